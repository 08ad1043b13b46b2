"""CPU: the C-ABI library loads without a GPU and exports every symbol the two headers under
include/ declare; gunrock.h's by-value structs have the layout a C compiler gives them."""
import ctypes as C
import os
import re
import subprocess

import pytest

from conftest import ROOT

INCLUDE = os.path.join(ROOT, "include")


def _declared(path):
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b((?:grb|gunrock)_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def lib():
    import gunrockinst_b200 as g
    if not os.path.exists(g._capi.LIB_PATH):
        g.build()
    return g._capi.lib()


def test_every_declared_symbol_is_exported(lib):
    import gunrockinst_b200 as g
    names = _declared(os.path.join(INCLUDE, "gunrock_b200.h")) + _declared(os.path.join(INCLUDE, "gunrock", "gunrock.h"))
    assert len(names) >= 39
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert not g._capi.MISSING


def test_loads_and_reports_without_gpu(lib):
    assert b"sm_100a" in lib.grb_version()
    assert lib.grb_device_count() >= 0          # 0 here; no compute call is made without a device


def test_reference_header_is_plain_c_and_struct_layout_matches(tmp_path):
    """gunrock.h must compile as C (shared_lib_tests/*.c include it) and the ctypes mirrors in
    _capi.py must agree with the compiler on size / field offsets (structs are passed by value)."""
    import gunrockinst_b200 as g
    src = tmp_path / "layout.c"
    src.write_text(r'''
#include <stdio.h>
#include <stddef.h>
#include <gunrock/gunrock.h>
int main(void) {
  printf("%zu %zu %zu %zu %zu %zu\n", sizeof(struct GunrockConfig), offsetof(struct GunrockConfig, src_node),
         offsetof(struct GunrockConfig, delta), offsetof(struct GunrockConfig, src_mode),
         sizeof(struct GunrockGraph), offsetof(struct GunrockGraph, edge_values));
  return 0; }''')
    exe = tmp_path / "layout"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", INCLUDE, str(src), "-o", str(exe)])
    got = [int(x) for x in subprocess.check_output([str(exe)]).split()]
    cfg, gg = g._capi.GunrockConfig, g._capi.GunrockGraph
    assert got == [C.sizeof(cfg), cfg.src_node.offset, cfg.delta.offset, cfg.src_mode.offset,
                   C.sizeof(gg), gg.edge_values.offset]


def test_product_package_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "gunrockinst_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in text and "from oracle" not in text and "liboracle" not in text, f


def test_malloced_result_is_handed_over_without_a_copy():
    """gunrock.h ownership rule (shared_lib_tests/test_bfs.c:58,66): node_values is malloc()ed by the
    library and freed by the caller.  The Python mirror wraps the block instead of copying it and must
    free() it exactly once, when the last view of the array is gone."""
    import gc
    import numpy as np
    import gunrockinst_b200 as g
    libc = C.CDLL(None)
    libc.malloc.restype = C.c_void_p
    libc.malloc.argtypes = [C.c_size_t]
    freed = []
    real_free = g._capi.free_host
    g._capi.free_host = lambda p: (freed.append(p), real_free(p))
    try:
        class Out:
            node_values = libc.malloc(4 * 10)
        C.memmove(Out.node_values, (C.c_int32 * 10)(*range(10)), 40)
        arr = g._take_node_values(Out, 10, np.uint32)
        assert arr.dtype == np.uint32 and arr.tolist() == list(range(10)) and arr.flags.writeable
        assert arr.ctypes.data == Out.node_values               # the library's block itself
        view = arr[2:5]
        del arr
        gc.collect()
        assert freed == [] and view.tolist() == [2, 3, 4]        # a view keeps the block alive
        del view
        gc.collect()
        assert freed == [Out.node_values]
        Out.node_values = None
        with pytest.raises(g.GunrockError):
            g._take_node_values(Out, 10, np.int32)
    finally:
        g._capi.free_host = real_free
