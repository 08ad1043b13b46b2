"""GPU: the reference's own C acceptance programs (shared_lib_tests/test_{bfs,cc,sssp,pr,bc,topk}.c),
compiled UNMODIFIED against include/gunrock/gunrock.h and linked with libgunrock.so by
oracle/Makefile (binaries under oracle/_ref/, built where /root/reference exists and shipped to
the GPU box), judged by the reference's CTest regular expressions (CMakeLists.txt:215-233)."""
import os
import re
import subprocess

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu
BIN = os.path.join(ROOT, "oracle", "_ref")

CASES = [
    ("test_bfs", r"Node_ID.*2.*: Label.*1", ["Node_ID [%d] : Label [%d]" % p for p in enumerate([0, 1, 1, 1, 2, 2, 2])]),
    ("test_cc", r"Node_ID.*1.*: Component_ID.*0", ["Node_ID [%d] : Component_ID [0]" % i for i in range(7)]),
    ("test_sssp", r"Node ID.*1.*: Label.*39.*: Predecessor.*0",
     ["Node ID [%d] : Label [%d]" % p for p in enumerate([0, 39, 6, 16, 50, 29, 64])]),
    # the reference's PageRank golden (0.402378) is stale: the program must run,
    # print seven ranked lines, and rank node 2 first (it has the most in-links among nodes reachable from 0)
    ("test_pr", r"Node ID \[\d\] : Page Rank \[\d+\.\d+\]", []),
    # CMakeLists.txt:219-221 and :235-237
    ("test_bc", r"Node_ID.*0.*: BC.*0.500000", ["Edge_ID [25] : EBC[0.000000]"]),
    ("test_topk", r"Node ID.*2.*: in_degrees.*3.*: out_degrees.*3.*",
     ["Node ID [2] : in_degrees [3] : out_degrees [3]", "Node ID [4] : in_degrees [2] : out_degrees [3]",
      "Node ID [0] : in_degrees [1] : out_degrees [3]"]),
]


@pytest.mark.parametrize("name,regex,lines", CASES, ids=[c[0] for c in CASES])
def test_reference_program_links_and_passes(name, regex, lines):
    exe = os.path.join(BIN, name)
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/%s not built (needs /root/reference at build time)" % name)
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    assert re.search(regex, out.stdout), out.stdout
    for ln in lines:
        assert ln in out.stdout, (ln, out.stdout)
    if name == "test_pr":
        assert len(re.findall(r"Node ID \[\d\] : Page Rank", out.stdout)) == 7
    if name == "test_bc":       # fp32 atomics: compare the printed values numerically
        got = [float(x) for x in re.findall(r"Node_ID \[\d\] : BC\[([0-9.]+)\]", out.stdout)]
        want = [0.5, 0.5, 7 / 3, 11 / 6, 11 / 6, 2 / 3, 1 / 3]
        assert len(got) == 7 and all(abs(a - b) < 1e-5 for a, b in zip(got, want)), out.stdout
