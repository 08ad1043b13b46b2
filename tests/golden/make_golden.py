"""Regenerates the fixtures in this directory.  Runs only where /root/reference exists.

For every MatrixMarket fixture of the reference (simple_example/bips98_606.mtx,
dataset/small/*.mtx) the REFERENCE's own reader + COO->CSR canonicaliser
(gunrock/graphio/market.cuh, gunrock/csr.cuh, compiled unmodified into
oracle/_ref/ref_graphio by oracle/Makefile) is run with --undirected and, for the
directed fixtures, also without; its CSR is stored as <name>[_dir].npz
(n, row_offsets int32, col int32, vals int32).  The reference's RandomNode sequence
(first 8 draws for n = 7 and n = 7135 from an unseeded process) goes to randomnode.json.

Also stored (derived by hand from the 7-vertex fixture of shared_lib_tests/test_*.c and
the CTest regexes in CMakeLists.txt:215-241): shared_lib.json.

usage: python tests/golden/make_golden.py
"""
import json
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle  # noqa: E402

REF = "/root/reference"
FIXTURES = {
    "bips98_606": "simple_example/bips98_606.mtx",
    "test_cc": "dataset/small/test_cc.mtx",
    "test_pr": "dataset/small/test_pr.mtx",
    "test_bc": "dataset/small/test_bc.mtx",
    "chesapeake": "dataset/small/chesapeake.mtx",
}


def main():
    oracle.build()
    for name, rel in FIXTURES.items():
        path = os.path.join(REF, rel)
        for undirected in (True, False):
            n, ro, col, vals = oracle.ref_graphio_market(path, undirected=undirected)
            out = os.path.join(HERE, name + ("" if undirected else "_dir") + ".npz")
            np.savez_compressed(out, n=np.int64(n), row_offsets=ro.astype(np.int32), col=col,
                                vals=vals if vals is not None else np.zeros(0, np.int32))
            print(out, n, col.size)
    # the text cache of the built CSR, written by the reference's own Csr::WriteToFile (gunrock/csr.cuh:140-176) next to
    # a scratch copy of the .mtx (graphio/market.cuh:254-326)
    import shutil
    import tempfile
    cdir = os.path.join(HERE, "csr_cache")
    os.makedirs(cdir, exist_ok=True)
    with tempfile.TemporaryDirectory() as tmp:
        for name in ("test_pr", "test_cc", "test_bc"):
            mtx = os.path.join(tmp, name + ".mtx")
            shutil.copy(os.path.join(REF, FIXTURES[name]), mtx)
            for undirected, kind in ((1, "undirected"), (0, "nonreversed")):
                subprocess.check_call([oracle.REF_GRAPHIO, "market", mtx, str(undirected), os.path.join(tmp, "out.bin")],
                                      stdout=subprocess.DEVNULL)
                shutil.copy(os.path.join(tmp, ".%s.mtx_%s_csr" % (name, kind)), os.path.join(cdir, "%s.mtx_%s_csr" % (name, kind)))
    draws = {}
    for n in (7, 7135):
        o = subprocess.check_output([oracle.REF_GRAPHIO, "randomnode", str(n), "8"]).decode().split()
        draws[str(n)] = [int(x) for x in o]
    json.dump(draws, open(os.path.join(HERE, "randomnode.json"), "w"))
    # the 7-vertex graph of shared_lib_tests/test_bfs.c:30-33 / test_sssp.c:33-35 and the
    # results its CTest regexes pin (CMakeLists.txt:215-229)
    shared = {
        "row_offsets": [0, 3, 6, 9, 11, 14, 15, 15],
        "col_indices": [1, 2, 3, 0, 2, 4, 3, 4, 5, 5, 6, 2, 5, 6, 6],
        "edge_values": [39, 6, 41, 51, 63, 17, 10, 44, 41, 13, 58, 43, 50, 59, 35],
        "bfs_labels": [0, 1, 1, 1, 2, 2, 2],
        "sssp_dist": [0, 39, 6, 16, 50, 29, 64],
        "cc_labels": [0, 0, 0, 0, 0, 0, 0],
        "regex": {"bfs": "Node_ID.*2.*: Label.*1", "cc": "Node_ID.*1.*: Component_ID.*0",
                  "sssp": "Node ID.*1.*: Label.*39.*: Predecessor.*0"},
    }
    json.dump(shared, open(os.path.join(HERE, "shared_lib.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
