"""The command-line rig (tests/rigs/gunrock_rig.c), counterpart of the reference's
tests/{bfs,dobfs,sssp,cc,pr}/test_*.cu programs: option parsing on the CPU, full runs with the
built-in CPU validation on the GPU (market and rmat graph types, the reference's flags)."""
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT

RIG = os.path.join(ROOT, "tests", "_bin", "gunrock_rig")


def _build():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "tests", "rigs")])


def test_rig_builds_and_prints_usage():
    _build()
    p = subprocess.run([RIG], capture_output=True, text=True)
    assert p.returncode == 1 and "market <file>" in p.stdout and "--traversal-mode" in p.stdout
    p = subprocess.run([RIG, "bfs", "nonsense"], capture_output=True, text=True)
    assert p.returncode == 1


def test_rig_answers_to_the_reference_binary_names():
    """tests/_bin/test_{bfs,dobfs,sssp,cc,pr}: the reference's per-primitive programs (`test_bfs market <file> ...`,
    tests/bfs/test_bfs.cu:58-89); the primitive comes from the name, the graph type is the first argument"""
    _build()
    for prim in ("bfs", "dobfs", "sssp", "cc", "pr"):
        exe = os.path.join(ROOT, "tests", "_bin", "test_" + prim)
        p = subprocess.run([exe], capture_output=True, text=True)
        assert p.returncode == 1 and "market <file>" in p.stdout
        p = subprocess.run([exe, "market", "/nonexistent/graph.mtx"], capture_output=True, text=True)
        assert p.returncode == 1 and "cannot read /nonexistent/graph.mtx" in (p.stdout + p.stderr)
        p = subprocess.run([exe, "nonsense", "x"], capture_output=True, text=True)
        assert p.returncode == 1 and "market <file>" in p.stdout


def _write_mtx(path, n, edges, values=None):
    with open(path, "w") as f:
        f.write("%%MatrixMarket matrix coordinate %s general\n" % ("integer" if values is not None else "pattern"))
        f.write("%d %d %d\n" % (n, n, len(edges)))
        for k, (a, b) in enumerate(edges):
            f.write("%d %d%s\n" % (a + 1, b + 1, "" if values is None else " %d" % values[k]))


@pytest.mark.gpu
@pytest.mark.parametrize("args", [
    ["bfs"], ["bfs", "--mark-pred"], ["bfs", "--idempotence=1"], ["bfs", "--traversal-mode=1", "--src=largestdegree"],
    ["bfs", "--traversal-mode=2", "--src=3", "--iteration-num=3"], ["dobfs"], ["dobfs", "--src=largestdegree", "--v"],
    ["sssp"], ["sssp", "--delta-factor=4", "--src=5"], ["cc"], ["pr"], ["pr", "--src=0", "--max-iter=50", "--error=1e-7"]])
def test_rig_market(tmp_path, args):
    _build()
    rng = np.random.RandomState(7)
    n = 600
    edges = [(int(a), int(b)) for a, b in rng.randint(0, n - 40, size=(3000, 2))]      # last 40 vertices stay isolated
    vals = [int(v) for v in rng.randint(1, 65, size=len(edges))]
    path = str(tmp_path / "g.mtx")
    _write_mtx(path, n, edges, vals if args[0] == "sssp" else None)
    p = subprocess.run([RIG, args[0], "market", path, "--undirected"] + args[1:], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert "TEST PASSED" in p.stdout and "INCORRECT" not in p.stdout
    if args[0] in ("bfs", "dobfs", "sssp") and not ("--idempotence=1" in args and "--mark-pred" in args):
        assert "Label Validity: CORRECT" in p.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("prim", ["bfs", "cc"])
def test_rig_under_the_reference_binary_name(prim):
    _build()
    exe = os.path.join(ROOT, "tests", "_bin", "test_" + prim)
    # the invocation of test_rig_rmat, minus the primitive argument
    p = subprocess.run([exe, "rmat", "--rmat_scale=14", "--rmat_edgefactor=8"], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert "TEST PASSED" in p.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("prim", ["bfs", "dobfs", "sssp", "cc", "pr"])
def test_rig_rmat(prim):
    _build()
    p = subprocess.run([RIG, prim, "rmat", "--rmat_scale=14", "--rmat_edgefactor=8"], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert "TEST PASSED" in p.stdout
    edges = int(p.stdout.split(" nodes, ")[1].split(" directed edges")[0])
    assert edges > (1 << 14) * 8        # a real graph was generated (duplicates and self loops removed)


@pytest.mark.gpu
def test_rig_directed_market_bfs(tmp_path):
    # no --undirected: the reader's (col, row) convention decides the edge direction (graphio/market.cuh:139-141)
    _build()
    path = str(tmp_path / "d.mtx")
    _write_mtx(path, 6, [(0, 1), (1, 2), (2, 3), (4, 5)])
    p = subprocess.run([RIG, "bfs", "market", path, "--src=3"], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0 and "Label Validity: CORRECT" in p.stdout, p.stdout + p.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("prim,flags", [("bfs", ["--undirected"]), ("sssp", ["--undirected"]), ("bfs", [])])
def test_rig_csr_cache(tmp_path, prim, flags):
    """--csr-cache-dir: the first run stores the reference's text cache of the built CSR (csr.cuh:140-176), the second one
    reads it instead of the .mtx (csr.cuh:182-232) -- same graph, same validated result"""
    _build()
    rng = np.random.RandomState(11)
    n = 300
    edges = [(int(a), int(b)) for a, b in rng.randint(0, n, size=(1500, 2))]
    vals = [int(v) for v in rng.randint(1, 65, size=len(edges))]
    path = str(tmp_path / "g.mtx")
    _write_mtx(path, n, edges, vals if prim == "sssp" else None)
    cdir = str(tmp_path / "cache")
    os.makedirs(cdir)
    cmd = [RIG, prim, "market", path, "--src=1", "--csr-cache-dir=" + cdir] + flags
    a = subprocess.run(cmd, capture_output=True, text=True, timeout=120)
    assert a.returncode == 0 and "Stored the CSR arrays" in a.stdout and "TEST PASSED" in a.stdout, a.stdout + a.stderr
    name = ".g.mtx_%s_csr" % ("undirected" if flags else "nonreversed")
    assert os.path.exists(os.path.join(cdir, name))
    os.rename(path, path + ".gone")                             # the second run cannot fall back to the .mtx
    b = subprocess.run(cmd, capture_output=True, text=True, timeout=120)
    assert b.returncode == 0 and "Reading directly from previously stored CSR arrays" in b.stdout and "TEST PASSED" in b.stdout, b.stdout + b.stderr
    line = [l for l in a.stdout.splitlines() if l.startswith("market graph:")]
    assert line and line == [l for l in b.stdout.splitlines() if l.startswith("market graph:")]
