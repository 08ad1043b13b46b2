"""CPU: the host loop of the 1D-partitioned BFS (gunrockinst_b200/partitioned.py) run at
world_size 1, 2 and 4 over gloo with a numpy stand-in for the device steps (tests/mg_sim.py);
depths must equal the oracle's serial BFS.  Also the host copy of the partition rule."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, scale, sources, dopt, alpha, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle
    from gunrockinst_b200 import partitioned as pt
    from mg_sim import NumpyEngine
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ro, co = oracle.rmat_csr(scale, 8)
        eng = NumpyEngine(ro, co, world.bit_length() - 1, rank)
        bfs = pt.PartitionedBfs(eng, direction_optimized=dopt, alpha=alpha)
        out = []
        for src in sources:
            st = bfs.run(src)
            labels = pt.gather_labels(eng)
            want = oracle.bfs(ro, co, src)
            kinds = "".join("P" if lv["kind"] == "pull" else "p" for lv in st["levels"])
            out.append((bool(np.array_equal(labels, want)), st["iterations"], int(want.max()) + 1, kinds,
                        st["edges_visited"], oracle.edges_visited(ro, want)))
        if rank == 0:
            q.put(out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("dopt", [False, True])
def test_partitioned_loop_matches_oracle(world, dopt):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    sources = [0, 5, 1023]
    procs = [ctx.Process(target=_worker, args=(r, world, port, 10, sources, dopt, 4.0, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    saw_pull = False
    for ok, iters, depth, kinds, ev, ev_want in res:
        assert ok
        assert ev == ev_want                                   # the GTEPS numerator (bfs_app.cu:82-91)
        assert iters in (depth, depth - 1) or depth == 0       # last level may or may not discover nothing
        saw_pull |= "P" in kinds
    assert saw_pull == dopt


def test_partition_rule_is_a_bijection():
    from gunrockinst_b200.partitioned import owner_of, global_ids
    for lp in (0, 1, 2, 3):
        n = 1 << 12
        seen = np.zeros(n, dtype=np.int32)
        for r in range(1 << lp):
            g = global_ids(n >> lp, lp, r)
            seen[g] += 1
            assert all(owner_of(int(v), lp) == r for v in g[:64])
        assert np.all(seen == 1)


def _pr_worker(rank, world, port, scale, undirected, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle
    from gunrockinst_b200.partitioned import global_ids
    from mg_sim import partitioned_pagerank
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ro, co = oracle.rmat_csr(scale, 4, undirected=undirected)
        lp = world.bit_length() - 1
        out = []
        for kw in (dict(), dict(src=0, threshold=0.01, max_iter=20), dict(max_iter=3)):
            mine, sweeps, prunes = partitioned_pagerank(ro, co, lp, rank, **kw)
            every = [None] * world
            dist.all_gather_object(every, mine)
            got = np.empty(ro.size - 1)
            for r, seg in enumerate(every):
                got[global_ids(seg.size, lp, r)] = seg
            want, it = oracle.pagerank(ro, co, kw.get("delta", 0.85), kw.get("threshold", 1e-6),
                                       kw.get("max_iter", 1000), kw.get("src", -1))
            out.append((sweeps, it, float(np.max(np.abs(got - want))), max(1.0, float(want.max())), prunes))
        if rank == 0:
            q.put(out)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("undirected", [True, False])
def test_partitioned_pagerank_restatement_matches_oracle(world, undirected):
    """the algorithm of grb_mgpr_* (replicated zero-degree map for the pruning, gather from replicated
    contributions) against the oracle's restatement of the reference's push formulation"""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_pr_worker, args=(r, world, port, 9, undirected, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for sweeps, it, err, scale, prunes in res:
        assert sweeps == it
        assert err <= 1e-6 * scale
    if not undirected:
        assert res[0][4] > 2                                   # the pruning cascaded over several sweeps


def _cc_worker(rank, world, port, scale, ef, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle
    from gunrockinst_b200.partitioned import global_ids
    from mg_sim import partitioned_cc
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ro, co = oracle.rmat_csr(scale, ef)
        lp = world.bit_length() - 1
        mine, iterations = partitioned_cc(ro, co, lp, rank)
        every = [None] * world
        dist.all_gather_object(every, mine)
        got = np.empty(ro.size - 1, dtype=np.int32)
        for r, seg in enumerate(every):
            got[global_ids(seg.size, lp, r)] = seg
        want, count = oracle.cc(ro, co)
        if rank == 0:
            q.put((bool(np.array_equal(got, want)), int((got == np.arange(got.size)).sum()) == count, iterations))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("ef", [8, 1])
def test_partitioned_cc_restatement_matches_oracle(world, ef):
    """the algorithm of grb_mgcc_* (minimum-label propagation with a best-sent replica and owner-side
    re-queueing) against the oracle's union-find labels"""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_cc_worker, args=(r, world, port, 9, ef, q)) for r in range(world)]
    for p in procs:
        p.start()
    same, count_ok, iterations = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert same and count_ok and iterations >= 2
