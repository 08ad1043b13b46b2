"""GPU: the partitioned BFS steps (grb_mgbfs_*) driven by gunrockinst_b200.partitioned.
Single process: world 1 through the same exchange code path (all segments local).  With more
than one device visible, `torchrun`-style workers over NCCL at world 2 (and 4, 8 when present)
on per-rank device-generated R-MAT slices; depths are compared with the oracle."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT, rmat_graph

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("scale", [10, 16])
@pytest.mark.parametrize("dopt", [False, True])
@pytest.mark.parametrize("exchange", ["nccl", "peer", "peer-persistent", "peer-tail"])
def test_world1_matches_oracle(gr, orc, scale, dopt, exchange, monkeypatch):
    """"peer-persistent": the whole search as one cooperative kernel (mgbfs_persistent.cuh, GRB_MG_PERSISTENT=1);
    "peer-tail": the host loop hands the last, short levels to the kernel's TAIL instance (GRB_MG_TAIL=1)"""
    from gunrockinst_b200 import partitioned as pt
    ro, co = rmat_graph(orc, scale)
    g = gr.Graph.from_csr(ro, co, symmetric=True)
    monkeypatch.delenv("GRB_MG_TAIL", raising=False)
    if exchange == "peer-tail":
        monkeypatch.setenv("GRB_MG_TAIL", "1")
        monkeypatch.delenv("GRB_MG_PERSISTENT", raising=False)
        exchange = "peer"
        persistent = False
    elif exchange == "peer-persistent":
        monkeypatch.setenv("GRB_MG_PERSISTENT", "1")
        exchange = "peer"
        persistent = True
    else:
        monkeypatch.delenv("GRB_MG_PERSISTENT", raising=False)
        persistent = False
    eng = pt.DeviceEngine(g)
    bfs = pt.PartitionedBfs(eng, direction_optimized=dopt, exchange=exchange)
    for src in (0, 3, ro.size - 2):
        st = bfs.run(src)
        want = orc.bfs(ro, co, src)
        assert np.array_equal(pt.gather_labels(eng), want)
        assert st["edges_visited"] == orc.edges_visited(ro, want)
        if dopt and scale >= 16 and src == 0:
            assert any(lv["kind"] == "pull" for lv in st["levels"])
        if persistent:
            assert st["gpu_launches"] <= 3             # two seeding kernels + the search
    eng.close()


@pytest.mark.parametrize("exchange", ["nccl", "peer", "peer-persistent"])
def test_world1_directed_graph_is_searched_top_down(gr, orc, exchange, monkeypatch):
    """a directed slice has no transpose: the partitioned search runs push levels only and still returns the oracle's depths"""
    from gunrockinst_b200 import partitioned as pt
    ro, co = rmat_graph(orc, 14, 8, False)
    g = gr.Graph.from_csr(ro, co, symmetric=False)
    if exchange == "peer-persistent":
        monkeypatch.setenv("GRB_MG_PERSISTENT", "1")
        exchange = "peer"
    else:
        monkeypatch.delenv("GRB_MG_PERSISTENT", raising=False)
    eng = pt.DeviceEngine(g)
    bfs = pt.PartitionedBfs(eng, direction_optimized=True, exchange=exchange)
    for src in (0, 7, ro.size - 2):
        st = bfs.run(src)
        assert np.array_equal(pt.gather_labels(eng), orc.bfs(ro, co, src))
        assert all(lv["kind"] == "push" for lv in st["levels"])
    eng.close()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, scale, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import gunrockinst_b200 as gr
    from gunrockinst_b200 import partitioned as pt
    import oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        lp = world.bit_length() - 1
        g = gr.Graph.rmat(scale, 16, device=rank, log_parts=lp, rank=rank)
        eng = pt.DeviceEngine(g)
        os.environ["GRB_MG_PERSISTENT"] = "1"                   # read when the context is created
        eng_p = pt.DeviceEngine(g)                              # the search as one cooperative kernel per device
        os.environ.pop("GRB_MG_PERSISTENT")
        os.environ["GRB_MG_TAIL"] = "1"
        eng_t = pt.DeviceEngine(g)                              # host loop + the TAIL instance for the last levels
        os.environ.pop("GRB_MG_TAIL")
        ro, co = oracle.rmat_csr(scale, 16)
        res = []
        for e, exchange in ((eng, "nccl"), (eng, "peer"), (eng_p, "peer"), (eng_t, "peer"), (eng, "nccl"), (eng, "peer"), (eng_p, "peer"),
                            (eng_t, "peer")):
            # alternating: either path starts from the other's leftovers
            for dopt in (True, False):
                bfs = pt.PartitionedBfs(e, direction_optimized=dopt, exchange=exchange)
                for src in (0, 1, 12345 % (1 << scale)):
                    st = bfs.run(src)
                    labels = pt.gather_labels(e)
                    want = oracle.bfs(ro, co, src)
                    res.append((bool(np.array_equal(labels, want)),
                                st["edges_visited"] == oracle.edges_visited(ro, want),
                                st["exchange_bytes"] > 0,
                                e is not eng_p or st["gpu_launches"] <= 3))
        # a directed graph (no per-device transpose): top-down only, both level loops
        gd = gr.Graph.rmat(scale, 16, undirected=False, device=rank, log_parts=lp, rank=rank)
        rod, cod = oracle.rmat_csr(scale, 16, undirected=False)
        for e in (pt.DeviceEngine(gd),):
            bfs = pt.PartitionedBfs(e, direction_optimized=True, exchange="peer")
            for src in (0, 5):
                st = bfs.run(src)
                res.append((bool(np.array_equal(pt.gather_labels(e), oracle.bfs(rod, cod, src))),
                            all(lv["kind"] == "push" for lv in st["levels"])))
        if rank == 0:
            q.put(res)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_multi_device_matches_oracle(gr, world):
    if gr.device_count() < world:
        pytest.skip("needs %d devices" % world)
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, 16, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = q.get(timeout=600)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert all(all(r) for r in res), res


# ---- partitioned SSSP over peer memory -------------------------------------------------------
def _check_sssp(ro, co, w, src, dist, preds, want):
    ok = bool(np.array_equal(dist, want))
    # predecessors: any valid parent (the reference's benign race) -- the tree
    # edge must exist and close the distance exactly; source and unreached vertices name themselves
    n = ro.size - 1
    v = np.arange(n)
    reached = (dist != 0xFFFFFFFF) & (v != src)
    ok &= bool(np.all(preds[~reached] == v[~reached]))
    for x in np.nonzero(reached)[0][:20000]:
        p = int(preds[x])
        nb = co[ro[p]:ro[p + 1]]
        hit = np.nonzero(nb == x)[0]
        ok &= hit.size > 0 and int(dist[p]) + int(w[ro[p] + hit[0]]) == int(dist[x])
    return ok


@pytest.mark.parametrize("scale", [10, 15])
def test_sssp_world1_matches_oracle(gr, orc, scale):
    from gunrockinst_b200 import partitioned as pt
    ro, co = rmat_graph(orc, scale)
    w = orc.edge_weights(ro, co, 7)
    g = gr.Graph.from_csr(ro, co, weights=w, symmetric=True)
    sp = pt.PartitionedSssp(g)
    for src in (0, 5, ro.size - 2):
        st = sp.run(src)
        want, _ = orc.sssp(ro, co, w, src)
        dist_, preds = sp.gather()
        assert _check_sssp(ro, co, w, src, dist_, preds, want)
        assert st["iterations"] >= 1 or ro[src + 1] == ro[src]
    sp.close()


def test_sssp_world1_directed_graph(gr, orc):
    """the partitioned SSSP only pushes along out-edges: directed slices need no transpose"""
    from gunrockinst_b200 import partitioned as pt
    ro, co = rmat_graph(orc, 13, 8, False)
    w = orc.edge_weights(ro, co, 3)
    g = gr.Graph.from_csr(ro, co, weights=w, symmetric=False)
    sp = pt.PartitionedSssp(g)
    for src in (0, 9):
        sp.run(src)
        want, _ = orc.sssp(ro, co, w, src)
        dist_, preds = sp.gather()
        assert _check_sssp(ro, co, w, src, dist_, preds, want)
    sp.close()


def _sssp_worker(rank, world, port, scale, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import gunrockinst_b200 as gr
    from gunrockinst_b200 import partitioned as pt
    import oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        lp = world.bit_length() - 1
        g = gr.Graph.rmat(scale, 16, device=rank, log_parts=lp, rank=rank, weight_seed=7)
        sp = pt.PartitionedSssp(g)
        ro, co = oracle.rmat_csr(scale, 16)
        w = oracle.edge_weights(ro, co, 7)
        res = []
        for src in (0, 1, 12345 % (1 << scale)):
            st = sp.run(src)
            dist_, preds = sp.gather()
            want, _ = oracle.sssp(ro, co, w, src)
            res.append(_check_sssp(ro, co, w, src, dist_, preds, want))
        if rank == 0:
            q.put(res)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sssp_multi_device_matches_oracle(gr, world):
    if gr.device_count() < world:
        pytest.skip("needs %d devices" % world)
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_sssp_worker, args=(r, world, port, 15, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = q.get(timeout=600)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert all(res), res


# ---- partitioned PageRank (grb_mgpr_*): same tolerance and iteration count as tests/test_pr_gpu.py ----
PR_TOL = 1e-6


def _check_pr(ranks, st, want, it):
    # per vertex, against the vertex's own value (no graph-wide scale): BASELINE.json north_star
    err = np.abs(ranks - want)
    worst = float(np.max(err - PR_TOL * np.maximum(1.0, np.abs(want)))) if want.size else -1.0
    return st["iterations"] == it and worst <= 0.0, (st["iterations"], it, float(err.max()) if want.size else 0.0)


@pytest.mark.parametrize("name", ["bips98_606", "chesapeake", "test_cc", "test_pr"])
def test_pagerank_world1_fixtures(gr, orc, name):
    from conftest import load_golden
    from gunrockinst_b200 import partitioned as pt
    n, ro, co, _ = load_golden(name)
    g = gr.Graph.from_csr(ro, co, symmetric=True)
    pr = pt.PartitionedPagerank(g)
    for kw in (dict(), dict(src=0, threshold=0.01, max_iter=20), dict(max_iter=3), dict(max_iter=0)):
        want, it = orc.pagerank(ro, co, kw.get("delta", 0.85), kw.get("threshold", 1e-6), kw.get("max_iter", 1000),
                                kw.get("src", -1))
        st = pr.run(**kw)
        ok, why = _check_pr(pr.gather(), st, want, it)
        assert ok, (kw, why)
    pr.close()


@pytest.mark.parametrize("scale", [10, 16])
def test_pagerank_world1_rmat(gr, orc, scale):
    from gunrockinst_b200 import partitioned as pt
    ro, co = rmat_graph(orc, scale)
    g = gr.Graph.from_csr(ro, co, symmetric=True)
    pr = pt.PartitionedPagerank(g)
    want, it = orc.pagerank(ro, co, 0.85, 1e-6, 1000, -1)
    for _ in range(2):                                  # a second run on the same context starts clean
        st = pr.run()
        ok, why = _check_pr(pr.gather(), st, want, it)
        assert ok, why
    assert st["iterations"] > 3 and st["valid_nodes"] == int(np.count_nonzero(np.diff(ro)))
    single, _, st1 = gr.pagerank(g)                     # the single-device push formulation, same graph
    assert st1["iterations"] == st["iterations"]
    assert np.all(np.abs(single - pr.gather()) <= PR_TOL * np.maximum(1.0, np.abs(want)))
    pr.close()


def test_pagerank_needs_symmetric_graph(gr, orc):
    from gunrockinst_b200 import partitioned as pt
    ro, co = rmat_graph(orc, 10, 4, False)
    g = gr.Graph.from_csr(ro, co, symmetric=False)
    with pytest.raises(gr.GunrockError):
        pt.PartitionedPagerank(g)


def _pr_worker(rank, world, port, scale, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import gunrockinst_b200 as gr
    from gunrockinst_b200 import partitioned as pt
    import oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        lp = world.bit_length() - 1
        g = gr.Graph.rmat(scale, 16, device=rank, log_parts=lp, rank=rank)
        pr = pt.PartitionedPagerank(g)
        ro, co = oracle.rmat_csr(scale, 16)
        res = []
        for kw in (dict(), dict(src=0, threshold=0.01, max_iter=20), dict(max_iter=3), dict()):
            want, it = oracle.pagerank(ro, co, kw.get("delta", 0.85), kw.get("threshold", 1e-6),
                                       kw.get("max_iter", 1000), kw.get("src", -1))
            st = pr.run(**kw)
            res.append(_check_pr(pr.gather(), st, want, it))
        if rank == 0:
            q.put(res)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_pagerank_multi_device_matches_oracle(gr, world):
    if gr.device_count() < world:
        pytest.skip("needs %d devices" % world)
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_pr_worker, args=(r, world, port, 15, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = q.get(timeout=600)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert all(ok for ok, _ in res), res


# ---- partitioned connected components (grb_mgcc_*): labels must be the oracle's min-vertex labels ----
@pytest.mark.parametrize("name", ["bips98_606", "chesapeake", "test_cc", "test_pr"])
def test_cc_world1_fixtures(gr, orc, name):
    from conftest import load_golden
    from gunrockinst_b200 import partitioned as pt
    n, ro, co, _ = load_golden(name)
    g = gr.Graph.from_csr(ro, co, symmetric=True)
    cc = pt.PartitionedCc(g)
    want, count = orc.cc(ro, co)
    for _ in range(2):                                  # a second run on the same context starts clean
        st = cc.run()
        labels, roots = cc.gather()
        assert np.array_equal(labels, want) and roots == count
        assert st["iterations"] >= 1
    cc.close()


@pytest.mark.parametrize("scale", [10, 16])
def test_cc_world1_rmat(gr, orc, scale):
    from gunrockinst_b200 import partitioned as pt
    ro, co = rmat_graph(orc, scale)
    g = gr.Graph.from_csr(ro, co, symmetric=True)
    cc = pt.PartitionedCc(g)
    cc.run()
    labels, roots = cc.gather()
    want, count = orc.cc(ro, co)
    assert np.array_equal(labels, want) and roots == count
    single, ncomp, _ = gr.cc(g)                         # the single-device hook / pointer-jump enactor
    assert np.array_equal(single, labels) and ncomp == roots
    cc.close()


def _cc_worker(rank, world, port, scale, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import gunrockinst_b200 as gr
    from gunrockinst_b200 import partitioned as pt
    import oracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        lp = world.bit_length() - 1
        res = []
        for ef in (16, 2):                              # one giant component + dust; many small components
            g = gr.Graph.rmat(scale, ef, device=rank, log_parts=lp, rank=rank)
            cc = pt.PartitionedCc(g)
            ro, co = oracle.rmat_csr(scale, ef)
            want, count = oracle.cc(ro, co)
            for _ in range(2):
                st = cc.run()
                labels, roots = cc.gather()
                res.append((bool(np.array_equal(labels, want)), roots == count, st["iterations"], st["elapsed_ms"]))
            cc.close()
            g.close()
        if rank == 0:
            q.put(res)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_cc_multi_device_matches_oracle(gr, world):
    if gr.device_count() < world:
        pytest.skip("needs %d devices" % world)
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_cc_worker, args=(r, world, port, 15, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = q.get(timeout=600)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert all(a and b for a, b, _, _ in res), res
