"""GPU: BFS through the C ABI (gunrock_bfs_func and the resident-graph entry points)
against the oracle -- bit-exact depths in every mode, every advance strategy."""
import numpy as np
import pytest

from conftest import load_golden, rmat_graph

pytestmark = pytest.mark.gpu


def check_preds(ro, co, labels, preds, src):
    n = ro.size - 1
    assert preds[src] == -1
    assert np.all(preds[labels == -1] == -2)            # bfs_problem.cuh:307: unvisited preds stay -2
    v = np.nonzero(labels > 0)[0]
    p = preds[v]
    assert np.all((p >= 0) & (p < n))
    assert np.all(labels[p] + 1 == labels[v])
    # (p, v) must be an edge of the graph
    for vv, pp in zip(v[:2000], p[:2000]):
        assert vv in co[ro[pp]:ro[pp + 1]]


def test_shared_lib_fixture_through_reference_api(gr, shared_lib):
    """shared_lib_tests/test_bfs.c: 7-vertex graph, src_mode randomize (first draw = 0)"""
    # the C program is a fresh process: libc rand() in its initial state (= srand(1)), first draw -> vertex 0
    import ctypes
    ctypes.CDLL(None).srand(1)
    labels = gr.gunrock_bfs_func(shared_lib["row_offsets"], shared_lib["col_indices"],
                                 src_mode=gr._capi.SRC_RANDOMIZE)
    assert labels.tolist() == shared_lib["bfs_labels"]


@pytest.mark.parametrize("name", ["bips98_606", "chesapeake", "test_cc", "test_bc_dir", "bips98_606_dir"])
@pytest.mark.parametrize("idem", [False, True])
def test_fixtures_reference_api(gr, orc, name, idem):
    n, ro, co, _ = load_golden(name)
    for src in (0, n - 1):
        got = gr.gunrock_bfs_func(ro, co, src_node=src, idempotence=idem)
        assert np.array_equal(got, orc.bfs(ro, co, src))


def test_largest_degree_source(gr, orc):
    n, ro, co, _ = load_golden("bips98_606")
    src = orc.largest_degree_node(ro)
    got = gr.gunrock_bfs_func(ro, co, src_mode=gr._capi.SRC_LARGEST_DEGREE)
    assert np.array_equal(got, orc.bfs(ro, co, src))


def test_unsupported_type_triple_returns_nothing(gr, shared_lib):
    import ctypes as C
    g, keep = gr._gunrock_graph(shared_lib["row_offsets"], shared_lib["col_indices"])
    out = gr._capi.GunrockGraph()
    dt = gr._capi.GunrockDataType(0, 0, gr._capi.VALUE_FLOAT)
    gr._capi.lib().gunrock_bfs_func(C.byref(out), C.byref(g), gr._config(), dt)
    assert not out.node_values


MODES = [
    dict(),                                         # auto TWC/LB
    dict(traversal_mode=1),                         # LB only
    dict(traversal_mode=2),                         # TWC only
    dict(direction_optimized=True),
    dict(direction_optimized=True, alpha=2.0, beta=2.0),
    dict(direction_optimized=True, alpha=1e9, beta=1e9),   # pull from the second level on
    dict(idempotence=True),
    dict(idempotence=True, traversal_mode=2),
    dict(mark_pred=True),
    dict(mark_pred=True, direction_optimized=True),
    dict(mark_pred=True, traversal_mode=2),
]
# the non-idempotent modes run as one cooperative kernel by default; the host-driven loop
# (one launch set + counter fetch per level, the reference's structure) must agree
MODES += [dict(m, host_loop=True) for m in MODES if not m.get("idempotence")]


@pytest.mark.parametrize("mode", MODES, ids=lambda m: ",".join("%s=%s" % kv for kv in m.items()) or "default")
@pytest.mark.parametrize("scale", [10, 16])
def test_rmat_all_modes(gr, orc, scale, mode):
    ro, co = rmat_graph(orc, scale)
    g = gr.Graph.from_csr(ro, co, symmetric=True)
    bfs = gr.Bfs(g, **mode)
    n = ro.size - 1
    rng = np.random.RandomState(scale)
    for src in [0, 1] + rng.randint(0, n, 3).tolist():
        labels, preds, st = bfs.run(src)
        want = orc.bfs(ro, co, src)
        assert np.array_equal(labels, want), "src %d mode %r" % (src, mode)
        assert st["edges_visited"] == orc.edges_visited(ro, want)
        if mode.get("mark_pred"):
            check_preds(ro, co, labels, preds, src)
    bfs.close()
    g.close()


def test_rmat_scale18_dobfs_and_64bit_offsets(gr, orc):
    ro, co = rmat_graph(orc, 18)
    want = orc.bfs(ro, co, 0)
    for fob in (4, 8):
        g = gr.Graph.from_csr(ro, co, symmetric=True, force_offset_bytes=fob)
        assert g.info.offset_bytes == fob
        for mode in (dict(direction_optimized=True, record_levels=True), dict(traversal_mode=1), dict(),
                     dict(direction_optimized=True, record_levels=True, host_loop=True), dict(host_loop=True)):
            bfs = gr.Bfs(g, **mode)
            labels, _, st = bfs.run(0)
            assert np.array_equal(labels, want)
            assert st["gpu_launches"] == 1 or mode.get("host_loop")
            if mode.get("direction_optimized"):
                assert st["pull_levels"] > 0
                lv = bfs.levels()
                assert len(lv) == st["iterations"]
                assert sum(r["discovered"] for r in lv) + 1 == int((want >= 0).sum())
            bfs.close()
        g.close()


def test_edge_cases(gr, orc):
    # a single vertex, an edgeless graph, an isolated source, a long path (many tiny levels)
    for ro, co, src in [([0, 0], [], 0), ([0, 0, 0, 0], [], 1), ([0, 1, 2, 2], [1, 0], 2)]:
        got = gr.gunrock_bfs_func(ro, co, src_node=src)
        assert np.array_equal(got, orc.bfs(ro, co, src))
    n = 3000
    rows = np.arange(n - 1)
    ro, co, _ = orc.coo_to_csr(n, np.concatenate([rows, rows + 1]), np.concatenate([rows + 1, rows]))
    for mode in (dict(), dict(direction_optimized=True), dict(idempotence=True), dict(host_loop=True)):
        g = gr.Graph.from_csr(ro, co, symmetric=True)
        labels, _, st = gr.Bfs(g, **mode).run(0)
        assert np.array_equal(labels, np.arange(n))
        assert st["iterations"] == n
    # a star: one hub with 100k leaves (CTA-cooperative gather, then a 100k-vertex frontier)
    k = 100000
    hub_rows = np.zeros(k, np.int64)
    leaves = np.arange(1, k + 1)
    ro, co, _ = orc.coo_to_csr(k + 1, np.concatenate([hub_rows, leaves]), np.concatenate([leaves, hub_rows]))
    for mode in (dict(traversal_mode=2), dict(traversal_mode=1), dict(direction_optimized=True)):
        g = gr.Graph.from_csr(ro, co, symmetric=True)
        for src in (0, 5):
            labels, _, _ = gr.Bfs(g, **mode).run(src)
            assert np.array_equal(labels, orc.bfs(ro, co, src))


def test_out_of_range_source_is_an_empty_search(gr):
    g = gr.Graph.from_csr([0, 1, 2], [1, 0])
    labels, _, st = gr.Bfs(g).run(7)
    assert labels.tolist() == [-1, -1] and st["iterations"] == 0


def test_repeated_runs_reuse_context(gr, orc):
    ro, co = rmat_graph(orc, 14)
    g = gr.Graph.from_csr(ro, co, symmetric=True)
    bfs = gr.Bfs(g, direction_optimized=True)
    for src in range(40):
        labels, _, _ = bfs.run(src)
        assert np.array_equal(labels, orc.bfs(ro, co, src))


@pytest.mark.parametrize("mode", [dict(direction_optimized=True), dict(), dict(mark_pred=True, direction_optimized=True),
                                  dict(direction_optimized=True, host_loop=True), dict(idempotence=True)],
                         ids=["dobfs", "push", "dobfs-pred", "dobfs-host-loop", "idempotent"])
def test_run_many_batches_match_single_runs(gr, orc, mode):
    """grb_bfs_run_many: a batch of searches inside one launch (persistent loop) gives, search by
    search, the statistics of separate runs, and leaves the last search's depths on the device"""
    import ctypes as C
    ro, co = rmat_graph(orc, 15)
    n = ro.size - 1
    g = gr.Graph.from_csr(ro, co, symmetric=True)
    bfs = gr.Bfs(g, **mode)
    bfs.run(0)                                              # first run sets the context up
    srcs = [0, 7, n - 1, 123, 0, 4242, n + 5, 99]           # one out-of-range source: an empty search
    many = bfs.run_many(srcs)
    for s, st in zip(srcs, many):
        want = orc.bfs(ro, co, s) if s < n else np.full(n, -1, np.int32)
        assert st["edges_visited"] == orc.edges_visited(ro, want), (s, mode)
        assert st["iterations"] == (int(want.max()) + 1 if s < n else 0)
    # depths of the last search are what the device holds
    labels = np.empty(n, np.int32)
    import torch
    ptr = bfs.device_labels()
    t = torch.as_tensor(gr.partitioned._RawCuda(ptr, n, "<i4"), device="cuda") if hasattr(gr, "partitioned") else None
    if t is None:
        from gunrockinst_b200 import partitioned
        t = torch.as_tensor(partitioned._RawCuda(ptr, n, "<i4"), device="cuda")
    assert np.array_equal(t.cpu().numpy(), orc.bfs(ro, co, srcs[-1]))
    bfs.close()
    g.close()


# ---- degree-ordered twin (grb_bfs_config.degree_order): same answers in the caller's ids ---------------
TWIN_MODES = [dict(), dict(direction_optimized=True), dict(direction_optimized=True, host_loop=True),
              dict(traversal_mode=1, host_loop=True), dict(mark_pred=True, direction_optimized=True),
              dict(idempotence=True), dict(traversal_mode=2)]


@pytest.mark.parametrize("mode", TWIN_MODES, ids=lambda m: ",".join("%s=%s" % kv for kv in m.items()) or "default")
@pytest.mark.parametrize("scale", [10, 16])
def test_degree_order_matches_oracle(gr, orc, scale, mode):
    ro, co = rmat_graph(orc, scale)
    g = gr.Graph.from_csr(ro, co, symmetric=True)
    bfs = gr.Bfs(g, degree_order=True, **mode)
    n = ro.size - 1
    rng = np.random.RandomState(100 + scale)
    for src in [0, 1, n - 1] + rng.randint(0, n, 3).tolist():
        labels, preds, st = bfs.run(src)
        want = orc.bfs(ro, co, src)
        assert np.array_equal(labels, want), "src %d mode %r" % (src, mode)
        assert st["edges_visited"] == orc.edges_visited(ro, want)
        if mode.get("mark_pred"):
            check_preds(ro, co, labels, preds, src)
    # a batch inside one launch translates its sources on the device
    srcs = [0, 5, n - 1, 77]
    many = bfs.run_many(srcs)
    for s, st in zip(srcs, many):
        assert st["edges_visited"] == orc.edges_visited(ro, orc.bfs(ro, co, s)), (s, mode)
    import torch
    from gunrockinst_b200 import partitioned
    t = torch.as_tensor(partitioned._RawCuda(bfs.device_labels(), n, "<i4"), device="cuda")
    assert np.array_equal(t.cpu().numpy(), orc.bfs(ro, co, srcs[-1]))
    bfs.close()
    g.close()


@pytest.mark.parametrize("name", ["bips98_606", "chesapeake", "test_cc", "test_bc_dir", "bips98_606_dir"])
def test_degree_order_fixtures(gr, orc, name):
    n, ro, co, _ = load_golden(name)
    sym = not name.endswith("_dir")
    g = gr.Graph.from_csr(ro, co, symmetric=sym)
    for mode in (dict(), dict(direction_optimized=True), dict(mark_pred=True, direction_optimized=True)):
        bfs = gr.Bfs(g, degree_order=True, **mode)
        for src in (0, n // 2, n - 1):
            labels, preds, _ = bfs.run(src)
            assert np.array_equal(labels, orc.bfs(ro, co, src)), (name, mode, src)
            if mode.get("mark_pred"):
                check_preds(ro, co, labels, preds, src)
        bfs.close()
    g.close()


# ---- the transpose: direction-optimising BFS on DIRECTED graphs (reference: the inverse graph of
# tests/dobfs/test_dobfs.cu:477-560, app/dobfs/dobfs_problem.cuh:211-260) --------------------------------
@pytest.mark.parametrize("name", ["bips98_606_dir", "test_bc_dir", "chesapeake_dir", "test_pr_dir"])
def test_dobfs_directed_fixtures_pull(gr, orc, name):
    n, ro, co, _ = load_golden(name)
    g = gr.Graph.from_csr(ro, co, symmetric=False)
    assert not g.check_symmetric()
    for mode in (dict(alpha=1e9, beta=1e9), dict(alpha=1e9, beta=1e9, host_loop=True),
                 dict(alpha=1e9, beta=1e9, mark_pred=True), dict(alpha=1e9, beta=1e9, degree_order=True)):
        bfs = gr.Bfs(g, direction_optimized=True, **mode)
        pulled = 0
        for src in range(0, n, max(1, n // 7)):
            labels, preds, st = bfs.run(src)
            want = orc.bfs(ro, co, src)
            assert np.array_equal(labels, want), (name, mode, src)
            pulled += st["pull_levels"]
            if mode.get("mark_pred"):
                check_preds(ro, co, labels, preds, src)
        assert pulled > 0, (name, mode)                 # the inverse graph exists: the search really pulls
        bfs.close()
    g.close()


@pytest.mark.parametrize("scale,ef", [(12, 8), (16, 16)])
def test_dobfs_directed_rmat_pull(gr, orc, scale, ef):
    ro, co = rmat_graph(orc, scale, ef, False)
    n = ro.size - 1
    g = gr.Graph.from_csr(ro, co, symmetric=False)
    assert not g.check_symmetric()
    for mode in (dict(), dict(host_loop=True), dict(degree_order=True), dict(alpha=1e9, beta=1e9)):
        bfs = gr.Bfs(g, direction_optimized=True, **mode)
        labels, _, st = bfs.run(0)
        want = orc.bfs(ro, co, 0)
        assert np.array_equal(labels, want), mode
        assert st["pull_levels"] > 0, mode
        assert st["edges_visited"] == orc.edges_visited(ro, want)
        bfs.close()
    g.close()


def test_transpose_and_symmetry_check(gr, orc):
    """grb_graph_build_transpose against a numpy transpose; grb_graph_check_symmetric on both kinds"""
    import ctypes as C
    ro, co = rmat_graph(orc, 12, 8, False)
    g = gr.Graph.from_csr(ro, co, symmetric=False)
    g.build_transpose()
    g.build_transpose()                                     # idempotent
    assert not g.check_symmetric()
    g.close()
    ro, co = rmat_graph(orc, 12)
    g = gr.Graph.from_csr(ro, co, symmetric=True)
    assert g.check_symmetric()
    g.close()
    # a symmetric promise that does not hold is detectable
    n, ro, co, _ = load_golden("test_bc_dir")
    g = gr.Graph.from_csr(ro, co, symmetric=True)
    assert not g.check_symmetric()
    g.close()


def test_cooperative_launch_failure_falls_back_to_host_loop(gr, orc, capfd):
    """when the cooperative launch of the persistent kernel is refused at run time the search
    (single or batched) must still be answered, by the host-driven loop"""
    import os
    ro, co = rmat_graph(orc, 14)
    n = ro.size - 1
    g = gr.Graph.from_csr(ro, co, symmetric=True)
    bfs = gr.Bfs(g, direction_optimized=True)
    labels, _, st = bfs.run(3)                              # sets the context up; persistent kernel
    assert st["gpu_launches"] == 1
    os.environ["GRB_FAIL_COOP_LAUNCH"] = "1"
    try:
        many = bfs.run_many([0, 9, 200])                    # batch refused -> search by search
        for s, st in zip([0, 9, 200], many):
            assert st["edges_visited"] == orc.edges_visited(ro, orc.bfs(ro, co, s))
            assert st["gpu_launches"] > 1                   # host loop
        labels, _, st = bfs.run(5)
        assert np.array_equal(labels, orc.bfs(ro, co, 5))
        bfs2 = gr.Bfs(g, direction_optimized=True)          # a fresh context: first fused search hits the refusal
        bfs2.run(1)
        labels, _, st = bfs2.run(7)
        assert np.array_equal(labels, orc.bfs(ro, co, 7)) and st["gpu_launches"] > 1
        bfs2.close()
    finally:
        del os.environ["GRB_FAIL_COOP_LAUNCH"]
    bfs.close()
    g.close()


# ---- streaming advance (oprtr/advance/lb_stream.cuh): TMA / cp.async staged tiles, hot visited words in shared
# memory.  stream_min_edges=2 forces it onto every LB level of these small graphs. ------------------------------
STREAM_MODES = [dict(traversal_mode=1), dict(), dict(direction_optimized=True), dict(idempotence=True),
                dict(idempotence=True, traversal_mode=1), dict(traversal_mode=1, degree_order=True),
                dict(direction_optimized=True, degree_order=True), dict(idempotence=True, degree_order=True)]


@pytest.mark.parametrize("mode", STREAM_MODES, ids=lambda m: ",".join("%s=%s" % kv for kv in m.items()) or "default")
@pytest.mark.parametrize("scale", [10, 16, 18])
def test_streaming_advance_matches_oracle(gr, orc, scale, mode):
    ro, co = rmat_graph(orc, scale)
    g = gr.Graph.from_csr(ro, co, symmetric=True)
    bfs = gr.Bfs(g, host_loop=True, stream_min_edges=2, record_levels=True, **mode)
    n = ro.size - 1
    rng = np.random.RandomState(7 + scale)
    streamed = filter_streamed = 0
    for src in [0, 1, n - 1] + rng.randint(0, n, 3).tolist():
        labels, _, st = bfs.run(src)
        want = orc.bfs(ro, co, src)
        assert np.array_equal(labels, want), "src %d mode %r" % (src, mode)
        assert st["edges_visited"] == orc.edges_visited(ro, want)
        lv = bfs.levels()
        streamed += sum(1 for r in lv if r["kind"] == "push-stream")
        filter_streamed += sum(1 for r in lv if r["kind"] == "filter-stream")
        if not mode.get("idempotence") and not mode.get("direction_optimized"):
            # a top-down search inspects every entry of the visited rows exactly once
            assert sum(r["edges"] for r in lv) == st["edges_visited"]
    # (an idempotent search keeps lb::ExpandKernel for its advance -- the streaming advance's claim IS the visited bit, which
    # an idempotent advance must leave to the filter -- and streams its FILTER instead: filter/stream_kernel.cuh)
    if not mode.get("idempotence") and (scale >= 16 or mode.get("traversal_mode") == 1):     # small graphs: only the LB-always modes reach the LB levels
        assert streamed > 0, mode
    if mode.get("idempotence"):
        assert filter_streamed > 0, mode
    bfs.close()
    g.close()


def test_streaming_advance_shapes(gr, orc):
    """tiles made of one long piece (a star), of thousands of one-entry pieces (a path, a perfect matching),
    pieces that start at every alignment, 64-bit offsets"""
    graphs = []
    k = 100000
    hub = np.zeros(k, np.int64)
    leaves = np.arange(1, k + 1)
    graphs.append(orc.coo_to_csr(k + 1, np.concatenate([hub, leaves]), np.concatenate([leaves, hub]))[:2])
    n = 20000                                                   # complete bipartite-ish: every row 67 long (odd: all alignments)
    rows = np.repeat(np.arange(300), 67)
    cols = 300 + (np.arange(300 * 67) * 7919) % (n - 300)
    graphs.append(orc.coo_to_csr(n, np.concatenate([rows, cols]), np.concatenate([cols, rows]))[:2])
    m = 50000                                                   # two hubs joined to a big matching: thousands of short pieces
    a = np.arange(2, 2 + m)
    graphs.append(orc.coo_to_csr(2 + m, np.concatenate([np.zeros(m, np.int64), a, np.ones(m, np.int64), a]),
                                 np.concatenate([a, np.zeros(m, np.int64), a, np.ones(m, np.int64)]))[:2])
    for ro, co in graphs:
        for fob in (4, 8):
            g = gr.Graph.from_csr(ro, co, symmetric=True, force_offset_bytes=fob)
            for mode in (dict(traversal_mode=1), dict(traversal_mode=1, degree_order=True), dict(idempotence=True)):
                bfs = gr.Bfs(g, host_loop=True, stream_min_edges=2, **mode)
                for src in (0, 5, ro.size - 2):
                    labels, _, _ = bfs.run(src)
                    assert np.array_equal(labels, orc.bfs(ro, co, src)), (ro.size, fob, mode, src)
                bfs.close()
            g.close()


@pytest.mark.parametrize("mode", [dict(traversal_mode=1, host_loop=True), dict(idempotence=True)],
                         ids=["topdown", "idempotent"])
def test_streaming_instances_at_scale_21(gr, orc, mode):
    """the streaming advance / filter where they are the default path (degree-ordered twin, levels of tens of millions of
    entries, several tiles per SM in flight): depths equal to the oracle's for several sources.  The ring's slot hand-back
    hazard of round 2 only showed at this size."""
    g = gr.Graph.rmat(21)
    ro, co = g.download()
    bfs = gr.Bfs(g, degree_order=True, record_levels=True, **mode)
    kinds = set()
    for src in (0, 1, 77777, (1 << 21) - 5):
        labels, _, st = bfs.run(src)
        assert np.array_equal(labels, orc.bfs(ro, co, src)), (src, mode)
        kinds |= {r["kind"] for r in bfs.levels()}
    assert ("filter-stream" if mode.get("idempotence") else "push-stream") in kinds
    bfs.close()
    g.close()
