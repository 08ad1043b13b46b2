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
