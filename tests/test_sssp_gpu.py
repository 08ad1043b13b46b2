"""GPU: SSSP through the C ABI against the oracle's Dijkstra -- bit-exact uint32 distances,
parents validated as tight tree edges (they are race-dependent in the reference)."""
import numpy as np
import pytest

from conftest import load_golden, rmat_graph

pytestmark = pytest.mark.gpu
UMAX = np.uint32(0xFFFFFFFF)


def check_preds(ro, co, w, dist, preds, src):
    n = ro.size - 1
    assert preds[src] == src                                   # sssp_problem.cuh:372: preds start as iota
    unreached = dist == UMAX
    assert np.array_equal(preds[unreached], np.nonzero(unreached)[0])
    v = np.nonzero(~unreached)[0]
    v = v[v != src]
    p = preds[v]
    assert np.all((p >= 0) & (p < n)) and not np.any(dist[p] == UMAX)
    # (p, v) is an edge whose weight closes the gap exactly
    for vv, pp in zip(v[:3000], p[:3000]):
        row = co[ro[pp]:ro[pp + 1]]
        ws = w[ro[pp]:ro[pp + 1]][row == vv]
        assert ws.size and (dist[pp].astype(np.int64) + ws.astype(np.int64) == int(dist[vv])).any()


def test_shared_lib_fixture_through_reference_api(gr, shared_lib):
    """shared_lib_tests/test_sssp.c; golden CMakeLists.txt:227-229: node 1 -> 39, predecessor 0"""
    # the C program is a fresh process: libc rand() in its initial state (= srand(1)), first draw -> vertex 0
    import ctypes
    ctypes.CDLL(None).srand(1)
    dist, preds = gr.gunrock_sssp_func(shared_lib["row_offsets"], shared_lib["col_indices"],
                                       shared_lib["edge_values"], src_mode=gr._capi.SRC_RANDOMIZE,
                                       mark_pred=True, delta_factor=1)
    assert dist.tolist() == shared_lib["sssp_dist"]
    assert dist[1] == 39 and preds[1] == 0 and preds[0] == 0


@pytest.mark.parametrize("name", ["bips98_606", "chesapeake", "test_cc", "test_bc_dir", "bips98_606_dir"])
def test_fixtures_reference_api(gr, orc, name):
    n, ro, co, vals = load_golden(name)
    w = (np.abs(vals.astype(np.int64)) % 64 + 1).astype(np.uint32)
    for src in (0, n - 1):
        want, _ = orc.sssp(ro, co, w, src)
        for mark in (True, False):
            dist, preds = gr.gunrock_sssp_func(ro, co, w, src_node=src, mark_pred=mark)
            assert np.array_equal(dist, want)
            if mark:
                check_preds(ro, co, w, dist, preds, src)


def test_unsupported_type_triple_returns_nothing(gr, shared_lib):
    import ctypes as C
    g, keep = gr._gunrock_graph(shared_lib["row_offsets"], shared_lib["col_indices"], shared_lib["edge_values"])
    out = gr._capi.GunrockGraph()
    preds = np.zeros(7, np.int32)
    dt = gr._capi.GunrockDataType(0, 0, gr._capi.VALUE_INT)
    gr._capi.lib().gunrock_sssp_func(C.byref(out), preds.ctypes.data_as(C.c_void_p), C.byref(g), gr._config(), dt)
    assert not out.node_values


@pytest.mark.parametrize("scale", [10, 16, 18])
@pytest.mark.parametrize("delta_factor", [1, 16])
def test_rmat_resident(gr, orc, scale, delta_factor):
    ro, co = rmat_graph(orc, scale)
    w = orc.edge_weights(ro, co, 7)
    g = gr.Graph.from_csr(ro, co, weights=w, symmetric=True, force_offset_bytes=8 if scale == 16 else 0)
    n = ro.size - 1
    rng = np.random.RandomState(scale)
    for src in [0] + rng.randint(0, n, 2).tolist():
        want, _ = orc.sssp(ro, co, w, src)
        for mark in (True, False):
            dist, preds, st = gr.sssp(g, src, delta_factor=delta_factor, mark_pred=mark)
            assert np.array_equal(dist, want), "src %d mark %s" % (src, mark)
            if mark:
                check_preds(ro, co, w, dist, preds, src)
    g.close()


def test_device_rmat_weights_and_large_weights(gr, orc):
    # graph and weights generated on the device (the C3 path), heavy-tailed weights force many buckets
    g = gr.Graph.rmat(15, weight_seed=99)
    ro, co, w = g.download(want_weights=True)
    want, _ = orc.sssp(ro, co, w, 0)
    dist, preds, st = gr.sssp(g, 0)
    assert np.array_equal(dist, want)
    g.close()
    rng = np.random.RandomState(3)
    w2 = (rng.pareto(1.0, co.size) * 50 + 1).clip(1, 1e6).astype(np.uint32)
    g = gr.Graph.from_csr(ro, co, weights=w2)
    want, _ = orc.sssp(ro, co, w2, 5)
    dist, preds, st = gr.sssp(g, 5)
    assert np.array_equal(dist, want)
    check_preds(ro, co, w2, dist, preds, 5)
    g.close()


def test_edge_cases(gr, orc):
    for ro, co, w, src in [([0, 0], [], [], 0), ([0, 0, 0, 0], [], [], 1), ([0, 1, 2, 2], [1, 0], [5, 5], 2),
                           ([0, 2, 3, 3], [1, 2, 2], [10, 1, 1], 0)]:
        dist, preds = gr.gunrock_sssp_func(ro, co, w, src_node=src, mark_pred=True)
        want, _ = orc.sssp(ro, co, np.asarray(w, np.uint32), src)
        assert np.array_equal(dist, want)
    # a weighted path: thousands of one-vertex iterations and bucket jumps
    n = 2000
    rows = np.arange(n - 1)
    ro, co, _ = orc.coo_to_csr(n, np.concatenate([rows, rows + 1]), np.concatenate([rows + 1, rows]))
    w = orc.edge_weights(ro, co, 5)
    g = gr.Graph.from_csr(ro, co, weights=w, symmetric=True)
    dist, preds, st = gr.sssp(g, 0)
    want, _ = orc.sssp(ro, co, w, 0)
    assert np.array_equal(dist, want)
    check_preds(ro, co, w, dist, preds, 0)
