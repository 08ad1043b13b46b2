"""GPU: betweenness centrality and top-k degree centrality through the C ABI (SURVEY.md section 8f
rank 1) against the oracle's restatement of the reference's CPU pass (tests/bc/test_bc.cu:235-320)
and of its three-sort top-k (topk_enactor.cuh:152-186).  BC accumulates fp32 atomics in an order
that differs run to run (as the reference's does): tolerance 1e-4 relative, stated here."""
import numpy as np
import pytest

from conftest import load_golden, rmat_graph

pytestmark = pytest.mark.gpu
RTOL = 1e-4

BC_RO = [0, 3, 6, 11, 15, 19, 23, 26]
BC_CO = [1, 2, 3, 0, 2, 4, 0, 1, 3, 4, 5, 0, 2, 5, 6, 1, 2, 5, 6, 2, 3, 4, 6, 3, 4, 5]


def test_bc_shared_lib_fixture_through_reference_api(gr, orc):
    """shared_lib_tests/test_bc.c; golden CMakeLists.txt:219-221: node 0 -> 0.5"""
    vals, ebc = gr.gunrock_bc_func(BC_RO, BC_CO)
    want, _, _ = orc.bc(BC_RO, BC_CO, -1)
    assert abs(vals[0] - 0.5) < 1e-6
    assert np.allclose(vals, want, rtol=RTOL, atol=1e-6)
    assert ebc.shape == (26,) and not ebc.any()          # the reference never accumulates edge values


@pytest.mark.parametrize("name", ["chesapeake", "test_cc", "test_bc", "test_pr"])
def test_bc_fixtures_all_sources(gr, orc, name):
    n, ro, co, _ = load_golden(name)
    want, _, _ = orc.bc(ro, co, -1)
    vals, _ = gr.gunrock_bc_func(ro, co)
    assert np.allclose(vals, want, rtol=RTOL, atol=1e-5)
    g = gr.Graph.from_csr(ro, co)
    got, _, st = gr.bc(g, -1)
    assert np.allclose(got, want, rtol=RTOL, atol=1e-5)
    g.close()


@pytest.mark.parametrize("scale", [10, 14])
def test_bc_rmat_single_source(gr, orc, scale):
    ro, co = rmat_graph(orc, scale)
    for fob in (4, 8):
        g = gr.Graph.from_csr(ro, co, symmetric=True, force_offset_bytes=fob)
        for src in (0, 5, (1 << scale) - 1):
            want, wsig, _ = orc.bc(ro, co, src)
            got, sig, st = gr.bc(g, src, want_sigmas=True)
            assert np.allclose(sig, wsig, rtol=RTOL)        # path counts (exact while below 2^24)
            assert np.allclose(got, want, rtol=1e-3, atol=1e-3)
            assert got[src] == 0
        g.close()


def test_bc_unsupported_type_triple_returns_nothing(gr):
    import ctypes as C
    g, keep = gr._gunrock_graph(BC_RO, BC_CO)
    out = gr._capi.GunrockGraph()
    dt = gr._capi.GunrockDataType(0, 0, gr._capi.VALUE_INT)
    gr._capi.lib().gunrock_bc_func(C.byref(out), C.byref(g), gr._config(src_node=-1), dt)
    assert not out.node_values


def transpose_offsets(ro, co):
    n = len(ro) - 1
    return np.concatenate([[0], np.cumsum(np.bincount(np.asarray(co, dtype=np.int64), minlength=n))])


def test_topk_shared_lib_fixture_through_reference_api(gr, orc):
    """shared_lib_tests/test_topk.c; golden CMakeLists.txt:235-237: node 2 -> in 3, out 3"""
    ro = [0, 3, 6, 9, 11, 14, 15, 15]
    co = [1, 2, 3, 0, 2, 4, 3, 4, 5, 5, 6, 2, 5, 6, 6]
    ci = [0, 1, 2, 5, 7, 9, 12, 15]
    ri = [1, 0, 0, 1, 4, 0, 2, 1, 2, 2, 3, 4, 3, 4, 5]
    ids, ind, outd = gr.gunrock_topk_func(ro, co, ci, ri, 3)
    assert ids.tolist() == [2, 4, 0] and ind.tolist() == [3, 2, 1] and outd.tolist() == [3, 3, 3]
    for k in (1, 7, 50):                                 # top_nodes is clamped to n (topk_app.cu:175-176)
        got = gr.gunrock_topk_func(ro, co, ci, ri, k)
        want = orc.topk(ro, ci, k)
        assert all(np.array_equal(a, b) for a, b in zip(got, want))


@pytest.mark.parametrize("scale", [10, 16])
def test_topk_rmat_directed(gr, orc, scale):
    ro, co = rmat_graph(orc, scale, 8, undirected=False)
    ci = transpose_offsets(ro, co)
    for k in (1, 10, 1000):
        got = gr.topk(ro, ci, k)
        want = orc.topk(ro, ci, k)
        assert all(np.array_equal(a, b) for a, b in zip(got, want))
