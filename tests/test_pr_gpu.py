"""GPU: PageRank through the C ABI against the oracle's fp64 restatement of the reference's
functor semantics.  Tolerance (BASELINE.json north_star): 1e-6 per vertex on the un-normalised
ranks -- |got[v] - want[v]| <= 1e-6 * max(1, |want[v]|), every vertex against its own value, no
graph-wide scale -- and the same number of power iterations must be executed."""
import numpy as np
import pytest

from conftest import load_golden, rmat_graph

pytestmark = pytest.mark.gpu
TOL = 1e-6


def assert_per_vertex(got, want, tol=TOL):
    """every vertex within tol of its own reference value (relative above 1, absolute below)"""
    err = np.abs(np.asarray(got, dtype=np.float64) - want)
    bound = tol * np.maximum(1.0, np.abs(want))
    bad = np.nonzero(err > bound)[0]
    assert bad.size == 0, (int(bad.size), int(bad[0]), float(err[bad[0]]), float(want[bad[0]]))


def _check(gr, orc, ro, co, symmetric, twin=False, **kw):
    src = kw.get("src", -1)
    want, it = orc.pagerank(ro, co, kw.get("delta", 0.85), kw.get("threshold", 1e-6), kw.get("max_iter", 1000), src)
    g = gr.Graph.from_csr(ro, co, symmetric=symmetric, force_offset_bytes=kw.get("offset_bytes", 0))
    if twin:
        g.build_twin()
        assert g.twin_entries() == g.edges          # same entries: grb_pagerank sweeps the twin
    ranks, srt, st = gr.pagerank(g, src=src, delta=kw.get("delta", 0.85), threshold=kw.get("threshold", 1e-6),
                                 max_iter=kw.get("max_iter", 1000), want_sorted=True)
    assert st["iterations"] == it
    assert_per_vertex(ranks, want)
    ids, sr = srt
    assert sorted(ids.tolist()) == list(range(ro.size - 1))
    assert np.all(np.diff(sr) <= 0)                                    # descending
    assert np.array_equal(sr, ranks[ids].astype(np.float32))
    ties = np.nonzero(np.diff(sr) == 0)[0]
    assert np.all(ids[ties] < ids[ties + 1])                           # stable: ties by ascending id
    g.close()
    return ranks, st


@pytest.mark.parametrize("name", ["bips98_606", "chesapeake", "test_cc", "test_pr", "test_pr_dir", "test_bc_dir",
                                  "bips98_606_dir"])
def test_fixtures(gr, orc, name):
    n, ro, co, _ = load_golden(name)
    sym = not name.endswith("_dir")
    _check(gr, orc, ro, co, sym)
    _check(gr, orc, ro, co, False, src=0, threshold=0.01, max_iter=20)      # test_pr.c's configuration
    _check(gr, orc, ro, co, False, max_iter=3)                               # stops on the iteration cap


def test_shared_lib_fixture_through_reference_api(gr, orc, shared_lib):
    """shared_lib_tests/test_pr.c: personalised (src 0), delta .85, error .01, max_iter 20, top 10"""
    ro, co = shared_lib["row_offsets"], shared_lib["col_indices"]
    want, _ = orc.pagerank(np.array(ro), np.array(co), 0.85, 0.01, 20, 0)
    ids, ranks = gr.gunrock_pr_func(ro, co, src_node=0, delta=0.85, error=0.01, max_iter=20, top_nodes=10)
    assert ids.size == 7                                               # min(n, top_nodes)
    order = np.lexsort((np.arange(7), -want.astype(np.float32)))
    assert ids.tolist() == order.tolist()
    assert np.allclose(ranks, want[order], atol=1e-6)
    ids3, _ = gr.gunrock_pr_func(ro, co, src_node=0, delta=0.85, error=0.01, max_iter=20, top_nodes=3)
    assert ids3.tolist() == order[:3].tolist()


@pytest.mark.parametrize("scale,ef,undirected", [(10, 16, True), (16, 16, True), (14, 4, False), (18, 8, True)])
def test_rmat(gr, orc, scale, ef, undirected):
    ro, co = rmat_graph(orc, scale, ef, undirected)
    ranks, st = _check(gr, orc, ro, co, undirected)
    assert st["iterations"] > 3
    if not undirected:
        assert st["pull_levels"] > 2                                   # pruning cascaded over several sweeps


@pytest.mark.parametrize("name", ["bips98_606", "chesapeake", "test_cc", "test_pr"])
def test_fixtures_on_degree_ordered_twin(gr, orc, name):
    """grb_graph_build_twin + grb_pagerank: the sweeps run on the relabelled graph, the source goes in and the ranks
    and the sorted pairs come out in the caller's ids"""
    n, ro, co, _ = load_golden(name)
    _check(gr, orc, ro, co, True, twin=True)
    _check(gr, orc, ro, co, True, twin=True, src=min(3, n - 1), threshold=0.01, max_iter=20)


@pytest.mark.parametrize("scale,ef", [(10, 16), (16, 16), (18, 8)])
def test_rmat_on_degree_ordered_twin(gr, orc, scale, ef):
    ro, co = rmat_graph(orc, scale, ef, True)
    _check(gr, orc, ro, co, True, twin=True)
    _check(gr, orc, ro, co, True, twin=True, src=7, max_iter=30)


@pytest.mark.parametrize("twin", [False, True])
def test_rmat_with_64_bit_offsets(gr, orc, twin):
    """the row-binned sweep instantiated for 64-bit row offsets (graphs past 2^31 entries use it)"""
    ro, co = rmat_graph(orc, 14, 16, True)
    _check(gr, orc, ro, co, True, twin=twin, offset_bytes=8, max_iter=40)


def test_twin_keeps_duplicate_entries(gr, orc):
    """a row that lists a neighbour twice counts it twice; the twin is the same multiset of entries (nothing is
    merged), so the sweeps on it give the oracle's ranks"""
    ro = np.array([0, 4, 6, 8], np.int64)
    co = np.array([1, 1, 2, 2, 0, 0, 0, 0], np.int32)       # edges 0-1 and 0-2, each listed twice on both sides
    _check(gr, orc, ro, co, True, twin=True)


def test_edge_cases(gr, orc):
    for ro, co in [([0, 0], []), ([0, 0, 0, 0], []), ([0, 1, 2, 2], [1, 0]), ([0, 1, 1], [1])]:
        ro, co = np.array(ro), np.array(co, dtype=np.int32)
        _check(gr, orc, ro, co, False)


def test_rmat_scale22_to_convergence(gr, orc):
    """BASELINE configs[1]'s graph, run to convergence (threshold 1e-6, ~130 sweeps): same sweep count as the fp64
    restatement and every vertex within 1e-6 of its own value.  The graph comes from the
    device generator (the same seeded CSR oracle.rmat_csr builds in minutes on one core)."""
    g = gr.Graph.rmat(22)
    ro, co = g.download()
    ranks, _, st = gr.pagerank(g, delta=0.85, threshold=1e-6, max_iter=1000)
    g.close()
    want, it = orc.pagerank(ro, co, 0.85, 1e-6, 1000, -1)
    assert st["iterations"] == it and it > 50
    assert_per_vertex(ranks, want)
    assert abs(float(ranks.sum()) - float(want.sum())) <= 1e-9 * float(want.sum())
