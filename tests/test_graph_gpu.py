"""GPU: resident graphs -- upload/download round trip and the device R-MAT builder against
the oracle's generator + canonicaliser (same seeded draws, bit-identical CSR)."""
import numpy as np
import pytest

from conftest import load_golden, rmat_graph

pytestmark = pytest.mark.gpu


def owner_of(v, log_parts):
    """Python restatement of graph::Partition::Owner (gunrockinst_b200/csrc/gunrock/graph/csr.cuh)"""
    v = np.asarray(v, dtype=np.int64)
    if log_parts == 0:
        return np.zeros_like(v)
    local = v >> log_parts
    mix = ((local.astype(np.uint64) * np.uint64(0x9E3779B1)) & np.uint64(0xFFFFFFFF)) >> np.uint64(32 - log_parts)
    return ((v & ((1 << log_parts) - 1)) ^ mix.astype(np.int64)).astype(np.int64)


def test_upload_download_roundtrip(gr):
    n, ro, co, vals = load_golden("bips98_606")
    for fob in (4, 8):
        g = gr.Graph.from_csr(ro, co, weights=vals.astype(np.uint32), force_offset_bytes=fob)
        assert (g.nodes, g.edges, g.info.offset_bytes, g.info.has_weights) == (n, co.size, fob, 1)
        ro2, co2, w2 = g.download(want_weights=True)
        assert np.array_equal(ro2, ro) and np.array_equal(co2, co) and np.array_equal(w2, vals.astype(np.uint32))
        g.close()


@pytest.mark.parametrize("scale,edgefactor", [(6, 4), (10, 16), (14, 16), (16, 8)])
@pytest.mark.parametrize("undirected", [True, False])
def test_device_rmat_matches_oracle(gr, orc, scale, edgefactor, undirected):
    want_ro, want_co = orc.rmat_csr(scale, edgefactor, undirected=undirected)
    g = gr.Graph.rmat(scale, edgefactor, undirected=undirected, weight_seed=99)
    ro, co, w = g.download(want_weights=True)
    assert g.nodes == 1 << scale and g.info.symmetric == int(undirected)
    assert np.array_equal(ro, want_ro) and np.array_equal(co, want_co)
    assert np.array_equal(w, orc.edge_weights(want_ro, want_co, 99))
    g.close()


@pytest.mark.parametrize("log_parts", [1, 2, 3])
def test_device_rmat_partitioned_slices(gr, orc, log_parts):
    scale = 13
    want_ro, want_co = rmat_graph(orc, scale)
    n = 1 << scale
    owner = owner_of(np.arange(n), log_parts)
    deg = np.diff(want_ro)
    seen = np.zeros(n, bool)
    total = 0
    for rank in range(1 << log_parts):
        g = gr.Graph.rmat(scale, log_parts=log_parts, rank=rank, force_offset_bytes=8 if rank == 0 else 0)
        assert g.local_nodes == n >> log_parts and (g.info.log_parts, g.info.rank) == (log_parts, rank)
        ro, co = g.download()
        mine = np.nonzero(owner == rank)[0]
        assert mine.size == g.local_nodes                   # the striping is a bijection
        assert np.array_equal(mine >> log_parts, np.arange(g.local_nodes))
        assert np.array_equal(np.diff(ro), deg[mine])
        for local in list(range(0, g.local_nodes, 97)) + [0, g.local_nodes - 1]:
            v = mine[local]
            assert np.array_equal(co[ro[local]:ro[local + 1]], want_co[want_ro[v]:want_ro[v + 1]])
        seen[mine] = True
        total += g.edges
        g.close()
    assert seen.all() and total == want_co.size
    # the hash evens out R-MAT's bit skew: no device holds far more than its share
    share = np.bincount(owner, weights=deg, minlength=1 << log_parts) / deg.sum()
    assert share.max() < 1.35 / (1 << log_parts)


def test_device_rmat_scale20_statistics(gr):
    g = gr.Graph.rmat(20)
    ro, co = g.download()
    n = g.nodes
    assert ro[0] == 0 and ro[-1] == co.size == g.edges
    src_of = np.repeat(np.arange(n, dtype=np.int64), np.diff(ro))
    key = src_of * n + co
    assert np.all(np.diff(key) > 0) and not np.any(src_of == co)
    # symmetric: the multiset of (col,row) keys equals the multiset of (row,col) keys
    rkey = np.sort(co.astype(np.int64) * n + src_of)
    assert np.array_equal(rkey, key)
    assert 0.9 * 32 * n < g.edges < 32 * n
    g.close()


@pytest.mark.parametrize("pinned", [False, True])
def test_stream_from_host_matches_resident_graph(gr, orc, pinned):
    """grb_graph_from_host_mapped (the reference's stream_from_host, problem_base.cuh:252-276): column ids
    and weights stay in host memory; every primitive must return what it returns on the uploaded copy."""
    ro, co = rmat_graph(orc, 14)
    w = orc.edge_weights(ro, co, 5)
    if pinned:                                          # memory the caller page-locked itself
        import torch
        co = torch.from_numpy(co).pin_memory().numpy()
        w = torch.from_numpy(w).pin_memory().numpy()
    g = gr.Graph.from_csr(ro, co, weights=w, symmetric=True, stream_from_host=True)
    assert (g.nodes, g.edges, g.info.has_weights) == (ro.size - 1, co.size, 1)
    ro2, co2, w2 = g.download(want_weights=True)
    assert np.array_equal(ro2, ro) and np.array_equal(co2, co) and np.array_equal(w2, w)
    for mode in (dict(), dict(direction_optimized=True), dict(idempotence=True), dict(mark_pred=True)):
        b = gr.Bfs(g, **mode)
        for src in (0, 77):
            labels, _, st = b.run(src)
            want = orc.bfs(ro, co, src)
            assert np.array_equal(labels, want), mode
            assert st["edges_visited"] == orc.edges_visited(ro, want)
        b.close()
    dist, _, _ = gr.sssp(g, 0)
    want, _ = orc.sssp(ro, co, w, 0)
    assert np.array_equal(dist, want)
    labels, ncomp, _ = gr.cc(g)
    want_l, want_n = orc.cc(ro, co)
    assert np.array_equal(labels, want_l) and ncomp == want_n
    ranks, _, st = gr.pagerank(g)
    want_r, it = orc.pagerank(ro, co, 0.85, 1e-6, 1000, -1)
    assert st["iterations"] == it and np.max(np.abs(ranks - want_r)) <= 1e-6 * max(1.0, float(want_r.max()))
    g.close()


def test_slice_from_host_round_trip(gr, orc):
    """grb_graph_slice_from_host: a rank's slice of a partitioned graph uploaded from host arrays is the slice the
    device generator built (same arrays back, same partition in the info block), and a world-1 'slice' is the graph"""
    import numpy as np
    for lp in (1, 2):
        for rank in range(1 << lp):
            a = gr.Graph.rmat(12, 8, log_parts=lp, rank=rank)
            ro, co = a.download()
            b = gr.Graph.from_slice(ro, co, a.nodes, lp, rank)
            ro2, co2 = b.download()
            assert np.array_equal(ro, ro2) and np.array_equal(co, co2)
            assert (b.nodes, b.local_nodes, b.edges) == (a.nodes, a.local_nodes, a.edges)
            assert (int(b.info.log_parts), int(b.info.rank)) == (lp, rank)
            a.close()
            b.close()
    ro, co = orc.rmat_csr(10, 8)
    g = gr.Graph.from_slice(ro, co, ro.size - 1, 0, 0)
    labels, _, _ = gr.Bfs(g, direction_optimized=True).run(0)
    assert np.array_equal(labels, orc.bfs(ro, co, 0))
    g.close()
    with pytest.raises(Exception):
        gr.Graph.from_slice(ro, co, ro.size, 1, 0)             # local_nodes << log_parts != global_nodes
