"""gunrockinst_b200 -- host-side mirror of the reference's interface for the advance /
filter hot path, on top of libgunrock.so (hand-written sm_100a CUDA, C ABI).

Two layers, both thin:

* the reference's own C API, argument for argument (gunrock/gunrock.h:106-151):
  ``gunrock_bfs_func / gunrock_cc_func / gunrock_sssp_func / gunrock_pr_func`` taking
  host CSR arrays -- used by the parity tests the way shared_lib_tests/test_*.c use it;
* the resident-graph API of include/gunrock_b200.h (``Graph``, ``Bfs``, ``sssp``, ``cc``,
  ``pagerank``): the graph stays in HBM between runs, which is what bench.py times.

numpy arrays are host buffers; nothing in this package computes on the CPU and nothing
imports oracle/.  If libgunrock.so is missing, every call raises ``GunrockError``.
"""
import ctypes as C
import numpy as np

from . import _capi
from ._capi import GunrockError, Stats, BfsConfig, GraphInfo, LevelRecord, build  # noqa: F401

RMAT_ABCD = (0.55, 0.2, 0.2, 0.05)     # reference defaults, gunrock/graphio/rmat.cuh:32-35


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def version():
    return _capi.lib().grb_version().decode()


def release_cached_memory():
    """hand the library's cached device / pinned blocks back to the driver (grb_release_cached_memory)"""
    _capi.lib().grb_release_cached_memory()


def device_count():
    return int(_capi.lib().grb_device_count())


class Graph:
    """A CSR graph resident in HBM (grb_graph_t)."""

    def __init__(self, handle):
        self._h = handle
        self.info = GraphInfo()
        _capi.check(_capi.lib().grb_graph_get_info(self._h, C.byref(self.info)), "grb_graph_get_info")

    # -- constructors ----------------------------------------------------------------
    @classmethod
    def from_csr(cls, row_offsets, col_indices, weights=None, symmetric=False, device=0,
                 force_offset_bytes=0, stream_from_host=False):
        """stream_from_host: upload the offsets only; column ids and weights stay in (page-locked,
        mapped) host memory and are read over PCIe by the kernels (problem_base.cuh:252-276).  The
        graph object keeps the host arrays alive."""
        ro = np.ascontiguousarray(row_offsets)
        if ro.dtype not in (np.int32, np.int64):
            ro = ro.astype(np.int64)
        co = np.ascontiguousarray(col_indices, dtype=np.int32)
        w = np.ascontiguousarray(weights, dtype=np.uint32) if weights is not None else None
        h = C.c_void_p()
        fn = "grb_graph_from_host_mapped" if stream_from_host else "grb_graph_from_host"
        _capi.check(getattr(_capi.lib(), fn)(
            device, ro.size - 1, co.size, _ptr(ro), ro.dtype.itemsize, _ptr(co), _ptr(w),
            int(bool(symmetric)), force_offset_bytes, C.byref(h)), fn)
        g = cls(h)
        if stream_from_host:
            g._host_arrays = (co, w)
        return g

    @classmethod
    def from_slice(cls, row_offsets, col_indices, global_nodes, log_parts, rank, weights=None, symmetric=True, device=0):
        """one rank's slice of a 1D-partitioned graph from host arrays (local rows, global neighbour ids): the arrays
        Graph.rmat(..., log_parts, rank).download() returns"""
        ro = np.ascontiguousarray(row_offsets)
        if ro.dtype not in (np.int32, np.int64):
            ro = ro.astype(np.int64)
        co = np.ascontiguousarray(col_indices, dtype=np.int32)
        w = np.ascontiguousarray(weights, dtype=np.uint32) if weights is not None else None
        h = C.c_void_p()
        _capi.check(_capi.lib().grb_graph_slice_from_host(
            device, int(global_nodes), int(log_parts), int(rank), ro.size - 1, co.size, _ptr(ro), ro.dtype.itemsize,
            _ptr(co), _ptr(w), int(bool(symmetric)), 0, C.byref(h)), "grb_graph_slice_from_host")
        return cls(h)

    @classmethod
    def rmat(cls, scale, edgefactor=16, seed=None, undirected=True, abcd=RMAT_ABCD, device=0,
             log_parts=0, rank=0, weight_seed=0, force_offset_bytes=0):
        """Seeded R-MAT generated, sorted and de-duplicated on the device; seed defaults to
        42 + scale."""
        if seed is None:
            seed = 42 + scale
        h = C.c_void_p()
        _capi.check(_capi.lib().grb_graph_rmat(
            device, scale, edgefactor << scale, seed, int(bool(undirected)), *abcd,
            log_parts, rank, weight_seed, force_offset_bytes, C.byref(h)), "grb_graph_rmat")
        return cls(h)

    # -- accessors ---------------------------------------------------------------------
    @property
    def nodes(self):
        return int(self.info.nodes)

    @property
    def local_nodes(self):
        return int(self.info.local_nodes)

    @property
    def edges(self):
        return int(self.info.edges)

    def download(self, want_weights=False):
        """-> (row_offsets, col_indices[, weights]) as numpy arrays (row offsets in the
        graph's own width)."""
        ro = np.empty(self.local_nodes + 1, dtype=np.int32 if self.info.offset_bytes == 4 else np.int64)
        co = np.empty(max(self.edges, 1), dtype=np.int32)
        w = np.empty(max(self.edges, 1), dtype=np.uint32) if (want_weights and self.info.has_weights) else None
        _capi.check(_capi.lib().grb_graph_download(self._h, _ptr(ro), _ptr(co), _ptr(w)), "grb_graph_download")
        if want_weights:
            return ro, co[:self.edges], (w[:self.edges] if w is not None else None)
        return ro, co[:self.edges]

    def build_transpose(self):
        """attach the transpose (CSC) of a directed graph: the inverse graph of a direction-optimising BFS"""
        _capi.check(_capi.lib().grb_graph_build_transpose(self._h), "grb_graph_build_transpose")

    def build_twin(self):
        """attach the degree-ordered twin (same graph, hubs first): searches with degree_order=True and PageRank
        sweeps of a symmetric graph run on it; results stay in the caller's vertex ids"""
        _capi.check(_capi.lib().grb_graph_build_twin(self._h), "grb_graph_build_twin")

    def twin_entries(self):
        """entries of the attached twin (== `edges`), -1 without one"""
        out = C.c_int64(-1)
        _capi.check(_capi.lib().grb_graph_twin_entries(self._h, C.byref(out)), "grb_graph_twin_entries")
        return int(out.value)

    def check_symmetric(self):
        """True iff the resident CSR equals its transpose (what `symmetric=True` promises)"""
        out = C.c_int(0)
        _capi.check(_capi.lib().grb_graph_check_symmetric(self._h, C.byref(out)), "grb_graph_check_symmetric")
        return bool(out.value)

    def device_ptrs(self):
        a, b, c = C.c_void_p(), C.c_void_p(), C.c_void_p()
        _capi.check(_capi.lib().grb_graph_device_ptrs(self._h, C.byref(a), C.byref(b), C.byref(c)),
                    "grb_graph_device_ptrs")
        return a.value, b.value, c.value

    def stream(self):
        """cudaStream_t (as int) the library issues this graph's work on"""
        s = C.c_void_p()
        _capi.check(_capi.lib().grb_graph_stream(self._h, C.byref(s)), "grb_graph_stream")
        return s.value or 0

    def close(self):
        if self._h:
            _capi.lib().grb_graph_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Bfs:
    """BFS context on a resident graph (Problem + Enactor, created once, run many times).

    Mirrors the knobs of the reference's BFS / DOBFS test rigs (tests/bfs/test_bfs.cu:58-89,
    tests/dobfs/test_dobfs.cu:533-534): mark_pred, idempotence, traversal_mode, alpha/beta.
    """

    def __init__(self, graph, mark_pred=False, idempotence=False, traversal_mode=0,
                 direction_optimized=False, alpha=0.0, beta=0.0, queue_sizing=0.0, record_levels=False,
                 host_loop=False, degree_order=False, stream_min_edges=0):
        self.graph = graph
        cfg = BfsConfig(int(mark_pred), int(idempotence), int(traversal_mode), int(direction_optimized),
                        float(alpha), float(beta), float(queue_sizing), int(record_levels), int(host_loop),
                        int(degree_order), int(stream_min_edges))
        self.mark_pred = bool(mark_pred)
        self._h = C.c_void_p()
        _capi.check(_capi.lib().grb_bfs_create(graph._h, C.byref(cfg), C.byref(self._h)), "grb_bfs_create")

    def run(self, src, want_labels=True, labels_out=None):
        """-> (labels or None, preds or None, stats dict)."""
        n = self.graph.local_nodes
        labels = preds = None
        if want_labels:
            labels = labels_out if labels_out is not None else np.empty(n, dtype=np.int32)
            if self.mark_pred:
                preds = np.empty(n, dtype=np.int32)
        st = Stats()
        _capi.check(_capi.lib().grb_bfs_run(self._h, int(src), _ptr(labels), _ptr(preds), C.byref(st)),
                    "grb_bfs_run")
        return labels, preds, st.as_dict()

    def run_many(self, sources):
        """One search per source, back to back inside the library; -> list of stats dicts."""
        srcs = np.ascontiguousarray(sources, dtype=np.int32)
        st = (Stats * max(srcs.size, 1))()
        _capi.check(_capi.lib().grb_bfs_run_many(self._h, _ptr(srcs), int(srcs.size), st), "grb_bfs_run_many")
        return [st[i].as_dict() for i in range(srcs.size)]

    def device_labels(self):
        p = C.c_void_p()
        _capi.check(_capi.lib().grb_bfs_device_labels(self._h, C.byref(p)), "grb_bfs_device_labels")
        return p.value

    def levels(self):
        cnt = C.c_int32()
        L = _capi.lib()
        _capi.check(L.grb_bfs_levels(self._h, None, 0, C.byref(cnt)), "grb_bfs_levels")
        arr = (LevelRecord * max(cnt.value, 1))()
        _capi.check(L.grb_bfs_levels(self._h, arr, cnt.value, C.byref(cnt)), "grb_bfs_levels")
        kinds = {0: "push-twc", 1: "pull", 2: "push-lb", 3: "push-small", 4: "filter", 5: "push-stream", 6: "filter-stream"}
        return [dict(level=r.level, kind=kinds.get(r.kind, str(r.kind)), frontier=r.frontier, edges=r.edges,
                     discovered=r.discovered, candidates=r.candidates, ms=r.ms, switch_ms=r.switch_ms) for r in arr[:cnt.value]]

    def close(self):
        if self._h:
            _capi.lib().grb_bfs_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def sssp(graph, src, delta_factor=1, mark_pred=True):
    """-> (dist uint32[n], preds int32[n] or None, stats)."""
    n = graph.local_nodes
    dist = np.empty(n, dtype=np.uint32)
    preds = np.empty(n, dtype=np.int32) if mark_pred else None
    st = Stats()
    _capi.check(_capi.lib().grb_sssp(graph._h, int(src), int(delta_factor), int(bool(mark_pred)),
                                     _ptr(dist), _ptr(preds), C.byref(st)), "grb_sssp")
    return dist, preds, st.as_dict()


def cc(graph):
    """-> (labels int32[n] = min vertex id of the component, num_components, stats)."""
    labels = np.empty(graph.local_nodes, dtype=np.int32)
    ncomp = C.c_int64()
    st = Stats()
    _capi.check(_capi.lib().grb_cc(graph._h, _ptr(labels), C.byref(ncomp), C.byref(st)), "grb_cc")
    return labels, int(ncomp.value), st.as_dict()


def pagerank(graph, src=-1, delta=0.85, threshold=1e-6, max_iter=1000, want_sorted=False):
    """-> (ranks float64[n] in vertex order, (sorted_ids, sorted_ranks) or None, stats)."""
    n = graph.local_nodes
    ranks = np.empty(n, dtype=np.float64)
    ids = np.empty(n, dtype=np.int32) if want_sorted else None
    sr = np.empty(n, dtype=np.float32) if want_sorted else None
    st = Stats()
    _capi.check(_capi.lib().grb_pagerank(graph._h, int(src), float(delta), float(threshold), int(max_iter),
                                         _ptr(ranks), _ptr(ids), _ptr(sr), C.byref(st)), "grb_pagerank")
    return ranks, ((ids, sr) if want_sorted else None), st.as_dict()


def bc(graph, src=-1, want_sigmas=False):
    """Betweenness centrality -> (bc float32[n] (x0.5; src < 0: every source), sigmas or None, stats)."""
    n = graph.local_nodes
    vals = np.empty(max(n, 1), dtype=np.float32)
    sig = np.empty(max(n, 1), dtype=np.float32) if want_sigmas else None
    st = Stats()
    _capi.check(_capi.lib().grb_bc(graph._h, int(src), _ptr(vals), _ptr(sig), None, C.byref(st)), "grb_bc")
    return vals[:n], (sig[:n] if want_sigmas else None), st.as_dict()


def topk(row_offsets, col_offsets, top_nodes, device=0):
    """Top-k degree centrality from the offsets of a graph and of its transpose
    -> (node_ids, in_degrees, out_degrees), int32[min(k, n)] each."""
    ro = np.ascontiguousarray(row_offsets, dtype=np.int32)
    ci = np.ascontiguousarray(col_offsets, dtype=np.int32)
    n = ro.size - 1
    k = max(min(int(top_nodes), n), 0)
    ids, ind, outd = (np.empty(max(k, 1), dtype=np.int32) for _ in range(3))
    _capi.check(_capi.lib().grb_topk(device, n, _ptr(ro), _ptr(ci), int(top_nodes), _ptr(ids), _ptr(ind),
                                     _ptr(outd), None), "grb_topk")
    return ids[:k], ind[:k], outd[:k]


# ---- the reference's C API, argument for argument ---------------------------------------
def _gunrock_graph(row_offsets, col_indices, edge_values=None):
    ro = np.ascontiguousarray(row_offsets, dtype=np.int32)
    co = np.ascontiguousarray(col_indices, dtype=np.int32)
    ev = np.ascontiguousarray(edge_values, dtype=np.uint32) if edge_values is not None else None
    g = _capi.GunrockGraph()
    g.num_nodes = ro.size - 1
    g.num_edges = co.size
    g.row_offsets = ro.ctypes.data
    g.col_indices = co.ctypes.data
    g.edge_values = ev.ctypes.data if ev is not None else None
    return g, (ro, co, ev)


def _config(**kw):
    c = _capi.GunrockConfig()
    c.device = kw.get("device", 0)
    c.mark_pred = kw.get("mark_pred", False)
    c.idempotence = kw.get("idempotence", False)
    c.src_mode = kw.get("src_mode", _capi.SRC_MANUALLY)
    c.src_node = kw.get("src_node", 0)
    c.queue_size = kw.get("queue_size", 1.0)
    c.delta = kw.get("delta", 0.85)
    c.error = kw.get("error", 0.01)
    c.max_iter = kw.get("max_iter", 20)
    c.top_nodes = kw.get("top_nodes", 0)
    c.delta_factor = kw.get("delta_factor", 1)
    return c


class _MallocBlock:
    """Owner of a buffer the library malloc()ed for the caller: free()s it when the last view goes."""

    def __init__(self, ptr):
        self.ptr = ptr

    def __del__(self):
        ptr, self.ptr = self.ptr, None
        if ptr and _capi is not None and getattr(_capi, "free_host", None) is not None:   # (module teardown)
            _capi.free_host(ptr)          # caller frees (shared_lib_tests/test_bfs.c:58,66)


def _take_node_values(out, n, dtype):
    """The library's result buffer as a numpy array, without a second host copy: the array keeps the
    malloc()ed block alive and releases it with free() when it is collected."""
    if not out.node_values:
        raise GunrockError("library returned no node_values (see stderr)")
    block = _MallocBlock(out.node_values)
    buf = (C.c_int32 * max(n, 1)).from_address(out.node_values)
    buf._block = block                      # numpy holds `buf` as the array's base, `buf` holds the block
    return np.frombuffer(buf, dtype=np.int32)[:n].view(dtype)


def gunrock_bfs_func(row_offsets, col_indices, **cfg):
    """gunrock_bfs_func on host CSR arrays -> int32 depths (gunrock.h:106-110)."""
    g, keep = _gunrock_graph(row_offsets, col_indices)
    out = _capi.GunrockGraph()
    dt = _capi.GunrockDataType(_capi.VTXID_INT, _capi.SIZET_INT, _capi.VALUE_INT)
    _capi.lib().gunrock_bfs_func(C.byref(out), C.byref(g), _config(**cfg), dt)
    return _take_node_values(out, int(g.num_nodes), np.int32)


def gunrock_cc_func(row_offsets, col_indices, **cfg):
    g, keep = _gunrock_graph(row_offsets, col_indices)
    out = _capi.GunrockGraph()
    dt = _capi.GunrockDataType(_capi.VTXID_INT, _capi.SIZET_INT, _capi.VALUE_INT)
    _capi.lib().gunrock_cc_func(C.byref(out), C.byref(g), _config(**cfg), dt)
    return _take_node_values(out, int(g.num_nodes), np.int32)


def gunrock_bc_func(row_offsets, col_indices, **cfg):
    """-> (bc float32[n], ebc float32[m]) (gunrock.h:112-116); src_node defaults to -1 = every source."""
    g, keep = _gunrock_graph(row_offsets, col_indices)
    out = _capi.GunrockGraph()
    dt = _capi.GunrockDataType(_capi.VTXID_INT, _capi.SIZET_INT, _capi.VALUE_FLOAT)
    cfg.setdefault("src_node", -1)
    _capi.lib().gunrock_bc_func(C.byref(out), C.byref(g), _config(**cfg), dt)
    m = int(g.num_edges)
    if not out.edge_values:
        raise GunrockError("library returned no edge_values (see stderr)")
    ebc = np.ctypeslib.as_array(C.cast(out.edge_values, C.POINTER(C.c_float)), shape=(max(m, 1),))[:m].copy()
    _capi.free_host(out.edge_values)
    return _take_node_values(out, int(g.num_nodes), np.float32), ebc


def gunrock_topk_func(row_offsets, col_indices, col_offsets, row_indices, top_nodes, **cfg):
    """-> (node_ids, in_degrees, out_degrees) int32[top_nodes] (gunrock.h:133-141)."""
    g, keep = _gunrock_graph(row_offsets, col_indices)
    ci = np.ascontiguousarray(col_offsets, dtype=np.int32)
    ri = np.ascontiguousarray(row_indices, dtype=np.int32)
    g.col_offsets = ci.ctypes.data
    g.row_indices = ri.ctypes.data
    n = int(g.num_nodes)
    k = max(min(int(top_nodes), n), 0)
    ids, ind, outd = (np.full(max(int(top_nodes), 1), -1, dtype=np.int32) for _ in range(3))
    out = _capi.GunrockGraph()
    dt = _capi.GunrockDataType(_capi.VTXID_INT, _capi.SIZET_INT, _capi.VALUE_INT)
    _capi.lib().gunrock_topk_func(C.byref(out), _ptr(ids), _ptr(ind), _ptr(outd), C.byref(g),
                                  _config(top_nodes=int(top_nodes), **cfg), dt)
    return ids[:k], ind[:k], outd[:k]


def gunrock_sssp_func(row_offsets, col_indices, edge_values, **cfg):
    """-> (dist uint32[n], preds int32[n]) (gunrock.h:126-131)."""
    g, keep = _gunrock_graph(row_offsets, col_indices, edge_values)
    n = int(g.num_nodes)
    out = _capi.GunrockGraph()
    preds = np.empty(max(n, 1), dtype=np.int32)
    dt = _capi.GunrockDataType(_capi.VTXID_INT, _capi.SIZET_INT, _capi.VALUE_UINT)
    _capi.lib().gunrock_sssp_func(C.byref(out), _ptr(preds), C.byref(g), _config(**cfg), dt)
    return _take_node_values(out, n, np.uint32), preds[:n]


def gunrock_pr_func(row_offsets, col_indices, **cfg):
    """-> (node_ids int32[k], ranks float32[k]) sorted by rank descending; k = top_nodes or n."""
    g, keep = _gunrock_graph(row_offsets, col_indices)
    n = int(g.num_nodes)
    out = _capi.GunrockGraph()
    ids = np.empty(max(n, 1), dtype=np.int32)
    ranks = np.empty(max(n, 1), dtype=np.float32)
    dt = _capi.GunrockDataType(_capi.VTXID_INT, _capi.SIZET_INT, _capi.VALUE_FLOAT)
    c = _config(**cfg)
    _capi.lib().gunrock_pr_func(C.byref(out), _ptr(ids), _ptr(ranks), C.byref(g), c, dt)
    k = min(n, c.top_nodes) if c.top_nodes > 0 else n
    return ids[:k], ranks[:k]
