"""numpy-facing wrapper around oracle/liboracle.so (the CPU restatement in oracle.c).

TEST INFRASTRUCTURE ONLY.  May be imported from tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs -- never from the
gunrockinst_b200 package (the product path has no CPU fallback).
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")
REF_GRAPHIO = os.path.join(_HERE, "_ref", "ref_graphio")


def build(force=False):
    """Compile liboracle.so (and oracle/_ref when /root/reference is present)."""
    src = os.path.join(_HERE, "oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])
    if os.path.isdir("/root/reference/gunrock") and not os.path.exists(REF_GRAPHIO):
        subprocess.call(["make", "-s", "-C", _HERE, "ref"])


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        p, i64, i32, u64, dbl = C.c_void_p, C.c_int64, C.c_int32, C.c_uint64, C.c_double
        L.orc_now_seconds.restype = dbl
        L.orc_coo_to_csr.restype = i64
        L.orc_coo_to_csr.argtypes = [i64, i64, p, p, p, p, p, p]
        L.orc_market_read.restype = C.c_int
        L.orc_market_read.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, p, p, p, p, p]
        L.orc_free.argtypes = [p]
        L.orc_csr_cache_write.restype = C.c_int
        L.orc_csr_cache_write.argtypes = [C.c_char_p, i64, i64, p, p, p]
        L.orc_csr_cache_read.restype = C.c_int
        L.orc_csr_cache_read.argtypes = [C.c_char_p, C.c_int, p, p, p, p, p]
        L.orc_random_node.restype = i32
        L.orc_random_node.argtypes = [i32]
        L.orc_largest_degree_node.restype = i32
        L.orc_largest_degree_node.argtypes = [i64, p]
        L.orc_bfs.restype = C.c_int
        L.orc_bfs.argtypes = [i64, p, p, i32, p, p]
        L.orc_bfs_sample.restype = C.c_int
        L.orc_bfs_sample.argtypes = [i64, p, p, i32, p, p, i64, C.POINTER(i64)]
        L.orc_edges_visited.restype = i64
        L.orc_edges_visited.argtypes = [i64, p, p]
        L.orc_sssp.restype = None
        L.orc_sssp.argtypes = [i64, p, p, p, i32, p, p]
        L.orc_cc.restype = i64
        L.orc_cc.argtypes = [i64, p, p, p]
        L.orc_pagerank.restype = C.c_int
        L.orc_pagerank.argtypes = [i64, p, p, dbl, dbl, C.c_int, i32, p]
        L.orc_bc.restype = None
        L.orc_bc.argtypes = [i64, p, p, i32, p, p, p]
        L.orc_topk.restype = None
        L.orc_topk.argtypes = [i64, p, p, i64, p, p, p]
        L.orc_rmat_draw.restype = u64
        L.orc_rmat_draw.argtypes = [u64, u64, C.c_uint32, C.c_uint32]
        L.orc_rmat_edges.restype = None
        L.orc_rmat_edges.argtypes = [C.c_int, u64, u64, u64, dbl, dbl, dbl, dbl, p, p]
        L.orc_edge_weight.restype = C.c_uint32
        L.orc_edge_weight.argtypes = [i32, i32, u64]
        L.orc_edge_weights.restype = None
        L.orc_edge_weights.argtypes = [i64, p, p, u64, p]
        L.orc_rmat_csr.restype = i64
        L.orc_rmat_csr.argtypes = [C.c_int, i64, u64, C.c_int, dbl, dbl, dbl, dbl, p, p]
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _offs(row_offsets):
    return np.ascontiguousarray(row_offsets, dtype=np.int64)


def _cols(col):
    return np.ascontiguousarray(col, dtype=np.int32)


RMAT_ABCD = (0.55, 0.2, 0.2, 0.05)     # reference defaults, graphio/rmat.cuh:32-35


def coo_to_csr(n, rows, cols, vals=None):
    rows = np.ascontiguousarray(rows, dtype=np.int32)
    cols = np.ascontiguousarray(cols, dtype=np.int32)
    k = rows.size
    ro = np.zeros(n + 1, dtype=np.int64)
    co = np.zeros(max(k, 1), dtype=np.int32)
    v_in = np.ascontiguousarray(vals, dtype=np.int64) if vals is not None else None
    vo = np.zeros(max(k, 1), dtype=np.int64) if vals is not None else None
    m = lib().orc_coo_to_csr(n, k, _ptr(rows), _ptr(cols), _ptr(v_in), _ptr(ro), _ptr(co), _ptr(vo))
    return ro, co[:m].copy(), (vo[:m].copy() if vo is not None else None)


def market_read(path, undirected=True, reversed_=False, load_values=True):
    """-> (n, row_offsets int64[n+1], col int32[m], values int64[m])"""
    L = lib()
    n, k = C.c_int64(), C.c_int64()
    pr, pc, pv = C.c_void_p(), C.c_void_p(), C.c_void_p()
    rc = L.orc_market_read(os.fsencode(path), int(undirected), int(reversed_), int(load_values),
                           C.byref(n), C.byref(k), C.byref(pr), C.byref(pc), C.byref(pv))
    if rc != 0:
        raise IOError("orc_market_read(%s) failed: %d" % (path, rc))
    kk = k.value
    rows = np.ctypeslib.as_array(C.cast(pr, C.POINTER(C.c_int32)), shape=(max(kk, 1),))[:kk].copy()
    cols = np.ctypeslib.as_array(C.cast(pc, C.POINTER(C.c_int32)), shape=(max(kk, 1),))[:kk].copy()
    vals = np.ctypeslib.as_array(C.cast(pv, C.POINTER(C.c_int64)), shape=(max(kk, 1),))[:kk].copy()
    for q in (pr, pc, pv):
        L.orc_free(q)
    ro, co, vo = coo_to_csr(n.value, rows, cols, vals)
    return n.value, ro, co, vo


def csr_cache_name(market_path, undirected=True, reversed_=False, directory=None):
    """the reference's cache file of a built graph (graphio/market.cuh:312-326): .<name>_{undirected,reversed,
    nonreversed}_csr next to the .mtx, or in `directory`"""
    d, f = os.path.split(os.path.abspath(market_path))
    kind = "undirected" if undirected else ("reversed" if reversed_ else "nonreversed")
    return os.path.join(directory or d, ".%s_%s_csr" % (f, kind))


def csr_cache_write(path, row_offsets, col, values=None):
    """the reference's text CSR (csr.cuh:140-176)"""
    ro, co = _offs(row_offsets), _cols(col)
    v = np.ascontiguousarray(values, dtype=np.int64) if values is not None else None
    rc = lib().orc_csr_cache_write(os.fsencode(path), ro.size - 1, co.size, _ptr(ro), _ptr(co), _ptr(v))
    if rc != 0:
        raise IOError("orc_csr_cache_write(%s) failed: %d" % (path, rc))


def csr_cache_read(path, load_values=False):
    """-> (n, row_offsets int64[n+1], col int32[m], values int64[m] or None) (csr.cuh:182-232)"""
    L = lib()
    n, m = C.c_int64(), C.c_int64()
    pr, pc, pv = C.c_void_p(), C.c_void_p(), C.c_void_p()
    rc = L.orc_csr_cache_read(os.fsencode(path), int(load_values), C.byref(n), C.byref(m), C.byref(pr), C.byref(pc), C.byref(pv))
    if rc != 0:
        raise IOError("orc_csr_cache_read(%s) failed: %d" % (path, rc))
    ro = np.ctypeslib.as_array(C.cast(pr, C.POINTER(C.c_int64)), shape=(n.value + 1,)).copy()
    co = np.ctypeslib.as_array(C.cast(pc, C.POINTER(C.c_int32)), shape=(max(m.value, 1),))[:m.value].copy()
    vo = np.ctypeslib.as_array(C.cast(pv, C.POINTER(C.c_int64)), shape=(max(m.value, 1),))[:m.value].copy() if pv.value else None
    for q in (pr, pc, pv):
        if q.value:
            L.orc_free(q)
    return n.value, ro, co, vo


def random_node(n):
    return int(lib().orc_random_node(n))


def largest_degree_node(row_offsets):
    ro = _offs(row_offsets)
    return int(lib().orc_largest_degree_node(ro.size - 1, _ptr(ro)))


def bfs(row_offsets, col, src, want_preds=False):
    ro, co = _offs(row_offsets), _cols(col)
    n = ro.size - 1
    labels = np.empty(n, dtype=np.int32)
    preds = np.empty(n, dtype=np.int32) if want_preds else None
    depth = lib().orc_bfs(n, _ptr(ro), _ptr(co), int(src), _ptr(labels), _ptr(preds))
    return (labels, preds, depth) if want_preds else labels


def bfs_timed(row_offsets, col, src):
    """-> (labels, seconds): brackets the traversal only, as the reference's
    cpu_timer does (tests/bfs/test_bfs.cu:282-317)."""
    ro, co = _offs(row_offsets), _cols(col)
    n = ro.size - 1
    labels = np.empty(n, dtype=np.int32)
    L = lib()
    t0 = L.orc_now_seconds()
    L.orc_bfs(n, _ptr(ro), _ptr(co), int(src), _ptr(labels), None)
    t1 = L.orc_now_seconds()
    return labels, t1 - t0


class BfsSampler:
    """Repeated (optionally edge-bounded) CPU BFS on one CSR for bench.py's CPU legs:
    arrays are converted once, scratch is preallocated, and each call is timed around
    the traversal only (tests/bfs/test_bfs.cu:282-317)."""

    def __init__(self, row_offsets, col):
        self.ro, self.co = _offs(row_offsets), _cols(col)
        self.n = self.ro.size - 1
        self.labels = np.empty(max(self.n, 1), dtype=np.int32)
        self.queue = np.empty(max(self.n, 1), dtype=np.int32)

    def run(self, src, max_edges=0):
        """-> (edges_scanned, seconds)"""
        L = lib()
        scanned = C.c_int64()
        t0 = L.orc_now_seconds()
        L.orc_bfs_sample(self.n, _ptr(self.ro), _ptr(self.co), int(src), _ptr(self.labels), _ptr(self.queue),
                         int(max_edges), C.byref(scanned))
        t1 = L.orc_now_seconds()
        return int(scanned.value), t1 - t0


def edges_visited(row_offsets, labels):
    ro = _offs(row_offsets)
    lb = np.ascontiguousarray(labels, dtype=np.int32)
    return int(lib().orc_edges_visited(ro.size - 1, _ptr(ro), _ptr(lb)))


def sssp(row_offsets, col, weights, src):
    ro, co = _offs(row_offsets), _cols(col)
    w = np.ascontiguousarray(weights, dtype=np.uint32)
    n = ro.size - 1
    dist = np.empty(n, dtype=np.uint32)
    preds = np.empty(n, dtype=np.int32)
    lib().orc_sssp(n, _ptr(ro), _ptr(co), _ptr(w), int(src), _ptr(dist), _ptr(preds))
    return dist, preds


def cc(row_offsets, col):
    ro, co = _offs(row_offsets), _cols(col)
    n = ro.size - 1
    labels = np.empty(n, dtype=np.int32)
    ncomp = lib().orc_cc(n, _ptr(ro), _ptr(co), _ptr(labels))
    return labels, int(ncomp)


def pagerank(row_offsets, col, delta=0.85, threshold=0.01, max_iter=20, src=-1):
    ro, co = _offs(row_offsets), _cols(col)
    n = ro.size - 1
    rank = np.empty(n, dtype=np.float64)
    it = lib().orc_pagerank(n, _ptr(ro), _ptr(co), float(delta), float(threshold),
                            int(max_iter), int(src), _ptr(rank))
    return rank, int(it)


def bc(row_offsets, col, src=-1):
    """-> (bc float64[n] (x0.5; src < 0: summed over every source), sigmas float64[n], depths int32[n]
    of the last source)."""
    ro, co = _offs(row_offsets), _cols(col)
    n = ro.size - 1
    out = np.empty(max(n, 1), dtype=np.float64)
    sig = np.empty(max(n, 1), dtype=np.float64)
    lab = np.empty(max(n, 1), dtype=np.int32)
    lib().orc_bc(n, _ptr(ro), _ptr(co), int(src), _ptr(out), _ptr(sig), _ptr(lab))
    return out[:n], sig[:n], lab[:n]


def topk(out_offsets, in_offsets, k):
    """-> (node_ids, in_degrees, out_degrees), int32[min(k, n)] each."""
    oo, io = _offs(out_offsets), _offs(in_offsets)
    n = oo.size - 1
    k = min(int(k), n)
    ids, ind, outd = (np.empty(max(k, 1), dtype=np.int32) for _ in range(3))
    lib().orc_topk(n, _ptr(oo), _ptr(io), k, _ptr(ids), _ptr(ind), _ptr(outd))
    return ids[:k], ind[:k], outd[:k]


def rmat_edges(scale, seed, e0, e1, abcd=RMAT_ABCD):
    k = int(e1 - e0)
    rows = np.empty(k, dtype=np.int32)
    cols = np.empty(k, dtype=np.int32)
    lib().orc_rmat_edges(scale, seed, e0, e1, *abcd, _ptr(rows), _ptr(cols))
    return rows, cols


def rmat_csr(scale, edgefactor=16, seed=None, undirected=True, abcd=RMAT_ABCD):
    """-> (row_offsets int64[n+1], col int32[m]); seed defaults to 42 + scale."""
    if seed is None:
        seed = 42 + scale
    n = 1 << scale
    edges = edgefactor << scale
    ro = np.zeros(n + 1, dtype=np.int64)
    co = np.zeros(max((2 if undirected else 1) * edges, 1), dtype=np.int32)
    m = lib().orc_rmat_csr(scale, edges, seed, int(undirected), *abcd, _ptr(ro), _ptr(co))
    return ro, co[:m].copy()


def edge_weights(row_offsets, col, seed):
    ro, co = _offs(row_offsets), _cols(col)
    w = np.empty(co.size, dtype=np.uint32)
    lib().orc_edge_weights(ro.size - 1, _ptr(ro), _ptr(co), seed, _ptr(w))
    return w


def ref_graphio_market(path, undirected=True, workdir=None):
    """Run the REFERENCE's own reader (oracle/_ref/ref_graphio) on a copy of
    `path` (it writes a cache file next to its input) -> (n, ro, col, vals)."""
    import shutil
    import tempfile
    if not os.path.exists(REF_GRAPHIO):
        raise FileNotFoundError(REF_GRAPHIO)
    d = workdir or tempfile.mkdtemp(prefix="refgio_")
    local = os.path.join(d, os.path.basename(path))
    shutil.copyfile(path, local)
    out = os.path.join(d, "out.bin")
    subprocess.check_call([REF_GRAPHIO, "market", local, "1" if undirected else "0", out],
                          stdout=subprocess.DEVNULL)
    return read_ref_bin(out)


def ref_graphio_rmat(scale, edgefactor, srand_seed, undirected=True):
    import tempfile
    d = tempfile.mkdtemp(prefix="refgio_")
    out = os.path.join(d, "out.bin")
    subprocess.check_call([REF_GRAPHIO, "rmat", str(scale), str(edgefactor), str(srand_seed),
                           "1" if undirected else "0", out], stdout=subprocess.DEVNULL)
    return read_ref_bin(out)


def read_ref_bin(path):
    with open(path, "rb") as f:
        n, m = np.frombuffer(f.read(16), dtype=np.int64)
        hv = int(np.frombuffer(f.read(4), dtype=np.int32)[0])
        ro = np.frombuffer(f.read(4 * (int(n) + 1)), dtype=np.int32).astype(np.int64)
        col = np.frombuffer(f.read(4 * int(m)), dtype=np.int32).copy()
        vals = np.frombuffer(f.read(4 * int(m)), dtype=np.int32).copy() if hv else None
    return int(n), ro, col, vals
