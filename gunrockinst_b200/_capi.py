"""ctypes binding of libgunrock.so (the C ABI declared in include/gunrock/gunrock.h and
include/gunrock_b200.h).

This module is the only place the Python host side touches native code.  There is
no CPU path behind it: if the library is missing or no CUDA device is usable, calls
raise -- nothing here imports oracle/.
"""
import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
LIB_PATH = os.path.join(_HERE, "lib", "libgunrock.so")
CSRC = os.path.join(_HERE, "csrc")


class GunrockError(RuntimeError):
    pass


def build(verbose=False):
    """Compile libgunrock.so for sm_100a (nvcc cross-compiles without a GPU)."""
    cmd = ["make", "-C", CSRC, "-j", str(min(8, os.cpu_count() or 1))]
    if not verbose:
        cmd.insert(1, "-s")
    subprocess.check_call(cmd)
    if not os.path.exists(LIB_PATH):
        raise GunrockError("build did not produce " + LIB_PATH)
    return LIB_PATH


# ---- struct mirrors --------------------------------------------------------------
class GraphInfo(C.Structure):
    _fields_ = [("nodes", C.c_int64), ("local_nodes", C.c_int64), ("edges", C.c_int64),
                ("offset_bytes", C.c_int32), ("symmetric", C.c_int32), ("has_weights", C.c_int32),
                ("device", C.c_int32), ("log_parts", C.c_int32), ("rank", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("elapsed_ms", C.c_double), ("reset_ms", C.c_double),
                ("iterations", C.c_int64), ("edges_visited", C.c_int64), ("nodes_visited", C.c_int64),
                ("edges_inspected", C.c_uint64), ("total_queued", C.c_uint64), ("algo_bytes", C.c_uint64),
                ("push_levels", C.c_int32), ("pull_levels", C.c_int32),
                ("gpu_launches", C.c_int32), ("reserved", C.c_int32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class LevelRecord(C.Structure):
    _fields_ = [("level", C.c_int32), ("kind", C.c_int32),
                ("frontier", C.c_uint64), ("edges", C.c_uint64), ("discovered", C.c_uint64),
                ("candidates", C.c_uint64), ("ms", C.c_float), ("switch_ms", C.c_float)]


class BfsConfig(C.Structure):
    _fields_ = [("mark_pred", C.c_int32), ("idempotence", C.c_int32), ("traversal_mode", C.c_int32),
                ("direction_optimized", C.c_int32), ("alpha", C.c_double), ("beta", C.c_double),
                ("queue_sizing", C.c_double), ("record_levels", C.c_int32), ("host_loop", C.c_int32),
                ("degree_order", C.c_int32), ("stream_min_edges", C.c_int32)]


class MgLevel(C.Structure):
    _fields_ = [("frontier_local", C.c_uint64), ("frontier_edges", C.c_uint64),
                ("edges_inspected", C.c_uint64), ("candidates", C.c_uint64),
                ("overflow", C.c_uint64), ("gpu_launches", C.c_uint64)]


class MgStats(C.Structure):
    _fields_ = [("elapsed_ms", C.c_double), ("iterations", C.c_int64), ("push_levels", C.c_int64),
                ("pull_levels", C.c_int64), ("gpu_launches", C.c_int64), ("edges_visited", C.c_uint64),
                ("edges_inspected", C.c_uint64), ("exchange_bytes", C.c_uint64), ("nodes_visited", C.c_uint64)]


class MgTrace(C.Structure):
    _fields_ = [("level", C.c_int32), ("kind", C.c_int32), ("frontier", C.c_uint64), ("discovered", C.c_uint64),
                ("edges", C.c_uint64), ("max_edges", C.c_uint64), ("exchange_bytes", C.c_uint64),
                ("candidates", C.c_uint64), ("compute_ms", C.c_float), ("exchange_ms", C.c_float)]


IPC_HANDLE_BYTES, MG_PEER_BUFFERS, MGSSSP_PEER_BUFFERS, MGPR_PEER_BUFFERS, MGCC_PEER_BUFFERS = 64, 4, 3, 3, 3


# gunrock.h (reference ABI, gunrock/gunrock.h:51-99)
class GunrockGraph(C.Structure):
    _fields_ = [("num_nodes", C.c_size_t), ("num_edges", C.c_size_t),
                ("row_offsets", C.c_void_p), ("col_indices", C.c_void_p),
                ("col_offsets", C.c_void_p), ("row_indices", C.c_void_p),
                ("node_values", C.c_void_p), ("edge_values", C.c_void_p)]


class GunrockDataType(C.Structure):
    _fields_ = [("VTXID_TYPE", C.c_int), ("SIZET_TYPE", C.c_int), ("VALUE_TYPE", C.c_int)]


class GunrockConfig(C.Structure):
    _fields_ = [("mark_pred", C.c_bool), ("idempotence", C.c_bool), ("src_node", C.c_int),
                ("device", C.c_int), ("max_iter", C.c_int), ("top_nodes", C.c_int),
                ("delta_factor", C.c_int), ("delta", C.c_float), ("error", C.c_float),
                ("queue_size", C.c_float), ("src_mode", C.c_int)]


SRC_MANUALLY, SRC_RANDOMIZE, SRC_LARGEST_DEGREE = 0, 1, 2
VTXID_INT, SIZET_INT, VALUE_INT, VALUE_UINT, VALUE_FLOAT = 0, 0, 0, 1, 2

_lib = None
MISSING = []


def lib():
    """Load libgunrock.so; raises GunrockError when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise GunrockError("%s not found: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)" % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    p, i32, i64, u64, dbl = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_double
    pp = C.POINTER(C.c_void_p)

    def sig(name, res, args):
        try:
            f = getattr(L, name)
        except AttributeError:      # header and library disagree: tests/test_abi.py fails on this
            MISSING.append(name)
            return
        f.restype = res
        f.argtypes = args

    sig("grb_version", C.c_char_p, [])
    sig("grb_device_count", C.c_int, [])
    sig("grb_release_cached_memory", None, [])
    sig("grb_graph_from_host", C.c_int, [C.c_int, i64, i64, p, C.c_int, p, p, C.c_int, C.c_int, pp])
    sig("grb_graph_from_host_mapped", C.c_int, [C.c_int, i64, i64, p, C.c_int, p, p, C.c_int, C.c_int, pp])
    sig("grb_graph_slice_from_host", C.c_int, [C.c_int, i64, C.c_int, C.c_int, i64, i64, p, C.c_int, p, p, C.c_int, C.c_int, pp])
    sig("grb_graph_rmat", C.c_int, [C.c_int, C.c_int, i64, u64, C.c_int, dbl, dbl, dbl, dbl,
                                    C.c_int, C.c_int, u64, C.c_int, pp])
    sig("grb_graph_build_transpose", C.c_int, [p])
    sig("grb_graph_build_twin", C.c_int, [p])
    sig("grb_graph_twin_entries", C.c_int, [p, C.POINTER(i64)])
    sig("grb_graph_check_symmetric", C.c_int, [p, C.POINTER(C.c_int)])
    sig("grb_graph_get_info", C.c_int, [p, C.POINTER(GraphInfo)])
    sig("grb_graph_download", C.c_int, [p, p, p, p])
    sig("grb_graph_device_ptrs", C.c_int, [p, pp, pp, pp])
    sig("grb_graph_stream", C.c_int, [p, pp])
    sig("grb_graph_destroy", None, [p])
    sig("grb_bfs_create", C.c_int, [p, C.POINTER(BfsConfig), pp])
    sig("grb_bfs_run", C.c_int, [p, i32, p, p, C.POINTER(Stats)])
    sig("grb_bfs_run_many", C.c_int, [p, p, i32, C.POINTER(Stats)])
    sig("grb_bfs_device_labels", C.c_int, [p, pp])
    sig("grb_bfs_levels", C.c_int, [p, C.POINTER(LevelRecord), i32, C.POINTER(i32)])
    sig("grb_bfs_destroy", None, [p])
    sig("grb_sssp", C.c_int, [p, i32, i32, i32, p, p, C.POINTER(Stats)])
    sig("grb_cc", C.c_int, [p, p, C.POINTER(i64), C.POINTER(Stats)])
    sig("grb_pagerank", C.c_int, [p, i32, dbl, dbl, i32, p, p, p, C.POINTER(Stats)])
    sig("grb_bc", C.c_int, [p, i32, p, p, p, C.POINTER(Stats)])
    sig("grb_topk", C.c_int, [C.c_int, i64, p, p, i32, p, p, p, C.POINTER(Stats)])
    sig("grb_test_advance", C.c_int, [p, C.c_int, C.c_int, p, i64, p, i64, C.POINTER(i64), p, p, p,
                                      C.POINTER(u64)])
    sig("grb_test_neighbor_reduce", C.c_int, [p, C.c_int, C.c_int, C.c_int, p, i64, p, p, p, i64, C.POINTER(i64)])
    sig("grb_test_filter", C.c_int, [C.c_int, i64, C.c_int, p, i64, p, i64, C.POINTER(i64), p])
    sig("grb_test_bisect", C.c_int, [C.c_int, i64, p, p, i64, C.c_int, C.c_uint32, C.c_uint32, p, p, i64, C.POINTER(i64),
                                     p, i64, C.POINTER(i64), C.POINTER(C.c_uint32)])
    sig("grb_test_row_bins", C.c_int, [p, p, i64, p, p, p, p, i64, p])
    # no allocation inside: every array argument is a device pointer (include/gunrock_b200.h)
    sig("grb_advance_lb_noalloc", C.c_int, [p, C.c_int, p, i64, p, i64, p, p, p, p, p])
    sig("grb_mgbfs_create", C.c_int, [p, pp])
    sig("grb_mgbfs_reset", C.c_int, [p, i32, C.POINTER(u64)])
    sig("grb_mgbfs_buffer", C.c_int, [p, C.c_int, pp, C.POINTER(u64)])
    sig("grb_mgbfs_push", C.c_int, [p, i32, u64, u64])
    sig("grb_mgbfs_absorb", C.c_int, [p, i32, C.c_int])
    sig("grb_mgbfs_pull", C.c_int, [p, i32, C.c_int])
    sig("grb_mgbfs_merge_visited", C.c_int, [p, C.c_int])
    sig("grb_mgbfs_queue_to_map", C.c_int, [p, C.c_int, u64])
    sig("grb_mgbfs_map_to_queue", C.c_int, [p, C.c_int])
    sig("grb_mgbfs_counters", C.c_int, [p, C.POINTER(MgLevel)])
    sig("grb_mgbfs_sync", C.c_int, [p])
    sig("grb_mgbfs_stream", C.c_int, [p, pp])
    sig("grb_mgbfs_labels", C.c_int, [p, p])
    sig("grb_mgbfs_destroy", None, [p])
    sig("grb_mgbfs_peer_export", C.c_int, [p, p])
    sig("grb_mgbfs_peer_import", C.c_int, [p, C.c_int, p, u64])
    sig("grb_mgbfs_run_peer", C.c_int, [p, i32, C.c_int, dbl, dbl, C.c_int, C.POINTER(MgStats)])
    sig("grb_mgbfs_peer_levels", C.c_int, [p, C.POINTER(MgTrace), i32, C.POINTER(i32)])
    sig("grb_mgsssp_create", C.c_int, [p, pp])
    sig("grb_mgsssp_peer_export", C.c_int, [p, p])
    sig("grb_mgsssp_peer_import", C.c_int, [p, C.c_int, p])
    sig("grb_mgsssp_run", C.c_int, [p, i32, C.POINTER(MgStats)])
    sig("grb_mgsssp_result", C.c_int, [p, p, p])
    sig("grb_mgsssp_destroy", None, [p])
    sig("grb_mgpr_create", C.c_int, [p, pp])
    sig("grb_mgpr_peer_export", C.c_int, [p, p])
    sig("grb_mgpr_peer_import", C.c_int, [p, C.c_int, p])
    sig("grb_mgpr_run", C.c_int, [p, i32, dbl, dbl, i32, C.POINTER(MgStats)])
    sig("grb_mgpr_result", C.c_int, [p, p])
    sig("grb_mgpr_destroy", None, [p])
    sig("grb_mgcc_create", C.c_int, [p, pp])
    sig("grb_mgcc_peer_export", C.c_int, [p, p])
    sig("grb_mgcc_peer_import", C.c_int, [p, C.c_int, p])
    sig("grb_mgcc_run", C.c_int, [p, C.POINTER(MgStats)])
    sig("grb_mgcc_result", C.c_int, [p, p, C.POINTER(i64)])
    sig("grb_mgcc_destroy", None, [p])
    for name in ("gunrock_bfs_func", "gunrock_cc_func", "gunrock_bc_func"):
        sig(name, None, [C.POINTER(GunrockGraph), C.POINTER(GunrockGraph), GunrockConfig, GunrockDataType])
    sig("gunrock_topk_func", None, [C.POINTER(GunrockGraph), p, p, p, C.POINTER(GunrockGraph),
                                    GunrockConfig, GunrockDataType])
    sig("gunrock_sssp_func", None, [C.POINTER(GunrockGraph), p, C.POINTER(GunrockGraph),
                                    GunrockConfig, GunrockDataType])
    sig("gunrock_pr_func", None, [C.POINTER(GunrockGraph), p, p, C.POINTER(GunrockGraph),
                                  GunrockConfig, GunrockDataType])
    _lib = L
    return L


def check(rc, what):
    if rc != 0:
        raise GunrockError("%s failed with CUDA error %d (see stderr)" % (what, rc))


def free_host(ptr):
    """free() a buffer the library malloc()ed for the caller (gunrock.h ownership rule)."""
    libc = C.CDLL(None)
    libc.free.argtypes = [C.c_void_p]
    libc.free(ptr)
