"""1D vertex-partitioned, direction-optimising BFS: one process per device, torch.distributed
for the exchange (NCCL over NVLink on B200s; gloo in the CPU tests of the host logic).

The reference declares the striping rule (gunrock/app/problem_base.cuh:186-210) and leaves
the multi-GPU path as a TODO (:336-338); this module is the host loop of that path.  The
device side of every step is libgunrock.so (grb_mgbfs_*, include/gunrock_b200.h,
csrc/gunrock/app/bfs/mgbfs_enactor.cuh) behind the small `engine` interface below; the test
suite drives the same loop with a numpy stand-in engine over gloo.

Per level (all bitmaps are global, 1 bit / vertex, owner-major so a device's bits are one
contiguous segment):

  push   engine.push -> all_to_all_single(next map -> recv map) -> engine.absorb
  pull   all_gather_into_tensor(frontier map, in place) -> engine.merge_visited -> engine.pull
  both   all_gather of every device's 64-byte counter block -> one host read: global frontier
         length, its degree sum (Beamer's direction rule, as in bfs_enactor.cuh), edges inspected

Payloads: push n/8 bytes per device per level (each device sends (P-1)/P of its next map),
pull n/8 bytes gathered per device, counters 64 B.  Exchange and compute are timed per level
with events on the stream the steps run on.
"""
import ctypes as C
import time

import numpy as np
import torch
import torch.distributed as dist

from . import _capi

CTR_WORDS = 8           # uint64 words of oprtr::LevelCounters
CTR_FRONTIER, CTR_DEGREE, CTR_EDGES, CTR_CANDS, CTR_OVERFLOW = 0, 1, 2, 4, 6


class _RawCuda:
    """wraps a raw device pointer so torch can view it (no copy, no ownership)"""

    def __init__(self, ptr, count, typestr):
        self.__cuda_array_interface__ = {"shape": (int(count),), "typestr": typestr,
                                         "data": (int(ptr), False), "version": 2}


class DeviceEngine:
    """The per-level device steps of one rank, over the C ABI (no CPU path)."""

    def __init__(self, graph):
        self.graph = graph
        L = _capi.lib()
        self._h = C.c_void_p()
        _capi.check(L.grb_mgbfs_create(graph._h, C.byref(self._h)), "grb_mgbfs_create")
        info = graph.info
        self.parts = 1 << info.log_parts
        self.rank = int(info.rank)
        self.nodes = int(info.nodes)
        self.local_nodes = int(info.local_nodes)
        self.local_edges = int(info.edges)
        self.symmetric = bool(info.symmetric)
        self.seg_words = self.local_nodes // 32
        self.device = torch.device("cuda", int(info.device))
        s = C.c_void_p()
        _capi.check(L.grb_mgbfs_stream(self._h, C.byref(s)), "grb_mgbfs_stream")
        self.stream = torch.cuda.ExternalStream(s.value or 0, device=self.device)
        words = self.seg_words * self.parts
        self.maps = [self._tensor(1, words, "<i4"), self._tensor(2, words, "<i4")]
        self.visited = self._tensor(0, words, "<i4")
        self.next = self._tensor(3, words, "<i4")
        self.recv = self._tensor(4, words, "<i4")
        self.counters = self._tensor(6, CTR_WORDS, "<i8")

    def _tensor(self, which, count, typestr):
        p, b = C.c_void_p(), C.c_uint64()
        _capi.check(_capi.lib().grb_mgbfs_buffer(self._h, which, C.byref(p), C.byref(b)), "grb_mgbfs_buffer")
        assert b.value >= count * int(typestr[2:]), (which, b.value, count)
        return torch.as_tensor(_RawCuda(p.value, count, typestr), device=self.device)

    def reset(self, src):
        d = C.c_uint64()
        _capi.check(_capi.lib().grb_mgbfs_reset(self._h, int(src), C.byref(d)), "grb_mgbfs_reset")
        return int(d.value)

    def push(self, level, queue_len, edges_hint):
        _capi.check(_capi.lib().grb_mgbfs_push(self._h, level, int(queue_len), int(edges_hint)), "grb_mgbfs_push")

    def absorb(self, level, cur_out):
        _capi.check(_capi.lib().grb_mgbfs_absorb(self._h, level, cur_out), "grb_mgbfs_absorb")

    def pull(self, level, cur):
        _capi.check(_capi.lib().grb_mgbfs_pull(self._h, level, cur), "grb_mgbfs_pull")

    def merge_visited(self, cur):
        _capi.check(_capi.lib().grb_mgbfs_merge_visited(self._h, cur), "grb_mgbfs_merge_visited")

    def queue_to_map(self, cur, queue_len):
        _capi.check(_capi.lib().grb_mgbfs_queue_to_map(self._h, cur, int(queue_len)), "grb_mgbfs_queue_to_map")

    def map_to_queue(self, cur):
        _capi.check(_capi.lib().grb_mgbfs_map_to_queue(self._h, cur), "grb_mgbfs_map_to_queue")

    def launches(self):
        lv = _capi.MgLevel()
        _capi.check(_capi.lib().grb_mgbfs_counters(self._h, C.byref(lv)), "grb_mgbfs_counters")
        return int(lv.gpu_launches)

    # ---- peer-memory exchange (include/gunrock_b200.h, grb_mgbfs_peer_*) -----------------------
    def enable_peer_exchange(self, group=None):
        """Exchange CUDA IPC handles of the buffers the peers write and map theirs; after this
        `run_peer` runs whole searches natively with NVLink stores instead of collectives.
        torch.distributed only ferries the handle bytes (and the global edge count) once."""
        L = _capi.lib()
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        if world == 1:
            self.peer_ready = True
            return
        nbytes = _capi.IPC_HANDLE_BYTES * _capi.MG_PEER_BUFFERS
        mine = (C.c_ubyte * nbytes)()
        _capi.check(L.grb_mgbfs_peer_export(self._h, mine), "grb_mgbfs_peer_export")
        every = [None] * world
        dist.all_gather_object(every, (bytes(mine), self.local_edges), group=group)
        blob = b"".join(h for h, _ in every)
        buf = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
        _capi.check(L.grb_mgbfs_peer_import(self._h, world, buf, sum(m for _, m in every)), "grb_mgbfs_peer_import")
        dist.barrier(group=group)           # nobody starts a search before every mapping exists
        self.peer_ready = True

    def run_peer(self, src, direction_optimized=True, alpha=14.0, beta=24.0, record_levels=False):
        st = _capi.MgStats()
        _capi.check(_capi.lib().grb_mgbfs_run_peer(self._h, int(src), int(direction_optimized), float(alpha),
                                                   float(beta), int(record_levels), C.byref(st)), "grb_mgbfs_run_peer")
        out = {k: getattr(st, k) for k, _ in st._fields_}
        out["levels"] = []
        if record_levels:
            rec = (_capi.MgTrace * 128)()
            cnt = C.c_int32()
            _capi.check(_capi.lib().grb_mgbfs_peer_levels(self._h, rec, 128, C.byref(cnt)), "grb_mgbfs_peer_levels")
            for r in rec[:cnt.value]:
                out["levels"].append(dict(level=r.level, kind="pull" if r.kind else "push", frontier=r.frontier,
                                          discovered=r.discovered, edges=r.edges, local_max_edges=r.max_edges,
                                          exchange_bytes=r.exchange_bytes, candidates=r.candidates, compute_ms=r.compute_ms,
                                          exchange_ms=r.exchange_ms))
        return out

    def labels(self, copy=True):
        """this rank's depths -> host.  The landing buffer is page-locked and allocated once (a fresh pageable array per
        call made the download crawl at 3-4 GB/s with eight ranks contending); copy=False returns a view of it that the
        next call overwrites."""
        if getattr(self, "_labels_pinned", None) is None:
            self._labels_pinned = torch.empty(max(self.local_nodes, 1), dtype=torch.int32, pin_memory=True)
        _capi.check(_capi.lib().grb_mgbfs_labels(self._h, C.c_void_p(self._labels_pinned.data_ptr())), "grb_mgbfs_labels")
        view = self._labels_pinned.numpy()[:self.local_nodes]
        return view.copy() if copy else view

    def device_labels(self):
        """torch view (no copy) of this rank's depths on the device"""
        return self._tensor(5, self.local_nodes, "<i4")

    # timing hooks (CUDA events on the stream every step and collective is ordered on)
    def stream_context(self):
        return torch.cuda.stream(self.stream)

    def mark(self):
        e = torch.cuda.Event(enable_timing=True)
        e.record(self.stream)
        return e

    @staticmethod
    def elapsed_ms(a, b):
        return a.elapsed_time(b)

    def synchronize(self):
        self.stream.synchronize()

    def close(self):
        if self._h:
            _capi.lib().grb_mgbfs_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def owner_of(v, log_parts):
    """the partition rule of csrc/gunrock/graph/csr.cuh (Partition::Owner), on the host"""
    if log_parts == 0:
        return 0
    local = (int(v) >> log_parts) & 0xFFFFFFFF
    mix = ((local * 0x9E3779B1) & 0xFFFFFFFF) >> (32 - log_parts)
    return (int(v) & ((1 << log_parts) - 1)) ^ mix


def global_ids(local_nodes, log_parts, rank):
    """global vertex id of every local row of `rank` (Partition::Global), vectorised"""
    local = np.arange(local_nodes, dtype=np.uint64)
    if log_parts == 0:
        return local.astype(np.int64)
    mix = ((local * np.uint64(0x9E3779B1)) & np.uint64(0xFFFFFFFF)) >> np.uint64(32 - log_parts)
    return ((local << np.uint64(log_parts)) | (np.uint64(rank) ^ mix)).astype(np.int64)


class PartitionedBfs:
    """Host loop of the partitioned DOBFS.  `engine` supplies the device steps and the tensors
    the collectives move; `group` is the torch.distributed process group (None = default)."""

    def __init__(self, engine, group=None, direction_optimized=True, alpha=14.0, beta=24.0, record_levels=True,
                 exchange="nccl"):
        """exchange: "nccl" -- collectives of torch.distributed between the device steps (the host loop
        below); "peer" -- the library's native loop with peer-memory stores (engine.run_peer)."""
        self.e = engine
        self.exchange = exchange
        if exchange == "peer" and not getattr(engine, "peer_ready", False):
            engine.enable_peer_exchange(group)
        self.group = group
        # a directed slice has no in-edges on this device: top-down only (the native loop takes the same decision)
        self.dopt = direction_optimized and bool(getattr(engine, "symmetric", True))
        self.alpha, self.beta = alpha, beta
        self.record_levels = record_levels
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        if self.world != engine.parts:
            raise ValueError("graph is partitioned %d ways but the group has %d ranks" % (engine.parts, self.world))
        self.all_counters = torch.zeros(self.world * CTR_WORDS, dtype=torch.int64, device=engine.counters.device)
        self.setup = torch.zeros(2, dtype=torch.int64, device=engine.counters.device)

    # ---- collectives ------------------------------------------------------------------------
    def _exchange_next(self):
        e = self.e
        if self.world == 1:
            e.recv.copy_(e.next)
        else:
            dist.all_to_all_single(e.recv, e.next, group=self.group)
        return e.next.numel() * 4 * (self.world - 1) // self.world

    def _gather_map(self, cur):
        e = self.e
        if self.world > 1:
            m = e.maps[cur]
            own = m[e.rank * e.seg_words:(e.rank + 1) * e.seg_words]
            dist.all_gather_into_tensor(m, own, group=self.group)
        return e.seg_words * 4 * (self.world - 1)

    def _gather_counters(self):
        e = self.e
        if self.world == 1:
            self.all_counters.copy_(e.counters)
        else:
            dist.all_gather_into_tensor(self.all_counters, e.counters, group=self.group)
        return self.all_counters.cpu().numpy().reshape(self.world, CTR_WORDS)     # the one host read of a level

    # ---- the search -------------------------------------------------------------------------------
    def run(self, src):
        """-> stats dict; labels stay on the devices (engine.labels())."""
        e = self.e
        if self.exchange == "peer":
            t_wall = time.perf_counter()
            st = e.run_peer(src, self.dopt, self.alpha, self.beta, self.record_levels)
            st["wall_ms"] = (time.perf_counter() - t_wall) * 1e3
            return st
        with e.stream_context():
            src_degree = e.reset(src)
            self.setup[0] = src_degree
            self.setup[1] = e.local_edges
            if self.world > 1:
                dist.all_reduce(self.setup, group=self.group)
            src_degree, m_global = (int(x) for x in self.setup.cpu())
            n = e.nodes
            local_q = 1 if (0 <= src < n and owner_of(src, e.parts.bit_length() - 1) == e.rank) else 0
            local_hint = src_degree if local_q else 0
            frontier, frontier_edges = (1 if 0 <= src < n else 0), src_degree
            unexplored = m_global - frontier_edges
            edges_visited = frontier_edges
            prev_len = 0
            pull = False
            cur, map_valid, gathered, queue_valid = 0, False, False, True
            level = 0
            levels, marks = [], []
            exch_bytes = inspected = 0
            t_wall = time.perf_counter()
            ev_start = e.mark()
            while frontier > 0:
                if self.dopt:
                    if not pull and frontier_edges > unexplored / self.alpha:
                        pull = True
                    elif pull and frontier < prev_len and frontier < n / self.beta:
                        pull = False
                m0 = e.mark() if self.record_levels else None
                lvl_bytes = 0
                if pull:
                    if not map_valid:
                        e.queue_to_map(cur, local_q)
                        map_valid = True
                    m1 = e.mark() if self.record_levels else None
                    if not gathered:
                        lvl_bytes += self._gather_map(cur)
                    m2 = e.mark() if self.record_levels else None
                    e.merge_visited(cur)
                    e.pull(level, cur)
                    cur ^= 1
                    queue_valid = False
                else:
                    if not queue_valid:
                        e.map_to_queue(cur)
                        queue_valid = True
                    e.push(level, local_q, local_hint)
                    m1 = e.mark() if self.record_levels else None
                    lvl_bytes += self._exchange_next()
                    m2 = e.mark() if self.record_levels else None
                    e.absorb(level, cur ^ 1)
                    cur ^= 1
                map_valid, gathered = True, False
                m3 = e.mark() if self.record_levels else None
                c = self._gather_counters()
                if int(c[:, CTR_OVERFLOW].sum()):
                    raise _capi.GunrockError("frontier queue overflow on a partitioned level")
                prev_len = frontier
                frontier = int(c[:, CTR_FRONTIER].sum())
                frontier_edges = int(c[:, CTR_DEGREE].sum())
                local_q = int(c[e.rank, CTR_FRONTIER])
                local_hint = int(c[e.rank, CTR_DEGREE])
                lvl_inspected = int(c[:, CTR_EDGES].sum())
                inspected += lvl_inspected
                unexplored = max(unexplored - frontier_edges, 0)
                edges_visited += frontier_edges
                exch_bytes += lvl_bytes
                if self.record_levels:
                    marks.append((m0, m1, m2, m3))
                    levels.append(dict(level=level, kind="pull" if pull else "push", frontier=prev_len,
                                       discovered=frontier, edges=lvl_inspected, exchange_bytes=lvl_bytes,
                                       candidates=int(c[:, CTR_CANDS].sum()) if pull else 0,
                                       local_max_edges=int(c[:, CTR_EDGES].max())))
                level += 1
            ev_end = e.mark()
            e.synchronize()
            wall_ms = (time.perf_counter() - t_wall) * 1e3
        for rec, (m0, m1, m2, m3) in zip(levels, marks):
            rec["exchange_ms"] = e.elapsed_ms(m1, m2)
            rec["compute_ms"] = e.elapsed_ms(m0, m1) + e.elapsed_ms(m2, m3)
        return dict(elapsed_ms=e.elapsed_ms(ev_start, ev_end), wall_ms=wall_ms, iterations=level,
                    gpu_launches=e.launches(),      # the context's counter restarts at every reset
                    edges_visited=edges_visited, edges_inspected=inspected, exchange_bytes=exch_bytes,
                    levels=levels, m_global=m_global)


class PartitionedSssp:
    """1D-partitioned SSSP over peer memory (include/gunrock_b200.h, grb_mgsssp_*): one instance per
    rank on that rank's slice of a weighted graph; every rank calls run() with the same source."""

    def __init__(self, graph, group=None):
        self.graph, self.group = graph, group
        L = _capi.lib()
        self._h = C.c_void_p()
        _capi.check(L.grb_mgsssp_create(graph._h, C.byref(self._h)), "grb_mgsssp_create")
        info = graph.info
        self.parts, self.rank = 1 << info.log_parts, int(info.rank)
        self.nodes, self.local_nodes = int(info.nodes), int(info.local_nodes)
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        if world != self.parts:
            raise ValueError("graph is partitioned %d ways but the group has %d ranks" % (self.parts, world))
        if world > 1:
            nbytes = _capi.IPC_HANDLE_BYTES * _capi.MGSSSP_PEER_BUFFERS
            mine = (C.c_ubyte * nbytes)()
            _capi.check(L.grb_mgsssp_peer_export(self._h, mine), "grb_mgsssp_peer_export")
            every = [None] * world
            dist.all_gather_object(every, bytes(mine), group=group)
            blob = b"".join(every)
            buf = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
            _capi.check(L.grb_mgsssp_peer_import(self._h, world, buf), "grb_mgsssp_peer_import")
            dist.barrier(group=group)

    def run(self, src):
        st = _capi.MgStats()
        _capi.check(_capi.lib().grb_mgsssp_run(self._h, int(src), C.byref(st)), "grb_mgsssp_run")
        return dict(elapsed_ms=st.elapsed_ms, iterations=st.iterations, gpu_launches=st.gpu_launches,
                    relaxations=st.edges_inspected, queued=st.nodes_visited)

    def result(self):
        """-> (dist uint32[local_nodes], preds int32[local_nodes]) of the owned rows"""
        d = np.empty(self.local_nodes, dtype=np.uint32)
        p = np.empty(self.local_nodes, dtype=np.int32)
        _capi.check(_capi.lib().grb_mgsssp_result(self._h, d.ctypes.data_as(C.c_void_p), p.ctypes.data_as(C.c_void_p)),
                    "grb_mgsssp_result")
        return d, p

    def gather(self):
        """global (dist, preds) in vertex order on every rank; for tests"""
        d, p = self.result()
        lp = self.parts.bit_length() - 1
        gd = np.empty(self.nodes, dtype=np.uint32)
        gp = np.empty(self.nodes, dtype=np.int32)
        if self.parts == 1:
            ids = global_ids(self.local_nodes, lp, 0)
            gd[ids], gp[ids] = d, p
            return gd, gp
        every = [None] * self.parts
        dist.all_gather_object(every, (d, p), group=self.group)
        for r, (dd, pp) in enumerate(every):
            ids = global_ids(self.local_nodes, lp, r)
            gd[ids], gp[ids] = dd, pp
        return gd, gp

    def close(self):
        if self._h:
            _capi.lib().grb_mgsssp_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PartitionedPagerank:
    """1D-partitioned PageRank over peer memory (include/gunrock_b200.h, grb_mgpr_*): one instance per
    rank on that rank's slice of a symmetric graph; every rank calls run() with the same arguments."""

    def __init__(self, graph, group=None):
        self.graph, self.group = graph, group
        L = _capi.lib()
        self._h = C.c_void_p()
        _capi.check(L.grb_mgpr_create(graph._h, C.byref(self._h)), "grb_mgpr_create")
        info = graph.info
        self.parts, self.rank = 1 << info.log_parts, int(info.rank)
        self.nodes, self.local_nodes = int(info.nodes), int(info.local_nodes)
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        if world != self.parts:
            raise ValueError("graph is partitioned %d ways but the group has %d ranks" % (self.parts, world))
        if world > 1:
            nbytes = _capi.IPC_HANDLE_BYTES * _capi.MGPR_PEER_BUFFERS
            mine = (C.c_ubyte * nbytes)()
            _capi.check(L.grb_mgpr_peer_export(self._h, mine), "grb_mgpr_peer_export")
            every = [None] * world
            dist.all_gather_object(every, bytes(mine), group=group)
            blob = b"".join(every)
            buf = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
            _capi.check(L.grb_mgpr_peer_import(self._h, world, buf), "grb_mgpr_peer_import")
            dist.barrier(group=group)

    def run(self, src=-1, delta=0.85, threshold=1e-6, max_iter=1000):
        st = _capi.MgStats()
        _capi.check(_capi.lib().grb_mgpr_run(self._h, int(src), float(delta), float(threshold), int(max_iter),
                                             C.byref(st)), "grb_mgpr_run")
        return dict(elapsed_ms=st.elapsed_ms, sweep_ms=st.edges_visited / 1e3, iterations=st.iterations,
                    prune_sweeps=st.push_levels, gpu_launches=st.gpu_launches, edges_inspected=st.edges_inspected,
                    valid_nodes=st.nodes_visited, exchange_bytes=st.exchange_bytes)

    def result(self):
        """-> fp64 ranks of the owned rows, local-row order"""
        r = np.empty(max(self.local_nodes, 1), dtype=np.float64)
        _capi.check(_capi.lib().grb_mgpr_result(self._h, r.ctypes.data_as(C.c_void_p)), "grb_mgpr_result")
        return r[:self.local_nodes]

    def gather(self):
        """global fp64 ranks in vertex order on every rank; for tests"""
        r = self.result()
        lp = self.parts.bit_length() - 1
        out = np.empty(self.nodes, dtype=np.float64)
        if self.parts == 1:
            out[global_ids(self.local_nodes, lp, 0)] = r
            return out
        every = [None] * self.parts
        dist.all_gather_object(every, r, group=self.group)
        for k, rr in enumerate(every):
            out[global_ids(self.local_nodes, lp, k)] = rr
        return out

    def close(self):
        if self._h:
            _capi.lib().grb_mgpr_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class PartitionedCc:
    """1D-partitioned connected components over peer memory (include/gunrock_b200.h, grb_mgcc_*): one
    instance per rank on that rank's slice of a symmetric graph; every rank calls run()."""

    def __init__(self, graph, group=None):
        self.graph, self.group = graph, group
        L = _capi.lib()
        self._h = C.c_void_p()
        _capi.check(L.grb_mgcc_create(graph._h, C.byref(self._h)), "grb_mgcc_create")
        info = graph.info
        self.parts, self.rank = 1 << info.log_parts, int(info.rank)
        self.nodes, self.local_nodes = int(info.nodes), int(info.local_nodes)
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        if world != self.parts:
            raise ValueError("graph is partitioned %d ways but the group has %d ranks" % (self.parts, world))
        if world > 1:
            nbytes = _capi.IPC_HANDLE_BYTES * _capi.MGCC_PEER_BUFFERS
            mine = (C.c_ubyte * nbytes)()
            _capi.check(L.grb_mgcc_peer_export(self._h, mine), "grb_mgcc_peer_export")
            every = [None] * world
            dist.all_gather_object(every, bytes(mine), group=group)
            blob = b"".join(every)
            buf = (C.c_ubyte * len(blob)).from_buffer_copy(blob)
            _capi.check(L.grb_mgcc_peer_import(self._h, world, buf), "grb_mgcc_peer_import")
            dist.barrier(group=group)

    def run(self):
        st = _capi.MgStats()
        _capi.check(_capi.lib().grb_mgcc_run(self._h, C.byref(st)), "grb_mgcc_run")
        return dict(elapsed_ms=st.elapsed_ms, iterations=st.iterations, gpu_launches=st.gpu_launches,
                    offers=st.edges_inspected, queued=st.nodes_visited)

    def result(self):
        """-> (labels int32[local_nodes] of the owned rows, number of owned component roots)"""
        lab = np.empty(max(self.local_nodes, 1), dtype=np.int32)
        roots = C.c_int64()
        _capi.check(_capi.lib().grb_mgcc_result(self._h, lab.ctypes.data_as(C.c_void_p), C.byref(roots)),
                    "grb_mgcc_result")
        return lab[:self.local_nodes], int(roots.value)

    def gather(self):
        """-> (global labels in vertex order, component count) on every rank; for tests"""
        lab, roots = self.result()
        lp = self.parts.bit_length() - 1
        out = np.empty(self.nodes, dtype=np.int32)
        if self.parts == 1:
            out[global_ids(self.local_nodes, lp, 0)] = lab
            return out, roots
        every = [None] * self.parts
        dist.all_gather_object(every, (lab, roots), group=self.group)
        total = 0
        for r, (ll, rr) in enumerate(every):
            out[global_ids(self.local_nodes, lp, r)] = ll
            total += rr
        return out, total

    def close(self):
        if self._h:
            _capi.lib().grb_mgcc_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def gather_labels(engine, group=None):
    """all ranks' labels -> one global int32[n] array in vertex order (on every rank); for tests"""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    mine = engine.labels()
    log_parts = engine.parts.bit_length() - 1
    out = np.empty(engine.nodes, dtype=np.int32)
    if world == 1:
        out[global_ids(engine.local_nodes, log_parts, 0)] = mine
        return out
    parts = [None] * world
    dist.all_gather_object(parts, mine, group=group)
    for r, lab in enumerate(parts):
        out[global_ids(engine.local_nodes, log_parts, r)] = lab
    return out


# ------------------------------------------------------------------------------------------
# bench.py --gpus N (N > 1): BASELINE config 5, one process per device under torchrun
# ------------------------------------------------------------------------------------------
def _pick_sources(graph, eng, count, seed, group=None):
    """vertex 0 first, then seeded random vertices with at least one edge -- the same list on every
    rank: each rank reports the degree of the candidates it owns, one all_reduce merges them"""
    lp = eng.parts.bit_length() - 1
    rng = np.random.RandomState(seed)
    cands = np.concatenate([[0], rng.randint(0, eng.nodes, 8 * count)]).astype(np.int64)
    ro_ptr, _, _ = graph.device_ptrs()
    osz = int(graph.info.offset_bytes)
    ro = torch.as_tensor(_RawCuda(ro_ptr, eng.local_nodes + 1, "<i4" if osz == 4 else "<i8"), device=eng.device)
    mine = np.array([owner_of(v, lp) == eng.rank for v in cands])
    rows = torch.as_tensor(cands >> lp, device=eng.device)
    deg = (ro[rows + 1] - ro[rows]).to(torch.int64) * torch.as_tensor(mine, device=eng.device)
    if dist.is_initialized() and eng.parts > 1:
        dist.all_reduce(deg, group=group)
    deg = deg.cpu().numpy()
    out = [int(v) for v, d in zip(cands, deg) if d > 0]
    seen, uniq = set(), []
    for v in out:
        if v not in seen:
            seen.add(v)
            uniq.append(v)
    return uniq[:count]


def bench(args, rank, world, workload_config, ClockSampler, pick_sources, measured_peaks, cpu_check=None):
    """One JSON line on rank 0: whole-job GTEPS of the 1D-partitioned DOBFS at `world` devices.
    cpu_check(row_offsets, col, src) -> depths: the harness's CPU checker (nothing in this package imports one)."""
    import json
    import os
    from . import Graph
    import sys
    local = int(os.environ.get("LOCAL_RANK", rank))
    # stdout carries exactly one JSON line: whatever libraries print meanwhile (NCCL announces its
    # version on stdout) goes to stderr
    sys.stdout.flush()
    saved_stdout = os.dup(1)
    os.dup2(2, 1)
    torch.cuda.set_device(local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29533")
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    t_setup = time.time()
    try:
        if world & (world - 1):
            raise SystemExit("bench.py: the partition needs a power-of-two device count")
        lp = world.bit_length() - 1
        scale = args.scale
        g = Graph.rmat(scale, 16, device=local, log_parts=lp, rank=rank)
        eng = DeviceEngine(g)
        exchange = getattr(args, "exchange", "peer")
        bfs = PartitionedBfs(eng, direction_optimized=True, record_levels=False, exchange=exchange)
        sources = _pick_sources(g, eng, 64, 1234 + scale)
        clocks = ClockSampler(local)
        warm = max(args.warmup, 3)
        dev = eng.device
        # ---- the same graph, unpartitioned, on ONE device (rank 0): the 1-device point every scaling figure of this line
        # refers to, and the labels the partitioned search has to reproduce
        single, ref_labels, oracle_parity = None, None, None
        if not getattr(args, "no_single_ref", False) and scale <= 27:
            if rank == 0:
                try:
                    from . import Bfs, release_cached_memory
                    t_s = time.time()
                    gfull = g if world == 1 else Graph.rmat(scale, 16, device=local)
                    sb = Bfs(gfull, direction_optimized=True)
                    s_stream = torch.cuda.ExternalStream(gfull.stream(), device=dev)
                    for i in range(3):
                        sb.run(sources[i % len(sources)], want_labels=False)
                    s_steps = max(1, min(args.steps, 64))
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    torch.cuda.synchronize()
                    a.record(s_stream)
                    s_stats = sb.run_many([sources[i % len(sources)] for i in range(s_steps)])
                    b.record(s_stream)
                    torch.cuda.synchronize()
                    s_ms = a.elapsed_time(b)
                    s_edges = sum(x["edges_visited"] for x in s_stats)
                    sb.run(sources[0], want_labels=False)
                    ref_labels = torch.as_tensor(_RawCuda(sb.device_labels(), eng.nodes, "<i4"), device=dev).clone()
                    single = {"gteps": s_edges / (s_ms * 1e-3) / 1e9, "ms_per_search": s_ms / s_steps, "searches": s_steps,
                              "graph": "the same seeded R-MAT scale %d, unpartitioned, resident on device 0" % scale,
                              "api": "grb_bfs_run_many (direction-optimising)", "setup_s": round(time.time() - t_s, 1)}
                    if scale <= 24 and cpu_check is not None:      # the caller's CPU checker (bench.py: the oracle)
                        ro, co = gfull.download()
                        oracle_parity = bool(np.array_equal(ref_labels.cpu().numpy(), cpu_check(ro, co, sources[0])))
                        del ro, co
                    sb.close()
                    if world > 1:
                        gfull.close()
                    release_cached_memory()
                except Exception as exc:                    # never lose the line over the reference point
                    single = {"error": "%s: %s" % (type(exc).__name__, str(exc)[:300])}
                    ref_labels = None
            dist.barrier()
        for i in range(warm):
            bfs.run(sources[i % len(sources)])
        dist.barrier()
        torch.cuda.synchronize()
        clocks.start()
        e0, e1 = eng.mark(), None
        edges = inspected = exch = 0
        levels = 0
        launches = 0
        for i in range(args.steps):
            st = bfs.run(sources[i % len(sources)])
            edges += st["edges_visited"]
            inspected += st["edges_inspected"]
            exch += st["exchange_bytes"]
            levels += st["iterations"]
            launches += st["gpu_launches"]
        e1 = eng.mark()
        dist.barrier()
        torch.cuda.synchronize()
        clocks.stop()
        t = torch.tensor([eng.elapsed_ms(e0, e1)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        cnt = torch.tensor([launches, eng.local_edges], dtype=torch.int64, device=dev)
        dist.all_reduce(cnt)
        launches_all, m_global = int(cnt[0].item()), int(cnt[1].item())

        # traced pass: per-level compute / exchange split (events on the stream), max over ranks
        tr = PartitionedBfs(eng, direction_optimized=True, record_levels=True, exchange=exchange)
        k = min(args.steps, 8)
        comp = torch.zeros(1, dtype=torch.float64, device=dev)
        exms = torch.zeros(1, dtype=torch.float64, device=dev)
        ex_bytes = 0
        lv_trace = None
        osz = int(g.info.offset_bytes)
        pull_bytes, pull_ms, pull_launches = 0, 0.0, 0
        for i in range(k):
            st = tr.run(sources[i % len(sources)])
            for l in st["levels"]:
                if l["kind"] == "pull":      # algorithmic bytes of the sweep on all devices
                    pull_bytes += l.get("candidates", 0) * 2 * osz + 5 * l["edges"] + 4 * l["discovered"] + 3 * (eng.nodes // 8)
                    pull_ms += l["compute_ms"]
                    pull_launches += 1
            comp += sum(l["compute_ms"] for l in st["levels"])
            exms += sum(l["exchange_ms"] for l in st["levels"])
            ex_bytes += st["exchange_bytes"]
            if i == 0:
                lv_trace = st["levels"]
        dist.all_reduce(comp, op=dist.ReduceOp.MAX)
        dist.all_reduce(exms, op=dist.ReduceOp.MAX)
        pms = torch.tensor([pull_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(pms, op=dist.ReduceOp.MAX)
        pull_ms = float(pms.item())

        # ---- parity of the partitioned search: depths of sources[0] gathered on rank 0 (device side, NCCL) against the
        # single-device search of the same graph (and, up to scale 24, that one against the CPU oracle)
        parity = None
        bfs.run(sources[0])
        mine = eng.device_labels()
        parts = [torch.empty_like(mine) for _ in range(world)] if rank == 0 else None
        if world > 1:
            dist.gather(mine, parts, dst=0)
        elif rank == 0:
            parts = [mine]
        if rank == 0:
            got = torch.empty(eng.nodes, dtype=torch.int32, device=dev)
            for r, lab in enumerate(parts):
                got[torch.as_tensor(global_ids(eng.local_nodes, lp, r), device=dev)] = lab
            reached = got >= 0
            hist = torch.bincount(got[reached].to(torch.int64)).cpu().tolist()
            parity = {"source": int(sources[0]), "reached": int(reached.sum().item()),
                      "sum_of_depths": int(got[reached].to(torch.int64).sum().item()), "depth_histogram": hist,
                      "equals_single_device_search": (bool(torch.equal(got, ref_labels)) if ref_labels is not None else None),
                      "single_device_search_equals_cpu_oracle": oracle_parity,
                      "how": "depths of every rank gathered on rank 0 over NCCL and compared on the device with the labels "
                             "of grb_bfs_run on the unpartitioned graph"}
            parity["ok"] = parity["equals_single_device_search"]
            del got, parts
        ref_labels = None

        # end to end for a resident partitioned graph: source in, this rank's labels out to the host (page-locked landing
        # buffer, reused)
        torch.cuda.synchronize()
        dist.barrier()
        e2e_steps = max(1, min(args.steps, args.e2e_steps))
        eng.labels(copy=False)
        t0 = time.perf_counter()
        e2e_edges = 0
        for i in range(e2e_steps):
            st = bfs.run(sources[i % len(sources)])
            eng.labels(copy=False)
            e2e_edges += st["edges_visited"]
        torch.cuda.synchronize()
        te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e_s = float(te.item())

        # ---- the same end to end with the partitioned CSR on the HOST: every rank uploads its slice from page-locked
        # memory, builds its context (peer mappings included), searches, downloads its labels -- all inside the timed region
        e2e_host = None
        if not getattr(args, "no_e2e_from_host", False):
            try:
                ro_h, co_h = g.download()
                ro_p = torch.from_numpy(ro_h).pin_memory()
                co_p = torch.from_numpy(co_h).pin_memory()
                del ro_h, co_h
                fh_steps, fh_edges, fh_s = max(1, min(3, e2e_steps)), 0, 0.0
                for i in range(-1, fh_steps):
                    torch.cuda.synchronize()
                    dist.barrier()
                    t0 = time.perf_counter()
                    g2 = Graph.from_slice(ro_p.numpy(), co_p.numpy(), eng.nodes, lp, rank, symmetric=True, device=local)
                    eng2 = DeviceEngine(g2)
                    bfs2 = PartitionedBfs(eng2, direction_optimized=True, record_levels=False, exchange=exchange)
                    st2 = bfs2.run(sources[max(i, 0) % len(sources)])
                    eng2.labels(copy=False)
                    torch.cuda.synchronize()
                    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
                    dist.all_reduce(dt, op=dist.ReduceOp.MAX)
                    eng2.close()
                    g2.close()
                    if i >= 0:
                        fh_s += float(dt.item())
                        fh_edges += st2["edges_visited"]
                e2e_host = {"value": fh_edges / max(fh_s, 1e-9) / 1e9, "unit": "GTEPS", "steps": fh_steps,
                            "ms_per_step": 1e3 * fh_s / fh_steps,
                            "h2d_bytes_per_step": int((ro_p.numel() * ro_p.element_size() + co_p.numel() * 4)) * world,
                            "d2h_bytes_per_step": 4 * eng.local_nodes * world,
                            "api": "per rank: grb_graph_slice_from_host (slice in page-locked host memory) + grb_mgbfs_create + "
                                   "peer mapping exchange + grb_mgbfs_run_peer + labels download; max over ranks"}
                del ro_p, co_p
            except Exception as exc:            # an extra figure: never lose the line over it
                e2e_host = {"error": "%s: %s" % (type(exc).__name__, str(exc)[:300])}

        if rank == 0:
            peak, peak_src = measured_peaks()
            value = edges / (ms * 1e-3) / 1e9
            cfg = workload_config(args, scale, eng.nodes, m_global)
            cfg["partition"] = "1D, owner = low bits XOR hash(local row), %d devices" % world
            cfg["level_loop"] = ("one cooperative kernel per device (mgbfs_persistent.cuh)" if os.environ.get("GRB_MG_PERSISTENT") == "1"
                                 else "host-driven (MGBFSEnactor::RunPeer)")
            cfg["exchange"] = ("push: all_to_all of the discovery bitmap; pull: all_gather of the frontier bitmap (NCCL)"
                               if exchange == "nccl" else
                               "peer memory over NVLink (CUDA IPC): pull sweeps store their frontier words on every device, "
                               "push levels store the non-empty lines of the discovery map into the owner's receive buffer, "
                               "counters and barriers are flag words in the peers' mailboxes; no collective call inside a search")
            comp_ms, ex_ms = float(comp.item()) / k, float(exms.item()) / k
            line = {
                "metric": "bfs_gteps_rmat", "value": value, "unit": "GTEPS", "n_gpus": world, "steps": args.steps,
                "warmup": warm, "ms_per_step": ms / max(args.steps, 1), "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "int32", "data": "synthetic", "config": cfg,
                "exchange": {"ms_per_search": ex_ms, "compute_ms_per_search": comp_ms,
                             "bytes_per_search_per_gpu": ex_bytes / k,
                             "achieved_gbs_per_gpu": (ex_bytes / k) / max(ex_ms * 1e-3, 1e-12) / 1e9,
                             "nvlink_peak_gbs_per_direction": 900.0,
                             "levels_per_search": levels / max(args.steps, 1), "first_search_levels": lv_trace},
                "roofline": {"bound": "hbm", "kernel": "advance::pull::Kernel (+ visited merge), all devices",
                             "achieved": (pull_bytes / (pull_ms * 1e-3) / 1e9) if pull_ms > 0 else None,
                             "peak": peak * world, "unit": "GB/s",
                             "frac": (pull_bytes / (pull_ms * 1e-3) / 1e9 / (peak * world)) if pull_ms > 0 else None,
                             "peak_source": peak_src + " x %d devices" % world, "traffic": None,
                             "launches_timed": pull_launches,
                             "note": "algorithmic bytes of the pull levels on all devices over their compute time "
                                     "(events on the stream, max over ranks); compute and exchange time per search "
                                     "are split in `exchange`"},
                "cpu_baseline": None,
                "single_device_same_graph": single,
                "speedup_vs_single_device": (value / single["gteps"]) if single and single.get("gteps") else None,
                "parity": parity,
                "e2e": {"value": e2e_edges / max(e2e_s, 1e-9) / 1e9, "unit": "GTEPS", "h2d_bytes_per_step": 4,
                        "d2h_bytes_per_step": 4 * eng.local_nodes * world, "steps": e2e_steps,
                        "api": "PartitionedBfs.run + labels download on every rank into a reused page-locked buffer "
                               "(graph resident, partitioned; no graph upload inside the step)"},
                "e2e_from_host": e2e_host,
                "gpu_launches": launches_all, "clocks": clocks.summary(),
                "graph500_gteps": value / 2.0, "edges_inspected_per_search": inspected / max(args.steps, 1),
                "setup_s": round(time.time() - t_setup, 1),
            }
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            print(json.dumps(line), flush=True)
            os.dup2(2, 1)
        eng.close()
        g.close()
    finally:
        dist.destroy_process_group()
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
        os.close(saved_stdout)
