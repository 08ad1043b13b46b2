"""A numpy stand-in for gunrockinst_b200.partitioned.DeviceEngine, used ONLY by the CPU tests of
the partitioned host loop (world_size 2 over gloo): it plays the part of one device's
grb_mgbfs_* steps on CPU tensors so that the driver's exchange / direction / termination logic
can be checked against the oracle without a GPU.  Test infrastructure, not a product path."""
import contextlib
import time

import numpy as np
import torch

from gunrockinst_b200.partitioned import CTR_WORDS, global_ids


def _bits(words_i32, n_bits):
    return np.unpackbits(words_i32.view(np.uint8), bitorder="little")[:n_bits].astype(bool)


def _words(bits):
    return np.packbits(bits, bitorder="little").view(np.int32)


class NumpyEngine:
    def __init__(self, row_offsets, col, log_parts, rank):
        self.parts, self.rank, self.log_parts = 1 << log_parts, rank, log_parts
        self.nodes = row_offsets.size - 1
        self.local_nodes = self.nodes >> log_parts
        assert self.local_nodes % 32 == 0
        self.seg_words = self.local_nodes // 32
        self.gid = [global_ids(self.local_nodes, log_parts, r) for r in range(self.parts)]
        self.pos = np.empty(self.nodes, dtype=np.int64)          # vertex -> owner-major bit position
        for r in range(self.parts):
            self.pos[self.gid[r]] = r * self.local_nodes + np.arange(self.local_nodes)
        mine = self.gid[rank]
        self.adj = [col[row_offsets[v]:row_offsets[v + 1]] for v in mine]
        self.deg = np.array([a.size for a in self.adj], dtype=np.int64)
        self.local_edges = int(self.deg.sum())
        words = self.seg_words * self.parts
        self.maps = [torch.zeros(words, dtype=torch.int32), torch.zeros(words, dtype=torch.int32)]
        self.visited = torch.zeros(words, dtype=torch.int32)
        self.next = torch.zeros(words, dtype=torch.int32)
        self.recv = torch.zeros(words, dtype=torch.int32)
        self.counters = torch.zeros(CTR_WORDS, dtype=torch.int64)
        self.lab = np.full(self.local_nodes, -1, dtype=np.int32)
        self.queue = np.zeros(0, dtype=np.int64)                 # local rows

    # -- helpers ------------------------------------------------------------------------------
    def _get(self, t):
        return _bits(t.numpy(), self.nodes)

    def _put(self, t, bits):
        t.copy_(torch.from_numpy(_words(bits).copy()))

    def _seg(self):
        return slice(self.rank * self.local_nodes, (self.rank + 1) * self.local_nodes)

    # -- the engine interface ----------------------------------------------------------------------
    def reset(self, src):
        for t in (*self.maps, self.visited, self.next, self.recv):
            t.zero_()
        self.counters.zero_()
        self.lab[:] = -1
        self.queue = np.zeros(0, dtype=np.int64)
        if not (0 <= src < self.nodes):
            return 0
        vis = self._get(self.visited)
        vis[self.pos[src]] = True
        self._put(self.visited, vis)
        p = self.pos[src]
        if p // self.local_nodes == self.rank:
            row = p % self.local_nodes
            self.lab[row] = 0
            self.queue = np.array([row], dtype=np.int64)
            return int(self.deg[row])
        return 0

    def push(self, level, queue_len, edges_hint):
        assert queue_len == self.queue.size
        self.counters.zero_()
        vis = self._get(self.visited)
        nxt = np.zeros(self.nodes, dtype=bool)
        inspected = 0
        for row in self.queue:
            nb = self.adj[row]
            inspected += nb.size
            p = self.pos[nb]
            nxt[p[~vis[p]]] = True
        self._put(self.next, nxt)
        self.counters[2] = inspected

    def absorb(self, level, cur_out):
        rec = self.recv.numpy().reshape(self.parts, self.seg_words)
        acc = np.bitwise_or.reduce(rec, axis=0)
        got = _bits(acc, self.local_nodes)
        vis = self._get(self.visited)
        new = got & ~vis[self._seg()]
        vis[self._seg()] |= new
        self._put(self.visited, vis)
        m = self._get(self.maps[cur_out])
        m[self._seg()] = new
        self._put(self.maps[cur_out], m)
        rows = np.nonzero(new)[0]
        self.lab[rows] = level + 1
        self.queue = rows
        self.counters[0] = rows.size
        self.counters[1] = int(self.deg[rows].sum())

    def pull(self, level, cur):
        self.counters.zero_()
        fin = self._get(self.maps[cur])
        vis = self._get(self.visited)
        out = self._get(self.maps[cur ^ 1])
        seg = self._seg()
        new = np.zeros(self.local_nodes, dtype=bool)
        inspected = cands = 0
        for row in np.nonzero(~vis[seg] & (self.deg > 0))[0]:
            cands += 1
            hit = np.nonzero(fin[self.pos[self.adj[row]]])[0]
            inspected += (hit[0] + 1) if hit.size else self.adj[row].size
            if hit.size:
                new[row] = True
        out[seg] = new
        vis[seg] |= new
        self._put(self.maps[cur ^ 1], out)
        self._put(self.visited, vis)
        rows = np.nonzero(new)[0]
        self.lab[rows] = level + 1
        self.counters[0] = rows.size
        self.counters[1] = int(self.deg[rows].sum())
        self.counters[2] = int(inspected)
        self.counters[4] = cands

    def merge_visited(self, cur):
        self.visited |= self.maps[cur]

    def queue_to_map(self, cur, queue_len):
        m = np.zeros(self.nodes, dtype=bool)
        m[self.rank * self.local_nodes + self.queue] = True
        self._put(self.maps[cur], m)

    def map_to_queue(self, cur):
        self.queue = np.nonzero(self._get(self.maps[cur])[self._seg()])[0]

    def labels(self):
        return self.lab.copy()

    def launches(self):
        return 0

    def stream_context(self):
        return contextlib.nullcontext()

    def mark(self):
        return time.perf_counter()

    @staticmethod
    def elapsed_ms(a, b):
        return (b - a) * 1e3

    def synchronize(self):
        pass
