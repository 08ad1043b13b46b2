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


def partitioned_pagerank(row_offsets, col, log_parts, rank, group=None, src=-1, delta=0.85, threshold=1e-6, max_iter=1000):
    """numpy restatement of the steps of csrc/gunrock/app/pr/mgpr_enactor.cuh for one rank (the
    replicas are filled by all_gather where the kernels store into peer memory, the counter blocks by
    all_reduce): pruning with the replicated "degree just reached zero" map, then gather sweeps over
    the in-edges of the owned rows from the replicated contributions.  Works on the in- and out-edge
    lists separately, so it also covers directed graphs (the device code requires a symmetric one).
    -> (fp64 ranks of the owned rows in local-row order, sweeps, pruning sweeps)"""
    import torch.distributed as dist
    import scipy.sparse as sp
    parts = 1 << log_parts
    n = row_offsets.size - 1
    ln = n >> log_parts
    gids = [global_ids(ln, log_parts, r) for r in range(parts)]
    pos = np.empty(n, dtype=np.int64)                         # vertex -> owner-major replica position
    for r in range(parts):
        pos[gids[r]] = r * ln + np.arange(ln)
    mine = gids[rank]
    a = sp.csr_matrix((np.ones(col.size, dtype=np.int8), col, row_offsets), shape=(n, n))
    at = a.T.tocsr()
    out_rows = [col[row_offsets[v]:row_offsets[v + 1]] for v in mine]
    in_rows = [at.indices[at.indptr[v]:at.indptr[v + 1]] for v in mine]

    def publish(segment):                                     # every device's segment -> the whole replica
        if parts == 1:
            return segment.copy()
        t = [torch.empty(ln, dtype=torch.from_numpy(segment).dtype) for _ in range(parts)]
        dist.all_gather(t, torch.from_numpy(np.ascontiguousarray(segment)), group=group)
        return np.concatenate([x.numpy() for x in t])

    def total(*vals):
        t = torch.tensor(vals, dtype=torch.int64)
        if parts > 1:
            dist.all_reduce(t, group=group)
        return [int(x) for x in t]

    deg = np.array([r.size for r in out_rows], dtype=np.int64)
    rank_v = np.full(ln, 1.0 - delta)
    zero = publish((deg == 0).astype(np.uint8))
    qlen, zeros = total(ln, int((deg == 0).sum()))
    prev, prune_sweeps = 0, 0
    while prev != qlen:
        prev = qlen
        prune_sweeps += 1
        if zeros == 0:
            break
        my_zero = zero[rank * ln:(rank + 1) * ln].astype(bool)
        for i, row in enumerate(out_rows):                    # advance: one off per out-neighbour that just hit zero
            deg[i] -= int(zero[pos[row]].sum())
        deg[my_zero] = -1                                     # rows that were at zero when the sweep began leave
        alive = deg >= 0
        new_zero = alive & (deg == 0)
        zero = publish(new_zero.astype(np.uint8))
        qlen, zeros = total(int(alive.sum()), int(new_zero.sum()))
    sweeps = 0
    if qlen > 0 and max_iter > 0:
        alive = deg > 0
        rank_v[~alive] = 0.0
        contrib = publish(np.where(alive, rank_v / np.maximum(deg, 1), 0.0))
        base = np.array([1.0 if (src == -1 or src == v) else 0.0 for v in mine])
        while True:
            acc = np.array([contrib[pos[row]].sum() if alive[i] else 0.0 for i, row in enumerate(in_rows)])
            nxt = delta * acc + (1.0 - delta) * base
            moving = int((np.abs(nxt - rank_v)[alive] > threshold).sum())
            rank_v[alive] = nxt[alive]
            contrib = publish(np.where(alive, rank_v / np.maximum(deg, 1), 0.0))
            sweeps += 1
            (moving,) = total(moving)
            if moving == 0 or sweeps >= max_iter:
                break
    return rank_v, sweeps, prune_sweeps


def partitioned_cc(row_offsets, col, log_parts, rank, group=None):
    """numpy restatement of csrc/gunrock/app/cc/mgcc_enactor.cuh for one rank: minimum-label
    propagation along the out-edges of the owned rows, offers to remote owners delivered by an
    all_gather of (vertex, label) pairs where the kernels issue atomicMin into peer memory.
    -> (labels of the owned rows in local-row order, iterations)"""
    import torch.distributed as dist
    parts = 1 << log_parts
    n = row_offsets.size - 1
    ln = n >> log_parts
    mine = global_ids(ln, log_parts, rank)
    local_of = {int(v): i for i, v in enumerate(mine)}
    label = mine.copy()
    queued = label.copy()
    best = np.arange(n, dtype=np.int64)                 # best label sent per global vertex
    frontier = list(range(ln))
    iterations = 0
    while True:
        offers = {}
        for i in frontier:
            v = int(mine[i])
            lab = int(label[i])
            for d in col[row_offsets[v]:row_offsets[v + 1]]:
                d = int(d)
                if lab < best[d]:
                    best[d] = lab
                    offers[d] = min(lab, offers.get(d, lab))
        if parts > 1:
            every = [None] * parts
            dist.all_gather_object(every, offers, group=group)
        else:
            every = [offers]
        touched = set()
        for off in every:
            for d, lab in off.items():
                if d in local_of and lab < label[local_of[d]]:
                    label[local_of[d]] = lab
                    touched.add(local_of[d])
        frontier = []
        for i in sorted(touched):
            if label[i] < queued[i]:
                queued[i] = label[i]
                frontier.append(i)
        total = torch.tensor([len(frontier)], dtype=torch.int64)
        if parts > 1:
            dist.all_reduce(total, group=group)
        iterations += 1
        if int(total) == 0:
            break
    return label.astype(np.int32), iterations
