"""GPU: the advance and filter operators on their own (grb_test_advance / grb_test_filter) against
a plain numpy restatement of the operator contract: which
(source, destination, CSR entry, input position) tuples the functor sees, what is emitted in
which output mode, and the counters."""
import ctypes as C

import numpy as np
import pytest

from conftest import rmat_graph

pytestmark = pytest.mark.gpu


def run_advance(gr, g, mode, output, queue, out_cap, n, m):
    L = gr._capi.lib()
    q = np.ascontiguousarray(queue, dtype=np.int32) if queue is not None else None
    out = np.full(max(out_cap, 1), -7, dtype=np.int32)
    eh = np.zeros(max(m, 1), np.int32); ss = np.zeros(max(n, 1), np.int64); dh = np.zeros(max(n, 1), np.int32)
    out_len, insp = C.c_int64(), C.c_uint64()
    p = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None
    rc = L.grb_test_advance(g._h, mode, output, p(q), (q.size if q is not None else n), p(out), out_cap,
                            C.byref(out_len), p(eh), p(ss), p(dh), C.byref(insp))
    assert rc == 0
    return out[:max(out_len.value, 0)], out_len.value, eh[:m], ss[:n], dh[:n], insp.value


def expect_advance(ro, co, queue, n):
    m = co.size
    eh = np.zeros(m, np.int64); ss = np.zeros(n, np.int64); dh = np.zeros(n, np.int64)
    slots = []
    for s in queue:
        if s < 0 or s >= n:
            continue
        for e in range(ro[s], ro[s + 1]):
            d = co[e]
            eh[e] += 1; ss[d] += s
            ok = (s + d) % 3 != 0
            if ok:
                dh[d] += 1
            slots.append(d if ok else -1)
    return eh, ss, dh, np.array(slots, dtype=np.int64)


@pytest.mark.parametrize("mode", [0, 1], ids=["twc", "lb"])
@pytest.mark.parametrize("scale", [8, 12])
def test_advance_contract(gr, orc, mode, scale):
    ro, co = rmat_graph(orc, scale, 8)
    n, m = ro.size - 1, co.size
    g = gr.Graph.from_csr(ro, co)
    rng = np.random.RandomState(scale)
    queues = [np.array([0]), np.arange(n), rng.randint(0, n, 300), np.array([], dtype=np.int64),
              np.array([5, -1, 5, n + 3, 0, -1]),                  # holes, duplicates, out-of-range ids are skipped
              None]                                               # NULL queue = identity frontier
    for q in queues:
        qq = np.arange(n) if q is None else q
        eh, ss, dh, slots = expect_advance(ro, co, qq, n)
        outputs = [0, 1] + ([2] if mode == 1 else [])
        for output in outputs:
            out, out_len, geh, gss, gdh, insp = run_advance(gr, g, mode, output, q, slots.size + 8, n, m)
            assert np.array_equal(geh, eh) and np.array_equal(gss, ss) and np.array_equal(gdh, dh)
            assert insp == slots.size
            if output == 1:                                      # compact: the passing destinations, any order
                assert out_len == int((slots >= 0).sum())
                assert np.array_equal(np.sort(out), np.sort(slots[slots >= 0]))
            elif output == 2:                                    # tombstone: one slot per edge, frontier order
                assert out_len == slots.size
                assert np.array_equal(out, slots)
    g.close()


@pytest.mark.parametrize("mode", [0, 1], ids=["twc", "lb"])
@pytest.mark.parametrize("atype", [1, 2, 3], ids=["V2E", "E2V", "E2E"])
def test_advance_frontier_types(gr, orc, mode, atype):
    """reference advance::TYPE (oprtr/advance/kernel_policy.cuh:54-60): an input EDGE id stands for the vertex
    the edge points to (edge_map_partitioned/kernel.cuh:356-359); an output edge frontier receives the CSR
    entry index of each successful edge (:409-413)"""
    ro, co = rmat_graph(orc, 10, 8)
    n, m = ro.size - 1, co.size
    g = gr.Graph.from_csr(ro, co)
    rng = np.random.RandomState(atype)
    in_edges, out_edges = atype in (2, 3), atype in (1, 3)
    queues = [rng.randint(0, m if in_edges else n, 500), np.array([3, -1, 3, (m if in_edges else n) + 5, 0]),
              np.arange(m if in_edges else n)[::7]]
    for q in queues:
        while True:     # the harness sizes the LB scratch for m expanded entries (a frontier without duplicates)
            verts = [(int(co[x]) if 0 <= x < m else -1) for x in q] if in_edges else list(q)
            if sum(int(ro[v + 1] - ro[v]) for v in verts if 0 <= v < n) <= m:
                break
            q = q[:len(q) // 2]
        eh, ss, dh, slots = expect_advance(ro, co, verts, n)
        eids = []                                                    # CSR entry behind every slot, frontier order
        for s in verts:
            if 0 <= s < n:
                eids.extend(range(ro[s], ro[s + 1]))
        eids = np.array(eids, dtype=np.int64)
        want = np.where(slots >= 0, eids, -1) if out_edges else slots
        for output in [0, 1] + ([2] if mode == 1 else []):
            out, out_len, geh, gss, gdh, insp = run_advance(gr, g, mode + 16 * atype, output, q, slots.size + 8, n, m)
            assert np.array_equal(geh, eh) and np.array_equal(gss, ss) and np.array_equal(gdh, dh)
            assert insp == slots.size
            if output == 1:
                assert np.array_equal(np.sort(out), np.sort(want[want >= 0]))
            elif output == 2:
                assert np.array_equal(out, want)
    g.close()


@pytest.mark.parametrize("op", [1, 3, 5, 6, 7, 8, 9], ids=["plus", "mul", "or", "and", "xor", "max", "min"])
@pytest.mark.parametrize("rtype", [1, 2], ids=["vertex", "edge"])
def test_neighbor_reduce(gr, orc, op, rtype):
    """reference: advance with R_OP / R_TYPE + SegReduceCsr (oprtr/advance/kernel.cuh:731-761,
    edge_map_partitioned/kernel.cuh:417-470): accepted edges contribute values[d] / values[e], rejected ones
    and empty lists the operator's identity"""
    ro, co = rmat_graph(orc, 10, 8)
    n, m = ro.size - 1, co.size
    g = gr.Graph.from_csr(ro, co)
    rng = np.random.RandomState(10 * op + rtype)
    values = (rng.randint(1, 3, m if rtype == 2 else n) if op == 3 else rng.randint(-1000, 1000, m if rtype == 2 else n)).astype(np.int32)
    ident = {1: 0, 3: 1, 5: 0, 6: -1, 7: 0, 8: -2**31, 9: 2**31 - 1}[op]
    fold = {1: lambda a, b: a + b, 3: lambda a, b: a * b, 5: lambda a, b: a | b, 6: lambda a, b: a & b,
            7: lambda a, b: a ^ b, 8: max, 9: min}[op]
    L = gr._capi.lib()
    p = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None
    for atype, q in ((0, np.concatenate([rng.randint(0, n, 400), [-1, n + 2, 0]])), (0, None), (2, rng.randint(0, m, 300))):
        items = np.arange(n) if q is None else q
        verts = [(int(co[x]) if 0 <= x < m else -1) for x in items] if atype == 2 else list(items)
        want, emitted = [], []
        for s in verts:
            acc = ident
            if 0 <= s < n:
                for e in range(ro[s], ro[s + 1]):
                    d = int(co[e])
                    if (s + d) % 3 != 0:
                        acc = fold(acc, int(values[e if rtype == 2 else d]))
                        if op == 3:
                            acc = ((acc + 2**31) % 2**32) - 2**31        # int32 wrap-around of the product
                        emitted.append(d)
            want.append(acc)
        qq = np.ascontiguousarray(q, dtype=np.int32) if q is not None else None
        got = np.zeros(len(items), np.int32)
        out = np.zeros(len(emitted) + 8, np.int32)
        out_len = C.c_int64()
        rc = L.grb_test_neighbor_reduce(g._h, op, rtype, atype, p(qq), len(items), p(values), p(got), p(out), out.size,
                                        C.byref(out_len))
        assert rc == 0
        assert np.array_equal(got.astype(np.int64), np.array(want, dtype=np.int64))
        assert out_len.value == len(emitted)
        assert np.array_equal(np.sort(out[:out_len.value]), np.sort(np.array(emitted, dtype=np.int32)))
    g.close()


def test_advance_reports_overflow(gr, orc):
    ro, co = rmat_graph(orc, 10, 8)
    n, m = ro.size - 1, co.size
    g = gr.Graph.from_csr(ro, co)
    for mode in (0, 1):
        _, out_len, *_ = run_advance(gr, g, mode, 1, np.arange(n), 10, n, m)
        assert out_len == -1


def run_filter(gr, n, compact, queue, out_cap):
    L = gr._capi.lib()
    q = np.ascontiguousarray(queue, dtype=np.int32) if queue is not None else None
    out = np.full(max(out_cap, 1), -7, dtype=np.int32)
    dh = np.zeros(max(n, 1), np.int32)
    out_len = C.c_int64()
    p = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None
    rc = L.grb_test_filter(0, n, compact, p(q), (q.size if q is not None else n), p(out), out_cap,
                           C.byref(out_len), p(dh))
    assert rc == 0
    return out[:max(out_len.value, 0)], out_len.value, dh[:n]


@pytest.mark.parametrize("size", [0, 1, 1000, 1024, 100001])
def test_filter_contract(gr, size):
    n = 5000
    rng = np.random.RandomState(size)
    q = rng.randint(-1, n, size)                                 # -1 = tombstones left by the advance
    keep = q[(q >= 0) & (q % 2 == 0)]
    want_hits = np.bincount(keep, minlength=n)
    out, out_len, dh = run_filter(gr, n, 1, q, size + 4)
    assert out_len == keep.size
    # compaction is stable inside a 1024-item tile (as the reference's raking scan, filter/cta.cuh:520-544);
    # tiles of different CTAs land in the order their atomics were served
    assert np.array_equal(np.sort(out), np.sort(keep))
    if size <= 1024:
        assert np.array_equal(out, keep)
    else:
        tiles = [q[i:i + 1024] for i in range(0, size, 1024)]
        runs = sorted(tuple(t[(t >= 0) & (t % 2 == 0)].tolist()) for t in tiles)
        bounds, got_runs, at = sorted(len(r) for r in runs), [], 0
        # every tile's survivors form one contiguous, ordered run of the output
        lens = {}
        for r in runs:
            lens.setdefault(r[:8], []).append(r)
        while at < out.size:
            cands = lens.get(tuple(out[at:at + 8].tolist()), [])
            hit = next((r for r in cands if tuple(out[at:at + len(r)].tolist()) == r), None)
            assert hit is not None, "output position %d does not start a tile's run" % at
            cands.remove(hit)
            at += len(hit)
    assert np.array_equal(dh, want_hits)
    # apply-only map (reference Kernel2): out = in, functor applied to the passing items
    out, out_len, dh = run_filter(gr, n, 0, q, size + 4)
    assert out_len == size and np.array_equal(out, q) and np.array_equal(dh, want_hits)


def test_filter_identity_frontier_and_overflow(gr):
    n = 3000
    out, out_len, dh = run_filter(gr, n, 1, None, n)
    assert np.array_equal(np.sort(out), np.arange(0, n, 2)) and dh.sum() == n // 2
    _, out_len, _ = run_filter(gr, n, 1, np.arange(n), 10)
    assert out_len == -1


# ---- priority queue: the near/far split stand-alone (reference: gunrock/priority_queue/kernel.cuh:404-503) ----------
def run_bisect(gr, n, scores, items, in_is_far, lower, upper, mark, cap=None):
    import ctypes as C
    from gunrockinst_b200 import _capi
    items = np.ascontiguousarray(items, dtype=np.int32)
    scores = np.ascontiguousarray(scores, dtype=np.uint32)
    mark = np.ascontiguousarray(mark, dtype=np.uint32).copy()
    cap = items.size + 4 if cap is None else cap
    near = np.empty(cap + 1, dtype=np.int32)
    far = np.empty(cap + 1, dtype=np.int32)
    nl, fl, mf = C.c_int64(), C.c_int64(), C.c_uint32()
    rc = _capi.lib().grb_test_bisect(0, n, scores.ctypes.data_as(C.c_void_p), items.ctypes.data_as(C.c_void_p), items.size,
                                     int(in_is_far), lower, upper, mark.ctypes.data_as(C.c_void_p),
                                     near.ctypes.data_as(C.c_void_p), cap, C.byref(nl),
                                     far.ctypes.data_as(C.c_void_p), cap, C.byref(fl), C.byref(mf))
    assert rc == 0
    return near[:max(nl.value, 0)], nl.value, far[:max(fl.value, 0)], fl.value, mark, mf.value


@pytest.mark.parametrize("size", [0, 1, 1023, 1024, 50000])
@pytest.mark.parametrize("in_is_far", [False, True])
def test_near_far_split_contract(gr, size, in_is_far):
    n = 7000
    rng = np.random.RandomState(31 * size + int(in_is_far))
    scores = rng.randint(0, 100, n).astype(np.uint32)
    lower, upper = 20, 60
    words = (n + 31) // 32
    if in_is_far:
        # the input is a far pile: distinct members, all marked; holes (-1) left by earlier splits
        items = rng.permutation(n)[:size].astype(np.int32)
        mark = np.zeros(words, np.uint32)
        np.bitwise_or.at(mark, items >> 5, (np.uint32(1) << (items & 31).astype(np.uint32)))
        if items.size > 10:
            items[rng.randint(0, items.size, items.size // 10)] = -1
    else:
        # a frontier with duplicates and holes; some vertices already sit in the far pile
        items = rng.randint(-1, n, size).astype(np.int32)
        already = rng.permutation(n)[:n // 5]
        mark = np.zeros(words, np.uint32)
        np.bitwise_or.at(mark, already >> 5, (np.uint32(1) << (already & 31).astype(np.uint32)))
    mark0 = mark.copy()
    near, nl, far, fl, mark_after, min_far = run_bisect(gr, n, scores, items, in_is_far, lower, upper, mark)
    live = items[items >= 0]
    sc = scores[live]
    want_near = live[(sc >= lower) & (sc < upper)]
    marked0 = ((mark0[live >> 5] >> (live & 31).astype(np.uint32)) & 1).astype(bool)
    far_cands = live[sc >= upper]
    if in_is_far:
        want_far = far_cands                                    # stays in the pile
        want_mark = mark0.copy()                                # ... and every item that left it clears its bit
        gone = live[sc < upper]
        np.bitwise_and.at(want_mark, gone >> 5, ~(np.uint32(1) << (gone & 31).astype(np.uint32)))
    else:
        want_far = np.unique(far_cands[~marked0[sc >= upper]])  # one copy of every vertex not in the pile yet
        want_mark = mark0.copy()
        np.bitwise_or.at(want_mark, want_far >> 5, (np.uint32(1) << (want_far & 31).astype(np.uint32)))
    assert nl == want_near.size and np.array_equal(np.sort(near), np.sort(want_near))
    assert fl == want_far.size and np.array_equal(np.sort(far), np.sort(want_far))
    assert np.array_equal(mark_after, want_mark)
    assert min_far == (int(scores[want_far].min()) if want_far.size else 0xFFFFFFFF)


def test_near_far_split_reports_overflow(gr):
    n = 4000
    scores = np.full(n, 5, np.uint32)
    _, nl, _, fl, _, _ = run_bisect(gr, n, scores, np.arange(n), False, 0, 10, np.zeros((n + 31) // 32, np.uint32), cap=100)
    assert nl == -1


# ---- rows binned by length, stand-alone (oprtr/row_bins.cuh: the fixed-row-set passes of PageRank and CC) -----------
def run_row_bins(gr, g, queue):
    from gunrockinst_b200 import _capi
    n, m = g.local_nodes, g.edges
    q = np.ascontiguousarray(queue, dtype=np.int32) if queue is not None else None
    qlen = q.size if q is not None else n
    small = np.full(qlen + 1, -1, np.int32); medium = np.full(qlen + 1, -1, np.int32); big = np.full(qlen + 1, -1, np.int32)
    cap = 2 * (m // 2048) + 2
    pieces = np.full((cap, 2), -1, np.int32)
    counts = np.zeros(8, np.int64)
    p = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None
    rc = _capi.lib().grb_test_row_bins(g._h, p(q), qlen, p(small), p(medium), p(big), p(pieces), cap, p(counts))
    assert rc == 0
    return small[:counts[0]], medium[:counts[1]], big[:counts[2]], pieces[:counts[3]], counts


def star_csr(n, hubs):
    """undirected: every hub is adjacent to every vertex; plus a ring, so that all three bins are populated"""
    rows = [set() for _ in range(n)]
    for h in hubs:
        for v in range(n):
            if v != h:
                rows[h].add(v); rows[v].add(h)
    for v in range(0, n, 3):
        for k in range(1, 1 + (v % 40)):
            w = (v + k) % n
            rows[v].add(w); rows[w].add(v)
    ro = np.zeros(n + 1, np.int64)
    ro[1:] = np.cumsum([len(r) for r in rows])
    co = np.concatenate([np.array(sorted(r), np.int32) for r in rows])
    return ro, co


@pytest.mark.parametrize("case", ["rmat14", "star", "star_subset", "empty_rows"])
def test_row_bins_contract(gr, orc, case):
    if case == "rmat14":
        ro, co = rmat_graph(orc, 14, 16)
        queue = None
    elif case in ("star", "star_subset"):
        ro, co = star_csr(5000, [0, 17, 4999])
        queue = None if case == "star" else np.random.RandomState(5).permutation(5000)[:3111]
    else:
        ro = np.array([0, 0, 3, 3, 3, 20, 20], np.int64)
        co = np.arange(20, dtype=np.int32) % 6
        queue = None
    n = ro.size - 1
    g = gr.Graph.from_csr(ro, co, symmetric=False)
    small, medium, big, pieces, counts = run_row_bins(gr, g, queue)
    g.close()
    rows = np.arange(n) if queue is None else np.asarray(queue)
    ln = (ro[1:] - ro[:-1])[rows]
    assert np.array_equal(np.sort(small), np.sort(rows[(ln >= 1) & (ln <= 16)]))
    assert np.array_equal(np.sort(medium), np.sort(rows[(ln >= 17) & (ln <= 2048)]))
    assert np.array_equal(np.sort(big), np.sort(rows[ln > 2048]))
    want_pieces = sorted((int(r), k) for r, l in zip(rows[ln > 2048], ln[ln > 2048]) for k in range((int(l) + 2047) // 2048))
    assert sorted(map(tuple, pieces.tolist())) == want_pieces
    assert counts[4] == int(ln.sum())
    # the CTAs of a pass cover every list: 8 pieces, 32 medium rows, 256 small rows per CTA
    assert counts[5] == (len(want_pieces) + 7) // 8 and counts[6] == (medium.size + 31) // 32 and counts[7] == (small.size + 255) // 256
    if case == "star":
        assert big.size == 3 and medium.size > 0 and small.size > 0
