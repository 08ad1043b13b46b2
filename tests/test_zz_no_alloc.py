"""GPU: grb_advance_lb_noalloc -- the LB advance for callers that cannot allocate (its scratch is a static
__device__ array inside the library, include/gunrock_b200.h) -- against the existing entry point
grb_test_advance(mode = LB) on the same graph and frontiers, through the C ABI, twice in a row and more:
the second call finds the first call's scratch contents and must not be disturbed by them.

Integer work: the comparison is exact (hit counts, source sums, counters; the compacted output frontier
as a multiset, as in test_operators_gpu.py -- its order depends on the schedule in both entry points).

The file sorts last on purpose (pytest -x): it has no reference counterpart and must not hide the parity
tests.  The device buffers the caller owns are torch tensors here; the entry point itself allocates nothing.
"""
import ctypes as C

import numpy as np
import pytest

from conftest import rmat_graph

pytestmark = pytest.mark.gpu

MAX_ITEMS = 262144          # GRB_NOALLOC_MAX_ITEMS


def existing_entry_point(gr, g, queue, out_cap, n, m):
    L = gr._capi.lib()
    q = np.ascontiguousarray(queue, dtype=np.int32)
    out = np.full(max(out_cap, 1), -7, dtype=np.int32)
    eh = np.zeros(max(m, 1), np.int32); ss = np.zeros(max(n, 1), np.int64); dh = np.zeros(max(n, 1), np.int32)
    out_len, insp = C.c_int64(), C.c_uint64()
    p = lambda a: a.ctypes.data_as(C.c_void_p)
    rc = L.grb_test_advance(g._h, 1, 1, p(q), q.size, p(out), out_cap, C.byref(out_len), p(eh), p(ss), p(dh), C.byref(insp))
    assert rc == 0
    return out[:max(out_len.value, 0)], out_len.value, eh[:m], ss[:n], dh[:n], insp.value


class DeviceBuffers:
    """what a caller of the no-alloc entry point owns; made once, reused by every call"""

    def __init__(self, torch, n, m, in_cap, out_cap):
        dev = "cuda:0"
        self.torch = torch
        self.q = torch.empty(max(in_cap, 1), dtype=torch.int32, device=dev)
        self.out = torch.empty(max(out_cap, 1), dtype=torch.int32, device=dev)
        self.eh = torch.empty(max(m, 1), dtype=torch.int32, device=dev)
        self.ss = torch.empty(max(n, 1), dtype=torch.int64, device=dev)
        self.dh = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
        self.counters = torch.empty(8, dtype=torch.int64, device=dev)
        self.out_cap = out_cap
        self.calls = 0

    def run(self, gr, g, queue, n, m, output=1):
        torch = self.torch
        L = gr._capi.lib()
        q = torch.from_numpy(np.ascontiguousarray(queue, dtype=np.int32))
        self.q[:q.numel()].copy_(q)
        self.out.fill_(-7); self.eh.zero_(); self.ss.zero_(); self.dh.zero_(); self.counters.fill_(-1)
        stream = torch.cuda.current_stream()
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]       # cudaMemGetInfo: sees any allocation, not only torch's
        rc = L.grb_advance_lb_noalloc(g._h, output, self.q.data_ptr(), q.numel(), self.out.data_ptr(), self.out_cap,
                                      self.eh.data_ptr(), self.ss.data_ptr(), self.dh.data_ptr(),
                                      self.counters.data_ptr(), C.c_void_p(stream.cuda_stream))
        assert rc == 0
        stream.synchronize()
        if self.calls:                                   # (the first launch loads the kernels: the driver's business)
            assert torch.cuda.mem_get_info()[0] >= free_before
        self.calls += 1
        c = self.counters.cpu().numpy().astype(np.uint64)
        out_len, insp, overflow = int(c[0]), int(c[2]), int(c[6])
        return (self.out[:out_len].cpu().numpy(), out_len, self.eh[:m].cpu().numpy(), self.ss[:n].cpu().numpy(),
                self.dh[:n].cpu().numpy(), insp, overflow)


def same(a, b):
    out_a, len_a, eh_a, ss_a, dh_a, insp_a = a
    out_b, len_b, eh_b, ss_b, dh_b, insp_b, overflow = b
    assert overflow == 0
    assert len_a == len_b and insp_a == insp_b
    assert np.array_equal(eh_a, eh_b) and np.array_equal(ss_a, ss_b) and np.array_equal(dh_a, dh_b)
    assert np.array_equal(np.sort(out_a), np.sort(out_b))


@pytest.mark.parametrize("offset_bytes", [4, 8], ids=["offsets32", "offsets64"])
def test_no_alloc_advance_matches_the_existing_entry_point(gr, orc, offset_bytes):
    import torch
    ro, co = rmat_graph(orc, 12, 8)
    n, m = ro.size - 1, co.size
    g = gr.Graph.from_csr(ro, co, force_offset_bytes=offset_bytes)
    rng = np.random.RandomState(5)
    # a big frontier first, smaller ones after it (they see its leftovers in the scratch), holes and duplicates,
    # an empty frontier, and the first one again
    queues = [np.arange(n), rng.randint(0, n, 300), np.array([5, -1, 5, n + 3, 0, -1]), np.array([0]),
              np.array([], dtype=np.int64), rng.randint(0, n, 3000), np.arange(n)]
    cap = int(m) * 2 + 64
    bufs = DeviceBuffers(torch, n, m, max(q.size for q in queues), cap)
    for q in queues:
        want = existing_entry_point(gr, g, q, cap, n, m)
        for _ in range(2):                      # two calls in a row on the same input
            same(want, bufs.run(gr, g, q, n, m))
    # output = none: the functor still runs, nothing is enqueued
    want = existing_entry_point(gr, g, queues[1], cap, n, m)
    got = bufs.run(gr, g, queues[1], n, m, output=0)
    assert got[1] == 0 and got[6] == 0 and got[5] == want[5]
    assert np.array_equal(got[2], want[2]) and np.array_equal(got[3], want[3]) and np.array_equal(got[4], want[4])
    g.close()


def test_no_alloc_advance_reports_a_full_output_queue(gr, orc):
    import torch
    ro, co = rmat_graph(orc, 8, 8)
    n, m = ro.size - 1, co.size
    g = gr.Graph.from_csr(ro, co)
    bufs = DeviceBuffers(torch, n, m, n, 16)
    got = bufs.run(gr, g, np.arange(n), n, m)
    assert got[6] != 0                          # overflow counter, as the existing entry point's out_len = -1
    # and the next call is clean again
    want = existing_entry_point(gr, g, np.array([0]), 4096, n, m)
    bufs2 = DeviceBuffers(torch, n, m, n, 4096)
    same(want, bufs2.run(gr, g, np.array([0]), n, m))
    g.close()


def test_no_alloc_advance_rejects_sizes_beyond_its_scratch(gr, orc, capfd):
    import torch
    ro, co = rmat_graph(orc, 8, 8)
    n, m = ro.size - 1, co.size
    g = gr.Graph.from_csr(ro, co)
    L = gr._capi.lib()
    bufs = DeviceBuffers(torch, n, m, MAX_ITEMS + 1, 64)
    bufs.q.zero_()
    bufs.counters.fill_(-1)
    torch.cuda.synchronize()
    rc = L.grb_advance_lb_noalloc(g._h, 1, bufs.q.data_ptr(), MAX_ITEMS + 1, bufs.out.data_ptr(), 64,
                                  bufs.eh.data_ptr(), bufs.ss.data_ptr(), bufs.dh.data_ptr(), bufs.counters.data_ptr(), None)
    assert rc != 0
    assert "GRB_NOALLOC_MAX_ITEMS" in capfd.readouterr().err
    torch.cuda.synchronize()
    assert int(bufs.counters[0].item()) == -1   # nothing ran
    # the largest size it accepts: every item is the same low-degree vertex
    deg = np.diff(ro)
    v = int(np.flatnonzero((deg > 0) & (deg <= 32))[0])
    bufs.q.fill_(v)
    rc = L.grb_advance_lb_noalloc(g._h, 0, bufs.q.data_ptr(), MAX_ITEMS, None, 0,
                                  bufs.eh.data_ptr(), bufs.ss.data_ptr(), bufs.dh.data_ptr(), bufs.counters.data_ptr(), None)
    assert rc == 0
    torch.cuda.synchronize()
    assert int(bufs.counters[2].item()) == MAX_ITEMS * int(deg[v]) and int(bufs.counters[6].item()) == 0
    g.close()
