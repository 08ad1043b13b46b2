#!/usr/bin/env python
"""bench.py -- BFS GTEPS on seeded synthetic R-MAT graphs (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--scale S]

One JSON line on stdout (rank 0).  A *step* is one breadth-first search from one
source of a fixed seeded source list over the resident graph.

  N = 1   BASELINE.json configs[1]: direction-optimising BFS, R-MAT scale 22,
          edgefactor 16, undirected, graph resident in HBM.
  N > 1   BASELINE.json configs[4]: R-MAT scale 27, 1D vertex-partitioned over N
          devices (one process per device), per-level frontier exchange over NCCL.

value   = sum(edges_visited) / device time of the K steps (CUDA events on the stream the
          library launches on, max over ranks); edges_visited = sum of out-degrees of the
          visited vertices, the reference's own numerator (gunrock/app/bfs/bfs_app.cu:82-91).
e2e     = same metric through the reference's C entry point gunrock_bfs_func
          (gunrock/gunrock.h:106-110) with the CSR in pinned HOST memory: upload, search and
          label download are all inside the timed region.
roofline, cpu_baseline: see DESIGN.md section "Measurement".

--impl reference times the reference's CPU BFS (tests/bfs/test_bfs.cu:258-322, restated in
oracle/oracle.c -- the reference TU itself does not compile) on the host cores.
"""
import argparse
import contextlib
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "bfs_gteps_rmat"
UNIT = "GTEPS"
N_SOURCES = 64


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


@contextlib.contextmanager
def quiet_stdout():
    """the reference API prints its statistics with printf (bfs_app.cu:103-118); keep this
    program's stdout to the one JSON line"""
    sys.stdout.flush()
    saved = os.dup(1)
    devnull = os.open(os.devnull, os.O_WRONLY)
    os.dup2(devnull, 1)
    try:
        yield
    finally:
        os.dup2(saved, 1)
        os.close(devnull)
        os.close(saved)


class ClockSampler:
    """SM clock / throttle reasons sampled through NVML while the timed region runs"""

    def __init__(self, index):
        self.samples, self.reasons, self.stop_flag, self.thread = [], set(), False, None
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4)}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.002)

    def start(self):
        if self.nv:
            self.stop_flag = False
            self.thread = threading.Thread(target=self._loop, daemon=True)
            self.thread.start()

    def stop(self):
        if self.thread:
            self.stop_flag = True
            self.thread.join()
            self.thread = None

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "samples": len(s), "reasons": sorted(self.reasons)}


def pick_sources(row_offsets, count, seed, global_ids=None):
    """vertex 0 (the R-MAT hub) first, then seeded random vertices with at least one edge"""
    import numpy as np
    deg = np.diff(row_offsets)
    rng = np.random.RandomState(seed)
    out = []
    if deg.size and deg[0] > 0:
        out.append(0)
    while len(out) < count and deg.size:
        v = int(rng.randint(0, deg.size))
        if deg[v] > 0:
            out.append(v)
    return out


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(kernel_key):
    """per-launch DRAM bytes of the dominant kernel from the committed ncu capture, if any"""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            rec = json.load(f).get(kernel_key)
        return (float(rec["bytes_per_launch"]), rec.get("source")) if rec else (None, None)
    except Exception:
        return None, None


def level_bytes(rec, offset_bytes, nodes):
    """algorithmic HBM bytes of one executed level"""
    F, E, D = rec["frontier"], rec["edges"], rec["discovered"]
    if rec["kind"] == "pull":
        return rec["candidates"] * 2 * offset_bytes + 5 * E + 4 * D + 3 * (nodes // 8)
    return F * (4 + 2 * offset_bytes) + 5 * E + 8 * D


# ----------------------------------------------------------------------------------------
# reference arm: the reference's CPU BFS on the host cores
# ----------------------------------------------------------------------------------------
def synthesize(scale, path):
    """`bench.py --synthesize SCALE PATH` (a child process of the reference arm): write the seeded R-MAT CSR
    -- the same arrays oracle.rmat_csr yields, built by the device generator in seconds where a GPU exists --
    to PATH.ro.npy / PATH.co.npy and exit.  The process that times the reference never loads libgunrock.so."""
    import numpy as np
    ro = co = None
    source = "oracle generator (CPU)"
    try:
        import gunrockinst_b200 as gr
        if gr.device_count() > 0:
            g = gr.Graph.rmat(scale)
            ro, co = g.download()
            g.close()
            source = "device generator in a child process that exited before timing (same seeded CSR as oracle.rmat_csr)"
    except Exception:
        ro = None
    if ro is None:
        import oracle
        oracle.build()
        ro, co = oracle.rmat_csr(scale)
    np.save(path + ".ro.npy", ro.astype(np.int64))
    np.save(path + ".co.npy", np.ascontiguousarray(co, dtype=np.int32))
    with open(path + ".src", "w") as f:
        f.write(source)


def load_synthetic_csr(scale):
    """-> (row_offsets, col, how): the reference arm's input, produced by a child process (see synthesize)"""
    import subprocess
    import tempfile
    import numpy as np
    base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else tempfile.gettempdir()
    path = os.path.join(base, "grb_bench_rmat_s%d_%d" % (scale, os.getpid()))
    try:
        subprocess.check_call([sys.executable, os.path.abspath(__file__), "--synthesize", str(scale), path],
                              stdout=subprocess.DEVNULL)
        ro = np.load(path + ".ro.npy", mmap_mode="r")
        co = np.load(path + ".co.npy", mmap_mode="r")
        how = open(path + ".src").read()
    finally:
        for ext in (".ro.npy", ".co.npy", ".src"):      # the mappings stay valid after the unlink
            try:
                os.unlink(path + ext)
            except OSError:
                pass
    return ro, co, how


def run_reference(args, rank, world):
    if rank != 0:
        return
    import numpy as np
    import oracle
    oracle.build()
    scale = args.scale
    t0 = time.time()
    ro, co, graph_source = load_synthetic_csr(scale)
    sources = pick_sources(ro, N_SOURCES, 1234 + scale)
    sampler = oracle.BfsSampler(ro, co)
    # whole searches where one takes well under a second (scale <= 24); above that a bounded sample: each
    # step scans at most `cap` neighbour entries of the search, and the line says which fraction that is
    cap = args.ref_edge_cap if args.ref_edge_cap >= 0 else (0 if scale <= 24 else 50_000_000)
    for i in range(args.warmup):
        sampler.run(sources[i % len(sources)], cap)
    edges = 0
    secs = 0.0
    for i in range(args.steps):
        e, t = sampler.run(sources[i % len(sources)], cap)
        edges += e
        secs += t
    value = edges / secs / 1e9 if secs > 0 else 0.0
    if cap:
        sample = "first <= %d scanned neighbour entries of each of %d searches (serial FIFO BFS)" % (cap, args.steps)
        fraction = min(1.0, cap / max(int(co.size), 1))
    else:
        sample = "%d complete searches (serial FIFO BFS, tests/bfs/test_bfs.cu:258-322 restated in oracle/oracle.c)" % args.steps
        fraction = 1.0
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / max(args.steps, 1),
        "higher_is_better": True, "scaling": "strong" if args.gpus > 1 else "weak", "vs_baseline": None,
        "dtype": "int32", "data": "synthetic",
        "config": workload_config(args, scale, int(ro.size - 1), int(co.size)),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": "port", "sample": sample,
                         "sample_fraction": fraction, "step_is_a_whole_search": not cap,
                         "cores_available": os.cpu_count(), "graph_source": graph_source},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "setup_s": round(time.time() - t0 - secs, 2),
    }
    print(json.dumps(line), flush=True)


def workload_config(args, scale, nodes, edges):
    if args.gpus == 1 and not getattr(args, "partitioned", False):
        w = "direction-optimising BFS, R-MAT scale %d edgefactor 16 undirected (BASELINE configs[1])" % scale
    else:
        exch = ("NCCL all-to-all / all-gather of the frontier bitmaps" if getattr(args, "exchange", "peer") == "nccl"
                else "stores into peer memory over NVLink from the producing kernels (the NCCL all-to-all is --exchange nccl)")
        w = ("BFS, R-MAT scale %d edgefactor 16 undirected, 1D vertex-partitioned over %d device%s, per-level frontier "
             "exchange: %s (BASELINE configs[4])" % (scale, args.gpus, "" if args.gpus == 1 else "s", exch))
    return {"workload": w, "scale": scale, "edgefactor": 16, "nodes": nodes, "directed_edges": edges,
            "rmat_abcd": [0.55, 0.2, 0.2, 0.05], "seed": 42 + scale, "sources": N_SOURCES,
            "l2_policy": "inputs larger than L2 (CSR %.2f GB vs 126 MB L2); labels/bitmaps are re-initialised "
                         "every step" % ((4.0 * edges + 4.0 * nodes) / 1e9)}


# ----------------------------------------------------------------------------------------
# the other single-device BASELINE configs at full size, each checked against the oracle
# ----------------------------------------------------------------------------------------
def run_full_size_configs(args, peak):
    """-> {"s25": {"dobfs": .., "topdown": ..}, "c3_sssp_s24": .., "c4_cc_s25": .., "c4_pr_s25": ..}: device time
    (CUDA events inside the library, best of 3), throughput, algorithmic bytes over that time against the measured HBM
    peak, and `parity` against the CPU oracle on the identical CSR (test infrastructure: the checker only)."""
    import numpy as np
    import gunrockinst_b200 as gr
    out = {}
    check = not args.no_cpu

    def guarded(name, fn):
        t0 = time.time()
        try:
            rec = fn()
        except Exception as exc:            # never lose the headline over an extra record
            rec = {"error": "%s: %s" % (type(exc).__name__, str(exc)[:300])}
        rec["wall_s"] = round(time.time() - t0, 1)
        out[name] = rec

    def best(run, reps=3):
        top = None
        for _ in range(reps):
            r = run()
            if top is None or r[-1]["elapsed_ms"] < top[-1]["elapsed_ms"]:
                top = r
        return top

    def rate(st, ms):
        return {"ms": ms, "algo_gbs": st["algo_bytes"] / ms / 1e6, "roofline_frac": st["algo_bytes"] / ms / 1e6 / peak}

    # ---- R-MAT scale 25: BFS (north star), CC and PageRank (configs[3]) share one graph ----------------------------
    g25 = ro = co = None
    try:
        g25 = gr.Graph.rmat(25)
        t_tw = time.time()
        g25.build_twin()                    # preprocessing (one more CSR, hubs first): the searches and PageRank below use it
        out["s25_twin_build_s"] = round(time.time() - t_tw, 3)
        if check:
            import oracle
            oracle.build()
            ro, co = g25.download()
    except Exception as exc:
        out["s25"] = {"error": str(exc)[:300]}

    def bfs25():
        rec = {"graph": {"nodes": g25.nodes, "directed_edges": g25.edges}, "source": 0,
               "layout": "degree-ordered twin of the graph (grb_bfs_config.degree_order); labels in the caller's ids"}
        want = oracle.bfs(ro, co, 0) if check else None
        for key, mode in (("dobfs", dict(direction_optimized=True, degree_order=True)),
                          ("topdown", dict(traversal_mode=1, host_loop=1, degree_order=True))):
            b = gr.Bfs(g25, **mode)
            labels, _, _ = b.run(0)
            _, _, st = best(lambda: b.run(0, want_labels=False))
            r = rate(st, st["elapsed_ms"])
            r.update(gteps=st["edges_visited"] / st["elapsed_ms"] / 1e6, edges_inspected=st["edges_inspected"],
                     levels=[st["push_levels"], st["pull_levels"]], gpu_launches=st["gpu_launches"],
                     parity=(bool(np.array_equal(labels, want)) if check else None))
            rec[key] = r
            b.close()
        return rec

    def cc25():
        gr.cc(g25)
        labels, k, st = best(lambda: gr.cc(g25))
        r = rate(st, st["elapsed_ms"])
        r.update(components=int(k), gpu_launches=st["gpu_launches"])
        if check:
            want, ncomp = oracle.cc(ro, co)
            r["parity"] = bool(np.array_equal(labels, want) and k == ncomp)
        return r

    def pr25():
        gr.pagerank(g25, max_iter=2)
        ranks, _, st = gr.pagerank(g25, delta=0.85, threshold=1e-6, max_iter=1000)
        r = rate(st, st["elapsed_ms"])
        r.update(sweeps=st["iterations"], ms_per_sweep=st["elapsed_ms"] / max(st["iterations"], 1),
                 rank_sum=float(ranks.sum()), rank_max=float(ranks.max()),
                 layout="row-binned sweep kernel on the degree-ordered twin of the graph (grb_graph_build_twin, "
                        "s25_twin_build_s); ranks in the caller's ids")
        # the same sweeps in the caller's vertex order, and through the LB advance (the routes this one replaced)
        for key, var, val in (("ms_per_sweep_callers_order", "GRB_PR_TWIN", "0"), ("ms_per_sweep_lb_advance", "GRB_PR_SWEEP", "lb")):
            os.environ["GRB_PR_TWIN"] = "0"
            os.environ[var] = val
            _, _, st8 = gr.pagerank(g25, delta=0.85, threshold=1e-6, max_iter=8)
            r[key] = st8["reserved"] / 1000.0 / max(st8["iterations"], 1)
            os.environ.pop("GRB_PR_TWIN", None); os.environ.pop(var, None)
        r.update(parity_note="PageRank is unpinned by the reference's own tests (stale golden, SURVEY.md section 4); "
                             "checked against the fp64 restatement of its functors: per-vertex 1e-6 after 5 sweeps here, "
                             "to convergence at scale 22 in tests/test_pr_gpu.py")
        if check:
            got, _, _ = gr.pagerank(g25, delta=0.85, threshold=1e-6, max_iter=5)
            want = oracle.pagerank(ro, co, 0.85, 1e-6, 5, -1)[0]
            err = np.abs(got - want)
            r["max_abs_diff_after_5_sweeps"] = float(err.max())
            r["parity"] = bool(np.all(err <= 1e-6 * np.maximum(1.0, np.abs(want))))
        return r

    if g25 is not None:
        guarded("s25", bfs25)
        guarded("c4_cc_s25", cc25)
        guarded("c4_pr_s25", pr25)
        g25.close()
        ro = co = None
        gr.release_cached_memory()

    def sssp24():
        g = gr.Graph.rmat(24, weight_seed=7)
        gr.sssp(g, 0)
        dist, _, st = best(lambda: gr.sssp(g, 0))
        r = rate(st, st["elapsed_ms"])
        r.update(graph={"nodes": g.nodes, "directed_edges": g.edges, "weights": "symmetric uint32 1..64"},
                 iterations=st["iterations"], relaxations=st["edges_inspected"])
        if check:
            r2, c2, w2 = g.download(want_weights=True)
            want, _ = oracle.sssp(r2, c2, w2, 0)
            r["parity"] = bool(np.array_equal(dist, want))
        g.close()
        return r

    guarded("c3_sssp_s24", sssp24)
    gr.release_cached_memory()
    return out


# ----------------------------------------------------------------------------------------
# B200 arm, one device
# ----------------------------------------------------------------------------------------
def run_single(args):
    import numpy as np
    import torch
    import gunrockinst_b200 as gr
    gr._capi.lib()                      # fails loudly if libgunrock.so is missing: no CPU path
    if gr.device_count() < 1 or not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(0)
    scale = args.scale
    t_setup = time.time()
    g = gr.Graph.rmat(scale)
    n, m, osz = g.nodes, g.edges, int(g.info.offset_bytes)
    ro, co = g.download()
    sources = pick_sources(ro, N_SOURCES, 1234 + scale)
    bfs = gr.Bfs(g, direction_optimized=True)
    stream = torch.cuda.ExternalStream(g.stream())
    clocks = ClockSampler(0)

    for i in range(max(args.warmup, 3)):
        bfs.run(sources[i % len(sources)], want_labels=False)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    clocks.start()
    edges = launches = algo_bytes = 0
    # the K searches run back to back inside the library (grb_bfs_run_many): every search still
    # ends with its own stream synchronisation (its summary is read on the host), but no Python
    # bookkeeping sits between two searches
    order = [sources[i % len(sources)] for i in range(args.steps)]
    e0.record(stream)
    stats = bfs.run_many(order)
    e1.record(stream)
    torch.cuda.synchronize()
    clocks.stop()
    enact_ms = sum(st["elapsed_ms"] + st["reset_ms"] for st in stats) / max(len(stats), 1)
    for st in stats:
        edges += st["edges_visited"]
        launches += st["gpu_launches"]
        algo_bytes += st["algo_bytes"]
    ms = e0.elapsed_time(e1)
    value = edges / (ms * 1e-3) / 1e9

    # ---- per-kernel roofline: one traced pass over the same sources (events around every level)
    tr = gr.Bfs(g, direction_optimized=True, record_levels=True, host_loop=True)
    for i in range(3):
        tr.run(sources[i % len(sources)], want_labels=False)
    agg = {}
    for i in range(min(args.steps, len(sources))):
        tr.run(sources[i % len(sources)], want_labels=False)
        for rec in tr.levels():
            a = agg.setdefault(rec["kind"], {"ms": 0.0, "bytes": 0, "launches": 0})
            a["ms"] += rec["ms"]
            a["bytes"] += level_bytes(rec, osz, n)
            a["launches"] += 1
    tr.close()
    dom = max(agg, key=lambda k: agg[k]["ms"])
    peak, peak_src = measured_peaks()
    kernel_names = {"pull": "advance::pull::Kernel", "push-lb": "advance::lb::ScanKernel+ExpandKernel",
                    "push-twc": "advance::twc::Kernel"}
    achieved = agg[dom]["bytes"] / (agg[dom]["ms"] * 1e-3) / 1e9 if agg[dom]["ms"] > 0 else 0.0
    roofline = {"bound": "hbm", "kernel": kernel_names.get(dom, dom), "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
                "traffic": ncu_traffic(dom)[0], "traffic_source": ncu_traffic(dom)[1],
                "traffic_note": "ncu capture of the launches of ONE search (source 0), per launch; algo_bytes_per_launch beside "
                                "it is the mean over the launches of every timed source: compare each with its own duration, "
                                "not with each other",
                "launches_timed": agg[dom]["launches"],
                "avg_launch_us": 1e3 * agg[dom]["ms"] / max(agg[dom]["launches"], 1),
                "algo_bytes_per_launch": agg[dom]["bytes"] / max(agg[dom]["launches"], 1),
                "share_of_kernel_time": agg[dom]["ms"] / sum(a["ms"] for a in agg.values()),
                "whole_step": {"algo_bytes_per_step": algo_bytes / max(args.steps, 1),
                               "achieved": algo_bytes / (ms * 1e-3) / 1e9,
                               "frac": algo_bytes / (ms * 1e-3) / 1e9 / peak}}

    # ---- the two operators on their bandwidth-heaviest launches (BASELINE metric: "advance/filter HBM GB/s
    # vs peak"): a top-down search inspects every edge (LB advance), an idempotent search pushes every
    # neighbour through the filter.  Best of 3 searches from sources[0], events around every launch.
    def best_launch(mode, kind, nbytes):
        b = gr.Bfs(g, record_levels=True, host_loop=1, **mode)
        top = None
        for _ in range(3):
            b.run(sources[0], want_labels=False)
            for rec in b.levels():
                if rec["kind"] in kind and rec["ms"] > 0:
                    gbs = nbytes(rec) / (rec["ms"] * 1e-3) / 1e9
                    if top is None or nbytes(rec) > top["algo_bytes"] or (nbytes(rec) == top["algo_bytes"] and gbs > top["achieved"]):
                        top = {"achieved": gbs, "frac": gbs / peak, "unit": "GB/s", "us": 1e3 * rec["ms"],
                               "algo_bytes": nbytes(rec), "items_in": rec["frontier"], "edges": rec["edges"],
                               "items_out": rec["discovered"], "kernel": rec["kind"]}
        b.close()
        return top
    operators = {
        "advance_lb": best_launch(dict(traversal_mode=1, degree_order=True), ("push-lb", "push-stream"), lambda r: level_bytes(r, osz, n)),
        "filter": best_launch(dict(idempotence=True, degree_order=True), ("filter", "filter-stream"), lambda r: 5 * r["frontier"] + 4 * r["discovered"]),
        "note": "largest launch of each operator in a top-down / an idempotent search from sources[0]; algorithmic "
                "bytes as in DESIGN.md section 3 (advance: F(4+2o)+5E+8D, filter: 5 in + 4 out).  Both run on the "
                "degree-ordered twin of the graph (grb_bfs_config.degree_order); the time covers the whole operator call: degree "
                "scan, mask snapshot, StreamExpandKernel and HarvestKernel when `kernel` is push-stream / filter-stream",
    }

    # ---- end to end through the reference's entry point, CSR in pinned host memory
    ro32 = torch.from_numpy(ro.astype(np.int32)).pin_memory()
    co32 = torch.from_numpy(co).pin_memory()
    e2e_steps = max(1, min(args.steps, args.e2e_steps))
    deg = np.diff(ro)
    e2e_edges, e2e_s = 0, 0.0
    with quiet_stdout():
        for i in range(2):
            gr.gunrock_bfs_func(ro32.numpy(), co32.numpy(), src_node=sources[0])
        for i in range(e2e_steps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            last = gr.gunrock_bfs_func(ro32.numpy(), co32.numpy(), src_node=sources[i % len(sources)])
            e2e_s += time.perf_counter() - t0       # the call returns after its own cudaDeviceSynchronize
            e2e_edges += int(deg[last >= 0].sum())   # host bookkeeping of the result, outside the timed call
    e2e_s = max(e2e_s, 1e-9)
    e2e = {"value": e2e_edges / e2e_s / 1e9, "unit": UNIT, "h2d_bytes_per_step": int(4 * (n + 1) + 4 * m),
           "d2h_bytes_per_step": int(4 * n), "ms_per_step": 1e3 * e2e_s / e2e_steps, "steps": e2e_steps,
           "api": "gunrock_bfs_func (gunrock.h:106-110): upload CSR from pinned host memory, top-down BFS, "
                  "download labels"}

    # ---- the same searches through the native entry points with the CSR left where it is: offsets
    # uploaded, column ids read in place over PCIe (grb_graph_from_host_mapped, the reference's
    # stream_from_host) by a direction-optimising search -- what a caller who knows the graph is
    # undirected can do; gunrock_bfs_func cannot (gunrock.h carries no such flag) and stays the `e2e` figure
    try:
        s_edges, s_s, lab = 0, 0.0, None
        for i in range(-2, e2e_steps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            gm = gr.Graph.from_csr(ro32.numpy(), co32.numpy(), symmetric=True, stream_from_host=True)
            bm = gr.Bfs(gm, direction_optimized=True)
            lab, _, _ = bm.run(sources[max(i, 0) % len(sources)])
            dt = time.perf_counter() - t0
            bm.close()
            gm.close()
            if i >= 0:
                s_s += dt
                s_edges += int(deg[lab >= 0].sum())
        e2e_streamed = {"value": s_edges / max(s_s, 1e-9) / 1e9, "unit": UNIT, "ms_per_step": 1e3 * s_s / e2e_steps,
                        "steps": e2e_steps, "h2d_bytes_per_step": int(4 * (n + 1)), "d2h_bytes_per_step": int(4 * n),
                        "labels_equal_e2e": bool(np.array_equal(lab, last)),
                        "api": "grb_graph_from_host_mapped + grb_bfs_create(direction_optimized) + grb_bfs_run per "
                               "step: offsets uploaded, column ids read in place from pinned host memory over PCIe "
                               "(not counted in h2d_bytes_per_step), labels downloaded"}
    except Exception as exc:          # an extra figure: never lose the line over it
        e2e_streamed = {"error": str(exc)[:200]}

    # ---- CPU baseline (the checker, timed on one host core) + parity of the timed path
    cpu = {"value": None, "unit": UNIT, "cores": 1, "kind": "port", "sample": "skipped"}
    parity = None
    if not args.no_cpu:
        import oracle
        oracle.build()
        sampler = oracle.BfsSampler(ro, co)
        ce, cs, k = 0, 0.0, 0
        while cs < args.cpu_seconds and k < len(sources):
            e, t = sampler.run(sources[k])
            if k == 0:
                labels, _, _ = bfs.run(sources[0])
                parity = bool(np.array_equal(labels, sampler.labels[:n]))
            ce += e
            cs += t
            k += 1
        cpu = {"value": ce / cs / 1e9, "unit": UNIT, "cores": 1, "cores_available": os.cpu_count(), "kind": "port",
               "sample": "%d complete searches of the same graph (sources[0:%d]), %.1f s of serial FIFO BFS "
                         "(tests/bfs/test_bfs.cu:258-322 restated in oracle/oracle.c)" % (k, k, cs)}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms / max(args.steps, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int32", "data": "synthetic", "config": workload_config(args, scale, n, m),
        "roofline": roofline, "operators": operators, "cpu_baseline": cpu, "e2e": e2e, "e2e_streamed": e2e_streamed, "gpu_launches": launches,
        "clocks": clocks.summary(), "parity_vs_oracle": parity,
        "graph500_gteps": value / 2.0, "device_ms_per_search": enact_ms, "setup_s": round(time.time() - t_setup, 1),
    }
    bfs.close()
    g.close()
    if not args.no_configs:
        gr.release_cached_memory()
        t_cfg = time.time()
        line["configs"] = run_full_size_configs(args, peak)
        line["configs"]["wall_s"] = round(time.time() - t_cfg, 1)
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=256)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scale", type=int, default=0, help="R-MAT scale (default: 22 on one device, 27 partitioned)")
    ap.add_argument("--e2e-steps", type=int, default=8)
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--ref-edge-cap", type=int, default=-1,
                    help="reference arm: scan at most this many neighbour entries per step (0: whole searches; "
                         "default: whole searches up to scale 24, 50 M entries above)")
    ap.add_argument("--synthesize", nargs=2, metavar=("SCALE", "PATH"), help=argparse.SUPPRESS)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-configs", action="store_true",
                    help="skip the full-size records of the other single-device BASELINE configs (scale-25 BFS / CC / "
                         "PageRank, scale-24 SSSP, each with its oracle check: several minutes)")
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="partitioned runs: peer-memory stores over NVLink (default) or NCCL collectives")
    ap.add_argument("--mg-persistent", action="store_true",
                    help="partitioned runs: the whole search as one cooperative kernel per device (mgbfs_persistent.cuh) "
                         "instead of the host-driven level loop")
    ap.add_argument("--no-e2e-from-host", action="store_true",
                    help="partitioned runs: skip the end-to-end figure that uploads every rank's slice from host memory")
    ap.add_argument("--no-single-ref", action="store_true",
                    help="partitioned runs: skip the single-device search of the same graph on rank 0 (reference point + parity)")
    ap.add_argument("--partitioned", action="store_true",
                    help="use the partitioned driver even on one device (the 1-device point of a scale-27 scaling table)")
    args = ap.parse_args()
    if args.synthesize:
        synthesize(int(args.synthesize[0]), args.synthesize[1])
        return
    if args.mg_persistent:
        os.environ["GRB_MG_PERSISTENT"] = "1"
    world = env_int("WORLD_SIZE", 1)
    rank = env_int("RANK", 0)
    if world > 1:
        args.gpus = world
    # launched by torchrun (the scaling run: N = 1, 2, 4, 8 back to back): every point is the SAME workload, the scale-27
    # partitioned search, so that the 1-device point is the one the others are compared with.  A plain `python bench.py`
    # stays BASELINE configs[1] (scale 22 on one device).
    if "RANK" in os.environ and "WORLD_SIZE" in os.environ and "LOCAL_RANK" in os.environ:
        args.partitioned = True             # (the reference arm times the same graph: its config has to match)
    if args.scale <= 0:
        args.scale = 22 if (args.gpus == 1 and not args.partitioned) else 27
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if args.gpus == 1 and not args.partitioned:
        run_single(args)
    else:
        from gunrockinst_b200 import partitioned
        cpu_check = None
        if not args.no_cpu and rank == 0:
            import oracle                   # the checker only: depths of the single-device search up to scale 24
            oracle.build()
            cpu_check = oracle.bfs
        partitioned.bench(args, rank, world, workload_config, ClockSampler, pick_sources, measured_peaks, cpu_check)


if __name__ == "__main__":
    main()
