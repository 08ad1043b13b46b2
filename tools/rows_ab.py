"""A/B of the row-binned passes (oprtr/row_bins.cuh) against the LB-advance routes they replace,
same process, same graph:  PageRank sweeps (GRB_PR_SWEEP=lb | rows1 = static grid | default = tickets) and the hook pass of connected
components (GRB_CC_HOOK=lb | rows1 = static grid | default = tickets).

usage: python tools/rows_ab.py [scale=25] [pr_sweeps=24] [runs=2] [cc]

One JSON line per primitive: time per route, results compared between the routes (CC: labels equal;
PageRank: same sweep count, max |diff| of the ranks), plus -- up to scale 20 -- the CPU oracle on the
identical CSR (test infrastructure, used as the checker only).  PageRank has a third route: the row-binned
sweeps on the degree-ordered twin of the graph (grb_graph_build_twin).
"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gunrockinst_b200 as gr


def route(var, value):
    if value is None:
        os.environ.pop(var, None)
    else:
        os.environ[var] = value


def main():
    scale = int(sys.argv[1]) if len(sys.argv) > 1 else 25
    sweeps = int(sys.argv[2]) if len(sys.argv) > 2 else 24
    runs = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    only = sys.argv[4] if len(sys.argv) > 4 else ""         # "cc": skip PageRank
    g = gr.Graph.rmat(scale)
    n, m = g.local_nodes, g.info.edges
    want_cc = want_pr = None
    if scale <= 20:
        import oracle
        ro, co = g.download()
        want_cc = oracle.cc(ro, co)[0]
        want_pr, want_it = oracle.pagerank(ro, co, 0.85, 1e-6, sweeps, -1)

    # ---- connected components
    rec = {"primitive": "cc", "scale": scale, "nodes": int(n), "entries": int(m)}
    labels = {}
    for name, val in (("lb", "lb"), ("rows1", "rows1"), ("rows", None)):
        route("GRB_CC_HOOK", val)
        best = None
        for _ in range(runs):
            lab, ncomp, st = gr.cc(g)
            if best is None or st["elapsed_ms"] < best:
                best = st["elapsed_ms"]
        labels[name] = lab
        rec[name] = {"ms": round(best, 4), "components": ncomp, "gpu_launches": st["gpu_launches"],
                     "entries_per_s": round(m / best / 1e6, 2)}
    route("GRB_CC_HOOK", None)
    rec["labels_equal"] = bool(np.array_equal(labels["lb"], labels["rows"]) and np.array_equal(labels["lb"], labels["rows1"]))
    if want_cc is not None:
        rec["rows_equal_oracle"] = bool(np.array_equal(labels["rows"], want_cc))
    rec["speedup"] = round(rec["lb"]["ms"] / rec["rows"]["ms"], 3)
    print(json.dumps(rec), flush=True)
    if only == "cc":
        g.close()
        return

    # ---- PageRank
    rec = {"primitive": "pagerank", "scale": scale, "nodes": int(n), "entries": int(m), "max_iter": sweeps}
    ranks = {}
    import time
    for name, val in (("lb", "lb"), ("rows1", "rows1"), ("rows", None), ("rows_twin", None)):
        route("GRB_PR_SWEEP", val)
        if name == "rows_twin":             # the same kernel on the degree-ordered twin of the graph
            t0 = time.time()
            g.build_twin()
            rec["twin_build_s"] = round(time.time() - t0, 3)
        best = None
        for _ in range(runs):
            r, _, st = gr.pagerank(g, max_iter=sweeps)
            per = st["reserved"] / 1000.0 / max(1, st["iterations"])
            if best is None or per < best:
                best = per
        ranks[name] = r
        rec[name] = {"ms_per_sweep": round(best, 4), "sweeps": st["iterations"], "elapsed_ms": round(st["elapsed_ms"], 3),
                     "entries_per_sweep": int(st["edges_inspected"] // max(1, st["iterations"])),
                     "gpu_launches": st["gpu_launches"], "rank_sum": float(r.sum())}
    route("GRB_PR_SWEEP", None)
    for name in ("rows1", "rows", "rows_twin"):
        d = np.abs(ranks["lb"] - ranks[name])
        rec[name]["max_abs_diff_vs_lb"] = float(d.max())
        rec[name]["within_1e-6_per_vertex"] = bool(np.all(d <= 1e-6 * np.maximum(1.0, np.abs(ranks["lb"]))))
        if want_pr is not None:
            rec[name]["vs_oracle_max_abs_diff"] = float(np.abs(ranks[name] - want_pr).max())
            rec[name]["sweeps_equal_oracle"] = bool(rec[name]["sweeps"] == want_it)
        rec[name]["speedup_vs_lb"] = round(rec["lb"]["ms_per_sweep"] / rec[name]["ms_per_sweep"], 3)
    print(json.dumps(rec), flush=True)
    g.close()


if __name__ == "__main__":
    main()
