"""Summarise `ncu --page source --print-source sass --csv` output: per launch, instruction totals by opcode,
the hottest instructions by executed count and by stall samples.  usage: ncu_sass_top.py file.csv [launch_index] [top]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
launches, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1][:120], "hdr": None, "ins": []}
        launches.append(cur)
    elif r and r[0] == "Address":
        cur["hdr"] = r
    elif cur and cur["hdr"] and len(r) == len(cur["hdr"]):
        cur["ins"].append(dict(zip(cur["hdr"], r)))
which = int(sys.argv[2]) if len(sys.argv) > 2 else -1
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
L = launches[which]
def f(x):
    try: return float(x)
    except Exception: return 0.0
tot = sum(f(i["Instructions Executed"]) for i in L["ins"])
samp = sum(f(i["# Samples"]) for i in L["ins"])
print(L["name"]); print("instructions", len(L["ins"]), "warp-instr executed", tot, "samples", samp)
by = collections.Counter()
for i in L["ins"]:
    by[i["Source"].split()[0] if not i["Source"].startswith("@") else i["Source"].split()[1]] += f(i["Instructions Executed"])
print("by opcode:", [(k, round(v / tot * 100, 1)) for k, v in by.most_common(25)])
print("-- top by executed")
for n, i in sorted(enumerate(L["ins"]), key=lambda t: -f(t[1]["Instructions Executed"]))[:top]:
    print("%5d %5.2f%% smp %5.2f%% thr %4.1f  %s" % (n, f(i["Instructions Executed"]) / tot * 100, f(i["# Samples"]) / max(samp, 1) * 100, f(i["Avg. Threads Executed"]), i["Source"][:100]))
print("-- top by samples")
for n, i in sorted(enumerate(L["ins"]), key=lambda t: -f(t[1]["# Samples"]))[:top]:
    st = {k: f(v) for k, v in i.items() if k.startswith("stall_") and "Not Issued" not in k and f(v) > 0}
    st = sorted(st.items(), key=lambda t: -t[1])[:3]
    print("%5d smp %5.2f%% exe %5.2f%%  %-70s %s" % (n, f(i["# Samples"]) / max(samp, 1) * 100, f(i["Instructions Executed"]) / tot * 100, i["Source"][:70], st))
