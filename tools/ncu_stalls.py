"""Stall-sample shares per source line of one profiled launch (ncu source page CSV, --print-source cuda,sass).
usage: python tools/ncu_stalls.py src.csv <function substring> <occurrence> [top N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
pat, occ, top = sys.argv[2], int(sys.argv[3]), int(sys.argv[4]) if len(sys.argv) > 4 else 30
sections, cur = [], None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur = {"file": r[1].split("/")[-1], "fn": None, "lines": []}
        sections.append(cur)
    elif r[0] == "Function Name":
        cur["fn"] = r[1]
    elif r[0] == "Line No":
        cur["hdr"] = r
    elif cur is not None and r[0].isdigit():
        cur["lines"].append(r)
launches, seen = [], None
for s in sections:
    if not launches or s["fn"] != launches[-1][0]["fn"] or s["file"] in seen:
        launches.append([]); seen = set()
    launches[-1].append(s); seen.add(s["file"])
sel = [l for l in launches if pat in l[0]["fn"]][occ]
NAMES = ("stall_barrier", "stall_short_sb", "stall_wait", "stall_not_selected", "stall_math", "stall_mio", "stall_long_sb", "stall_lg",
         "stall_branch_resolving", "stall_no_inst", "stall_dispatch", "stall_selected", "stall_drain", "stall_membar", "stall_misc", "stall_tex", "stall_sleep")
out = []
for s in sel:
    h = s["hdr"]
    ie, si, te = h.index("Instructions Executed"), h.index("# Samples"), h.index("Thread Instructions Executed")
    ex, wf = h.index("L1 Wavefronts Shared Excessive"), h.index("L1 Wavefronts Shared")
    stl = {n: h.index(n) for n in NAMES if n in h}
    for r in s["lines"]:
        try:
            d = dict(i=int(r[ie]), s=int(r[si]), t=int(r[te]), f=s["file"], ln=int(r[0]), src=r[1].strip()[:90], ex=int(r[ex] or 0), wf=int(r[wf] or 0))
            d["st"] = {n: int(r[j] or 0) for n, j in stl.items()}
            out.append(d)
        except (ValueError, IndexError):
            pass
ti, ts = sum(o["i"] for o in out), sum(o["s"] for o in out)
print(f"{sel[0]['fn'][:70]}: {ti/1e9:.3f} G warp instructions, {ts} samples")
agg = {}
for o in out:
    for n, v in o["st"].items():
        agg[n] = agg.get(n, 0) + v
print("stall shares of all samples (%):", {k[6:]: round(100 * v / ts, 1) for k, v in sorted(agg.items(), key=lambda x: -x[1]) if v})
for o in sorted(out, key=lambda o: -o["s"])[:top]:
    t3 = sorted(o["st"].items(), key=lambda x: -x[1])[:3]
    print("%5.2f%% smp %5.2f%% ins thr %4.1f %s:%d %s | %s" % (100 * o["s"] / ts, 100 * o["i"] / ti, o["t"] / max(1, o["i"]), o["f"], o["ln"], o["src"][:64],
                                                          " ".join("%s %.1f" % (k[6:], 100 * v / ts) for k, v in t3)))
tex, twf = sum(o["ex"] for o in out), sum(o["wf"] for o in out)
print(f"shared-memory wavefronts {twf/1e9:.2f} G, of which excessive (bank conflicts) {tex/1e9:.2f} G")
for o in sorted(out, key=lambda o: -o["ex"])[:8]:
    if o["ex"]:
        print("  %5.1f%% of the excess: %s:%d %s" % (100 * o["ex"] / max(1, tex), o["f"], o["ln"], o["src"][:80]))
