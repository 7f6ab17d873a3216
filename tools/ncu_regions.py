"""Warp instructions per 32 instances by region of count_kernel, from an ncu source-page export
(`ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv`).
usage: python tools/ncu_regions.py src.csv <occurrence of count_kernel> <instances of that launch> [markers.json]
Regions are found from marker comments in kc_count.cuh (so they follow the code as it moves)."""
import csv, sys, os, re
rows = list(csv.reader(open(sys.argv[1])))
occ, inst = int(sys.argv[2]), float(sys.argv[3])
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
sel = [l for l in launches if "count_kernel" in l[0]["fn"]][occ]
src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "mhm2_proxy_b200", "csrc", "kc_count.cuh")).read().split("\n")
MARK = [("count_kernel(const CountArgs", "item setup / scheduling"), ("auto update = ", "counters (update)"),
        ("auto find_or_claim = ", "slow path (find_or_claim)"), ("auto drain = ", "miss queue: drain"),
        ("for (u32 sg0 = 0", "round setup (segment cache, ranges)"), ("auto fetch = ", "a. stage records"),
        ("---- a. one supermer record", "a. stage records"), ("---- b. scan of the k-mer counts", "b. scan + instance list"),
        ("---- c. every k-mer instance", "c. loop control"), ("if (t < total) {", "c. fast path: list entry, lookup, key compare"),
        ("const u32 mm = __ballot_sync", "c. miss queue: push"), ("const bool stop = __any_sync", "chunk end"),
        ("what is still waiting", "miss queue: drain"), ("const bool over = s_over", "item end: status, filter, compaction, output")]
region, cur_r = {}, "helpers"
for i, line in enumerate(src, 1):
    for pat, name in MARK:
        if pat in line:
            cur_r = name
    region[i] = cur_r
agg, tot, thr = {}, 0.0, {}
for s in sel:
    h = s["hdr"]
    ie, te = h.index("Instructions Executed"), h.index("Thread Instructions Executed")
    for r in s["lines"]:
        try:
            n, t, ln = float(r[ie]), float(r[te]), int(r[0])
        except (ValueError, IndexError):
            continue
        f = s["file"]
        if f == "kc_count.cuh":
            p = region.get(ln, "?")
        elif f == "kc_extract.cuh":
            p = "c. extract (k-mer, revcomp, canonical, extensions)"
        elif f == "kc_device.cuh":
            p = "hash" if 50 <= ln <= 72 else ("c. extract (k-mer, revcomp, canonical, extensions)" if 130 <= ln <= 142 or ln < 50 or 86 <= ln <= 130 else "item end: status, filter, compaction, output")
        else:
            p = "library intrinsics (" + f + ")"
        agg[p] = agg.get(p, 0) + n
        thr[p] = thr.get(p, 0) + t
        tot += n
it = inst / 32
print(f"{sel[0]['fn'][:70]}: {tot/1e9:.2f} G warp instructions, {tot/it:.1f} per 32 instances")
for p, n in sorted(agg.items(), key=lambda x: -x[1]):
    print(f"{100*n/tot:6.2f}%  {n/it:7.1f} per 32 inst  {thr[p]/max(n,1):5.1f} lanes  {p}")
