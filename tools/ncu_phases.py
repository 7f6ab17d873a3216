"""Warp-instruction / stall-sample share per phase of count_kernel from an ncu source-page export
(`ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv`; `--print-source cuda > text.csv` gives the
embedded source text, whose comment lines carry the phase markers). usage: ncu_phases.py src.csv text.csv <occurrence>"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
occ = int(sys.argv[3]) if len(sys.argv) > 3 else 0
text = {int(r[0]): r[1] for r in csv.reader(open(sys.argv[2])) if len(r) >= 2 and r[0].isdigit()}
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
MARK = [("---- a.", "a. stage records (global -> shared)"), ("---- b.", "b. scan + instance list"), ("---- c.", "c. extract k-mer, revcomp, canonical"),
        ("murmur64_words", "c. hash (MurmurHash3)"), ("---- find or claim", "d. find or claim (index probe, key compare, CAS)"),
        ("claimed ||", "d. counters (shared atomics)"), ("group_barrier_or", "chunk end (barrier, bookkeeping)"), ("---- e.", "e. filter + compaction + output")]
g = {}
def add(name, i, t, s):
    d = g.setdefault(name, [0, 0, 0]); d[0] += i; d[1] += t; d[2] += s
for s in sel:
    h = s["hdr"]
    ie, te, si = h.index("Instructions Executed"), h.index("Thread Instructions Executed"), h.index("# Samples")
    if s["file"] == "kc_count.cuh":  # phase of every source line, from the markers in the embedded text
        phase_of, phase, inside = {}, "helpers (barriers, smem loads)", False
        for ln in sorted(text):
            src = text[ln]
            if "count_kernel(const CountArgs" in src:
                phase, inside = "item setup (clear table, prefetch)", True
            if inside and "build_items_kernel" in src:
                inside, phase = False, "other"
            if inside:
                for m, name in MARK:
                    if m in src:
                        phase = name
            phase_of[ln] = phase
    for r in s["lines"]:
        if s["file"] == "kc_count.cuh":
            name = phase_of.get(int(r[0]), "other")
        elif s["file"] == "kc_device.cuh":
            ln = int(r[0])
            name = "c. hash (MurmurHash3)" if ln <= 50 else "c. extract k-mer, revcomp, canonical" if ln <= 145 else "e. filter + compaction + output"
        elif "atomic" in s["file"]:
            name = "d. counters (shared atomics)"
        else:
            name = "other (" + s["file"] + ")"
        try:
            add(name, int(r[ie]), int(r[te]), int(r[si]))
        except (ValueError, IndexError):
            pass
ti = sum(v[0] for v in g.values()); ts = sum(v[2] for v in g.values())
print(f"{sel[0]['fn'][:70]}: {ti/1e9:.2f} G warp-instructions")
for k, v in sorted(g.items(), key=lambda kv: -kv[1][0]):
    print(f"| {k} | {100*v[0]/ti:.1f} % | {100*v[2]/max(1,ts):.1f} % | {v[1]/max(1,v[0]):.1f} |")
