"""Per-source-line totals from `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > src.csv`.
usage: python tools/ncu_lines.py src.csv <function substring> <occurrence index> [top N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
pat, occ, top = sys.argv[2], int(sys.argv[3]), int(sys.argv[4]) if len(sys.argv) > 4 else 40
# the export repeats (file, function) sections per profiled launch, in launch order
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
# group sections into launches: a new launch starts when the function name changes or the file list restarts
launches, seen = [], None
for s in sections:
    if not launches or s["fn"] != launches[-1][0]["fn"] or s["file"] in seen:
        launches.append([])
        seen = set()
    launches[-1].append(s)
    seen.add(s["file"])
sel = [l for l in launches if pat in l[0]["fn"]][occ]
out = []
for s in sel:
    h = s["hdr"]
    ie, te, si = h.index("Instructions Executed"), h.index("Thread Instructions Executed"), h.index("# Samples")
    for r in s["lines"]:
        try:
            out.append((int(r[ie]), int(r[te]), int(r[si]), s["file"], int(r[0]), r[1].strip()[:110]))
        except (ValueError, IndexError):
            pass
tot = sum(o[0] for o in out); tots = sum(o[2] for o in out)
print(f"{sel[0]['fn'][:60]}  warp-instr {tot/1e6:.1f} M, samples {tots}")
for i, t, s_, f, ln, src in sorted(out, reverse=True)[:top]:
    print(f"{100*i/tot:5.2f}% instr {100*s_/max(1,tots):5.2f}% smp thr {t/max(1,i):4.1f} {f}:{ln}: {src}")
