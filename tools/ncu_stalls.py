"""Per-source-line warp-stall breakdown of an ncu report (source page): python tools/ncu_stalls.py report.ncu-rep [top_n]
Lines sorted by stall samples; for each the share of every stall reason above 8 %."""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows, fname, hdr = [], None, None
tot = {}
for r in csv.reader(io.StringIO(txt)):
    if not r: continue
    if r[0] in ("File Name", "File Path"): fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or not r[0].isdigit(): continue
    d = dict(zip(hdr[4:], r[4:]))
    try: smp = int(d["# Samples"]); inst = int(d["Instructions Executed"])
    except Exception: continue
    st = {}
    for k, v in d.items():
        if k.startswith("stall_") and "Not Issued" not in k:
            try: st[k[6:]] = int(v)
            except Exception: pass
    for k, v in st.items(): tot[k] = tot.get(k, 0) + v
    rows.append((smp, inst, fname, int(r[0]), r[1].strip()[:90], st))
S = sum(r[0] for r in rows); I = sum(r[1] for r in rows)
print("all lines: " + "  ".join(f"{k} {100*v/max(S,1):.1f}%" for k, v in sorted(tot.items(), key=lambda kv: -kv[1]) if v * 100 > S))
rows.sort(key=lambda r: -r[0])
for smp, inst, f, ln, src, st in rows[:top]:
    why = "  ".join(f"{k} {100*v/max(smp,1):.0f}%" for k, v in sorted(st.items(), key=lambda kv: -kv[1]) if v * 12 > smp)
    print(f"{100*smp/S:5.1f}% smp {100*inst/I:5.1f}% inst  {f}:{ln}  [{why}]  {src}")
