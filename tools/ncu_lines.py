"""Per-source-line summary of an ncu report's source page (cuda,sass view):
python tools/ncu_lines.py report.ncu-rep [top_n]  -> lines sorted by warp instructions executed."""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows, fname, hdr = [], None, None
for r in csv.reader(io.StringIO(txt)):
    if not r: continue
    if r[0] in ("File Name", "File Path"): fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or r[0] == "" or not r[0].isdigit(): continue
    d = dict(zip(hdr[4:], r[4:]))
    try:
        inst = int(d["Instructions Executed"]); thr = int(d["Thread Instructions Executed"]); smp = int(d["# Samples"])
    except Exception: continue
    rows.append((inst, thr, smp, fname, int(r[0]), r[1].strip()[:110]))
tot_i = sum(r[0] for r in rows); tot_t = sum(r[1] for r in rows); tot_s = sum(r[2] for r in rows)
print(f"total warp-inst {tot_i:.4g} thread-inst {tot_t:.4g} avg threads {tot_t/max(tot_i,1):.2f} samples {tot_s}")
rows.sort(reverse=True)
for inst, thr, smp, f, ln, src in rows[:top]:
    print(f"{100*inst/tot_i:5.1f}% inst {100*smp/max(tot_s,1):5.1f}% smp  thr/inst {thr/max(inst,1):5.1f}  {f}:{ln}  {src}")
# optional: group by named line ranges  CATS="name:file:lo-hi,..."
import os
cats = os.environ.get("CATS")
if cats:
    agg = {}
    for spec in cats.split(","):
        name, f, rng = spec.split(":"); lo, hi = map(int, rng.split("-"))
        for inst, thr, smp, ff, ln, src in rows:
            if ff == f and lo <= ln <= hi:
                a = agg.setdefault(name, [0, 0, 0]); a[0] += inst; a[1] += thr; a[2] += smp
    print("--- categories")
    for name, (i, t, sm) in agg.items():
        print(f"{name:24s} {100*i/tot_i:5.1f}% inst  {100*sm/max(tot_s,1):5.1f}% smp  thr/inst {t/max(i,1):5.1f}")
