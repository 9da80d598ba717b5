"""Per-source-line summary of an ncu report's source page (cuda,sass view):
python tools/ncu_lines.py report.ncu-rep [top_n]  -> lines sorted by warp instructions executed."""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows, fname, hdr = [], None, None
for r in csv.reader(io.StringIO(txt)):
    if not r: continue
    if r[0] == "File Name": fname = r[1].split("/")[-1]; continue
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
