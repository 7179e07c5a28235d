"""Per-source-line instruction and stall-sample totals of an ncu report (run here, no GPU):  python tools/ncu_lines.py rep [top]"""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
h = rows[hi]
ie = h.index("Instructions Executed"); ss = h.index("Warp Stall Sampling (All Samples)"); te = h.index("Thread Instructions Executed")
lines = []
for r in rows[hi + 1:]:
    if len(r) > ie and r[0]:
        try:
            lines.append((int(r[ie]), int(r[ss]), int(r[te]), int(r[0]), r[1].strip()))
        except ValueError:
            pass
tot = sum(l[0] for l in lines); tots = max(1, sum(l[1] for l in lines))
print(f"total warp-instr {tot:.4g}")
for inst, smp, thr, no, src in sorted(lines, reverse=True)[:top]:
    print(f"{100*inst/tot:5.1f}% inst {100*smp/tots:5.1f}% smp  thr {thr/max(1,inst):4.1f}  L{no:<5d} {src[:110]}")
