"""Per-CUDA-source-line instruction / stall-sample totals of an ncu report captured with --import-source on."""
import csv, io, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Line No')
h = rows[hi]
ie = h.index('Instructions Executed'); te = h.index('Thread Instructions Executed'); ss = h.index('Warp Stall Sampling (All Samples)')
lines = []
for r in rows[hi + 1:]:
    if r and r[0].isdigit():
        try:
            lines.append((int(r[0]), r[1].strip(), int(r[ie]), int(r[te]), int(r[ss])))
        except Exception:
            pass
tot = sum(l[2] for l in lines); tots = max(1, sum(l[4] for l in lines))
print(f"total warp-instr {tot:.4g}")
for ln, text, i, t, s in sorted(lines, key=lambda l: -l[2])[:top]:
    print(f"{ln:5d} {100*i/tot:5.1f}% inst {100*s/tots:5.1f}% smp thr {t/max(1,i):4.1f} | {text[:110]}")
