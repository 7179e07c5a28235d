"""One table of the ptxas -v output of every kernel (registers, spill bytes, static shared memory): python tools/ptxas_summary.py build/ptxas"""
import glob, os, re, subprocess, sys
d = sys.argv[1] if len(sys.argv) > 1 else "build/ptxas"
rows = []
for f in sorted(glob.glob(os.path.join(d, "*.log"))):
    txt = open(f).read().split("ptxas info    : Compiling entry function ")
    for blk in txt[1:]:
        name = blk.split("'")[1]
        regs = re.search(r"Used (\d+) registers", blk)
        smem = re.search(r"(\d+) bytes smem", blk)
        sp = re.search(r"(\d+) bytes spill stores, (\d+) bytes spill loads", blk)
        try:
            dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip().split("(")[0]
        except Exception:
            dem = name
        rows.append((os.path.basename(f)[:-4], dem.replace("void ", "").replace("ccg::", ""), int(regs.group(1)) if regs else -1,
                     int(sp.group(1)) if sp else 0, int(sp.group(2)) if sp else 0, int(smem.group(1)) if smem else 0))
print(f"{'file':12s} {'kernel':70s} {'regs':>5s} {'spill st':>9s} {'spill ld':>9s} {'smem B':>7s}")
for r in rows:
    print(f"{r[0]:12s} {r[1][:70]:70s} {r[2]:5d} {r[3]:9d} {r[4]:9d} {r[5]:7d}")
