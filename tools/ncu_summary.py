"""Summarise an ncu report (run here, no GPU): headline counters + SASS opcode mix of the captured kernel."""
import csv, re, collections, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, u = rows[0], rows[1]
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'sm__inst_executed.avg.per_cycle_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'launch__grid_size', 'launch__waves_per_multiprocessor',
        'lts__t_bytes.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum']
for v in rows[2:]:
    print("kernel:", v[h.index('Kernel Name')][:90])
    for i, n in enumerate(h):
        if n in keys:
            print(f"  {n} [{u[i]}] = {v[i]}")
    stalls = [(float(v[i]), n) for i, n in enumerate(h) if 'smsp__average_warps_issue_stalled' in n and n.endswith('_per_issue_active.ratio') and v[i]]
    for val, n in sorted(stalls, reverse=True)[:6]:
        print(f"  stall {n.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')}: {val:.2f}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
h = rows[hi]; ix = {n: i for i, n in enumerate(h)}
ie, te, ss = ix['Instructions Executed'], ix['Thread Instructions Executed'], ix['Warp Stall Sampling (All Samples)']
data = []
for r in rows[hi + 1:]:
    try:
        data.append((r[1], int(r[ie]), int(r[te]), int(r[ss])))
    except Exception:
        pass
tot = sum(d[1] for d in data); tots = max(1, sum(d[3] for d in data))
ops = collections.Counter(); opss = collections.Counter(); opt = collections.Counter()
for s, i, t, sm in data:
    m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)', s)
    op = m.group(2).split('.')[0] if m else s[:10]
    ops[op] += i; opss[op] += sm; opt[op] += t
print(f"total warp-instr {tot:.4g}, avg active threads {sum(d[2] for d in data)/max(1,tot):.1f}")
for op, c in ops.most_common(14):
    print(f"  {op:8s} {100*c/tot:5.1f}% inst  {100*opss[op]/tots:5.1f}% samples  thr/inst {opt[op]/max(1,c):.1f}")
