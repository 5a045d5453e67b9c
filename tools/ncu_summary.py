"""Summarise an ncu report: key raw metrics + top stall lines (run here, no GPU needed)."""
import csv, subprocess, sys, collections
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units, val = rows[0], rows[1], rows[2]
keep = ['Kernel Name', 'gpu__time_duration.sum', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'l1tex__m_xbar2l1tex_read_bytes.sum',
        'l1tex__m_xbar2l1tex_read_bytes.sum.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed',
        'launch__registers_per_thread', 'sm__cycles_elapsed.avg.per_second', 'sm__cycles_elapsed.avg',
        'launch__cluster_size', 'launch__occupancy_cluster_max_active', 'sm__ctas_launched.sum', 'launch__grid_size',
        'sm__cycles_active.avg', 'l1tex__data_bank_reads.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_writes.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed']
for h, u, v in zip(hdr, units, val):
    if h in keep or 'cluster' in h.lower():
        print(f"{h:90s} {v:>24s} {u}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hdr = rows[1]
data = [r for r in rows[2:] if len(r) == len(hdr)]
S, SRC, EX = hdr.index('Warp Stall Sampling (All Samples)'), hdr.index('Source'), hdr.index('Instructions Executed')
cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
tot = sum(float(r[S] or 0) for r in data)
print("\nstall reasons (all warps):")
for i in cols:
    t = sum(float(r[i] or 0) for r in data)
    if t / tot > 0.01:
        print(f"  {hdr[i]:24s} {t/tot*100:5.1f}%")
print("\ntop sampled SASS lines:")
for k in sorted(sorted(range(len(data)), key=lambda k: -float(data[k][S] or 0))[:int(sys.argv[2]) if len(sys.argv) > 2 else 25]):
    r = data[k]
    rs = sorted([(float(r[i] or 0), hdr[i]) for i in cols], reverse=True)[:1]
    print(f"{float(r[S])/tot*100:5.1f}% line{k:5d} ex={r[EX]:>10s} {r[SRC][:72]:72s} {rs[0][1]}")
