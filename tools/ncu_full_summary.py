"""Summarise an `ncu --set full` report of the production kernels (tools/gpu_profile.sh culled) as a markdown table.
usage: ncu -i gpurun_out/r02_culled.ncu-rep --page raw --csv > /tmp/culled_raw.csv; python tools/ncu_full_summary.py /tmp/culled_raw.csv"""
import csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
g = lambda r, n: r[hdr.index(n)] if n in hdr else ''
to_ms = lambda v, u: float(v) / 1e6 if u in ('ns', 'nsecond') else float(v) / 1e3 if u in ('us', 'usecond') else float(v)
to_mb = lambda v, u: float(v) * {'byte': 1e-6, 'Kbyte': 1e-3, 'Mbyte': 1.0, 'Gbyte': 1e3}[u]
tu, ru, wu = (units[hdr.index(n)] for n in ('gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum'))
st = [h for h in hdr if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio')]
print("| kernel | ms | FMA pipe % | issue active % | registers | warps active % | threads / instr | DRAM MB | L1 hit % | L2 hit % | top stalls (cycles per issued instruction) |\n|---|---|---|---|---|---|---|---|---|---|---|")
for r in rows[2:]:
    k = g(r, "Kernel Name"); m = re.search(r"(k_\w+(<[^>]*>)?)", k); k = m.group(1).replace('(bool)', '') if m else k[:30]
    vals = sorted([(float(r[hdr.index(h)] or 0), h.replace('smsp__average_warps_issue_stalled_', '').replace('_per_issue_active.ratio', '')) for h in st], reverse=True)
    vals = [(v, n) for v, n in vals if n != 'selected'][:3]
    print(f"| `{k}` | {to_ms(g(r,'gpu__time_duration.sum'),tu):.3f} | {float(g(r,'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active')):.0f} | {float(g(r,'smsp__issue_active.avg.pct_of_peak_sustained_active')):.0f} | "
          f"{g(r,'launch__registers_per_thread')} | {float(g(r,'sm__warps_active.avg.pct_of_peak_sustained_active')):.0f} | {float(g(r,'smsp__thread_inst_executed_per_inst_executed.ratio')):.1f} | "
          f"{to_mb(g(r,'dram__bytes_read.sum'),ru)+to_mb(g(r,'dram__bytes_write.sum'),wu):.0f} | {float(g(r,'l1tex__t_sector_hit_rate.pct') or 0):.0f} | {float(g(r,'lts__t_sector_hit_rate.pct') or 0):.0f} | " + ", ".join(f"{n} {v:.1f}" for v, n in vals) + " |")
