#!/bin/bash
# usage: profiles/ncu_report.sh <tag>   (reads gpurun_out/prof_<tag>.ncu-rep, writes profiles/ncu_<tag>_{summary,lines}.txt)
TAG=$1
cd "$(dirname "$0")/.."
ncu -i gpurun_out/prof_$TAG.ncu-rep --page raw --csv > gpurun_out/raw_$TAG.csv 2>/dev/null
ncu -i gpurun_out/prof_$TAG.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/src_$TAG.csv 2>/dev/null
python - "$TAG" <<'PY'
import csv, sys
tag = sys.argv[1]
rows = list(csv.reader(open('gpurun_out/raw_%s.csv' % tag)))
hdr, units, vals = rows[0], rows[1], rows[2]
keep = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__block_size', 'launch__grid_size',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed', 'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed',
        'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_bytes.sum', 'lts__t_sectors_srcunit_tex_op_read.sum']
with open('profiles/ncu_%s_summary.txt' % tag, 'w') as f:
    f.write('# ncu --set full --clock-control none, kvq_scan_kernel, bench.py config 2 (3M x 150 bp reads, 0.98 GB), 1 launch\n')
    for h, u, v in zip(hdr, units, vals):
        if h in keep or ('issue_stalled' in h and 'per_issue_active' in h):
            f.write('%s %s %s\n' % (h, u, v))
print(open('profiles/ncu_%s_summary.txt' % tag).read())
PY
python profiles/ncu_lines.py gpurun_out/src_$TAG.csv 60 > profiles/ncu_${TAG}_lines.txt
python - "$TAG" <<'PY'
import csv, sys, re
tag = sys.argv[1]
rows = [r for r in csv.reader(open('gpurun_out/launches_%s.csv' % tag)) if len(r) > 10 and r[0].isdigit()]
tot = {}
with open('profiles/ncu_%s_launches.txt' % tag, 'w') as f:
    f.write('# ncu --metrics gpu__time_duration.sum --clock-control none; python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1\n')
    for r in rows:
        name = re.sub(r'<unnamed>::|void ', '', r[4].split('(')[0]); v = float(r[-1].replace(',', ''))
        f.write('%s %s grid=%s block=%s %s %s\n' % (r[0], name, r[8], r[7], r[-1], r[-2]))
        tot.setdefault(name, [0, 0.0]); tot[name][0] += 1; tot[name][1] += v
    for k, (n, v) in tot.items():
        f.write('# total %s: %d launches, %.0f ns, share %.1f%%\n' % (k, n, v, 100 * v / sum(x[1] for x in tot.values())))
PY
tail -3 profiles/ncu_${TAG}_launches.txt
