"""Summarise an ncu report: key raw metrics + SASS hot segments.  usage: ncu_summary.py raw.csv src.csv"""
import csv, sys
raw, src = sys.argv[1], sys.argv[2]
rows = list(csv.reader(open(raw)))
hdr, units, vals = rows[0], rows[1], rows[2]
want = ['gpu__time_duration.sum', 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum', 'l1tex__t_bytes.sum',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'smsp__warps_eligible.avg.per_cycle_active', 'sm__cycles_elapsed.max',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__pcsamp_warps_issue_stalled_long_scoreboard', 'smsp__pcsamp_warps_issue_stalled_wait', 'smsp__pcsamp_warps_issue_stalled_math_pipe_throttle',
        'smsp__pcsamp_warps_issue_stalled_not_selected', 'smsp__pcsamp_warps_issue_stalled_branch_resolving', 'smsp__pcsamp_warps_issue_stalled_short_scoreboard',
        'smsp__pcsamp_warps_issue_stalled_lg_throttle', 'smsp__pcsamp_warps_issue_stalled_no_instructions', 'smsp__pcsamp_warps_issue_stalled_selected',
        'smsp__pcsamp_warps_issue_stalled_barrier', 'smsp__pcsamp_warps_issue_stalled_mio_throttle', 'smsp__pcsamp_warps_issue_stalled_dispatch_stall',
        'local_load', 'smsp__inst_executed_op_local']
for i, h in enumerate(hdr):
    if h in want or any(h == w for w in want):
        print('%-80s %s %s' % (h, vals[i], units[i]))
rows = list(csv.reader(open(src)))
hdr = rows[1]
ia, isrc, ie, ns, it = (hdr.index(x) for x in ('Address', 'Source', 'Instructions Executed', '# Samples', 'Avg. Threads Executed'))
data = [(r[ia], r[isrc], int(r[ie] or 0), int(r[ns] or 0), r[it]) for r in rows[2:]
        if len(r) > ie and (r[ie] or '0').isdigit() and (r[ns] or '0').isdigit()]
for j in range(1, len(data)):          # several instances of the kernel in one report: keep the first
    if data[j][0] == data[0][0]:
        data = data[:j]
        break
tot = sum(d[2] for d in data); tots = sum(d[3] for d in data)
print('total inst', tot, 'samples', tots, 'n sass', len(data))
seglen = int(sys.argv[3]) if len(sys.argv) > 3 else 80
for s in range(0, len(data), seglen):
    seg = data[s:s + seglen]
    e = sum(d[2] for d in seg); sm = sum(d[3] for d in seg)
    if e / max(tot,1) < 0.003 and sm / max(tots,1) < 0.003: continue
    ops = {}
    for d in seg:
        t = d[1].split()
        op = t[1] if t[0].startswith('@') else t[0]
        ops[op] = ops.get(op, 0) + d[2]
    top = sorted(ops.items(), key=lambda x: -x[1])[:6]
    print('%4d-%4d inst %5.1f%% samp %5.1f%% thr %s  %s' % (s, s + len(seg), 100 * e / tot, 100 * sm / tots, seg[len(seg) // 2][4],
          ' '.join('%s:%.1f' % (k, 100 * v / tot) for k, v in top)))
