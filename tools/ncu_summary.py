"""Summarise an .ncu-rep: key metrics per kernel + SASS opcode mix and top stall lines of one kernel.
usage: ncu_summary.py REPORT [kernel-regex]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
rx = sys.argv[2] if len(sys.argv) > 2 else None
WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct', 'sm__cycles_elapsed.avg']
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    print('-----', r[hdr.index('Kernel Name')][:90])
    for w in WANT:
        if w in hdr:
            print("  %-72s %s %s" % (w, r[hdr.index(w)], units[hdr.index(w)]))
    stalls = [(float(r[i]), h) for i, h in enumerate(hdr) if 'warp_issue_stalled' in h and h.endswith('per_warp_active.pct') and r[i]]
    for v, h in sorted(stalls, reverse=True)[:6]:
        print("  stall %-60s %.1f" % (h.replace('smsp__warp_issue_stalled_', '').replace('_per_warp_active.pct', ''), v))
if rx:
    src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + rx], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    h = next(i for i, r in enumerate(rows) if 'Source' in r)
    hdr = rows[h]
    si, ie, ss = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('# Samples')
    ops, samp, tot, lines = collections.Counter(), collections.Counter(), 0, []
    for r in rows[h + 1:]:
        if len(r) <= ie or not r[ie].isdigit():
            continue
        toks = r[si].split()
        op = toks[1] if toks[0].startswith('@') else toks[0]
        op = op.split('.')[0]
        ops[op] += int(r[ie]); samp[op] += int(r[ss]); tot += int(r[ie])
        lines.append((int(r[ss]), int(r[ie]), r[si]))
    print('total warp-instructions', tot)
    for k, v in ops.most_common(22):
        print("  %-10s %12d %5.1f%%  samples %d" % (k, v, 100 * v / tot, samp[k]))
    print('top stall lines:')
    for s_, e_, t_ in sorted(lines, reverse=True)[:14]:
        print("  %6d %10d  %s" % (s_, e_, t_[:100]))
