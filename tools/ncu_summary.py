"""Summarises ncu outputs into the small text files kept under profiles/:  launch list -> per-kernel share table,
.ncu-rep -> selected raw metrics per kernel.   python tools/ncu_summary.py launches <csv> <out> ; ... rep <ncu-rep> <out>"""
import collections
import csv
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum', 'sm__cycles_elapsed.max',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_pipe_tc_wavefronts_mem_shared.sum',
        'l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_red.sum',
        'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum',
        'SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts.avg', 'SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts_mem_lgds.avg',
        'SM_A.TriageCompute.l1tex__data_pipe_lsu_wavefronts_mem_shared.avg', 'smsp__sass_inst_executed_op_utcmma.sum',
        'smsp__sass_inst_executed_op_tmem_ldt.sum',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_membar_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio']


def launches(src, dst, title):
    lines = [l for l in open(src) if not l.startswith('==')]
    agg, tot = collections.OrderedDict(), 0.0
    for row in csv.DictReader(lines):
        a = agg.setdefault(row['Kernel Name'], [0, 0.0])
        a[0] += 1
        a[1] += float(row['Metric Value'])
        tot += float(row['Metric Value'])
    out = ['# ' + title, '# gpu__time_duration.sum per launch under ncu (cold cache, serialised): compare SHARES', 'kernel,launches,total_ms,share']
    for k, (n, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append('"%s",%d,%.3f,%.4f' % (k[:100], n, v / 1e6, v / tot))
    open(dst, 'w').write('\n'.join(out) + '\n')
    print('\n'.join(out[:10]))


def rep(src, dst, title):
    raw = subprocess.run(['ncu', '-i', src, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    out = ['# ' + title]
    for r in rows[2:]:
        out.append('')
        out.append('kernel: ' + r[idx['Kernel Name']])
        for k in KEYS:
            if k in idx:
                out.append('  %-92s %s %s' % (k, r[idx[k]], units[idx[k]]))
    open(dst, 'w').write('\n'.join(out) + '\n')
    print('\n'.join(out))


if __name__ == '__main__':
    {'launches': launches, 'rep': rep}[sys.argv[1]](sys.argv[2], sys.argv[3], sys.argv[4] if len(sys.argv) > 4 else '')
