"""Summarise an ncu raw/source CSV pair: python tools/ncu_summary.py <raw.csv> [<src.csv>]"""
import collections
import csv
import sys

KEYS = ['gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'sm__inst_executed.avg.per_cycle_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'lts__t_sectors_srcunit_tex_op_read.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum',
        'l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum', 'l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum', 'smsp__inst_executed.sum',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warp_latency_per_inst_issued.ratio', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed']


def raw(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    ni = hdr.index('Kernel Name')
    print('kernels:', [r[ni][:50] for r in rows[2:]])
    for k in KEYS:
        if k in hdr:
            i = hdr.index(k)
            print(f'{k} [{units[i]}]', [r[i] for r in rows[2:]])
    for i, h in enumerate(hdr):
        if 'warp_issue_stalled' in h and h.endswith('_per_warp_active.pct'):
            vals = [float(r[i]) for r in rows[2:]]
            if max(vals) > 3:
                print('  stall', h.replace('smsp__warp_issue_stalled_', '').replace('_per_warp_active.pct', ''), [f'{v:.1f}' for v in vals])


def src(path, top=14):
    rows = list(csv.reader(open(path)))
    blocks, cur = [], None
    for r in rows:
        if r and r[0] == 'Kernel Name':
            cur = {'name': r[1], 'rows': []}; blocks.append(cur); continue
        if cur is not None:
            cur['rows'].append(r)
    seen = set()
    for b in blocks:
        hdr, data = b['rows'][0], b['rows'][1:]
        si, ie, ws = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('Warp Stall Sampling (All Samples)')
        tot = sum(int(r[ie] or 0) for r in data if len(r) > ie)
        if (b['name'], tot) in seen:
            continue
        seen.add((b['name'], tot))
        print(b['name'][:60], 'sass lines', len(data), 'warp instr', tot)
        ops, samp = collections.Counter(), collections.Counter()
        for r in data:
            if len(r) <= ie or not r[si]:
                continue
            t = r[si].split()
            op = t[1] if t[0].startswith('@') and len(t) > 1 else t[0]
            op = op.split('.')[0]
            ops[op] += int(r[ie] or 0); samp[op] += int(r[ws] or 0)
        ts = max(1, sum(samp.values()))
        print('   ' + '  '.join(f'{op} {100*c/tot:.1f}%/{100*samp[op]/ts:.1f}%' for op, c in ops.most_common(top)))


if __name__ == '__main__':
    raw(sys.argv[1])
    if len(sys.argv) > 2:
        src(sys.argv[2])
