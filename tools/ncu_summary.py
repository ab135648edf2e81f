#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU): key metrics, stall reasons, hottest source lines.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [--lines 25]"""
import csv
import io
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_warps',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed.sum',
        'smsp__inst_executed.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed',
        'smsp__inst_executed.min', 'smsp__inst_executed.max', 'smsp__inst_executed.avg',
        'smsp__issue_active.min', 'smsp__issue_active.max', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__thread_inst_executed_per_inst_executed.ratio',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum',
        'lts__t_sectors_srcunit_tex_op_read.sum', 'lts__t_sectors_srcunit_tex_op_write.sum',
        'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct']


def page(rep, name, extra=()):
    out = subprocess.run(['ncu', '-i', rep, '--page', name, '--csv'] + list(extra), capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    nlines = int(sys.argv[sys.argv.index('--lines') + 1]) if '--lines' in sys.argv else 25
    rows = page(rep, 'raw')
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print('==', r[hdr.index('Kernel Name')])
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print('  %-62s %s %s' % (k, r[i], units[i]))
        st = [(hdr[i], r[i]) for i in range(len(hdr))
              if 'smsp__average_warp' in hdr[i] and 'issue_stalled' in hdr[i] and 'not_issued' not in hdr[i]]
        st.sort(key=lambda t: -float(t[1].replace(',', '') or 0))
        for name, v in st[:8]:
            print('  stall %-50s %s' % (name.split('issue_stalled_')[1].split('_per_')[0], v))
    src = page(rep, 'source', ['--print-source', 'cuda,sass'])
    h = None
    cur = line = None
    agg = {}
    for r in src:
        if len(r) >= 2 and r[0] == 'File Name':
            cur = r[1].split('/')[-1]
            continue
        if len(r) > 6 and r[0] == 'Line No':
            h = r
            continue
        if h is None or len(r) <= 6:
            continue
        if r[0] != '':
            line = (cur, r[0], r[1].strip())
            agg.setdefault(line, [0.0, 0.0])
            continue
        if r[2] in ('...', ''):
            continue
        cs, cx = h.index('# Samples'), h.index('Instructions Executed')

        def num(x):
            try:
                return float(x.replace(',', ''))
            except Exception:
                return 0.0
        agg[line][0] += num(r[cs])
        agg[line][1] += num(r[cx])
    if agg:
        tot = sum(v[0] for v in agg.values()) or 1.0
        toti = sum(v[1] for v in agg.values()) or 1.0
        print('== CUDA lines by warp instructions executed (total %d warp-instr, %d stall samples)' % (toti, tot))
        for k in sorted(agg, key=lambda k: -agg[k][1])[:nlines]:
            v = agg[k]
            print('  %5.1f%% inst %5.1f%% smp  %s:%s %s' % (100 * v[1] / toti, 100 * v[0] / tot, k[0], k[1], k[2][:90]))
        print('== CUDA lines by stall samples')
        for k in sorted(agg, key=lambda k: -agg[k][0])[:nlines]:
            v = agg[k]
            print('  %5.1f%% smp %5.1f%% inst  %s:%s %s' % (100 * v[0] / tot, 100 * v[1] / toti, k[0], k[1], k[2][:90]))


if __name__ == '__main__':
    main()
