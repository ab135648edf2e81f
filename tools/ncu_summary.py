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
    src = page(rep, 'source', ['--print-source', 'cuda'])
    src = [r for r in src if len(r) > 6]
    if len(src) > 2:
        h = src[0]
        ci, cs, cx = h.index('Source'), h.index('# Samples'), h.index('Instructions Executed')
        cl = h.index('Address') if 'Address' in h else 0

        def num(x):
            try:
                return float(x.replace(',', ''))
            except Exception:
                return 0.0
        body = src[1:]
        tot = sum(num(r[cs]) for r in body) or 1.0
        toti = sum(num(r[cx]) for r in body) or 1.0
        print('== source lines by stall samples (total %d samples, %d warp-instr)' % (tot, toti))
        order = sorted(range(len(body)), key=lambda k: -num(body[k][cs]))
        for k in order[:nlines]:
            r = body[k]
            print('  %5.1f%% smp %5.1f%% inst  L%-4s %s' % (100 * num(r[cs]) / tot, 100 * num(r[cx]) / toti, r[cl], r[ci].strip()[:100]))
        print('== source lines by instructions executed')
        order = sorted(range(len(body)), key=lambda k: -num(body[k][cx]))
        for k in order[:nlines]:
            r = body[k]
            print('  %5.1f%% inst %5.1f%% smp  L%-4s %s' % (100 * num(r[cx]) / toti, 100 * num(r[cs]) / tot, r[cl], r[ci].strip()[:100]))


if __name__ == '__main__':
    main()
