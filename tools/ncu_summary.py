#!/usr/bin/env python3
"""Turns ncu outputs brought back in gpurun_out/ into the small text summaries committed under profiles/.
  python tools/ncu_summary.py launches <launches.csv>            per-kernel share of the step
  python tools/ncu_summary.py raw <report.ncu-rep>               key metrics per profiled launch
  python tools/ncu_summary.py sass <report.ncu-rep> <kernel-regex>   opcode mix of the first matching launch
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct', 'smsp__inst_executed.sum',
        'smsp__sass_thread_inst_executed_op_ffma_pred_on.sum', 'smsp__sass_thread_inst_executed_op_fadd_pred_on.sum',
        'smsp__sass_thread_inst_executed_op_fmul_pred_on.sum']


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = None
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows:
        if r[0] == 'ID':
            hdr = r
            continue
        if hdr is None:
            continue
        d = dict(zip(hdr, r))
        v = float(d['Metric Value'].replace(',', ''))
        v *= {'ns': 1.0, 'us': 1e3, 'ms': 1e6, 's': 1e9}.get(d['Metric Unit'], 1.0)
        k = d['Kernel Name'].split('(')[0]
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"# per-kernel device time from `ncu --metrics gpu__time_duration.sum --clock-control none` ({path}); total {tot / 1e6:.3f} ms")
    print(f"{'kernel':50s} {'launches':>8s} {'ms':>10s} {'share':>7s}")
    for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{k:50s} {v[0]:8d} {v[1] / 1e6:10.3f} {100 * v[1] / tot:6.1f}%")


def raw(rep):
    txt = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units = rows[0], rows[1]
    print(f"# key metrics per profiled launch (ncu --set full --clock-control none), from {rep}")
    for r in rows[2:]:
        print('----', r[hdr.index('Kernel Name')][:90])
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f"  {k:72s} {r[i][:24]:>24s} {units[i]}")


def sass(rep, pat):
    txt = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass', '--kernel-name', 'regex:' + pat,
                          '--launch-count', '1'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr = rows[1]
    iE, iS, iT = hdr.index('Instructions Executed'), hdr.index('Source'), hdr.index('Thread Instructions Executed')
    ops = collections.Counter(); tot = 0; thr = 0
    for r in rows[2:]:
        if r[0] == 'Kernel Name':
            break
        t = r[iS].split()
        op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
        ops[op] += int(r[iE]); tot += int(r[iE]); thr += int(r[iT])
    print(f"# SASS opcode mix (warp-level instructions executed) of {rows[0][1][:80]}; total {tot}; avg active threads {thr / max(tot, 1):.2f}")
    for k, v in ops.most_common(24):
        print(f"  {k:10s} {v:13d} {100 * v / tot:5.1f}%")


if __name__ == '__main__':
    {'launches': launches, 'raw': raw, 'sass': sass}[sys.argv[1]](*sys.argv[2:])
