#!/usr/bin/env python
"""Turn one ``ncu --set full --import-source on`` report into the markdown summary kept under profiles/.

    python scripts/summarize_ncu.py gpurun_out/ros4.ncu-rep "title line" [steps_total] > profiles/<name>.md

Runs ``ncu -i <rep> --page raw --csv`` and ``--page source --csv --print-source sass`` (no GPU needed) and prints
the launch metrics that the roofline discussion in DESIGN.md uses plus the SASS opcode mix.  ``steps_total`` (sum of
the per-system step counters of the launch) adds "warp instructions per warp-step".
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = [
    'gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
    'sass__inst_executed_local_loads', 'sass__inst_executed_local_stores', 'sass__inst_executed_shared_loads',
    'sm__cycles_elapsed.avg', 'sm__cycles_elapsed.avg.per_second',
    'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
    'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
    'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__inst_executed.sum',
    'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__sass_average_branch_targets_threads_uniform.pct',
    'smsp__sass_thread_inst_executed_op_dfma_pred_on.sum.per_cycle_elapsed',
    'smsp__sass_thread_inst_executed_op_dmul_pred_on.sum.per_cycle_elapsed',
    'smsp__sass_thread_inst_executed_op_dadd_pred_on.sum.per_cycle_elapsed',
    'dram__bytes_read.sum', 'dram__bytes_write.sum',
] + ['smsp__average_warps_issue_stalled_%s_per_issue_active.ratio' % k for k in (
    'wait', 'not_selected', 'math_pipe_throttle', 'dispatch_stall', 'short_scoreboard', 'long_scoreboard',
    'branch_resolving', 'no_instruction')]


def ncu_csv(rep, *args):
    out = subprocess.run(['ncu', '-i', rep] + list(args) + ['--csv'], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep, title = sys.argv[1], sys.argv[2]
    steps = float(sys.argv[3]) if len(sys.argv) > 3 else None
    raw = ncu_csv(rep, '--page', 'raw')
    hdr, units, vals = raw[0], raw[1], raw[2]
    m = dict(zip(hdr, zip(units, vals)))
    print('# ' + title + '\n')
    print('Kernel: `%s`\n' % m.get('Kernel Name', ('', '?'))[1])
    print('| metric | unit | value |\n|---|---|---|')
    for k in KEYS:
        if k in m:
            print('| %s | %s | %s |' % (k, m[k][0], m[k][1]))
    src = ncu_csv(rep, '--page', 'source', '--print-source', 'sass')
    hdr = src[1]
    data = src[2:]
    isrc, iex, ith, ismp = (hdr.index(c) for c in ('Source', 'Instructions Executed', 'Thread Instructions Executed',
                                                    '# Samples'))
    tot = sum(int(r[iex]) for r in data)
    tth = sum(int(r[ith]) for r in data)
    ops, oth, smp = collections.Counter(), collections.Counter(), collections.Counter()
    for r in data:
        t = r[isrc].split()
        op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
        ops[op] += int(r[iex])
        oth[op] += int(r[ith])
        smp[op] += int(r[ismp])
    ts = max(1, sum(smp.values()))
    mx = max(int(r[iex]) for r in data)
    hot = [r for r in data if int(r[iex]) >= 0.5*mx]
    hi, hth = sum(int(r[iex]) for r in hot), sum(int(r[ith]) for r in hot)
    print('\n## SASS (source page)\n')
    print('* warp instructions executed: %d, mean active lanes %.2f / 32' % (tot, tth/tot))
    print('* step loop (instructions executed >= half of the maximum count): %d SASS instructions, %.1f %% of the '
          'issued warp instructions, %.2f lanes active; everything else (refill, retire, rare paths): %.1f %% at '
          '%.2f lanes' % (len(hot), 100*hi/tot, hth/hi, 100*(tot - hi)/tot, (tth - hth)/max(1, tot - hi)))
    if steps:
        wsteps = steps/(hth/hi)
        print('* %.3g attempted steps in the launch -> %.0f warp instructions per warp-step' % (steps, tot/wsteps))
    print('\n| opcode | % of instructions | mean active lanes | % of stall samples |\n|---|---|---|---|')
    for op, c in ops.most_common(20):
        print('| %s | %.2f | %.1f | %.2f |' % (op, 100*c/tot, oth[op]/c, 100*smp[op]/ts))


if __name__ == '__main__':
    main()
