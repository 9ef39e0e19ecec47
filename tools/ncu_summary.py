#!/usr/bin/env python
"""Summarise one kernel of an .ncu-rep (made with --set full --import-source on) as text for profiles/.

usage: python tools/ncu_summary.py <report.ncu-rep> [rows of 32 canvas pixels per launch] > profiles/<name>.txt
Reads the report with `ncu -i ... --page raw/source --csv` (no GPU needed)."""
import collections
import csv
import io
import subprocess
import sys

WANT = [
    'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread',
    'launch__grid_size', 'launch__block_size', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
    'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
    'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__cycles_active.avg', 'sm__cycles_elapsed.avg',
    'sm__cycles_active.min', 'sm__cycles_active.max',
    'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
    'l1tex__data_pipe_lsu_wavefronts.sum', 'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed',
    'smsp__warps_eligible.avg.per_cycle_active', 'smsp__warps_active.avg.per_cycle_active',
    'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio',
    'smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio',
    'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active', 'lts__t_bytes.sum', 'l1tex__m_xbar2l1tex_read_bytes.sum',
]


def page(rep, which):
    out = subprocess.run(['ncu', '-i', rep, '--page', which, '--csv'], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    rows32 = float(sys.argv[2]) if len(sys.argv) > 2 else 7680 * 2160 / 32
    raw = page(rep, 'raw')
    hdr, units, data = raw[0], raw[1], raw[2:]
    for w in WANT:
        if w in hdr:
            i = hdr.index(w)
            print(f"{w:86s} {units[i]:16s} {[r[i] for r in data]}")
    src = page(rep, 'source')
    h = None
    for k, r in enumerate(src):
        if 'Source' in r and 'Instructions Executed' in r:
            h = k
            break
    if h is None:
        return
    hdr = src[h]
    isrc, iex, ismp = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('# Samples')
    stall_cols = [(i, c) for i, c in enumerate(hdr) if c.startswith('stall_')]
    ops, smp, tot = collections.Counter(), collections.Counter(), 0
    stalls = collections.Counter()
    buckets = collections.Counter()
    for r in src[h + 1:]:
        if len(r) <= iex or not r[iex].isdigit():
            continue
        n = int(r[iex])
        tot += n
        t = r[isrc].split()
        op = t[1] if t[0].startswith('@') else t[0]
        op = op.split('.')[0]
        ops[op] += n
        smp[op] += int(r[ismp] or 0)
        buckets[n // 32] += 1
        for i, c in stall_cols:
            if r[i].isdigit():
                stalls[c] += int(r[i])
    print(f"total warp instructions {tot} = {tot / rows32:.2f} per 32 canvas pixels")
    print("  ".join(f"{op}:{n / rows32:.2f}" for op, n in ops.most_common(30)))
    ts = sum(stalls.values()) or 1
    print("stall samples: " + ", ".join(f"{c[6:]}:{100 * n / ts:.1f}%" for c, n in stalls.most_common(12)))


if __name__ == '__main__':
    main()
