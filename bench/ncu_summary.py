"""Summarise ncu artefacts into small text files under profiles/ (the .ncu-rep files stay in gpurun_out/).

    python bench/ncu_summary.py launches gpurun_out/launches.csv profiles/step_launches.md
    python bench/ncu_summary.py report   gpurun_out/prof_top.ncu-rep profiles/top_kernels_ncu.md
"""
import collections
import csv
import re
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'launch__shared_mem_per_block_dynamic', 'smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_membar_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio']


def short(name: str) -> str:
    return re.sub(r'\(.*', '', name).replace('void ', '').replace('llmt::', '')[:100]


def launches(src: str, dst: str) -> None:
    lines = [l for l in open(src) if not l.startswith('==')]
    agg = collections.defaultdict(lambda: [0, 0.0])
    total = 0.0
    for row in csv.DictReader(lines):
        if row.get('Metric Name') != 'gpu__time_duration.sum':
            continue
        v = float(row['Metric Value'].replace(',', ''))
        v = v / 1e3 if row['Metric Unit'] in ('ns', 'nsecond') else (v * 1e3 if row['Metric Unit'] in ('ms', 'msecond') else v)
        k = re.sub(r'<.*', '', short(row['Kernel Name']))
        agg[k][0] += 1
        agg[k][1] += v
        total += v
    with open(dst, 'w') as f:
        f.write(f'# Kernel launch list of one training step (ncu gpu__time_duration, serialised: compare shares)\n\n')
        f.write(f'total {total / 1e3:.2f} ms over {sum(c for c, _ in agg.values())} launches\n\n')
        f.write('| share | time (us) | launches | kernel |\n|---|---|---|---|\n')
        for k, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f'| {100 * v / total:.1f}% | {v:.1f} | {c} | `{k}` |\n')
    print(open(dst).read())


def report(src: str, dst: str) -> None:
    raw = subprocess.run(['ncu', '-i', src, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    seen = collections.OrderedDict()
    for r in rows[2:]:
        seen.setdefault(short(r[idx['Kernel Name']]), r)      # first instance of every distinct kernel
    with open(dst, 'w') as f:
        f.write(f'# ncu --set full summary ({src}); one instance per distinct kernel\n')
        for name, r in seen.items():
            f.write(f'\n## `{name}`\n\n')
            for k in KEYS:
                if k in idx:
                    f.write(f'- {k}: {r[idx[k]]} {units[idx[k]]}\n')
    print(open(dst).read())


if __name__ == '__main__':
    {'launches': launches, 'report': report}[sys.argv[1]](sys.argv[2], sys.argv[3])
