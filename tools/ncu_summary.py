#!/usr/bin/env python
"""Condense ncu output into the small text summaries kept under profiles/.

    python tools/ncu_summary.py launches gpurun_out/<tag>_launches.csv  profiles/<name>.csv
    python tools/ncu_summary.py full     gpurun_out/<tag>.ncu-rep       profiles/<name>.csv [layer names...]

`launches`: the `--metrics gpu__time_duration.sum` launch list -> id, kernel, grid, block, microseconds,
plus the share of the itmo kernels' total per kernel family at the bottom.
`full`: the `--set full` report read through `ncu -i ... --page raw --csv` -> one row per launch with the
metrics DESIGN.md and bench.py's roofline.traffic quote (duration, DRAM bytes, tensor-pipe and shared-memory
activity, L2/DRAM throughput, registers, shared memory).
"""
from __future__ import annotations

import csv
import io
import re
import subprocess
import sys

FULL_METRICS = [
    ('gpu__time_duration.sum', 'time_us'),
    ('dram__bytes_read.sum', 'dram_read_MB'),
    ('dram__bytes_write.sum', 'dram_write_MB'),
    ('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed', 'tensor_pipe_pct'),
    ('sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active', 'hmma_pct_active'),
    ('sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm_throughput_pct'),
    ('l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed', 'smem_tc_wavefront_pct'),
    ('l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex_pct'),
    ('lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l2_pct'),
    ('dram__throughput.avg.pct_of_peak_sustained_elapsed', 'dram_pct'),
    ('launch__registers_per_thread', 'regs'),
    ('launch__shared_mem_per_block_dynamic', 'smem_dyn_KB'),
    ('launch__grid_size', 'grid'),
    ('launch__block_size', 'block'),
]


def short_name(k: str) -> str:
    m = re.search(r'itmo::(\w+)(<[^>]*>)?', k)
    return m.group(0) if m else 'torch:' + re.sub(r'\s+', ' ', k)[:48]


def launches(src: str, dst: str) -> None:
    txt = open(src).read()
    rows = list(csv.DictReader(io.StringIO(txt[txt.index('"ID","Process ID"'):])))
    fam: dict[str, list[float]] = {}
    with open(dst, 'w', newline='') as f:
        w = csv.writer(f)
        w.writerow(['id', 'kernel', 'grid', 'block', 'time_us'])
        for r in rows:
            us = float(r['Metric Value'].replace(',', '')) / 1e3
            name = short_name(r['Kernel Name'])
            w.writerow([r['ID'], name, r['Grid Size'], r['Block Size'], f'{us:.2f}'])
            if name.startswith('itmo::'):
                fam.setdefault(name.split('<')[0], []).append(us)
        tot = sum(sum(v) for v in fam.values()) or 1.0
        w.writerow([])
        w.writerow(['# itmo kernel family', 'launches', 'total_us', 'share_of_itmo_time'])
        for k, v in sorted(fam.items(), key=lambda kv: -sum(kv[1])):
            w.writerow(['# ' + k, len(v), f'{sum(v):.1f}', f'{sum(v) / tot:.4f}'])


def full(src: str, dst: str, names: list[str]) -> None:
    # src: the .ncu-rep, or the raw page already exported on the GPU box (`ncu -i rep --page raw --csv > x.csv`)
    if src.endswith('.csv'):
        raw = open(src).read()
    else:
        raw = subprocess.run(['ncu', '-i', src, '--page', 'raw', '--csv'], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    kcol = col['Kernel Name']
    with open(dst, 'w', newline='') as f:
        w = csv.writer(f)
        w.writerow(['launch', 'layer', 'kernel'] + [f'{alias}' for _, alias in FULL_METRICS])
        w.writerow(['', '', ''] + [units[col[m]] if m in col else '' for m, _ in FULL_METRICS])
        for i, r in enumerate(body):
            vals = [r[col[m]] if m in col else '' for m, _ in FULL_METRICS]
            w.writerow([r[col['ID']], names[i] if i < len(names) else '', short_name(r[kcol])] + vals)


if __name__ == '__main__':
    if len(sys.argv) < 4:
        sys.exit(__doc__)
    if sys.argv[1] == 'launches':
        launches(sys.argv[2], sys.argv[3])
    else:
        full(sys.argv[2], sys.argv[3], sys.argv[4:])
