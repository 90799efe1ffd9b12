#!/usr/bin/env python
"""Summaries of ncu output kept under profiles/ (run in the build container; no GPU needed).

    python profiles/summarize_ncu.py launches gpurun_out/x_launches.csv  > profiles/x_launches.txt
    python profiles/summarize_ncu.py full gpurun_out/x_prof.ncu-rep      > profiles/x_prof.txt
"""
import collections
import csv
import re
import subprocess
import sys

FULL_METRICS = [
    'gpu__time_duration.sum', 'launch__grid_size', 'launch__block_size', 'launch__registers_per_thread',
    'launch__shared_mem_per_block_dynamic', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
    'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
    'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
    'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__cycles_elapsed.max',
    'l1tex__t_bytes.sum', 'lts__t_bytes.sum', 'smsp__inst_executed.sum',
]


def launches(path):
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    r = csv.reader(lines)
    hdr = next(r)
    ki, vi, ui = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
    agg = collections.OrderedDict()
    n = 0
    for row in r:
        v = float(row[vi].replace(',', ''))
        v = v / 1000 if row[ui] == 'ns' else (v * 1000 if row[ui] == 'ms' else v)
        short = re.sub(r'<.*', '', re.sub(r'\(.*', '', row[ki])).replace('void ', '')
        a = agg.setdefault(short, [0, 0.0])
        a[0] += 1
        a[1] += v
        n += 1
    tot = sum(v[1] for v in agg.values())
    print('# ncu --metrics gpu__time_duration.sum --clock-control none  (cold-cache, serialised: compare SHARES)')
    print('# %s: %d launches, %.1f us summed kernel time' % (path, n, tot))
    print('%-44s %7s %12s %8s' % ('kernel', 'count', 'us/launch', 'share'))
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print('%-44s %7d %12.2f %7.1f%%' % (k[-44:], v[0], v[1] / v[0], 100 * v[1] / tot))


def full(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    print('# ncu --set full --clock-control none --import-source on : %s' % path)
    ki = hdr.index('Kernel Name')
    for j, row in enumerate(data):
        print('launch %d: %s' % (j, row[ki][:110]))
        for m in FULL_METRICS:
            if m in hdr:
                i = hdr.index(m)
                print('    %-66s %16s %s' % (m, row[i], units[i]))


def fullcsv(path, limit=40):
    """Compact table from the raw CSV page of an `ncu --set full` report (written on the GPU box:
    `ncu -i x.ncu-rep --page raw --csv`): one line per launch."""
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ki = hdr.index('Kernel Name')

    def col(name):
        return hdr.index(name) if name in hdr else None
    cols = [('us', 'gpu__time_duration.sum'), ('grid', 'launch__grid_size'), ('block', 'launch__block_size'),
            ('regs', 'launch__registers_per_thread'), ('tensor%', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active'),
            ('dramR_MB', 'dram__bytes_read.sum'), ('dramW_MB', 'dram__bytes_write.sum'),
            ('dram%', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'),
            ('l2%', 'lts__throughput.avg.pct_of_peak_sustained_elapsed'), ('sm%', 'sm__throughput.avg.pct_of_peak_sustained_elapsed'),
            ('warps%', 'sm__warps_active.avg.pct_of_peak_sustained_active')]
    scale = {'byte': 1e-6, 'Kbyte': 1e-3, 'Mbyte': 1.0, 'Gbyte': 1e3, 'ns': 1e-3, 'us': 1.0, 'ms': 1e3, 'usecond': 1.0, 'nsecond': 1e-3, 'msecond': 1e3}
    print('# ncu --set full --clock-control none (raw page): %s' % path)
    print('%-46s' % 'kernel' + ''.join('%10s' % c[0] for c in cols))
    for row in data[:limit]:
        name = re.sub(r'<.*', '', re.sub(r'\(.*', '', row[ki])).replace('void ', '')[-46:]
        out = '%-46s' % name
        for label, metric in cols:
            i = col(metric)
            if i is None or row[i] == '':
                out += '%10s' % '-'
                continue
            v = float(row[i].replace(',', ''))
            if label in ('us', 'dramR_MB', 'dramW_MB'):
                v *= scale.get(units[i], 1.0)
            out += '%10.2f' % v if label not in ('grid', 'block', 'regs') else '%10d' % v
        print(out)


def traffic(path, workload, family, calls):
    """dram bytes (read + write) of every launch in an `ncu --set full` capture that covers `calls`
    library calls of `family`, averaged per call -> profiles/traffic.json (read by bench.py)."""
    import json
    import os
    if path.endswith('.csv'):      # `ncu -i x.ncu-rep --page raw --csv` written on the GPU box (reports > 64 MiB stay there)
        out = open(path).read()
    else:
        out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    if len(sys.argv) > 6:          # optional: kernel-name regex and number of matching launches to keep
        ki = hdr.index('Kernel Name')
        data = [r for r in data if re.search(sys.argv[6], r[ki])]
        if len(sys.argv) > 7:
            data = data[:int(sys.argv[7])]
    scale = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
    total = 0.0
    for m in ('dram__bytes_read.sum', 'dram__bytes_write.sum'):
        i = hdr.index(m)
        total += sum(float(r[i].replace(',', '')) * scale[units[i]] for r in data)
    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'traffic.json')
    table = json.load(open(dst)) if os.path.exists(dst) else {}
    table.setdefault(workload, {})[family] = round(total / float(calls), 1)
    table.setdefault('_source', {})['%s/%s' % (workload, family)] = '%s: %d kernel launches over %s calls' % (
        os.path.basename(path), len(data), calls)
    json.dump(table, open(dst, 'w'), indent=1, sort_keys=True)
    print('%s %s: %.1f dram bytes per call' % (workload, family, total / float(calls)))


if __name__ == '__main__':
    if sys.argv[1] == 'traffic':
        traffic(sys.argv[2], sys.argv[3], sys.argv[4], sys.argv[5])
    else:
        {'launches': launches, 'full': full, 'fullcsv': fullcsv}[sys.argv[1]](sys.argv[2])
