"""Digest of an `ncu --set full` capture: headline metrics (raw page) and executed instructions / shared-memory
wavefronts / stall samples per opcode (source page).  usage: ncu_digest.py raw.csv [sass.csv] [units-per-launch]"""
import collections
import csv
import re
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts.avg', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.avg', 'l1tex__data_pipe_lsu_wavefronts_mem_lgds.avg',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'sm__cycles_elapsed.avg', 'derived__memory_l1_wavefronts_shared_excessive']
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    d = dict(zip(hdr, r))
    print('==', d.get('Kernel Name', '')[:90])
    for k in KEYS:
        if k in d:
            print(f'  {k:72s} {d[k]} {units[hdr.index(k)]}')
    for k in hdr:
        if 'issue_stalled' in k and k.endswith('per_issue_active.ratio') and float(d[k] or 0) >= 0.2:
            print(f'  {k:72s} {d[k]}')
if len(sys.argv) > 2:
    nunits = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
    rows = list(csv.reader(open(sys.argv[2])))
    hdr = rows[1]
    ix = {c: i for i, c in enumerate(hdr)}
    byop, samples, wf, wfi = (collections.Counter() for _ in range(4))
    tot = 0.0
    for r in rows[2:]:
        if len(r) < len(hdr) or r[0] in ('Address', 'Kernel Name'):
            continue
        m = re.match(r'(@!?U?P\d+\s+)?([A-Z0-9_.]+)', r[ix['Source']].strip())
        op = (m.group(2) if m else r[ix['Source']]).split('.')[0]
        try:
            n = float(r[ix['Instructions Executed']])
        except ValueError:
            continue
        tot += n
        byop[op] += n
        samples[op] += float(r[ix['# Samples']])
        wf[op] += float(r[ix['L1 Wavefronts Shared']])
        wfi[op] += float(r[ix['L1 Wavefronts Shared Ideal']])
    ts = sum(samples.values())
    print(f'SASS lines {len(rows) - 2}; instructions executed {tot:.0f} = {tot / nunits:.1f} per unit')
    for op, n in byop.most_common(24):
        print(f'  {op:12s} {n / tot * 100:5.1f}%  {n / nunits:8.1f}/unit  samples {samples[op] / ts * 100:5.1f}%  smem wavefronts {wf[op] / nunits:7.1f} (ideal {wfi[op] / nunits:7.1f})')
