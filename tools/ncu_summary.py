"""Summarise an .ncu-rep (here, without a GPU): key raw metrics, stall-reason totals and the hottest
instructions of each profiled kernel.   python tools/ncu_summary.py file.ncu-rep [top_n]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 12
KEEP = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size',
        'launch__block_size', 'launch__waves_per_multiprocessor', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'smsp__inst_executed.sum', 'lts__t_sector_hit_rate.pct',
        'sm__cycles_active.avg']
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
for r in rows[2:]:
    print('==', r[idx['Kernel Name']][:90])
    for k in KEEP:
        if k in idx:
            print('   %-66s %s %s' % (k, r[idx[k]], units[idx[k]]))
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
sections, cur = [], None
for r in rows:
    if r and r[0] == 'Kernel Name':
        cur = {'name': r[1], 'hdr': None, 'data': []}
        sections.append(cur)
    elif r and r[0] == 'Address' and cur is not None:
        cur['hdr'] = r
    elif cur is not None and cur['hdr'] and len(r) >= len(cur['hdr']):
        cur['data'].append(r)
for s in sections:
    idx = {h: i for i, h in enumerate(s['hdr'])}
    data = s['data']
    print('== source:', s['name'][:90])
    stalls = [h for h in s['hdr'] if h.startswith('stall_') and 'Not Issued' not in h]
    tot = {n: sum(int(r[idx[n]] or 0) for r in data) for n in stalls}
    allsum = sum(tot.values()) or 1
    print('   stalls:', ', '.join('%s %.1f%%' % (n[6:], 100.0 * v / allsum) for n, v in sorted(tot.items(), key=lambda x: -x[1]) if v * 50 > allsum))
    ops = {}
    for r in data:
        op = r[idx['Source']].split()
        op = [t for t in op if not t.startswith('@')]
        name = op[0].split('.')[0] if op else '?'
        ops[name] = ops.get(name, 0) + int(r[idx['Instructions Executed']] or 0)
    total = sum(ops.values()) or 1
    print('   warp instr %d:' % total, ', '.join('%s %.1f%%' % (k, 100.0 * v / total) for k, v in sorted(ops.items(), key=lambda x: -x[1])[:14]))
    ns = sum(int(r[idx['# Samples']] or 0) for r in data) or 1
    for r in sorted(data, key=lambda r: -int(r[idx['# Samples']] or 0))[:top_n]:
        print('   %5.1f%% %9s  %s' % (100.0 * int(r[idx['# Samples']]) / ns, r[idx['Instructions Executed']], r[idx['Source']].strip()[:80]))
