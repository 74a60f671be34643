#!/usr/bin/env python
"""Turns the raw ncu exports of profiles/r02_ncu.sh (gpurun_out/r02_mega_<streams>_raw.csv, `ncu --page raw --csv` of one
`--set full` capture of mega_generate_kernel over 8 tokens) into the committed evidence:
  profiles/r02_mega_ncu_<streams>_selected.csv   the counters the roofline discussion uses (name, unit, value)
  profiles/r02_mega_ncu_traffic.json             dram bytes per token step per operating point -- what bench.py prints as roofline.traffic
    python profiles/ncu_traffic.py [gpurun_out]"""
import csv, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, 'gpurun_out')
TOKENS = 8
KEEP = ('dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__time_duration.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'dram__cycles_active.avg', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic', 'lts__t_bytes.sum', 'lts__t_sector_hit_rate.pct',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__cycles_elapsed.max')
out = {}
for streams in (1024, 128):
    path = os.path.join(src, f'r02_mega_{streams}_raw.csv')
    if not os.path.exists(path):
        continue
    rows = list(csv.reader(open(path)))
    names, units, vals = rows[0], rows[1], rows[2]
    d = {n: (u, v) for n, u, v in zip(names, units, vals)}
    with open(os.path.join(ROOT, 'profiles', f'r02_mega_ncu_{streams}_selected.csv'), 'w', newline='') as f:
        w = csv.writer(f)
        w.writerow(['metric', 'unit', 'value'])
        w.writerow(['kernel', '', d['Kernel Name'][1]])
        for k in KEEP:
            if k in d:
                w.writerow([k, d[k][0], d[k][1]])
    scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
    rd = float(d['dram__bytes_read.sum'][1]) * scale[d['dram__bytes_read.sum'][0]]
    wr = float(d['dram__bytes_write.sum'][1]) * scale[d['dram__bytes_write.sum'][0]]
    out[str(streams)] = {'dram_bytes_per_token': (rd + wr) / TOKENS, 'dram_read_bytes_per_token': rd / TOKENS, 'dram_write_bytes_per_token': wr / TOKENS,
                         'tokens_in_capture': TOKENS, 'kernel_us_under_ncu': float(d['gpu__time_duration.sum'][1]),
                         'source': f'profiles/r02_mega_ncu_{streams}_selected.csv (ncu --set full, one launch of mega_generate_kernel, {TOKENS} tokens, profiles/r02_ncu.sh)'}
json.dump(out, open(os.path.join(ROOT, 'profiles', 'r02_mega_ncu_traffic.json'), 'w'), indent=1)
print(json.dumps(out, indent=1))
