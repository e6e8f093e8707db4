"""profiles/r1_dram_traffic.json from the ncu DRAM pass of one bench step (scripts_dev/gpu.sh dram):
per-kernel launches, DRAM bytes (read + write) and duration of the LAST step in the launch list."""
import csv, json, re, sys, collections
src, dst = sys.argv[1], sys.argv[2]
rows = [r for r in csv.reader(open(src)) if len(r) > 10]
hdr = rows[0]
ID, K, M, V, U = hdr.index('ID'), hdr.index('Kernel Name'), hdr.index('Metric Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
launch = collections.OrderedDict()
for r in rows[1:]:
    d = launch.setdefault(r[ID], {'name': r[K], 'bytes': 0.0, 'ns': 0.0})
    v = float(r[V].replace(',', ''))
    if r[M].startswith('dram__bytes'):
        d['bytes'] += v * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[r[U]]
    elif r[M].startswith('gpu__time'):
        d['ns'] += v * {'ns': 1, 'us': 1e3, 'ms': 1e6, 'nsecond': 1, 'usecond': 1e3, 'msecond': 1e6}[r[U]]
L = list(launch.values())
starts = [i for i, d in enumerate(L) if 'minmax_init' in d['name']]
step = [d for d in L[starts[-1]:] if 'at::' not in d['name']]
def short(n):
    m = re.search(r'(\w+)(<[^>]*>)?\(', n)
    return (m.group(1) + (m.group(2) or '')) if m else n[:40]
per = collections.OrderedDict()
for d in step:
    e = per.setdefault(short(d['name']), {'launches': 0, 'dram_bytes': 0.0, 'time_us': 0.0})
    e['launches'] += 1; e['dram_bytes'] += d['bytes']; e['time_us'] += d['ns'] / 1e3
gemm = [k for k in per if k.startswith(('conv_tc_kernel', 'conv_dwpw_kernel', 'conv_halo_kernel', 'head_taps_kernel'))]
out = {'source': 'ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none '
                 '--cache-control none over bench.py --steps 1 --warmup 3 (scripts_dev/gpu.sh dram); last step of the list',
       'launches_in_step': len(step), 'gemm_launches': sum(per[k]['launches'] for k in gemm),
       'gemm_dram_bytes_per_step': sum(per[k]['dram_bytes'] for k in gemm),
       'all_dram_bytes_per_step': sum(e['dram_bytes'] for e in per.values()),
       'step_time_us_serialised': sum(e['time_us'] for e in per.values()), 'per_kernel': per}
json.dump(out, open(dst, 'w'), indent=1)
print(json.dumps({k: out[k] for k in ('launches_in_step', 'gemm_launches', 'gemm_dram_bytes_per_step', 'all_dram_bytes_per_step', 'step_time_us_serialised')}))
