"""Copy the evidence of one GPU session from gpurun_out/ into profiles/ under a
round tag: bench lines, the ncu launch list, key rows of the `ncu --set full`
captures (one csv per kernel), per-source-line stall tables, and
profiles/traffic.json (dram bytes per launch, read by bench.py's roofline.traffic).

  python tools/save_profiles.py r1f [--envs=N] setup_legpar_kernel solve_warp_kernel finish_kernel
"""
import csv, io, json, os, shutil, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT, PROF = os.path.join(ROOT, 'gpurun_out'), os.path.join(ROOT, 'profiles')
tag, kernels = sys.argv[1], sys.argv[2:]
envs = 4096
if kernels and kernels[0].startswith('--envs='):
  envs = int(kernels[0].split('=')[1]); kernels = kernels[1:]
WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'smsp__cycles_active.avg', 'sm__cycles_elapsed.max', 'launch__occupancy_limit', 'launch__waves_per_multiprocessor',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'smsp__average_warps_issue_stalled',
        'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'launch__shared_mem_per_block', 'sm__inst_executed_pipe']
for f, g in (('bench_4096.json', 'bench_4096'), ('bench_65536.json', 'bench_65536'), ('bench_ref.json', 'bench_reference_arm'),
             ('launches_c.csv', 'launches_timed_region')):
  src = os.path.join(OUT, f)
  if os.path.exists(src):
    shutil.copy(src, os.path.join(PROF, '%s_%s%s' % (tag, g, os.path.splitext(f)[1])))
traffic = {}
insts = {}
for k in kernels:
  rep = os.path.join(OUT, '%s_%s.ncu-rep' % (k, tag))
  if not os.path.exists(rep):
    continue
  raw = subprocess.check_output(['ncu', '-i', rep, '--page', 'raw', '--csv'], text=True, stderr=subprocess.DEVNULL)
  rows = list(csv.reader(io.StringIO(raw)))
  hdr, units, vals = rows[0], rows[1], rows[2]
  d = {}
  with open(os.path.join(PROF, '%s_%s_ncu_full.csv' % (tag, k)), 'w') as f:
    f.write('metric,unit,value\n')
    for h, u, v in zip(hdr, units, vals):
      if h == 'Kernel Name' or (any(w in h for w in WANT) and 'per_second' not in h):
        f.write('%s,%s,"%s"\n' % (h, u, v))
      d[h] = (u, v)
  def byt(name):
    u, v = d[name]
    return float(v.replace(',', '')) * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[u]
  traffic[k] = byt('dram__bytes_read.sum') + byt('dram__bytes_write.sum')
  insts[k] = float(d['smsp__inst_executed.sum'][1].replace(',', ''))
  sub = {'setup_kernel': 'setup_kernel+ILi2E', 'setup_legpar_kernel': 'setup_legpar_kernel+ILi2E', 'solve_warp_kernel': 'solve_warp_kernel+ILi2ELb0ELi14' if envs >= 10240 else 'solve_warp_kernel+ILi2ELb0ELi0', 'finish_kernel': 'finish_kernel+ILi2E',
         'solve_gram_kernel': 'solve_gram_kernel+ILi2ELb0'}.get(k, k)
  txt = subprocess.run([sys.executable, os.path.join(ROOT, 'tools', 'ncu_lines.py'), rep, sub, '40'], text=True,
                       capture_output=True).stdout
  open(os.path.join(PROF, '%s_%s_lines.txt' % (tag, k)), 'w').write(txt)
path = os.path.join(PROF, 'traffic.json')
try:
  allt = json.load(open(path))
except Exception:
  allt = {}
by = allt.get('by_envs', {})
cur = by.get(str(envs), {'dram_bytes_per_launch': {}, 'warp_instructions_per_launch': {}})
cur['round'] = tag
cur['dram_bytes_per_launch'].update(traffic)
cur['warp_instructions_per_launch'].update(insts)
by[str(envs)] = cur
allt = {'how': 'ncu --set full --clock-control none, one launch inside bench.py (cold caches: ncu flushes between passes); '
               'per batch size (environments per GPU)', 'by_envs': by}
json.dump(allt, open(path, 'w'), indent=1)
print(envs, traffic)
