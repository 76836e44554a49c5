"""Per-source-line view of an ncu capture: joins the SASS rows of
`ncu -i X.ncu-rep --page source --csv` with the line table of the library's
cubin (`nvdisasm -g`), matching instructions by order inside the kernel.

  python tools/ncu_lines.py <report.ncu-rep> <kernel substring> [top N]

Prints, per source line: stall samples, warp instructions executed, the
dominant stall reasons.
"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, 'gym_solo_b200', 'libsolo8_b200.so')


def sass_lines(kernel_sub):
  """[(file, line)] per SASS instruction of the first function whose mangled
  name contains every '+'-separated piece of kernel_sub."""
  tmp = tempfile.mkdtemp()
  subprocess.check_call(['cuobjdump', '-xelf', 'all', LIB], cwd=tmp, stdout=subprocess.DEVNULL)
  cubin = [f for f in os.listdir(tmp) if f.endswith('.cubin')][0]
  txt = subprocess.check_output(['nvdisasm', '-g', '-c', os.path.join(tmp, cubin)], text=True)
  pieces = kernel_sub.split('+')
  out, cur, active = [], ('?', 0), False
  for ln in txt.splitlines():
    m = re.match(r'\s*\.text\.(\S+):', ln)
    if m:
      if active and out:
        break
      active = all(p in m.group(1) for p in pieces)
      continue
    if not active:
      continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
      cur = (os.path.basename(m.group(1)), int(m.group(2)))
      continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+\S', ln):
      out.append((cur, ln.split('*/', 1)[1].strip().rstrip(';')))
  return out


def main():
  rep, ksub = sys.argv[1], sys.argv[2]
  top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
  # reports with several kernels: keep the rows of the one asked for
  raw = subprocess.check_output(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name',
                                 'regex:' + ksub.split('+')[0]], text=True, stderr=subprocess.DEVNULL)
  rows = list(csv.reader(io.StringIO(raw)))
  hi = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
  hdr = rows[hi]
  body = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
  lines = sass_lines(ksub)
  if len(lines) != len(body):
    print('warning: %d SASS rows in the report vs %d in the cubin (rebuilt since the capture?)'
          % (len(body), len(lines)), file=sys.stderr)
  col = {h: i for i, h in enumerate(hdr)}
  stall_cols = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
  agg = collections.defaultdict(lambda: dict(samples=0, inst=0, stalls=collections.Counter(), n=0))
  total = 0
  for (loc, _), r in zip(lines, body):
    a = agg[loc]
    s = int(r[col['# Samples']] or 0)
    a['samples'] += s
    total += s
    a['inst'] += int(r[col['Instructions Executed']] or 0)
    a['n'] += 1
    for h in stall_cols:
      v = int(r[col[h]] or 0)
      if v:
        a['stalls'][h[6:]] += v
  print('total samples %d over %d SASS instructions' % (total, len(body)))
  print('%-28s %8s %6s %10s %5s  %s' % ('file:line', 'samples', '%', 'warp-inst', 'sass', 'top stalls'))
  for loc, a in sorted(agg.items(), key=lambda kv: -kv[1]['samples'])[:top]:
    st = ', '.join('%s %d' % kv for kv in a['stalls'].most_common(3))
    print('%-28s %8d %6.2f %10d %5d  %s' % ('%s:%d' % loc, a['samples'], 100.0 * a['samples'] / max(1, total),
                                          a['inst'], a['n'], st))


if __name__ == '__main__':
  main()
