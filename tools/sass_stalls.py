"""Static single-warp timing of a SASS address range: sum of the per-instruction stall fields
(bits 105..108 of every 128-bit instruction; /opt/skills/guides/B300_MICROARCH.md "single-warp
issue model"), with the scoreboard slots shown.  For a warp that runs alone -- the tail of
solve_warp_kernel at 4096 environments -- the sum of stall counts along a path is its time,
up to the variable-latency waits.

  python tools/sass_stalls.py <lib.so> <mangled-kernel-substring> <start-hex> <end-hex>
"""
import re, subprocess, sys

def load(lib, kernel):
  out = subprocess.run(['cuobjdump', '-sass', lib], capture_output=True, text=True).stdout
  blocks = out.split('Function : ')
  for b in blocks:
    if kernel in b.split('\n', 1)[0]:
      return b
  raise SystemExit('kernel not found')

def parse(text):
  ins = []
  lines = text.splitlines()
  i = 0
  pat = re.compile(r'/\*([0-9a-f]{4,})\*/\s+(.*?);\s*/\* (0x[0-9a-f]{16}) \*/')
  while i < len(lines):
    m = pat.search(lines[i])
    if m and i + 1 < len(lines):
      m2 = re.search(r'/\* (0x[0-9a-f]{16}) \*/', lines[i + 1])
      if m2:
        addr = int(m.group(1), 16); hi = int(m2.group(1), 16)
        stall = (hi >> 41) & 0xf; yld = (hi >> 45) & 1; wbar = (hi >> 46) & 7; rbar = (hi >> 49) & 7; wait = (hi >> 52) & 0x3f
        ins.append((addr, m.group(2).strip(), stall, yld, wbar, rbar, wait))
        i += 2
        continue
    i += 1
  return ins

if __name__ == '__main__':
  lib, kern, a, b = sys.argv[1], sys.argv[2], int(sys.argv[3], 16), int(sys.argv[4], 16)
  ins = parse(load(lib, kern))
  tot = 0
  for addr, txt, stall, yld, wbar, rbar, wait in ins:
    if a <= addr <= b:
      tot += stall
      print('%05x st%2d %s w%d r%d wm%02x  %s' % (addr, stall, 'Y' if yld else ' ', wbar, rbar, wait, txt[:70]))
  print('instructions in range, sum of stalls:', sum(1 for x in ins if a <= x[0] <= b), tot)
