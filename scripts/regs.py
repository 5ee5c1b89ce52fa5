#!/usr/bin/env python
"""registers / spills / stack per kernel from the nvcc -Xptxas -v log of a system's build directory:
    python scripts/regs.py robertson [filter]"""
import re, subprocess, sys
sys.path.insert(0, '.')
from pyodesys_b200 import systems
name = sys.argv[1]
flt = sys.argv[2] if len(sys.argv) > 2 else ''
s = getattr(systems, name)(compile=True)
log = open(s._native.build_dir + '/nvcc.log').read()
cur = None
for line in log.splitlines():
    m = re.search(r"Compiling entry function '([^']+)'", line)
    if m:
        cur = m.group(1)
        if cur.startswith('_Z'):
            cur = subprocess.run(['c++filt', cur], capture_output=True, text=True).stdout.strip()
        cur = cur.replace('pyodesys_cuda::', '').replace('(KernelArgs)', '')
        stack = ''
    m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
    if m:
        stack = 'stack %s spill st %s ld %s' % m.groups()
    m = re.search(r"Used (\d+) registers", line)
    if m and cur and flt in cur:
        print('%4s regs  %-28s %s' % (m.group(1), stack, cur))
