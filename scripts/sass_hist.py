#!/usr/bin/env python
"""Static SASS opcode histogram of one kernel of a built system library.
usage: sass_hist.py <lib.so> <kernel-substring>"""
import collections, re, subprocess, sys
so, pat = sys.argv[1], sys.argv[2]
txt = subprocess.run(['cuobjdump', '-sass', so], capture_output=True, text=True).stdout
cur, hist = None, collections.Counter()
for ln in txt.splitlines():
    m = re.search(r'Function : (\S+)', ln)
    if m:
        cur = m.group(1)
        continue
    if cur and pat in cur:
        m = re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)', ln)
        if m:
            hist[m.group(2).split('.')[0]] += 1
tot = sum(hist.values())
print('total', tot)
print(' '.join('%s:%d' % kv for kv in hist.most_common(25)))
