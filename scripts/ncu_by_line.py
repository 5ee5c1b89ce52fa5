#!/usr/bin/env python
"""Attribute the executed warp instructions of one profiled kernel to source lines (no GPU needed).

    python scripts/ncu_by_line.py <report.ncu-rep> <lib.so> <kernel-mangled-substring> [top]

ncu's ``--page source --print-source sass`` lists every SASS instruction of the kernel (callees included) with its
executed count; ``nvdisasm -g`` of the cubin inside <lib.so> gives the (innermost) file:line of the same instructions
in the same order.  Printed: executed warp instructions per file, per line (top N) and per opcode class.
"""
import collections, csv, io, os, re, subprocess, sys, tempfile

rep, so, pat = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
opfilter = sys.argv[5] if len(sys.argv) > 5 else None  # e.g. LDL: lines of that opcode only
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
H = rows[hdr]
ci, cs, ct = H.index('Instructions Executed'), H.index('Source'), H.index('Thread Instructions Executed')
cw = H.index('# Samples') if '# Samples' in H else None
cwv = H.index('L1 Wavefronts Shared') if 'L1 Wavefronts Shared' in H else None
cwx = H.index('L1 Wavefronts Shared Excessive') if 'L1 Wavefronts Shared Excessive' in H else None


def _i(v):
    try:
        return int(v)
    except ValueError:
        return 0


sass = [(r[cs].strip(), _i(r[ci]), _i(r[ct]), _i(r[cw]) if cw is not None else 0, _i(r[cwv]) if cwv else 0,
         _i(r[cwx]) if cwx else 0) for r in rows[hdr + 1:] if len(r) == len(H)]
tmp = tempfile.mkdtemp()
subprocess.run(['cuobjdump', '-xelf', 'all', os.path.abspath(so)], cwd=tmp, capture_output=True)
cub = [f for f in os.listdir(tmp) if f.endswith('.cubin')][0]
dis = subprocess.run(['nvdisasm', '-g', '-c', os.path.join(tmp, cub)], capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(dis) if l.startswith('.text.') and pat in l)
lines, cur = [], ('?', 0)
for l in dis[start + 1:]:
    if l.startswith('//-----') or (l.startswith('.text.') and pat not in l):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r'\s+/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
    if m:
        lines.append((cur, m.group(2).strip()))
if len(lines) != len(sass):
    print('warning: %d disassembled vs %d profiled instructions' % (len(lines), len(sass)))
n = min(len(lines), len(sass))
tot = sum(s[1] for s in sass[:n])
byfile, byline, byop, samples, wav, wavx = (collections.Counter() for _ in range(6))
for (loc, txt), (src, cnt, thr, smp, wv, wx) in zip(lines[:n], sass[:n]):
    wav[loc] += wv
    wavx[loc] += wx
    if opfilter and not (txt.split()[1] if txt.startswith('@') else txt.split()[0]).startswith(opfilter):
        continue
    byfile[loc[0]] += cnt
    byline[loc] += cnt
    samples[loc] += smp
    op = txt.split()[1] if txt.startswith('@') else txt.split()[0]
    byop[op.split('.')[0]] += cnt
print('total executed warp instructions: %.4g' % tot)
for f, c in byfile.most_common():
    print('%6.2f %%  %s' % (100.0*c/tot, f))
print('--- lines')
for (f, ln), c in byline.most_common(top):
    print('%6.2f %%  samples %6d  %s:%d' % (100.0*c/tot, samples[(f, ln)], f, ln))
ts = sum(samples.values())
if ts:
    print('--- lines by stall samples (share of %d samples ~ share of the time)' % ts)
    for (f, ln), c in samples.most_common(top):
        print('%6.2f %%  instr %5.2f %%  %s:%d' % (100.0*c/ts, 100.0*byline[(f, ln)]/tot, f, ln))
print('--- opcodes')
for op, c in byop.most_common(25):
    print('%6.2f %%  %s' % (100.0*c/tot, op))
tw = sum(wav.values())
if tw:
    print('--- shared-memory / shuffle wavefronts: %.4g, excessive (bank conflicts) %.4g' % (tw, sum(wavx.values())))
    for loc, v in wav.most_common(top//2):
        print('%6.2f %%  (%5.1f %% excessive)  %s:%d' % (100.0*v/tw, 100.0*wavx[loc]/max(1, v), loc[0], loc[1]))
