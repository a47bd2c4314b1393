"""Summarise a GZ_TC_TIMELINE dump (front-warp marks, per-CTA wall times, MMA-warp marks)."""
import re
import statistics as st
import sys

txt = open(sys.argv[1]).read().strip().split('launch ')
blk = [b for b in txt if b.startswith('B=')][-1]
front, rest = blk.split('cta_ns')
cta, mma = rest.split('mma')
mma = mma.split('back')[0]
F = [list(map(int, l.split())) for l in front.splitlines()[1:] if len(l.split()) == 4]
M = [list(map(int, l.split())) for l in mma.splitlines() if len(l.split()) == 7]
med = lambda x: int(st.median(x))
print(blk.splitlines()[0])
print('group period (cycles)', med([F[i + 1][0] - F[i][0] for i in range(5, len(F) - 2)]))
print('front: wait conv2 %d | ep2 + stage next %d | ep1 next %d | idle %d' % (
    med([f[1] - f[0] for f in F[5:-2]]), med([f[2] - f[1] for f in F[5:-2]]), med([f[3] - f[2] for f in F[5:-2]]),
    med([F[i + 1][0] - F[i][3] for i in range(5, len(F) - 2)])))
names = ['issue conv2', 'wait conv3 accum free', 'barrier A0 (ep2 quarter 0)', 'conv3 st 0-3 + barriers A1, B',
         'conv1 next + conv3 st 4-17', 'barrier C (ep1 next)']
for k in range(6):
    print('mma  %-28s %6d' % (names[k], med([m[k + 1] - m[k] for m in M[5:-2]])))
vals = [(int(a), int(b)) for a, b in re.findall(r'(-?\d+):(-?\d+)', cta.split('duration)')[1])]
d = [v[1] for v in vals]
print('per-CTA wall time ns: min %d median %d max %d (n=%d)' % (min(d), st.median(d), max(d), len(d)))
