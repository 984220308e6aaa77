"""Opcode histogram per kernel of the built library: python tools/sass_summary.py [libnfcore.so] > profiles/rNN_sass_summary.txt
Evidence for the Blackwell-native instructions (UTCHMMA = tcgen05.mma, LDTM/STTM = tcgen05.ld/st, UTMALDG = TMA loads,
FFMA2/FMUL2/FADD2 = packed f32x2 arithmetic, MUFU.* = transcendental unit)."""
import collections, os, re, subprocess, sys
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'vi-normflows_b200', 'lib', 'libnfcore.so')
out = subprocess.run(['cuobjdump', '-sass', lib], capture_output=True, text=True).stdout
kern, hist = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.match(r'\s+Function : (\S+)', line)
    if m:
        kern = subprocess.run(['c++filt', m.group(1)], capture_output=True, text=True).stdout.strip()
        hist[kern] = collections.Counter()
        continue
    m = re.match(r'\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)', line)
    if m and kern:
        op = m.group(1)
        key = op if op.startswith(('MUFU', 'UTC', 'UTMA', 'LDTM', 'STTM', 'UBLKCP')) else op.split('.')[0]
        hist[kern][key] += 1
WATCH = ['UTCHMMA', 'LDTM', 'STTM', 'UTMALDG', 'FFMA2', 'FMUL2', 'FADD2', 'FFMA', 'MUFU.EX2', 'MUFU.RCP', 'MUFU.LG2', 'MUFU.SIN', 'MUFU.COS', 'SHFL', 'LDS', 'STS', 'LDG', 'STG', 'ATOM', 'ATOMG', 'RED']
tot = collections.Counter()
print('%-110s %7s  %s' % ('kernel', 'instr', 'notable opcodes'))
for k, h in hist.items():
    n = sum(h.values())
    tot.update(h)
    agg = collections.Counter()
    for op, c in h.items():
        for w in WATCH:
            if op == w or op.startswith(w + '.'):
                agg[w] += c
    note = ' '.join('%s=%d' % (w, agg[w]) for w in WATCH if agg[w])
    print('%-110s %7d  %s' % (k[:110], n, note))
print()
print('library totals:', ' '.join('%s=%d' % (w, sum(c for op, c in tot.items() if op == w or op.startswith(w + '.'))) for w in WATCH))
