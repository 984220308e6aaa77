"""Turn the JSON lines of tools/scale_c5.py runs at 1, 2, 4, 8 GPUs into the scaling table (stdin -> stdout)."""
import json, sys
rows = [json.loads(l) for l in sys.stdin if l.startswith('{')]
by = {}
for r in rows:
    by.setdefault(r['D'], {})[r['n_gpus']] = r
print('C5: K=64, 2^%d samples in total, strong scaling (ms/step: device time of objective launches + exchange; ms/call: through the NumPy-facing call)' % rows[0]['log2_samples_total'])
print('%4s %5s %12s %12s %14s %9s %11s %9s  %s' % ('D', 'GPUs', 'ms/step', 'ms/call', 'samples/s', 'speed-up', 'efficiency', 'exch.share', 'exchange'))
for D in sorted(by):
    base = by[D].get(1)
    for g in sorted(by[D]):
        r = by[D][g]
        sp = base['ms_per_step'] / r['ms_per_step'] if base else float('nan')
        print('%4d %5d %12.3f %12.3f %14.4e %9.2f %11.3f %9.3f  %s' % (D, g, r['ms_per_step'], r['ms_per_call'], r['samples_per_s'], sp, sp / g, r['exchange_share'], r['exchange']))
