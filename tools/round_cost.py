"""Cost of one round of the packed-pair kernel as a function of the samples per thread (tuning aid):
N = 148*512*S samples for S = 2, 4, 6, 8 (one round each) and two/three full rounds."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sweep
T = 148 * 512
for S, label in ((2, '1 round of 2'), (4, '1 round of 4'), (6, '1 round of 6'), (8, '1 round of 8'), (10, '8 + 2'), (12, '8 + 4'), (14, '8 + 6'),
                 (16, '8 + 8'), (24, '8 + 8 + 8'), (32, '4 x 8')):
    sweep.run('%-12s' % label, K=32, N=T * S, steps=200)
