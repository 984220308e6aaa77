"""Run one probe configuration (for sweeps and ncu): python tools/one.py CFG K logN steps [target]
CFG = "warps,samples per thread,of which in tensor memory" of the packed-pair kernel, "default", or "off" (scalar kernel)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
cfg = sys.argv[1]
if cfg == 'off':
    os.environ['NF_PAIR'] = '0'
elif cfg != 'default':
    os.environ['NF_PAIR_CFG'] = cfg
import sweep
sweep.run('cfg=' + cfg, K=int(sys.argv[2]), N=1 << int(sys.argv[3]), steps=int(sys.argv[4]), target=sys.argv[5] if len(sys.argv) > 5 else 'p1')
