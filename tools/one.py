"""Run one probe configuration (for ncu): python tools/one.py CFG K logN steps"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
os.environ['NF_ELBO_CFG'] = sys.argv[1]
import sweep
sweep.run('cfg=' + sys.argv[1], K=int(sys.argv[2]), N=1 << int(sys.argv[3]), steps=int(sys.argv[4]))
