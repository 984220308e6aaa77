import os, sys
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import sweep
for lg in (17, 18, 19, 20, 21, 22, 24):
    sweep.run('default', N=1 << lg, steps=100 if lg < 23 else 10)
