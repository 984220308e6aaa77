"""normflows/normflows/config.py: paths and the package-global RandomState(101)."""
from pathlib import Path

import numpy.random as npr

root = Path(__file__).parent.parent.parent
data = root / 'data'
mnist = data / 'mnist'
notebooks = root / 'notebooks'
figs = root / 'figures'
models = root / 'models'
results = root / 'results'

figname = str(figs / '{}_flows_iter_{}.png')
rs = npr.RandomState(101)
