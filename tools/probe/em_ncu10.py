#!/usr/bin/env python
"""Short EM run for an ncu launch list: W = 10, 17 and 50 PWMs, 3 iterations each (k_em_accumulate_all, k_em_reduce_update)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import _pkg

pkg = _pkg.load()
g = pkg.PengCuda(0)
g.ingest_synthetic(1_000_000, 200, 1234, 0, "TGCTGASTCAGCA", 0.3, 0.0)
v64 = g.background_model(2, precision=pkg.F64)
g.count(10, pkg.STRAND_BOTH)
g.stats(None, v64, K=2)
for P in (17, 50):
    pw = np.random.RandomState(1).dirichlet(np.ones(4), (P, 10))
    out, its = g.em(pw, pkg.F64, 1e4, 0.0, 3)
print("ok")
g.close()
