#!/usr/bin/env python
"""Short EM run for ncu: W = 12, 50 PWMs, 2 iterations of the FP64 arm (k_em_accumulate_all)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import _pkg

pkg = _pkg.load()
g = pkg.PengCuda(0)
g.ingest_synthetic(1_000_000, 200, 1234, 0, "TGCTGASTCAGCA", 0.3, 0.0)
v64 = g.background_model(2, precision=pkg.F64)
g.count(12, pkg.STRAND_BOTH)
g.stats(None, v64, K=2)
pw = np.random.RandomState(1).dirichlet(np.ones(4), (50, 12))
out, its = g.em(pw, pkg.F64, 1e4, 0.0, 2)
print("ok", its.sum())
g.close()
