#!/usr/bin/env python
"""Where the time of one peng_cuda_em call goes: the protocol entry points timed one by one (host clock, synchronised)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import _pkg

pkg = _pkg.load()
g = pkg.PengCuda(0)
g.ingest_synthetic(1_000_000, 200, 1234, 0, "TGCTGASTCAGCA", 0.3, 0.0)
v64 = g.background_model(2, precision=pkg.F64)
for W in (10, 12):
    g.count(W, pkg.STRAND_BOTH)
    g.stats(None, v64, K=2)
    for P in (17, 50):
        pw = np.random.RandomState(W).dirichlet(np.ones(4), (P, W))
        for rep in range(3):
            t = [time.perf_counter()]
            g.em_begin(pw, 1e4, 0.0, 10); g.synchronize(); t.append(time.perf_counter())
            for it in range(4):
                g.em_accumulate(); g.synchronize(); t.append(time.perf_counter())
                g.em_update(False); g.synchronize(); t.append(time.perf_counter())
            g.em_end(); t.append(time.perf_counter())
            d = np.diff(t) * 1e3
            print(f"W={W} P={P} rep={rep}: begin {d[0]:.3f} ms | acc/upd " + " ".join(f"{d[1 + 2 * i]:.3f}/{d[2 + 2 * i]:.3f}" for i in range(4)) +
                  f" | end {d[-1]:.3f} ms", flush=True)
        t0 = time.perf_counter(); g.em(pw, pkg.F64, 1e4, 0.0, 10); print(f"   em() 10 iterations: {1e3 * (time.perf_counter() - t0):.3f} ms", flush=True)
g.close()
