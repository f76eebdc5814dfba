#!/usr/bin/env python
"""EM PWM-iterations/s of the FP64 arm over pattern lengths (BASELINE metric 2; configs[4] asks for the long-k-mer table).
Counts come from a synthetic set counted on the device; every PWM runs all iterations (threshold 0).

    python tools/em_bench.py [--W 10 12 14] [--pwms 17 50]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--W", type=int, nargs="+", default=[10, 12, 14])
    ap.add_argument("--pwms", type=int, nargs="+", default=[17, 50])
    ap.add_argument("--iterations", type=int, default=10)
    args = ap.parse_args()
    import _pkg

    pkg = _pkg.load()
    g = pkg.PengCuda(0)
    g.ingest_synthetic(1_000_000, 200, 1234, 0, "TGCTGASTCAGCA", 0.3, 0.0)
    v32 = g.background_model(2, precision=pkg.F32REF)
    v64 = g.background_model(2, precision=pkg.F64)
    out = []
    for W in args.W:
        g.count(W, pkg.STRAND_PLUS if W > 12 else pkg.STRAND_BOTH)
        g.stats(None, v64, K=2)  # FP64 tables only
        rng = np.random.RandomState(W)
        for P in args.pwms:
            pw = rng.dirichlet(np.ones(4), (P, W)) * 0.3
            for m in range(P):
                for p in range(W):
                    pw[m, p, "ACGT".index("TGCTGACTCAGCA"[(m + p) % 13])] += 0.7
            pw /= pw.sum(2, keepdims=True)
            g.em(pw, pkg.F64, 1e4, 0.0, 2)
            g.synchronize()
            t0 = time.perf_counter()
            _, its = g.em(pw, pkg.F64, 1e4, 0.0, args.iterations)
            dt = time.perf_counter() - t0
            bins = 4 ** W
            rec = {"W": W, "pwms": P, "iterations": int(its.sum()), "seconds": dt, "pwm_iterations_per_s": float(its.sum()) / dt,
                   "table_bytes_per_iteration": 16 * bins, "table_GBps": 16 * bins * args.iterations / dt / 1e9,
                   "fp64_gflops_est": (W + 8) * bins * float(its.sum()) / dt / 1e9}
            print(json.dumps(rec), flush=True)
            out.append(rec)
    g.close()


if __name__ == "__main__":
    main()
