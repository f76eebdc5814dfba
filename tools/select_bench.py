#!/usr/bin/env python
"""Seed selection (peng_cuda_select_base_patterns): device partition levels vs the whole-table host replay.

    python tools/select_bench.py [--W 10 12]
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
    ap.add_argument("--W", type=int, nargs="+", default=[10, 12])
    args = ap.parse_args()
    import _pkg

    pkg = _pkg.load()
    for W in args.W:
        res = {}
        for mode in ("device", "host"):
            if mode == "host":
                os.environ["PENG_SELECT_HOST"] = "1"
            else:
                os.environ.pop("PENG_SELECT_HOST", None)
            g = pkg.PengCuda(0)
            g.ingest_synthetic(1_000_000, 200, 1234, 0, "TGCTGASTCAGCA", 0.3, 0.0)
            v32 = g.background_model(2, precision=pkg.F32REF)
            g.count(W, pkg.STRAND_BOTH)
            g.stats(v32, None, K=2)
            g.select_base_patterns(10.0, 3)
            t = []
            for _ in range(5):
                t0 = time.perf_counter()
                sel = g.select_base_patterns(10.0, 3)
                t.append(time.perf_counter() - t0)
            res[mode] = (min(t), sel)
            g.close()
        assert np.array_equal(res["device"][1], res["host"][1])
        print(json.dumps({"W": W, "seeds": int(len(res["device"][1])), "device_ms": 1e3 * res["device"][0], "host_ms": 1e3 * res["host"][0]}),
              flush=True)


if __name__ == "__main__":
    main()
