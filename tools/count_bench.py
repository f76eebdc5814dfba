"""Counting-step timing over shapes / pattern lengths / strands (CUDA events inside the library, L2 flushed).
usage: count_bench.py [n_seq seq_len W strand(BOTH|PLUS)]...   e.g.  count_bench.py 1000000 200 10 BOTH 10000000 300 14 PLUS"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import _pkg
pkg = _pkg.load()
args = sys.argv[1:] or ["1000000", "200", "10", "BOTH"]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
last = None
g = None
for k in range(0, len(args), 4):
    n, L, W, strand = int(args[k]), int(args[k + 1]), int(args[k + 2]), args[k + 3]
    if (n, L) != last:
        if g is not None:
            g.close()
        g = pkg.PengCuda(0)
        g.set_stream(torch.cuda.current_stream().cuda_stream)
        g.ingest_synthetic(n, L, 1234, 0, "TGCTGASTCAGCA", 0.3, 0.0)
        g.enable_timing(True)
        g.set_count_path(int(os.environ.get("PENG_COUNT_PATH", "0")))  # 0 auto, 1 L2 atomics, 2 partition
        last = (n, L)
    rows, ev = [], []
    for i in range(6):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        ltot = g.count(W, pkg.STRAND_BOTH if strand == "BOTH" else pkg.STRAND_PLUS)
        b.record()
        torch.cuda.synchronize()
        ev.append(a.elapsed_time(b))
        rows.append([g.get_timing(p) for p in range(7)])
    r = np.median(np.array(rows[2:]), axis=0)
    ms = float(np.median(ev[2:]))
    print(json.dumps({"n_seq": n, "seq_len": L, "W": W, "strand": strand, "path": g.last_count_path, "ltot": int(ltot),
                      "step_ms": round(ms, 4), "gbases_per_s": round(n * L / ms / 1e6, 2),
                      "phases_ms": {"mark": round(float(r[0]), 4), "count": round(float(r[1]), 4), "fold": round(float(r[2]), 4),
                                    "P": round(float(r[3]), 4), "slow": round(float(r[4]), 4), "H": round(float(r[5]), 4), "P2": round(float(r[6]), 4)}}),
          flush=True)
