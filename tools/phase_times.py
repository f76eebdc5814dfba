"""Phase timing of the counting step on the BASELINE configs[2] shape (CUDA events inside the library)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import _pkg
pkg = _pkg.load()
g = pkg.PengCuda(0)
g.set_stream(torch.cuda.current_stream().cuda_stream)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
g.ingest_synthetic(n, 200, 1234, 0, "TGCTGASTCAGCA", 0.3, 0.0)
g.enable_timing(True)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for path in (1, 2):
    g.set_count_path(path)
    for both in (1, 0):
        rows = []
        for i in range(8):
            flush.zero_()
            g.count(10, both)
            rows.append([g.get_timing(p) for p in range(6)])
        r = np.median(np.array(rows[3:]), axis=0)
        print(f"path {path} both={both}: mark {r[0]:.3f}  count {r[1]:.3f}  fold {r[2]:.3f} | P {r[3]:.3f}  slow {r[4]:.3f}  H {r[5]:.3f} ms")
