"""Host-buffer ingest timing per mode: raw bytes + device packing / host packing / hybrid (peng_cuda_set_ingest_mode)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import _pkg
pkg = _pkg.load()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
g = pkg.PengCuda(0)
if os.environ.get("PENG_TOOL_STREAM") == "legacy":  # what a caller on torch's default stream hands over
    g.set_stream(torch.cuda.current_stream().cuda_stream)
elif os.environ.get("PENG_TOOL_STREAM") == "torch":
    _s = torch.cuda.Stream()
    g.set_stream(_s.cuda_stream)
g.ingest_synthetic(n, 200, 1234, 0, "TGCTGASTCAGCA", 0.3, 0.0)
codes, offs = g.export_codes()
pin = torch.from_numpy(codes).pin_memory()
for mode, frac in [(1, -1), (2, -1), (0, -1), (0, 0.0), (0, 0.15), (0, 0.25), (0, 0.35), (0, 0.5), (0, 0.65), (0, 0.8), (0, 0.9), (0, 0.95)]:
    g.set_ingest_mode(mode, frac)
    ts = []
    for i in range(8):
        t0 = time.perf_counter(); g.ingest_codes(pin.numpy(), offs); g.synchronize(); ts.append(time.perf_counter() - t0)
    print(f"mode {mode} fraction {frac}: {1e3 * np.median(ts[2:]):.2f} ms  ({n * 200 / np.median(ts[2:]) / 1e9:.1f} Gbases/s), host-packed {g.last_ingest_host_packed}", flush=True)

# ingest + count with and without the count prefetch (stage P inside the ingest), same box, same buffers
g.set_ingest_mode(0, -1)
for arm in (0, 10, 0, 10):
    g.set_count_prefetch(arm, 1)
    ti, tc = [], []
    for i in range(10):
        t0 = time.perf_counter(); g.ingest_codes(pin.numpy(), offs); g.synchronize(); t1 = time.perf_counter()
        g.count(10, 1, want_ltot=False); g.synchronize(); t2 = time.perf_counter()
        ti.append(t1 - t0); tc.append(t2 - t1)
    print(f"prefetch {'on ' if arm else 'off'}: ingest {1e3 * np.median(ti[2:]):.2f} ms + count {1e3 * np.median(tc[2:]):.2f} ms = {1e3 * (np.median(ti[2:]) + np.median(tc[2:])):.2f} ms", flush=True)
g.set_count_prefetch(0)
