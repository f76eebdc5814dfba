import sys, time, numpy as np
sys.path.insert(0, '/root/repo')
import _pkg
pkg = _pkg.load()
MOTIF = "TGCTGASTCAGCA"
def run(W, both, world, paths):
    hs = [pkg.PengCuda(0) for _ in range(world)]
    blobs = np.stack([h.peer_export(W) for h in hs])
    for r, h in enumerate(hs): h.peer_attach(r, world, blobs)
    n = 30000
    for step in range(3):
        cuts = [n * r // world for r in range(world + 1)]
        for r, h in enumerate(hs):
            h.set_count_path(paths[(r + step) % len(paths)])
            h.ingest_synthetic(cuts[r + 1] - cuts[r], 200, 32, cuts[r], MOTIF, 0.3, 0.01)
            h.count_local32(W, int(both))
        t0 = time.perf_counter()
        for h in hs: h.count_exchange32(want_ltot=False)
        res = []
        for h in hs:
            try: res.append(h.count_finish())
            except Exception as e: res.append(str(e)[-110:])
        print(W, both, world, paths, 'step', step, f'{1e3*(time.perf_counter()-t0):.1f} ms', res, flush=True)
    for h in hs: h.close()
for paths in ([0, 0], [2, 2], [0, 2]):
    run(10, False, 2, paths)
run(10, True, 2, [0, 2])
run(12, True, 3, [0, 0]); run(12, True, 3, [2, 2])
