import os, subprocess, sys, tempfile, time
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tools')
import numpy as np
import _pkg
from e2e_cli_large import write_fasta, MOTIF
pkg = _pkg.load()
g = pkg.PengCuda(0)
g.ingest_synthetic(1_000_000, 200, 1234, 0, MOTIF, 0.3, 0.0)
codes, _ = g.export_codes(); g.close()
d = tempfile.mkdtemp()
fa = os.path.join(d, 'set.fa'); write_fasta(codes, 1_000_000, 200, fa)
ours = '/root/repo/peng-motif_b200/host/peng_motif'
for env in ({}, {'PENG_FASTA_SERIAL': '1'}):
    for i in range(3):
        t0 = time.perf_counter()
        r = subprocess.run([ours, fa, '-w', '10', '-o', os.path.join(d, 'o.meme'), '--timings'], capture_output=True, text=True, env={**os.environ, 'PENG_INGEST_DEBUG': '1', **env})
        print(env, round(time.perf_counter() - t0, 3), [l for l in r.stderr.splitlines() if 'ingest' in l])
