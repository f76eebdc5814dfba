"""The fused cross-GPU sum + fold + mirror over peer memory (peng_cuda_peer_* / peng_cuda_count_exchange32,
k_peer_exchange32) on ONE GPU: a peer group of one rank, and peer groups of two and three ranks whose handles live in this
process on the same device (their kernels meet through the same flag words and peer pointers as over NVLink; real
multi-GPU runs are checked by `bench.py --gpus N`, which compares the exchange with the NCCL path bit for bit).
Expected tables: the oracle on the whole set.  Bit-exact."""
import numpy as np
import pytest

import _oracle as O

pytestmark = pytest.mark.gpu
MOTIF = "TGCTGASTCAGCA"


def expected(gpu, n, W, both, seed):
    gpu.ingest_synthetic(n, 200, seed, 0, MOTIF, 0.3, 0.01)
    codes, offs = gpu.export_codes()
    return O.count(codes, offs, W, both)


@pytest.mark.parametrize("W,both", [(8, True), (10, True), (10, False), (12, True), (12, False)])
def test_group_of_one(pkg, gpu, W, both):
    ref, lt = expected(gpu, 20_000, W, both, 31)
    g = pkg.PengCuda(0)
    try:
        g.peer_attach(0, 1, g.peer_export(W))
        for path in (0, 2):  # canonical bins (L2 / shared-memory kernels) and forward-id bins (partition paths)
            g.set_count_path(path)
            g.ingest_synthetic(20_000, 200, 31, 0, MOTIF, 0.3, 0.01)
            g.count_local32(W, int(both))
            assert g.count_exchange32() == lt
            assert np.array_equal(g.get_counts(), ref), (W, both, path)
        with pytest.raises(pkg.PengCudaError):
            g.count(W + 2, int(both))  # the tables are shared with the group: detach first
        g.peer_detach()
        g.count(W + 2, int(both))
    finally:
        g.close()


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("W,both", [(10, True), (10, False), (12, True)])
def test_ranks_in_one_process(pkg, gpu, W, both, world):
    """`world` handles on device 0, each with its shard; every rank ends with the whole set's finished table and ltot.
    The exchanges are launched without waiting (ltot = NULL) on the handles' own streams and meet on the device."""
    n = 30_000
    ref, lt = expected(gpu, n, W, both, 32)
    hs = [pkg.PengCuda(0) for _ in range(world)]
    try:
        blobs = np.stack([h.peer_export(W) for h in hs])
        for r, h in enumerate(hs):
            h.peer_attach(r, world, blobs)
        for step in range(3):  # the flag epochs advance: several exchanges in a row
            cuts = [n * r // world for r in range(world + 1)]
            for r, h in enumerate(hs):
                h.set_count_path(2 if (r + step) % 2 else 0)  # mixed bin conventions in one sum
                h.ingest_synthetic(cuts[r + 1] - cuts[r], 200, 32, cuts[r], MOTIF, 0.3, 0.01)
                h.count_local32(W, int(both))
            for h in hs:
                h.count_exchange32(want_ltot=False)
            for h in hs:
                assert h.count_finish() == lt
                assert np.array_equal(h.get_counts(), ref), (W, both, world, step)
    finally:
        for h in hs:
            h.close()


def test_missing_rank_times_out(pkg, monkeypatch):
    """A rank that never calls the exchange must not hang the others: the wait gives up and the finish reports it."""
    monkeypatch.setenv("PENG_PEER_TIMEOUT_MS", "500")  # read by peng_cuda_create (default 20 s)
    hs = [pkg.PengCuda(0) for _ in range(2)]
    try:
        blobs = np.stack([h.peer_export(10) for h in hs])
        for r, h in enumerate(hs):
            h.peer_attach(r, 2, blobs)
            h.ingest_synthetic(5_000, 200, 1, r * 5_000, MOTIF, 0.3, 0.0)
        hs[0].count_local32(10, 1)
        with pytest.raises(pkg.PengCudaError) as e:
            hs[0].count_exchange32()
        assert "did not arrive" in str(e.value)
    finally:
        for h in hs:
            h.close()
