"""The N > 1 path on the CPU: world_size-2 gloo run of the multi-GPU protocol —
shard sequences by bases, build per-rank UNMIRRORED histograms, one all_reduce(sum) of histogram + ltot,
mirror afterwards — with the kernel bodies standing in for the GPU (tests/sim).  The result must be
bit-identical to the single-rank table."""
import os
import sys

import numpy as np
import pytest

import _cases
import _oracle as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, W, both, q):
    import torch
    import torch.distributed as dist

    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _pkg
    import _sim as S

    pkg = _pkg.load()
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    codes, offs, _, _ = O.read_fasta(os.path.join(_cases.GOLDEN, "data", "MafK_100seqs.fasta"))
    first, last = pkg.shard_sequences(offs, rank, world)
    sub_offs = (offs[first:last + 1] - offs[first]).astype(np.uint64)
    sub_codes = codes[int(offs[first]):int(offs[last])]
    st = S.pack(sub_codes, sub_offs)
    hist, ltot = S.count(st, W, both, 4, mirror=False)
    bgc = S.bg_counts(st, 2)
    buf = torch.from_numpy(np.concatenate([hist, [np.uint64(ltot)], bgc]).astype(np.int64))
    dist.all_reduce(buf)  # the one collective of the path
    tot = buf.numpy().astype(np.uint64)
    counts = tot[: 4 ** W].copy()
    if both:
        S.lib.sim_mirror(counts.ctypes.data_as(S.u64p), W)
    if rank == 0:
        q.put((counts, int(tot[4 ** W]), tot[4 ** W + 1:].copy(), (first, last)))
    dist.destroy_process_group()


@pytest.mark.parametrize("W,both", [(8, True), (6, False)])
def test_two_rank_histogram_allreduce(W, both):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000) + W
    procs = [ctx.Process(target=_worker, args=(r, 2, port, W, both, q)) for r in range(2)]
    for p in procs:
        p.start()
    counts, ltot, bgc, shard0 = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    codes, offs, _, _ = O.read_fasta(os.path.join(_cases.GOLDEN, "data", "MafK_100seqs.fasta"))
    ref, ref_ltot = O.count(codes, offs, W, both)
    assert ltot == ref_ltot and np.array_equal(counts, ref)
    _, n = O.f32.background_model(codes, offs)
    assert np.array_equal(bgc.astype(np.int64), n)
    assert 0 < shard0[1] < len(offs) - 1  # rank 0 really owned a strict subset


def test_shard_sequences_partition(pkg):
    rng = np.random.RandomState(0)
    for world in (1, 2, 3, 8):
        lens = rng.randint(1, 500, 1000)
        offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
        cuts = [pkg.shard_sequences(offs, r, world) for r in range(world)]
        assert cuts[0][0] == 0 and cuts[-1][1] == 1000
        for a, b in zip(cuts, cuts[1:]):
            assert a[1] == b[0]
        bases = [int(offs[l] - offs[f]) for f, l in cuts]
        assert max(bases) - min(bases) <= 2 * 500
    assert pkg.shard_synthetic(10, 0, 3) == (0, 3) and pkg.shard_synthetic(10, 2, 3) == (6, 4)
