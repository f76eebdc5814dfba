"""The partition path of the W = 10 counter (k_count_part -> k_count_slow -> k_hist_part -> k_fold_canon) against the
oracle and against the L2-atomic path, through the C ABI.  Bit-exact, as for every count."""
import os

import numpy as np
import pytest

import _cases
import _oracle as O
from test_kernel_logic_sim import palindrome_rich_sequences

pytestmark = pytest.mark.gpu
DATA = os.path.join(_cases.GOLDEN, "data")
W = 10


@pytest.fixture()
def part(gpu):
    gpu.set_count_path(2)
    yield gpu
    gpu.set_count_path(0)


def check(gpu, codes, offs):
    gpu.ingest_codes(codes, offs)
    for both in (True, False):
        ltot = gpu.count(W, int(both))
        assert gpu.last_count_path == 2
        ref, lt = O.count(codes, offs, W, both)
        assert ltot == lt, both
        assert np.array_equal(gpu.get_counts(), ref), both


@pytest.mark.parametrize("seed", range(16))
def test_dirty_and_palindrome_rich(part, seed):
    seqs = _cases.dirty_sequences(seed) if seed % 2 else palindrome_rich_sequences(seed)
    check(part, *O.encode(seqs))


def test_mafk(part):
    codes, offs, _, _ = O.read_fasta(os.path.join(DATA, "MafK.fasta"))
    check(part, codes, offs)


def test_empty_short_and_invalid(part):
    check(part, *O.encode(["A", "ACGTACGTA", "ACGTACGTAC", "ACGTACGTACG", "N" * 40, "ACGTNACGTACGTACGTAC"]))


@pytest.mark.parametrize("kind", ["homopolymer", "tandem", "one_class", "skewed_random"])
def test_adversarial_queue_pressure(part, kind):
    """Inputs that overflow a class queue (every W-mer of a round in one id range) or consist of slow units only."""
    rng = np.random.RandomState(5)
    if kind == "homopolymer":
        seqs = ["A" * 5000, "T" * 3000, "C" * 777]
    elif kind == "tandem":
        seqs = ["ACG" * 2000, "AACCGGTT" * 900, "AT" * 1500]
    elif kind == "one_class":
        # all W-mers end in 'AA' (digits 8, 9 = A): a single class takes every key of every round; the varying
        # prefix keeps most of them out of the conflict candidates
        seqs = []
        for _ in range(300):
            s = []
            for _ in range(40):
                s.append("".join(rng.choice(list("CGT"), 8)) + "AA")
            seqs.append("".join(s))
    else:
        seqs = ["".join(rng.choice(list("ACGT"), 400, p=[0.85, 0.05, 0.05, 0.05])) for _ in range(400)]
    check(part, *O.encode(seqs))


def test_paths_agree_on_synthetic(gpu, pkg):
    """BASELINE configs[2] shape at reduced size + a dirty variant: partition path == L2-atomic path == oracle."""
    for nfrac in (0.0, 0.01):
        gpu.ingest_synthetic(60_000, 200, 99, 0, "TGCTGASTCAGCA", 0.3, nfrac)
        codes, offs = gpu.export_codes()
        for both in (True, False):
            gpu.set_count_path(1)
            l1 = gpu.count(W, int(both)); c1 = gpu.get_counts()
            assert gpu.last_count_path == 1
            gpu.set_count_path(2)
            l2 = gpu.count(W, int(both)); c2 = gpu.get_counts()
            assert gpu.last_count_path == 2
            gpu.set_count_path(0)
            l0 = gpu.count(W, int(both)); c0 = gpu.get_counts()
            assert gpu.last_count_path == 2  # 12 M positions: automatic choice is the partition path
            ref, lt = O.count(codes, offs, W, both)
            assert l1 == l2 == l0 == lt
            assert np.array_equal(c1, ref) and np.array_equal(c2, ref) and np.array_equal(c0, ref)


def test_sharded_sum_matches_whole(part, pkg):
    """count_local leaves canonical (unmirrored) bins: the sum over shards + finish equals the whole-set table."""
    import torch

    part.ingest_synthetic(40_000, 200, 7, 0, "TGCTGASTCAGCA", 0.3, 0.0)
    ltot = part.count(W, pkg.STRAND_BOTH)
    whole = part.get_counts()
    acc, lt = np.zeros(4 ** W + 1, np.uint64), 0
    for r in range(4):
        part.ingest_synthetic(10_000, 200, 7, r * 10_000, "TGCTGASTCAGCA", 0.3, 0.0)
        part.count_local(W, pkg.STRAND_BOTH)
        part.synchronize()
        acc += torch.as_tensor(part.count_buffer(), device="cuda").cpu().numpy().view(np.uint64)
    assert int(acc[-1]) == ltot
    rc = np.array([O.revcomp(i, W) for i in range(0, 4 ** W, 4 ** W // 64)])  # spot check of the mirror relation
    ids = np.arange(0, 4 ** W, 4 ** W // 64)
    canon = np.minimum(ids, rc)
    assert np.array_equal(acc[:-1][canon], whole[ids])


def test_composition_skew_at_scale(gpu, pkg):
    """An AT-rich set (and an extreme one) large enough for the automatic partition paths: the classes are far from
    equally filled, bucket regions of the heavy classes may run full (their blocks then go straight to the table)."""
    rng = np.random.RandomState(17)
    for p in ([0.32, 0.18, 0.18, 0.32], [0.7, 0.1, 0.1, 0.1]):
        n, L = 120_000, 200
        codes = (rng.choice(4, n * L, p=p) + 1).astype(np.uint8)
        offs = (np.arange(n + 1) * L).astype(np.uint64)
        gpu.ingest_codes(codes, offs)
        for Wt, both in ((10, True), (10, False), (12, False)):
            gpu.set_count_path(1)
            l1 = gpu.count(Wt, int(both)); c1 = gpu.get_counts()
            gpu.set_count_path(2)
            l2 = gpu.count(Wt, int(both)); c2 = gpu.get_counts()
            assert gpu.last_count_path == 2 and l1 == l2 and np.array_equal(c1, c2), (p, Wt, both)
    gpu.set_count_path(0)
    ref, lt = O.count(codes, offs, 10, True)
    assert gpu.count(10, 1) == lt and np.array_equal(gpu.get_counts(), ref)


@pytest.mark.parametrize("shard_path", [0, 2], ids=["auto", "partition"])
@pytest.mark.parametrize("W_,both,n", [(10, True, 40_000), (10, False, 40_000), (8, True, 4_000), (12, True, 4_000), (12, False, 4_000),
                                       (14, True, 100_000), (14, False, 100_000)])
def test_sharded_sum32_matches_whole(gpu, pkg, W_, both, n, shard_path):
    """The 32-bit exchange: unfolded uint32 tables (+ ltot in word 4^W) of 4 shards summed as the all-reduce would, written
    back, folded once by count_finish32 — equals the whole-set table.  shard_path 0: the shards are small and take the
    L2-atomic / shared-memory kernels (canonical bins); shard_path 2: the shards are FORCED onto the partition / deep
    partition paths (fast W-mers in their forward-id bin), the combination bench.py times at N > 1."""
    import torch

    gpu.set_count_path(0)
    gpu.ingest_synthetic(n, 200, 7, 0, "TGCTGASTCAGCA", 0.3, 0.01)
    ltot = gpu.count(W_, int(both))
    ids = np.random.RandomState(1).randint(0, 4 ** W_, 1 << 16, dtype=np.int64).astype(np.uint64)
    whole, total = gpu.gather_counts(ids), gpu.counts_total()
    acc = None
    gpu.set_count_path(shard_path)
    try:
        for r in range(4):
            gpu.ingest_synthetic(n // 4, 200, 7, r * (n // 4), "TGCTGASTCAGCA", 0.3, 0.01)
            gpu.count_local32(W_, int(both))
            if shard_path == 2:
                assert gpu.last_count_path == (2 if W_ >= 10 else 1)
            gpu.synchronize()  # the handle works on its own stream, torch on another
            buf = torch.as_tensor(gpu.count_buffer32(), device="cuda")
            acc = buf.clone() if acc is None else acc + buf
            torch.cuda.synchronize()  # torch's stream has read the table before the handle's stream reuses it
            with pytest.raises(pkg.PengCudaError):
                gpu.count_finish()  # the 64-bit finish must refuse an unfolded table
    finally:
        gpu.set_count_path(0)
    buf.copy_(acc)  # what the all-reduce leaves on every rank
    torch.cuda.synchronize()
    assert gpu.count_finish32() == ltot
    assert np.array_equal(gpu.gather_counts(ids), whole) and gpu.counts_total() == total


@pytest.mark.parametrize("W_,last_on_partition", [(10, False), (10, True), (12, False), (12, True)])
def test_sum32_of_mixed_paths(gpu, pkg, W_, last_on_partition):
    """Ranks choose their counting path from their OWN shard size, so base-balanced shards that straddle a threshold mix
    the two bin conventions in one summed table (partition paths: fast W-mers in the forward-id bin; L2 / shared-memory
    kernels: everything in the canonical bin).  count_finish32 must fold such a sum correctly whichever path the
    finishing rank took itself: one partition-path shard + one L2-path shard, BOTH strands, against the oracle."""
    import torch

    n = 6_000
    gpu.ingest_synthetic(2 * n, 200, 21, 0, "TGCTGASTCAGCA", 0.3, 0.01)
    codes, offs = gpu.export_codes()
    ref, lt = O.count(codes, offs, W_, True)
    order = [(0, 2), (1, 1)] if not last_on_partition else [(1, 1), (0, 2)]  # (shard, path); the last one finishes
    acc = None
    try:
        for shard, path in order:
            gpu.set_count_path(path)
            gpu.ingest_synthetic(n, 200, 21, shard * n, "TGCTGASTCAGCA", 0.3, 0.01)
            gpu.count_local32(W_, pkg.STRAND_BOTH)
            assert gpu.last_count_path == path
            gpu.synchronize()
            buf = torch.as_tensor(gpu.count_buffer32(), device="cuda")
            acc = buf.clone() if acc is None else acc + buf
            torch.cuda.synchronize()  # torch's stream has read the table before the handle's stream reuses it
    finally:
        gpu.set_count_path(0)
    buf.copy_(acc)
    torch.cuda.synchronize()
    assert gpu.count_finish32() == lt
    assert np.array_equal(gpu.get_counts(), ref)
