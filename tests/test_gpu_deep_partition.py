"""The deep partition path of the long pattern lengths (W = 12, 14, 16: k_count_part -> k_count_slow -> k_part2 ->
k_hist_leaf) against the oracle, the dict-based restatement and the L2-atomic path, through the C ABI.  Bit-exact."""
import numpy as np
import pytest

import _cases
import _oracle as O
from test_kernel_logic_sim import long_period_sequences, palindrome_rich_sequences

pytestmark = pytest.mark.gpu


@pytest.fixture()
def deep(gpu):
    gpu.set_count_path(2)
    yield gpu
    gpu.set_count_path(0)


def check_sparse(gpu, seqs, W, both, path):
    """Tables too large to copy: the expected bins, random other bins, and the sum over all bins."""
    gpu.set_count_path(path)
    ltot = gpu.count(W, int(both))
    assert gpu.last_count_path == 2
    ref, lt = _cases.py_count_sparse(seqs, W, both)
    assert ltot == lt, (W, both, path)
    exp = {}
    for k, c in ref.items():
        exp[k] = c
        if both:
            exp[O.revcomp(k, W)] = c
    keys = np.array(sorted(exp), np.uint64)
    assert np.array_equal(gpu.gather_counts(keys), np.array([exp[int(k)] for k in keys], np.uint64)), (W, both, path)
    others = np.random.RandomState(W).randint(0, 4 ** W, 4096, dtype=np.int64).astype(np.uint64)
    others = others[~np.isin(others, keys)]
    assert not gpu.gather_counts(others).any()
    assert gpu.counts_total() == sum(exp.values()), (W, both, path)


@pytest.mark.parametrize("seed", range(6))
def test_small_sets_all_long_lengths(deep, seed):
    seqs = (_cases.dirty_sequences(seed), palindrome_rich_sequences(seed), long_period_sequences(seed))[seed % 3]
    codes, offs = O.encode(seqs)
    deep.ingest_codes(codes, offs)
    for W in (12, 14, 16):
        for both in (True, False):
            for path in (2, 3):
                check_sparse(deep, seqs, W, both, path)


@pytest.mark.parametrize("seed", range(4))
def test_w12_full_table_against_oracle(deep, seed):
    seqs = _cases.dirty_sequences(seed, n_max=40, len_max=3000) + long_period_sequences(seed)
    codes, offs = O.encode(seqs)
    deep.ingest_codes(codes, offs)
    for both in (True, False):
        ref, lt = O.count(codes, offs, 12, both)
        for path in (2, 3):
            deep.set_count_path(path)
            assert deep.count(12, int(both)) == lt and deep.last_count_path == 2
            assert np.array_equal(deep.get_counts(), ref), (both, path)


@pytest.mark.parametrize("kind", ["homopolymer", "tandem", "one_class", "skewed_random"])
def test_adversarial_queue_pressure(deep, kind):
    """Inputs that overflow the class queues of stage P and stage P2 (most ids in one class) or are slow units only."""
    rng = np.random.RandomState(5)
    if kind == "homopolymer":
        seqs = ["A" * 5000, "T" * 3000, "C" * 777]
    elif kind == "tandem":
        seqs = ["ACG" * 2000, "AACCGGTTACGTTGCA"[:15] * 900, "AT" * 1500, "ACGTTGCATGCAAG" * 300]
    elif kind == "one_class":
        seqs = ["".join("".join(rng.choice(list("CGT"), 9)) + "AAAAAAA" for _ in range(60)) for _ in range(300)]
    else:
        seqs = ["".join(rng.choice(list("ACGT"), 700, p=[0.85, 0.05, 0.05, 0.05])) for _ in range(500)]
    codes, offs = O.encode(seqs)
    deep.ingest_codes(codes, offs)
    for both in (True, False):
        ref, lt = O.count(codes, offs, 12, both)
        deep.set_count_path(2)
        assert deep.count(12, int(both)) == lt
        assert np.array_equal(deep.get_counts(), ref), (kind, both)
    deep.set_count_path(1)
    for W in (14, 16):
        for both in (True, False):
            deep.set_count_path(1)
            l1 = deep.count(W, int(both)); t1 = deep.counts_total()
            ids = np.random.RandomState(W).randint(0, 4 ** W, 1 << 16, dtype=np.int64).astype(np.uint64)
            # bins that are certainly occupied: the first W-mers of the sequences
            first = np.array([sum((int(c) - 1) << (2 * p) for p, c in enumerate(codes[int(o):int(o) + W]))
                              for o in offs[:-1] if (codes[int(o):int(o) + W] > 0).all()], np.uint64)
            ids = np.concatenate([ids, first])
            g1 = deep.gather_counts(ids)
            deep.set_count_path(2)
            l2 = deep.count(W, int(both)); t2 = deep.counts_total()
            assert deep.last_count_path == 2
            assert l1 == l2 and t1 == t2 and np.array_equal(g1, deep.gather_counts(ids)), (kind, W, both)


def test_paths_agree_on_synthetic(gpu, pkg):
    """BASELINE configs[4] shape at reduced size (+ a dirty variant): deep path == L2-RED variant == L2-atomic path,
    W = 12 against the oracle table, W = 14 / 16 on 2^20 random bins, the bins of the planted motif and the total."""
    for nfrac in (0.0, 0.01):
        gpu.ingest_synthetic(40_000, 300, 99, 0, "TGCTGASTCAGCA", 0.3, nfrac)
        codes, offs = gpu.export_codes()
        for both in (False, True):
            ref, lt = O.count(codes, offs, 12, both)
            for path in (0, 2, 3):
                gpu.set_count_path(path)
                assert gpu.count(12, int(both)) == lt
                assert gpu.last_count_path == (2 if path else 1)  # 12 M positions: below the W = 12 crossover of 2^25
                assert np.array_equal(gpu.get_counts(), ref), (both, path)
            for W in (14, 16):
                ids = np.random.RandomState(W).randint(0, 4 ** W, 1 << 20, dtype=np.int64).astype(np.uint64)
                first = np.array([sum((int(c) - 1) << (2 * p) for p, c in enumerate(codes[int(o):int(o) + W]))
                                  for o in offs[:2000] if (codes[int(o):int(o) + W] > 0).all()], np.uint64)
                ids = np.concatenate([ids, first])
                res = []
                for path in (1, 2, 3):
                    gpu.set_count_path(path)
                    lt_w = gpu.count(W, int(both))
                    assert gpu.last_count_path == (1 if path == 1 else 2)
                    res.append((lt_w, gpu.counts_total(), gpu.gather_counts(ids)))
                assert res[0][0] == res[1][0] == res[2][0] and res[0][1] == res[1][1] == res[2][1], (W, both)
                assert np.array_equal(res[0][2], res[1][2]) and np.array_equal(res[0][2], res[2][2]), (W, both)
                assert res[0][2][-10:].all()
    gpu.set_count_path(0)
