"""Tiled fold / mirror of the long pattern lengths (k_fold_canon_tiled, k_mirror_tiled: W >= 12) against the
one-thread-per-id kernels (PENG_FOLD_UNTILED=1), through the C ABI: whole tables at W = 12, sampled bins + totals at
W = 14, single-chunk (written table, fused and separate mirror) and multi-chunk (accumulated table) sets.  Bit-exact."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture()
def pair(pkg):
    """(tiled, untiled) handles; the switch is read when a handle is created."""
    a = pkg.PengCuda(0)
    os.environ["PENG_FOLD_UNTILED"] = "1"
    try:
        b = pkg.PengCuda(0)
    finally:
        del os.environ["PENG_FOLD_UNTILED"]
    yield a, b
    a.close()
    b.close()


def ingest_both(pair, n_seq, seq_len, n_fraction=0.0):
    for g in pair:
        g.ingest_synthetic(n_seq, seq_len, 4321, 0, "TGCTGASTCAGCA", 0.3, n_fraction)
        g.set_count_path(2)


@pytest.mark.parametrize("n_fraction", [0.0, 0.02])
def test_single_chunk_w12_whole_table(pair, n_fraction):
    a, b = pair
    ingest_both(pair, 150_000, 300, n_fraction)
    for strand in (1, 0):  # BOTH, PLUS (PLUS never takes the pair fold: the two handles must agree anyway)
        la, lb = a.count(12, strand), b.count(12, strand)
        assert a.last_count_path == 2 and b.last_count_path == 2
        ta, tb = a.get_counts(), b.get_counts()
        assert la == lb and np.array_equal(ta, tb), strand
        if strand == 1:
            assert ta.sum() > la  # mirrored: the pair counts appear at both ids
        # separate mirror pass (what a rank runs after the all-reduce): local + finish == count
        a.count_local(12, strand)
        assert a.count_finish() == la and np.array_equal(a.get_counts(), ta), strand
        b.count_local(12, strand)
        assert b.count_finish() == lb and np.array_equal(b.get_counts(), ta), strand


def test_single_chunk_w14_sampled(pair):
    a, b = pair
    ingest_both(pair, 150_000, 300)
    codes, offs = a.export_codes()
    ids = np.random.RandomState(14).randint(0, 4 ** 14, 1 << 20, dtype=np.int64).astype(np.uint64)
    first = np.array([sum((int(c) - 1) << (2 * p) for p, c in enumerate(codes[int(o):int(o) + 14])) for o in offs[:3000]
                      if (codes[int(o):int(o) + 14] > 0).all()], np.uint64)
    ids = np.concatenate([ids, first])
    for fused in (True, False):
        res = []
        for g in (a, b):
            if fused:
                lt = g.count(14, 1)
            else:
                g.count_local(14, 1)
                lt = g.count_finish()
            res.append((lt, g.counts_total(), g.gather_counts(ids)))
        assert res[0][0] == res[1][0] and res[0][1] == res[1][1], fused
        assert np.array_equal(res[0][2], res[1][2]) and res[0][2][-10:].all(), fused


@pytest.mark.parametrize("n_seq", [4_000_000, 14_400_000])
def test_multi_chunk_w12(pair, n_seq):
    """1.2 G positions = two chunks folded once at the end (the table is written, mirror fused); 4.35 G positions > 2^32:
    the deep path folds before a uint32 bin could overflow, that fold and the last one ADD to the 64-bit table."""
    a, b = pair
    ingest_both(pair, n_seq, 300)
    la, lb = a.count(12, 1), b.count(12, 1)
    assert a.last_count_path == 2 and la == lb == n_seq * 289
    ta = a.get_counts()
    assert np.array_equal(ta, b.get_counts())
    assert int(ta.sum()) > la
    ids = np.random.RandomState(1).randint(0, 4 ** 12, 4096, dtype=np.int64).astype(np.uint64)
    rc = np.array([sum((3 - ((int(i) >> (2 * p)) & 3)) << (2 * (11 - p)) for p in range(12)) for i in ids], np.uint64)
    assert np.array_equal(ta[ids.astype(np.int64)], ta[rc.astype(np.int64)])  # mirrored
    # PLUS: the written (first) fold without pairs
    assert a.count(12, 0) == b.count(12, 0) and np.array_equal(a.get_counts(), b.get_counts())
