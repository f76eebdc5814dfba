"""BASELINE configs[3] code path: W = 10 on MORE than 2^30 stream positions per GPU (several chunks of the partition
path: per-chunk fold, final accumulating fold) and the 64-bit exchange that a 25 Gbp set needs (2.5e10 > 2^32).

The reference cannot count such a set in one piece (`unsigned int j`, src/base_pattern.cpp:335-336 of the reference), so
the expected table is SURVEY §8c's recipe: the sum of the oracle's tables over shards of <= 1 Gbp — exact, because
neither the non-overlap rule nor the resume rule ever crosses a sequence boundary (src/base_pattern.cpp:331-393).
Bit-exact, through the C ABI."""
import numpy as np
import pytest

import _oracle as O

pytestmark = pytest.mark.gpu
W = 10
MOTIF = "TGCTGASTCAGCA"
N_SEQ, L, SEED, NFRAC = 7_000_000, 200, 4242, 0.01  # 1.414e9 stream positions: two chunks of the partition path
N_RANK0 = 6_000_000                                  # "rank 0" of the exchange test holds > 2^30 positions by itself
ORACLE_SHARDS = 5


@pytest.fixture(scope="module")
def expected(gpu):
    """{both: (table, ltot)} = sum of the oracle over 5 shards of 280 Mbp (exported from the device generator)."""
    out = {}
    per = N_SEQ // ORACLE_SHARDS
    acc = {b: [np.zeros(4 ** W, np.uint64), 0] for b in (True, False)}
    for s in range(ORACLE_SHARDS):
        gpu.ingest_synthetic(per, L, SEED, s * per, MOTIF, 0.3, NFRAC)
        codes, offs = gpu.export_codes()
        for b in (True, False):
            c, lt = O.count(codes, offs, W, b)
            acc[b][0] += c
            acc[b][1] += lt
    for b in (True, False):
        out[b] = (acc[b][0], acc[b][1])
    return out


@pytest.mark.parametrize("both", [True, False], ids=["BOTH", "PLUS"])
def test_multichunk_count_matches_shard_sum(gpu, expected, both):
    gpu.set_count_path(0)
    gpu.ingest_synthetic(N_SEQ, L, SEED, 0, MOTIF, 0.3, NFRAC)
    assert N_SEQ * (L + 2) > 2 ** 30
    ltot = gpu.count(W, int(both))
    assert gpu.last_count_path == 2  # automatic choice: the partition path, in two chunks
    ref, lt = expected[both]
    assert ltot == lt
    assert np.array_equal(gpu.get_counts(), ref)


@pytest.mark.parametrize("both", [True, False], ids=["BOTH", "PLUS"])
def test_multichunk_u64_exchange_matches_shard_sum(gpu, expected, both):
    """count_local -> count_buffer (summed over two 'ranks' as ncclAllReduce would) -> count_finish; rank 0's shard runs
    in two chunks, rank 1's in one: the unmirrored canonical uint64 tables + ltot add up to the whole set's."""
    import torch

    gpu.set_count_path(0)
    gpu.ingest_synthetic(N_SEQ - N_RANK0, L, SEED, N_RANK0, MOTIF, 0.3, NFRAC)
    gpu.count_local(W, int(both))
    gpu.synchronize()
    part1 = torch.as_tensor(gpu.count_buffer(), device="cuda").clone()
    gpu.ingest_synthetic(N_RANK0, L, SEED, 0, MOTIF, 0.3, NFRAC)
    assert N_RANK0 * (L + 2) > 2 ** 30
    gpu.count_local(W, int(both))
    gpu.synchronize()
    buf = torch.as_tensor(gpu.count_buffer(), device="cuda")
    buf += part1  # what the all-reduce leaves on every rank
    torch.cuda.synchronize()
    ref, lt = expected[both]
    assert gpu.count_finish() == lt
    assert np.array_equal(gpu.get_counts(), ref)


@pytest.mark.parametrize("W_,both", [(2, True), (4, True), (8, False)])
def test_small_W_more_than_2_31_positions(gpu, W_, both):
    """W <= 8 counts into CTA-private shared-memory bins that are merged into the uint32 table: with few bins one of them
    can pass 2^32 on a multi-Gbp set, so the launches are cut at 2^31 positions with a fold between them.  11 M x 200 bp =
    2.2e9 positions (two launches) against the sum of three single-launch shards."""
    n = 11_100_000
    gpu.ingest_synthetic(n, 200, 11, 0, MOTIF, 0.3, 0.0)
    ltot = gpu.count(W_, int(both))
    whole = gpu.get_counts()
    acc, lt = np.zeros(4 ** W_, np.uint64), 0
    per = n // 3
    for s in range(3):
        gpu.ingest_synthetic(per, 200, 11, s * per, MOTIF, 0.3, 0.0)
        lt += gpu.count(W_, int(both))
        acc += gpu.get_counts()
    assert ltot == lt == n * (200 - W_ + 1)
    assert np.array_equal(whole, acc)
