"""peng_cuda_ingest_codes from host buffers: raw bytes + device packing, host packing (host_pack.cpp) and the hybrid of the
two must leave bit-identical device buffers — checked through everything that reads them: export, k-mer counts (codes +
validity), background counts (sequence starts)."""
import os

import numpy as np
import pytest

import _cases
import _oracle as O

pytestmark = pytest.mark.gpu


@pytest.fixture()
def g(gpu):
    yield gpu
    gpu.set_ingest_mode(0, -1.0)


def fingerprint(gpu, codes, offs, W=8):
    c2, o2 = gpu.export_codes()
    assert np.array_equal(o2, offs) and np.array_equal(c2, np.where(codes > 4, 0, codes))
    ltot = gpu.count(W, 1)
    counts = gpu.get_counts()
    _, bg = gpu.background_model(2, return_counts=True)
    return ltot, counts, bg


def check_modes(gpu, codes, offs, modes):
    gpu.set_ingest_mode(1, -1.0)
    gpu.ingest_codes(codes, offs)
    assert gpu.last_ingest_host_packed == 0
    ref = fingerprint(gpu, codes, offs)
    r2, lt = O.count(codes, offs, 8, True)
    assert ref[0] == lt and np.array_equal(ref[1], r2)
    for mode, frac in modes:
        gpu.set_ingest_mode(mode, frac)
        gpu.ingest_codes(codes, offs)
        if mode == 2 or frac >= 0.5:
            assert gpu.last_ingest_host_packed > 0, (mode, frac)
        got = fingerprint(gpu, codes, offs)
        assert got[0] == ref[0] and np.array_equal(got[1], ref[1]) and np.array_equal(got[2], ref[2]), (mode, frac)


@pytest.mark.parametrize("seed", range(6))
def test_small_dirty_sets(g, seed):
    codes, offs = O.encode(_cases.dirty_sequences(seed, n_max=30, len_max=900))
    check_modes(g, codes, offs, [(2, -1.0), (0, 0.5), (0, 0.05), (0, 0.97)])


def test_ragged_lengths_many_phases(g):
    rng = np.random.RandomState(11)
    lens = rng.randint(1, 600, 400_000)  # 120 M bases: 3 staging phases when packed entirely on the host
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    codes = rng.randint(0, 5, int(offs[-1])).astype(np.uint8)
    check_modes(g, codes, offs, [(2, -1.0), (0, 0.77), (0, 0.31), (0, -1.0)])


def test_fixed_length_set_and_automatic_choice(g, pkg):
    g.ingest_synthetic(300_000, 200, 5, 0, "TGCTGASTCAGCA", 0.3, 0.01)
    codes, offs = g.export_codes()
    check_modes(g, codes, offs, [(2, -1.0), (0, 0.6)])
    g.set_ingest_mode(0, -1.0)  # automatic: hybrid with >= 4 host threads and AVX-512, raw otherwise — same buffers either way
    g.ingest_codes(codes, offs)
    assert g.count(10, 1) > 0
    ref, lt = O.count(codes, offs, 10, True)
    assert np.array_equal(g.get_counts(), ref)


def test_implied_validity(g):
    """Phases without letters other than ACGT send no validity words (k_valid_implied rebuilds them from the offsets, the
    listed blocks — share boundaries, blocks holding another letter — are written over them): a clean set, a set with rare
    other letters, and a set whose later phases are dirty (those send their validity words as before)."""
    rng = np.random.RandomState(21)
    lens = rng.randint(1, 600, 400_000)  # 120 M bases: 3 staging phases
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    total = int(offs[-1])
    clean = rng.randint(1, 5, total).astype(np.uint8)
    check_modes(g, clean, offs, [(2, -1.0), (0, 0.7)])
    g.set_ingest_mode(2, -1.0)
    g.ingest_codes(clean, offs)
    packed_all = g.last_ingest_h2d_bytes
    assert packed_all < 0.26 * total + offs.nbytes + (1 << 20), packed_all  # 2-bit codes + offsets + a few listed blocks
    rare = clean.copy()
    rare[rng.choice(total, total // 20_000, replace=False)] = 0
    check_modes(g, rare, offs, [(2, -1.0), (0, 0.4)])
    g.set_ingest_mode(2, -1.0)
    g.ingest_codes(rare, offs)
    assert packed_all < g.last_ingest_h2d_bytes < packed_all + 32 * (total // 20_000)
    late = clean.copy()
    late[int(0.6 * total):] = rng.randint(0, 5, total - int(0.6 * total))
    check_modes(g, late, offs, [(2, -1.0), (0, 0.9)])
    g.set_ingest_mode(2, -1.0)
    g.ingest_codes(late, offs)
    assert 0.26 * total < g.last_ingest_h2d_bytes - offs.nbytes < 0.375 * total


def test_fasta_parallel_equals_streaming(g, monkeypatch):
    """peng_cuda_ingest_fasta: the parallel path (file cut at header lines, records read in place from the mapping by the
    host packer, wrapped records joined in an arena) against the streaming path (PENG_FASTA_SERIAL=1: one scanning thread,
    raw letters to the device) and the oracle's reader — a 3 MB file of single-line and wrapped records, lower case, N, CR
    line ends, blank lines, records without sequence; whole and as shards."""
    from test_kernel_logic_sim import _random_fasta

    rng = np.random.RandomState(77)
    text = _random_fasta(rng, 20_000)
    path = _cases.write_fasta(text)
    try:
        assert os.path.getsize(path) > (2 << 20)  # 12 chunks
        codes, offs, mn, mx = O.read_fasta(path)
        info = g.ingest_fasta(path)
        assert g.last_ingest_host_packed == len(offs) - 1  # the parallel path packs on the host
        assert info == dict(n_seq=len(offs) - 1, total_bases=len(codes), min_len=mn, max_len=mx)
        c, o = g.export_codes()
        assert np.array_equal(o, offs) and np.array_equal(c, codes)
        par = fingerprint(g, codes, offs)
        for shard_parallel in (True, False):  # shards default to the streaming reader (one host thread per device calls them)
            if shard_parallel:
                monkeypatch.setenv("PENG_FASTA_SHARD_PARALLEL", "1")
            else:
                monkeypatch.delenv("PENG_FASTA_SHARD_PARALLEL")
            shards = []
            for p in range(3):
                g.ingest_fasta_shard(path, p, 3)
                assert shard_parallel or g.last_ingest_host_packed == 0
                shards.append(g.export_codes())
            assert np.array_equal(np.concatenate([s[0] for s in shards]), codes)
            assert np.array_equal(np.concatenate([np.diff(s[1]) for s in shards]), np.diff(offs))
        monkeypatch.setenv("PENG_FASTA_SERIAL", "1")
        info2 = g.ingest_fasta(path)
        assert g.last_ingest_host_packed == 0 and info2 == info
        ser = fingerprint(g, codes, offs)
        assert par[0] == ser[0] and np.array_equal(par[1], ser[1]) and np.array_equal(par[2], ser[2])
    finally:
        os.unlink(path)


def test_long_sequence_falls_back(g):
    rng = np.random.RandomState(2)
    offs = np.array([0, 1000, 50_001_000, 50_002_000], np.uint64)  # 50 M letters in one record: more than a staging phase
    codes = rng.randint(1, 5, int(offs[-1])).astype(np.uint8)
    g.set_ingest_mode(2, -1.0)
    g.ingest_codes(codes, offs)
    assert g.last_ingest_host_packed == 0  # plain path
    c2, _ = g.export_codes()
    assert np.array_equal(c2, codes)


def chain_sequences(seed):
    """Runs of exactly-W-long segments separated by single Ns (the parity chain of the k-mer start rule walks back through
    them), tandem repeats and ordinary sequence, in lengths that put sequence boundaries anywhere in a 64-position block."""
    rng = np.random.RandomState(seed)
    seqs = []
    for _ in range(rng.randint(40, 120)):
        kind = rng.randint(0, 4)
        if kind == 0:
            seqs.append("N".join("".join(rng.choice(list("ACGT"), 10)) for _ in range(rng.randint(1, 30))))
        elif kind == 1:
            u = "".join(rng.choice(list("ACGT"), rng.randint(1, 7)))
            seqs.append((u * 60)[: rng.randint(10, 300)])
        else:
            seqs.append("".join(rng.choice(list("ACGT"), rng.randint(1, 400))))
    return seqs


@pytest.mark.parametrize("seed", range(6))
def test_count_prefetch_small_sets(g, pkg, seed):
    """peng_cuda_set_count_prefetch: stage P of the W = 10 partition path runs INSIDE ingest_codes, unit range by unit range
    (the raw tail first, then every packed phase, then the units around the split) — the table equals the oracle's for every
    split position, both strands, and a following count with other parameters is not disturbed."""
    codes, offs = O.encode(chain_sequences(seed))
    g.set_count_path(2)
    try:
        for both in (True, False):
            ref, lt = O.count(codes, offs, 10, both)
            for mode, frac in ((0, 0.5), (0, 0.07), (0, 0.93), (2, -1.0), (0, 0.31)):
                g.set_ingest_mode(mode, frac)
                g.set_count_prefetch(10, int(both))
                g.ingest_codes(codes, offs)
                assert g.count(10, int(both)) == lt and g.last_count_path == 2
                assert np.array_equal(g.get_counts(), ref), (seed, both, mode, frac)
                assert np.array_equal(g.get_counts32(), ref.astype(np.uint32))
                assert g.count(10, int(both)) == lt and np.array_equal(g.get_counts(), ref)  # second count: from scratch
                g.ingest_codes(codes, offs)  # armed for `both`, counted with the other strand and another W: prefetch dropped
                r2, l2 = O.count(codes, offs, 10, not both)
                assert g.count(10, int(not both)) == l2 and np.array_equal(g.get_counts(), r2)
                g.ingest_codes(codes, offs)
                r3, l3 = O.count(codes, offs, 8, both)
                assert g.count(8, int(both)) == l3 and np.array_equal(g.get_counts(), r3)
    finally:
        g.set_count_prefetch(0)
        g.set_count_path(0)


def test_count_prefetch_many_phases(g, pkg):
    """Ragged 120 M bases: several staging phases, tapered tail, automatic partition path; prefetch == plain count == oracle."""
    rng = np.random.RandomState(12)
    lens = rng.randint(1, 600, 400_000)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    codes = rng.randint(0, 5, int(offs[-1])).astype(np.uint8)
    ref, lt = O.count(codes, offs, 10, True)
    try:
        for mode, frac in ((0, -1.0), (0, 0.77), (2, -1.0), (1, -1.0)):
            g.set_ingest_mode(mode, frac)
            g.set_count_prefetch(10, 1)
            g.ingest_codes(codes, offs)
            assert g.count(10, 1) == lt
            assert np.array_equal(g.get_counts(), ref), (mode, frac)
    finally:
        g.set_count_prefetch(0)


@pytest.mark.parametrize("seed", range(3))
def test_ingest_sequences_per_sequence_arrays(g, seed):
    """peng_cuda_ingest_sequences: one pageable array per sequence (what Sequence::getSequence() / getL() give) — the same
    device buffers as the concatenated ingest, in every mode (host packing from the pointers, gather + plain path)."""
    seqs = _cases.dirty_sequences(seed, n_max=60, len_max=700)
    codes, offs = O.encode(seqs)
    arrays = [np.ascontiguousarray(codes[int(offs[i]):int(offs[i + 1])].copy()) for i in range(len(seqs))]
    rng = np.random.RandomState(seed)
    keep = [arrays[i] for i in rng.permutation(len(arrays))]  # separate allocations, not one buffer in order
    ptrs = np.array([a.ctypes.data for a in arrays], np.uint64)
    lens = np.array([len(a) for a in arrays], np.uint32)
    g.set_ingest_mode(1, -1.0)
    g.ingest_codes(codes, offs)
    ref = fingerprint(g, codes, offs)
    for mode in (0, 1, 2):
        g.set_ingest_mode(mode, -1.0)
        g.ingest_sequences(ptrs, lens)
        got = fingerprint(g, codes, offs)
        assert got[0] == ref[0] and np.array_equal(got[1], ref[1]) and np.array_equal(got[2], ref[2]), mode
    del keep


def test_nibble_staging_opt_in(pkg, monkeypatch):
    """PENG_NIBBLE=1: the raw part of the hybrid ingest crosses PCIe as nibbles (host: nibble_pack, device: k_pack<2>) —
    same device buffers as the raw-byte path, for several splits (codes above 4: tests/test_kernel_logic_sim.py)."""
    monkeypatch.setenv("PENG_NIBBLE", "1")
    h = pkg.PengCuda(0)
    try:
        rng = np.random.RandomState(3)
        lens = rng.randint(1, 600, 200_000)
        offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
        codes = rng.randint(0, 5, int(offs[-1])).astype(np.uint8)
        check_modes(h, codes, offs, [(0, 0.5), (0, 0.05), (0, 0.93), (0, -1.0)])
    finally:
        h.close()
