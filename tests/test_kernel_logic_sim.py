"""Kernel LOGIC on the CPU: the per-thread bodies of the CUDA kernels (peng-motif_b200/csrc/core_*.cuh) are
compiled as host code by tests/sim and driven thread by thread with the launch geometry of the real
kernels, then compared with the oracle.  This is a GPU-less debugging aid; the parity tests proper are the
`-m gpu` suite, which goes through the C ABI on a B200."""
import os
from collections import Counter

import numpy as np
import pytest

import _cases
import _oracle as O
import _sim as S

DATA = os.path.join(_cases.GOLDEN, "data")


@pytest.mark.parametrize("case", [c for c in _cases.golden_cases() if c["exit"] == 0], ids=lambda c: c["tag"])
def test_count_golden_cases(case):
    path = _cases.write_fasta(case["fasta"])
    codes, offs, _, _ = O.read_fasta(path)
    os.unlink(path)
    W, both = case["W"], case["strand"] == "BOTH"
    st = S.pack(codes, offs)
    assert np.array_equal(S.unpack(st, offs), codes)
    for wpc in (1, 4, 16):  # thread chunk geometry must not matter
        hist, ltot = S.count(st, W, both, wpc)
        assert ltot == case["ltot"]
        assert np.array_equal(hist, _cases.case_counts(case))
    _, n = O.f32.background_model(codes, offs)
    assert np.array_equal(S.bg_counts(st, 2).astype(np.int64), n)


def test_pack_ascii_equals_codes():
    text = open(os.path.join(DATA, "MafK_100seqs.fasta")).read()
    codes, offs, _, _ = O.read_fasta(os.path.join(DATA, "MafK_100seqs.fasta"))
    letters = np.frombuffer("".join(l for l in text.split("\n") if not l.startswith(">")).encode(), np.uint8)
    a, b = S.pack(letters, offs, ascii=True), S.pack(codes, offs)
    assert np.array_equal(a.codes, b.codes) and np.array_equal(a.valid, b.valid) and np.array_equal(a.sstart, b.sstart)


@pytest.mark.parametrize("seed", range(12))
def test_count_dirty_random(seed):
    seqs = _cases.dirty_sequences(seed)
    codes, offs = O.encode(seqs)
    st = S.pack(codes, offs)
    rng = np.random.RandomState(seed)
    for W in (2, 4, 6, 8, 10):
        for both in (True, False):
            hist, ltot = S.count(st, W, both, int(rng.choice([1, 2, 16])))
            ref, lt = O.count(codes, offs, W, both)
            assert lt == ltot and np.array_equal(hist, ref), (W, both)
    for W in (12, 14, 16):  # no 4^W table on the test box: compare the multiset of counted keys
        for both in (True, False):
            keys, ltot = S.count_keys(st, W, both, int(rng.choice([1, 2, 16])))
            ref, lt = _cases.py_count_sparse(seqs, W, both)
            assert lt == ltot and Counter(keys.tolist()) == ref, (W, both)


@pytest.mark.parametrize("W,both", [(10, True), (10, False), (12, True)])
def test_count_mafk(W, both):
    codes, offs, _, _ = O.read_fasta(os.path.join(DATA, "MafK.fasta"))
    st = S.pack(codes, offs)
    hist, ltot = S.count(st, W, both, 16)
    ref, lt = O.count(codes, offs, W, both)
    assert lt == ltot and np.array_equal(hist, ref)


def test_synthetic_generator_properties():
    motif = [3, 2, 1, 3, 2, 0, 4, 3, 1, 0, 2, 1, 0]  # TGCTGASTCAGCA
    st = S.synth(2000, 200, 1234, 0, motif, 0.3, 0.0)
    offs = (np.arange(2001) * 200).astype(np.uint64)
    codes = S.unpack(st, offs)
    assert codes.min() >= 1 and codes.max() <= 4
    frac = np.bincount(codes, minlength=5)[1:] / codes.size
    assert np.all(np.abs(frac - 0.25) < 0.01)
    s = "".join("NACGT"[c] for c in codes)
    planted = sum(("TGCTGACTCAGCA" in s[i * 200:(i + 1) * 200]) or ("TGCTGAGTCAGCA" in s[i * 200:(i + 1) * 200])
                  for i in range(2000))  # the motif is its own reverse complement up to S
    assert 0.25 * 2000 < planted < 0.35 * 2000
    # shards of one global set are reproducible: sequences [500, 700) generated alone equal the slice
    part = S.synth(200, 200, 1234, 500, motif, 0.3, 0.0)
    assert np.array_equal(S.unpack(part, offs[:201]), codes[500 * 200:700 * 200])
    dirty = S.unpack(S.synth(500, 100, 7, 0, None, 0.0, 0.02), (np.arange(501) * 100).astype(np.uint64))
    assert 0.01 < (dirty == 0).mean() < 0.03


@pytest.mark.parametrize("name,W,both", [("cfg1_W8_BOTH", 8, True), ("cfg1_W8_PLUS", 8, False)])
def test_tables(name, W, both):
    g = np.load(os.path.join(_cases.GOLDEN, name + ".npz"))
    codes, offs, _, _ = O.read_fasta(os.path.join(DATA, "MafK_100seqs.fasta"))
    counts, ltot = O.count(codes, offs, W, both)
    for dt, suf in [(np.float32, "f32"), (np.float64, "f64")]:
        for k in range(3):
            assert np.array_equal(S.bgprob(g["v_" + suf], W, k, both, dt), g[f"bgprob_k{k}_{suf}"])
        e, z, lp = S.stats(counts, g[f"bgprob_k2_{suf}"], ltot, W)
        assert np.array_equal(e, g["expected_" + suf]) and np.array_equal(z, g["z_" + suf])
        assert np.array_equal(lp, g["logp_" + suf])  # same libm on the host; on the GPU log() may differ by 1 ulp


@pytest.mark.parametrize("W", [4, 6, 8])
def test_iupac_enumeration(W):
    for iid in _cases.random_iupac_ids(W, W, 120):
        assert np.array_equal(S.iupac_list(iid, W, 0).astype(np.uint64), O.iupac_expand(iid, W, False))
        reps = S.iupac_list(iid, W, 2)
        canon = np.array(sorted(min(int(x), O.revcomp(int(x), W)) for x in reps), np.uint64)
        assert np.array_equal(canon, np.unique(O.iupac_expand(iid, W, True)))


def test_em_chain_order_is_reference_order():
    g = np.load(os.path.join(_cases.GOLDEN, "cfg1_W8_BOTH.npz"))
    codes, offs, _, _ = O.read_fasta(os.path.join(DATA, "MafK_100seqs.fasta"))
    counts, _ = O.count(codes, offs, 8, True)
    bg = g["bgprob_k2_f32"]
    for pwm in g["adv_pwm_f32"]:
        acc = S.em_acc_f32(pwm, 8, counts, bg)
        s = ((acc[:, 0] + acc[:, 1]) + acc[:, 2]) + acc[:, 3]
        assert np.array_equal((acc / s[:, None]).astype(np.float32), O.f32.em_iteration(pwm, 8, counts, bg))


# ---- partition path: fast / slow split by bit-parallel conflict candidates ---------------------------------
def palindrome_rich_sequences(seed):
    """Reverse-complement palindromes of every length 4..40 and periodic stretches next to random sequence: the
    inputs that must reach the slow path of the partition kernel."""
    rng = np.random.RandomState(1000 + seed)
    comp = {"A": "T", "C": "G", "G": "C", "T": "A"}
    seqs = []
    for _ in range(rng.randint(2, 6)):
        parts = []
        for _ in range(rng.randint(3, 12)):
            kind = rng.randint(0, 4)
            if kind == 0:
                parts.append("".join(rng.choice(list("ACGT"), rng.randint(1, 40))))
            elif kind == 1:
                half = "".join(rng.choice(list("ACGT"), rng.randint(2, 21)))
                parts.append(half + "".join(comp[c] for c in reversed(half)))
            elif kind == 2:
                unit = "".join(rng.choice(list("ACGT"), rng.randint(1, 10)))
                parts.append((unit * 30)[: rng.randint(8, 60)])
            else:
                parts.append("N" * rng.randint(1, 3))
        seqs.append("".join(parts))
    return seqs


@pytest.mark.parametrize("seed", range(16))
def test_partition_path_split_is_exact(seed):
    seqs = _cases.dirty_sequences(seed) if seed % 2 else palindrome_rich_sequences(seed)
    codes, offs = O.encode(seqs)
    st = S.pack(codes, offs)
    for W in (4, 6, 8, 10):
        for both in (True, False):
            hist, ltot, stt = S.count_part(st, W, both)
            ref, lt = O.count(codes, offs, W, both)
            assert stt["missed"] == 0, (W, both, stt)
            assert lt == ltot and np.array_equal(hist, ref), (W, both, stt)


def long_period_sequences(seed):
    """Tandem repeats with units of 1..15 letters, long reverse-complement palindromes and random stretches: the
    conflict shapes of the long pattern lengths (lags up to W - 1 = 15)."""
    rng = np.random.RandomState(2000 + seed)
    comp = {"A": "T", "C": "G", "G": "C", "T": "A"}
    seqs = []
    for _ in range(rng.randint(2, 5)):
        parts = []
        for _ in range(rng.randint(4, 14)):
            kind = rng.randint(0, 4)
            if kind == 0:
                parts.append("".join(rng.choice(list("ACGT"), rng.randint(1, 60))))
            elif kind == 1:
                half = "".join(rng.choice(list("ACGT"), rng.randint(4, 30)))
                parts.append(half + "".join(comp[c] for c in reversed(half)))
            elif kind == 2:
                unit = "".join(rng.choice(list("ACGT"), rng.randint(1, 16)))
                parts.append((unit * 40)[: rng.randint(12, 90)])
            else:
                parts.append("N" * rng.randint(1, 3))
        seqs.append("".join(parts))
    return seqs


@pytest.mark.parametrize("seed", range(12))
def test_partition_path_split_long_patterns(seed):
    """W = 12, 14, 16 (deep partition path): the fast/slow split emits exactly the keys of the exact ring logic."""
    seqs = _cases.dirty_sequences(seed) if seed % 3 == 0 else long_period_sequences(seed)
    codes, offs = O.encode(seqs)
    st = S.pack(codes, offs)
    for W in (12, 14, 16):
        for both in (True, False):
            keys, ltot, stt = S.count_part_keys(st, W, both)
            ref, lt = S.count_keys(st, W, both)
            assert stt["missed"] == 0, (W, both, stt)
            assert lt == ltot and np.array_equal(keys, np.sort(ref)), (W, both, stt)
    hist, ltot, stt = S.count_part(st, 12, True)
    ref, lt = O.count(codes, offs, 12, True)
    assert stt["missed"] == 0 and lt == ltot and np.array_equal(hist, ref)


@pytest.mark.parametrize("seed", range(8))
def test_bit_plane_scan_equals_two_bit_scan(seed):
    """conflict_candidates_planes (what stage P runs) flags exactly the W-mers conflict_candidates flags, every W."""
    seqs = _cases.dirty_sequences(seed) if seed % 2 == 0 else long_period_sequences(seed)
    codes, offs = O.encode(seqs)
    st = S.pack(codes, offs)
    for W in (4, 6, 8, 10, 12, 14, 16):
        for both in (True, False):
            bad, flagged = S.cand_planes_mismatch(st, W, both)
            assert bad == 0, (W, both, bad)
            assert flagged > 0 or seed % 2 == 0, (W, both)


@pytest.mark.parametrize("case", [c for c in _cases.golden_cases() if c["exit"] == 0 and c["W"] in (4, 6, 8, 10)],
                         ids=lambda c: c["tag"])
def test_partition_path_golden_cases(case):
    path = _cases.write_fasta(case["fasta"])
    codes, offs, _, _ = O.read_fasta(path)
    os.unlink(path)
    hist, ltot, stt = S.count_part(S.pack(codes, offs), case["W"], case["strand"] == "BOTH")
    assert stt["missed"] == 0 and ltot == case["ltot"]
    assert np.array_equal(hist, _cases.case_counts(case))


@pytest.mark.parametrize("both", [True, False])
def test_partition_path_mafk_and_candidate_rate(both):
    codes, offs, _, _ = O.read_fasta(os.path.join(DATA, "MafK.fasta"))
    hist, ltot, stt = S.count_part(S.pack(codes, offs), 10, both)
    ref, lt = O.count(codes, offs, 10, both)
    assert stt["missed"] == 0 and lt == ltot and np.array_equal(hist, ref)
    # the candidate test is a superset; it must stay a small one or the slow path would dominate
    assert stt["needless_slow"] < 0.01 * ltot, stt
    # random sequence (BASELINE configs[2] shape): ~1 looked-at W-mer in 1000 may go to the slow path
    st = S.synth(3000, 200, 1234, 0, [3, 2, 1, 3, 2, 0, 4, 3, 1, 0, 2, 1, 0], 0.3, 0.0)
    o = (np.arange(3001) * 200).astype(np.uint64)
    c = S.unpack(st, o)
    hist, ltot, stt = S.count_part(st, 10, both)
    ref, lt = O.count(c, o, 10, both)
    assert stt["missed"] == 0 and lt == ltot and np.array_equal(hist, ref)
    assert ltot - stt["fast"] < 0.02 * ltot, stt


# ---- FASTA record splitting of the streaming ingest (host_ingest.h: scan_fasta) ---------------------------
@pytest.mark.parametrize("case", _cases.golden_cases(), ids=lambda c: c["tag"])
def test_scan_fasta_small_cases(case):
    path = _cases.write_fasta(case["fasta"])
    try:
        if case["exit"] != 0:
            with pytest.raises(ValueError):
                S.scan_fasta(path)
            return
        codes, offs = S.scan_fasta(path)
        ref_codes, ref_offs, _, _ = O.read_fasta(path)
    finally:
        os.unlink(path)
    assert np.array_equal(offs, ref_offs) and np.array_equal(codes, ref_codes[: len(codes)])


@pytest.mark.parametrize("text", ["", "\n\n", ">a\n", ">a\nACGT", ">a\nACGT\n>b\n", ">\nAC\nGT\n\n>x y\nNNAC\r\n", ">a\n>b\n>c\nA\n",
                                  ">a\nACGT\n>b\nAC"])
def test_scan_fasta_edge_files(text):
    """empty file, blank lines only, header without sequence, unterminated last line, CR, empty header, trailing record cut"""
    path = _cases.write_fasta(text)
    try:
        codes, offs = S.scan_fasta(path)
        ref_codes, ref_offs, _, _ = O.read_fasta(path)
    finally:
        os.unlink(path)
    assert np.array_equal(offs, ref_offs) and np.array_equal(codes, ref_codes[: len(codes)])


def _random_fasta(rng, n_rec):
    """records with wrapped and single-line sequences, blank lines, empty records, CR line ends, lower case, N"""
    out = []
    for r in range(n_rec):
        out.append(">r%d some text\n" % r if rng.rand() < 0.9 else ">\n")
        kind = rng.randint(0, 10)
        if kind == 0:
            continue  # record without sequence: ignored with a warning
        L = int(rng.choice([1, 5, 59, 60, 61, 200, 700]))
        seq = "".join(rng.choice(list("ACGTacgtNn"), L, p=[0.22, 0.22, 0.22, 0.22, 0.02, 0.02, 0.02, 0.02, 0.02, 0.02]))
        if kind < 5:
            out.append(seq + "\n")
        else:
            w = int(rng.choice([10, 60, 61]))
            for i in range(0, L, w):
                out.append(seq[i:i + w] + ("\r\n" if kind == 9 else "\n"))
                if rng.rand() < 0.05:
                    out.append("\n")
    return "".join(out)


@pytest.mark.parametrize("seed", range(5))
def test_scan_fasta_chunked_equals_serial(seed):
    """The parallel ingest cuts the file at header lines and scans the chunks independently (ChunkedScan): same records,
    same warnings, same 'last record' as one pass — for any number of chunks, and sharded like peng_cuda_ingest_fasta_shard."""
    rng = np.random.RandomState(300 + seed)
    text = _random_fasta(rng, 400)
    if seed == 1:
        text += ">tail\nACGT"  # unterminated last line: dropped, its record ignored
    if seed == 2:
        text += ">empty\n"
    path = _cases.write_fasta(text)
    try:
        ref_l, ref_o, ref_w, ref_last = S.scan_fasta_chunked(path, 1)
        c0, o0 = S.scan_fasta(path)
        assert np.array_equal(o0, ref_o) and len(c0) == len(ref_l)
        for n in (2, 3, 7, 16, 64, 500):
            l, o, w, last = S.scan_fasta_chunked(path, n)
            assert np.array_equal(o, ref_o) and np.array_equal(l, ref_l) and w == ref_w and last == ref_last, n
        for n_parts in (2, 3):
            ls, lens, ws = [], [], 0
            for part in range(n_parts):
                l1, o1, w1, _ = S.scan_fasta_chunked(path, 1, part, n_parts)
                l, o, w, _ = S.scan_fasta_chunked(path, 9, part, n_parts)
                assert np.array_equal(o, o1) and np.array_equal(l, l1) and w == w1
                ls.append(l)
                lens.append(np.diff(o))
            assert np.array_equal(np.concatenate(ls), ref_l) and np.array_equal(np.concatenate(lens), np.diff(ref_o))
    finally:
        os.unlink(path)


def test_scan_fasta_chunked_errors():
    """A sequence line before any header / a space in a sequence line is fatal in whichever chunk it lies."""
    good = "".join(">r%d\nACGTACGTAC\n" % i for i in range(200))
    for bad in ("ACGT\n" + good, good + ">x\nAC GT\n", good[:1500] + ">y\nA C\n" + good[1500:]):
        path = _cases.write_fasta(bad)
        try:
            for n in (1, 4, 50):
                with pytest.raises(ValueError):
                    S.scan_fasta_chunked(path, n)
        finally:
            os.unlink(path)


def test_scan_fasta_missing_file():
    with pytest.raises(ValueError):
        S.scan_fasta("/nonexistent/file.fa")


# ---- seed selection: the pruned introsort must leave the ids with z >= threshold exactly where std::sort leaves them ----
def _z_tables(seed, n):
    """tie-heavy z tables: reverse-complement-like exact pairs, plateaus of equal values, a few large values"""
    rng = np.random.RandomState(seed)
    kind = seed % 4
    if kind == 0:
        z = rng.standard_normal(n).astype(np.float32)
    elif kind == 1:
        z = np.round(rng.standard_normal(n) * 3).astype(np.float32)  # ~20 distinct values: massive ties
    elif kind == 2:
        half = rng.standard_normal(n // 2).astype(np.float32) * 4
        z = np.concatenate([half, half[::-1]])  # every value exactly twice
    else:
        z = np.sort(rng.standard_normal(n).astype(np.float32) * 5)[:: -1 if seed % 8 == 3 else 1].copy()  # sorted / reversed
    hot = rng.choice(n, 200, replace=False)
    z[hot] = np.round(rng.uniform(5, 60, 200), 1).astype(np.float32)
    z[hot[:50]] = z[hot[50:100]]  # exact ties among the large values too
    return z


@pytest.mark.parametrize("seed", range(16))
def test_pruned_sort_matches_std_sort(seed):
    n = 4 ** 8 if seed % 2 else 4 ** 9
    z = _z_tables(seed, n)
    for thr in (10.0, 3.0, 0.0, -1.0, 1e9, -1e9):
        lead, a, b = S.sort_leading(z, thr)
        k = int((z >= thr).sum())
        assert lead >= k, (thr, lead, k)
        assert np.array_equal(a[:k], b[:k]), thr
        assert (z[a[lead:].astype(np.int64)] < thr).all()


@pytest.mark.parametrize("seed", range(6))
def test_select_matches_oracle(seed):
    W = 8
    n = 4 ** W
    z = _z_tables(seed, n)
    counts = np.random.RandomState(100 + seed).randint(0, 8, n).astype(np.uint64)
    for thr, ct, single, fn in ((10.0, 3, False, True), (3.0, 1, True, False), (0.5, 2, False, False)):
        assert np.array_equal(S.select(z, counts, W, thr, ct, single, fn), O.f32.select(z, counts, W, thr, ct, single, fn))


@pytest.mark.parametrize("seed", range(16))
def test_partition_pass_list_formulation_equals_libstdcxx(seed):
    """The data-parallel restatement of a partition pass (two position lists, pairwise swaps, cut = min(A[m], B[m-1])) that
    the device runs for seed selection leaves the same array and the same cut as std::__unguarded_partition_pivot, level
    after level down the left spine — ties, plateaus, sorted and reversed inputs included."""
    for n in (17, 100, 4 ** 6, 4 ** 8):
        z = _z_tables(seed, max(n, 512))[:n].copy()
        assert S.partition_level_agrees(z), (seed, n)
    assert S.partition_level_agrees(np.zeros(1000, np.float32))  # every element ties with the pivot
    assert S.partition_level_agrees(np.arange(1000, dtype=np.float32))
    assert S.partition_level_agrees(-np.arange(1000, dtype=np.float32))


@pytest.mark.parametrize("seed", range(8))
def test_select_by_device_levels_matches_oracle(seed):
    """Seed selection with the top partition levels run in the device's formulation and the rest on (z, id) pairs
    (host_ingest.h: sort_leading_by_levels) == the oracle's std::sort-based list."""
    W = 8 if seed % 2 else 9  # 4^8 and 4^9 ids
    n = 4 ** W
    z = _z_tables(seed, n)
    counts = np.random.RandomState(100 + seed).randint(0, 8, n).astype(np.uint64)
    for thr, ct, single, fn in ((10.0, 3, False, True), (3.0, 1, True, False), (0.5, 2, False, False), (-100.0, 0, True, False)):
        for stop in (16, 64, 5000):
            got, levels = S.select_levels(z, counts, W, thr, ct, single, fn, stop)
            assert levels >= 1
            assert np.array_equal(got, O.f32.select(z, counts, W, thr, ct, single, fn)), (thr, stop)


def test_select_with_nan_falls_back_to_plain_sort():
    W = 8
    z = _z_tables(1, 4 ** W)
    z[123] = np.nan
    lead, a, b = S.sort_leading(z, 10.0)
    assert lead == len(z) and np.array_equal(a, b)


# ---- packed ingest: the host threads write the device layout themselves (host_pack.cpp) ------------------------------------
def _host_pack_matches(src, offs, ascii, n_threads, max_blocks):
    st = S.pack(src, offs, ascii)
    codes, valid, nw, phases = S.host_pack(src, offs, ascii, n_threads, max_blocks)
    assert phases > 0, phases
    q = ((nw * 32 + 63) // 64) * 64  # positions written (whole 64-position blocks)
    assert np.array_equal(codes[: q // 16][: 2 * nw], st.codes[: 2 * nw]), (n_threads, max_blocks)
    assert np.array_equal(valid[: q // 32][:nw], st.valid[:nw]), (n_threads, max_blocks)
    assert not codes[2 * nw: q // 16].any() and not valid[nw: q // 32].any()
    return phases


@pytest.mark.parametrize("seed", range(10))
def test_host_pack_equals_device_pack(seed):
    seqs = _cases.dirty_sequences(seed, n_max=30, len_max=700)
    codes, offs = O.encode(seqs)
    for n_threads, max_blocks in ((1, 1 << 14), (3, 1 << 14), (4, 40), (7, 23)):
        _host_pack_matches(codes, offs, False, n_threads, max_blocks)
    text = np.frombuffer("".join(seqs).encode(), np.uint8)
    lower = np.frombuffer("".join(seqs).lower().replace("n", "x").encode(), np.uint8)
    _host_pack_matches(text, offs, True, 3, 64)
    _host_pack_matches(lower, offs, True, 2, 1 << 14)


def test_host_pack_many_phases_and_threads():
    rng = np.random.RandomState(3)
    lens = rng.randint(1, 300, 5000)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    codes = rng.randint(0, 5, int(offs[-1])).astype(np.uint8)
    assert _host_pack_matches(codes, offs, False, 8, 512) > 10
    _host_pack_matches(codes, offs, False, 16, 1 << 16)
    # fixed length 200 (BASELINE configs[2] shape), lengths around the 64-letter chunk size, and an empty set
    offs = (np.arange(3001) * 200).astype(np.uint64)
    _host_pack_matches(rng.randint(1, 5, 600000).astype(np.uint8), offs, False, 5, 2000)
    for L in (1, 2, 61, 62, 63, 64, 65, 126, 127, 128, 129):
        offs = (np.arange(41) * L).astype(np.uint64)
        _host_pack_matches(rng.randint(0, 5, 40 * L).astype(np.uint8), offs, False, 3, 9)
    _host_pack_matches(np.zeros(0, np.uint8), np.zeros(1, np.uint64), False, 4, 64)


@pytest.mark.parametrize("trial", range(6))
def test_host_pack_block_edges(trial):
    """The AVX-512 packer works on output blocks (head / whole blocks / tail of every sequence): lengths around the block
    size, empty sequences, sequences whose separators end a block, many sequences inside one block, codes > 4."""
    rng = np.random.RandomState(90 + trial)
    pool = [0, 1, 2, 5, 9] if trial % 3 == 0 else [0, 1, 2, 3, 30, 60, 61, 62, 63, 64, 65, 66, 67, 124, 125, 126, 127, 128, 190,
                                                   191, 192, 193, 200, 700]
    lens = rng.choice(pool, 3000 if trial % 3 == 0 else 400)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    codes = rng.randint(0, 7, int(offs[-1])).astype(np.uint8)
    for n_threads, max_blocks in ((1, 1 << 14), (5, 64), (3, 33)):
        _host_pack_matches(codes, offs, False, n_threads, max_blocks)


def test_host_pack_implied_validity():
    """A phase whose letters are all ACGT sends no validity words: the device rebuilds them from the offsets
    (core_seq.cuh: implied_valid_word) and only the listed blocks travel (share boundaries, blocks holding another letter).
    `S.host_pack` rebuilds every phase that way and compares with the words the packer wrote."""
    rng = np.random.RandomState(5)
    lens = rng.randint(1, 400, 4000)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    clean = rng.randint(1, 5, int(offs[-1])).astype(np.uint8)
    phases = _host_pack_matches(clean, offs, False, 4, 2000)
    n_blocks = (int(offs[-1]) + 2 * len(lens) + 64) // 64
    assert S.host_pack.last_exceptions <= phases * (4 * 4 * 2 + 2) < n_blocks // 20  # only the share boundaries
    dirty = clean.copy()
    hit = rng.choice(len(dirty), 300, replace=False)
    dirty[hit] = 0
    _host_pack_matches(dirty, offs, False, 4, 2000)
    assert 250 <= S.host_pack.last_exceptions <= 300 + phases * 34


@pytest.mark.timeout(300, method="thread")
def test_host_pack_pool_shared_by_several_callers():
    """`peng_motif --gpus N` ingests its shards from one host thread per device, and all of them pack through the
    process-wide pool: jobs of different callers must take turns (Pool::start .. finish), each caller with two phases in
    flight as in ingest_host_hybrid.  Without that the workers' job state is torn and the callers hang."""
    rng = np.random.RandomState(6)
    lens = rng.randint(1, 400, 20000)
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    codes = rng.randint(0, 5, int(offs[-1])).astype(np.uint8)
    st = S.pack(codes, offs, False)
    for n_callers, n_threads in ((4, 4), (3, 1), (2, 8)):
        assert S.host_pack_concurrent(codes, offs, n_callers, n_threads, 10, 300, st.codes) == 0, (n_callers, n_threads)


def test_host_pack_scalar_path_and_oversized_sequence():
    rng = np.random.RandomState(4)
    offs = np.array([0, 50, 5000, 5100], np.uint64)
    codes = rng.randint(0, 5, 5100).astype(np.uint8)
    assert S.host_pack(codes, offs, False, 2, 16)[3] == -1  # the 4950-letter sequence needs 78 blocks
    _host_pack_matches(codes, offs, False, 2, 200)


@pytest.mark.parametrize("seed", range(6))
def test_nibble_path_equals_byte_path(seed):
    """The half-way packing of the hybrid ingest (host: nibble_pack over the pool's threads, device: pack_word<2>) leaves the
    same code / validity / sequence-start words as the one-byte-per-base path, for every (even) split position; letter codes
    above 4 count as 'other' in both."""
    rng = np.random.RandomState(seed)
    lens = rng.randint(1, 700, rng.randint(5, 400))
    offs = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint64)
    src = rng.randint(0, 7, int(offs[-1])).astype(np.uint8)  # 5, 6: not letters
    ref = S.pack(src, offs, False)
    for s_split in (0, 1, len(lens) // 2, len(lens) - 1):
        base = (int(offs[s_split]) - min(int(offs[s_split]), 128)) & ~1
        st, fw = S.pack_nibbles(src, offs, base, n_threads=1 + seed % 4)
        assert fw <= (64 + int(offs[s_split]) + 2 * s_split + 31) // 32 + 1
        assert np.array_equal(st.codes[2 * fw:], ref.codes[2 * fw:]) and np.array_equal(st.valid[fw:], ref.valid[fw:])
        assert np.array_equal(st.sstart[fw:], ref.sstart[fw:])
