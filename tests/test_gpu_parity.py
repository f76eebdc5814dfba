"""Parity tests proper: the CUDA path, called through the C ABI (include/peng_cuda.h), against the oracle on the
same inputs, against the committed golden fixtures recorded from the unmodified reference, and — at
BASELINE.json's full sizes — through size-independent properties.

Bars: bit-exact for all counts, ltot, IUPAC count sums, selected pattern lists and every PENG_F32REF table
(float32 reference arithmetic); <= 1e-9 relative for PENG_F64 z-scores / expected counts / PWMs against
the FP64 arm (tolerance written below).  One documented exception: the float32 log p-value table goes
through the device's double log(), which is not bit-identical to glibc's; it is compared at float32
resolution and must be bit-identical for all but a handful of entries.
"""
import os
from collections import Counter

import numpy as np
import pytest

import _cases
import _oracle as O
from tests_hash import sha

pytestmark = pytest.mark.gpu
DATA = os.path.join(_cases.GOLDEN, "data")
RTOL64 = 1e-9  # north_star: "within 1e-9 relative error in FP64"


def rel_ok(a, b, rtol=RTOL64):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    both_inf = np.isinf(a) & np.isinf(b) & (np.sign(a) == np.sign(b))
    both_nan = np.isnan(a) & np.isnan(b)
    with np.errstate(invalid="ignore"):
        close = np.abs(a - b) <= rtol * np.maximum(np.abs(a), np.abs(b))
    return bool(np.all(close | both_inf | both_nan))


# ---------------------------------------------------------------------------------------------------------
# (a) ingest
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("case", _cases.golden_cases(), ids=lambda c: c["tag"])
def test_ingest_and_count_golden_cases(gpu, pkg, case):
    path = _cases.write_fasta(case["fasta"])
    try:
        if case["exit"] != 0:
            with pytest.raises(pkg.PengCudaError) as e:
                gpu.ingest_fasta(path)
            assert e.value.code == -4  # PENG_E_FORMAT where the reference exits(1)
            return
        info = gpu.ingest_fasta(path)
    finally:
        os.unlink(path)
    assert (info["n_seq"], info["min_len"], info["max_len"]) == (case["n_seq"], case["min_len"], case["max_len"])
    W, both = case["W"], case["strand"] == "BOTH"
    ltot = gpu.count(W, pkg.STRAND_BOTH if both else pkg.STRAND_PLUS)
    assert ltot == case["ltot"]
    assert np.array_equal(gpu.get_counts(), _cases.case_counts(case))
    # background model (device-counted (k+1)-mers incl. the N quirks) -> v bit patterns
    v = gpu.background_model(2, precision=pkg.F32REF)
    assert [int(x) for x in v.view(np.uint32)] == case["v_bits"]
    # tables + seed list
    gpu.stats(v, v.astype(np.float64), K=2)
    k = min(W - 1, 2)
    assert sha(gpu.get_table(pkg.TABLE_BGPROB, pkg.F32REF, k)) == case["bgprob_sha256"]
    assert sha(gpu.get_table(pkg.TABLE_EXPECTED, pkg.F32REF)) == case["expected_sha256"]
    assert sha(gpu.get_table(pkg.TABLE_ZSCORE, pkg.F32REF)) == case["z_sha256"]
    assert gpu.select_base_patterns(10.0, 3).tolist() == case["seeds"]


def test_ingest_codes_roundtrip_and_fasta_equivalence(gpu, pkg):
    codes, offs, mn, mx = O.read_fasta(os.path.join(DATA, "MafK.fasta"))
    gpu.ingest_codes(codes, offs)
    c2, o2 = gpu.export_codes()
    assert np.array_equal(c2, codes) and np.array_equal(o2, offs)
    info = gpu.ingest_fasta(os.path.join(DATA, "MafK.fasta"))
    assert info == dict(n_seq=len(offs) - 1, total_bases=len(codes), min_len=mn, max_len=mx)
    c3, o3 = gpu.export_codes()
    assert np.array_equal(c3, codes) and np.array_equal(o3, offs)


def test_ingest_empty_and_ragged(gpu, pkg):
    # ragged lengths incl. sequences shorter than W and of exactly W
    seqs = ["A", "ACGT", "ACGTACGTAC", "ACGTACGTACG", "N", "NNNNNNNNNNNN", "ACGTNACGTACGTACGTAC", "T" * 300]
    codes, offs = O.encode(seqs)
    gpu.ingest_codes(codes, offs)
    for W in (2, 4, 10, 12):
        for both in (True, False):
            ltot = gpu.count(W, int(both))
            ref, lt = O.count(codes, offs, W, both)
            assert ltot == lt and np.array_equal(gpu.get_counts(), ref)
    # an empty set: zero k-mers everywhere
    gpu.ingest_codes(np.zeros(0, np.uint8), np.zeros(1, np.uint64))
    assert gpu.count(4, pkg.STRAND_BOTH) == 0 and not gpu.get_counts().any()


# ---------------------------------------------------------------------------------------------------------
# (b) counting
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("seed", range(8))
def test_count_dirty_random_all_lengths(gpu, pkg, seed):
    seqs = _cases.dirty_sequences(seed)
    codes, offs = O.encode(seqs)
    gpu.ingest_codes(codes, offs)
    for W in (2, 4, 6, 8, 10, 12):
        for both in (True, False):
            ltot = gpu.count(W, int(both))
            ref, lt = O.count(codes, offs, W, both)
            assert ltot == lt, (W, both)
            assert np.array_equal(gpu.get_counts(), ref), (W, both)


@pytest.mark.parametrize("seed", range(2))
def test_count_maximum_pattern_lengths(gpu, pkg, seed):
    """W = 14 (2 GiB table) and the formal maximum W = 16 (32 GiB table): compare the sparse multiset of
    counted keys with an independent dict-based restatement."""
    seqs = _cases.dirty_sequences(100 + seed)
    codes, offs = O.encode(seqs)
    gpu.ingest_codes(codes, offs)
    for W in (14, 16):
        for both in (True, False):
            ltot = gpu.count(W, int(both))
            ref, lt = _cases.py_count_sparse(seqs, W, both)
            assert ltot == lt
            if both:  # table holds the pair count at both ids
                exp = {}
                for k, c in ref.items():
                    exp[k] = c
                    exp[O.revcomp(k, W)] = c
            else:
                exp = dict(ref)
            # the 4^W table (2 GiB / 32 GiB) stays on the device: gather the expected bins plus random
            # other bins, and check the sum over ALL bins so that no stray increment can hide
            keys = np.array(sorted(exp), np.uint64)
            assert np.array_equal(gpu.gather_counts(keys), np.array([exp[int(k)] for k in keys], np.uint64)), (W, both)
            others = np.random.RandomState(W).randint(0, 4 ** W, 4096, dtype=np.int64).astype(np.uint64)
            others = others[~np.isin(others, keys)]
            assert not gpu.gather_counts(others).any()
            assert gpu.counts_total() == sum(exp.values()), (W, both)


@pytest.mark.parametrize("name,fasta,W,both", [("cfg1_W8_BOTH", "MafK_100seqs.fasta", 8, True),
                                               ("cfg1_W8_PLUS", "MafK_100seqs.fasta", 8, False),
                                               ("cfg2_W10_BOTH", "MafK.fasta", 10, True)])
def test_configs_against_reference_goldens(gpu, pkg, name, fasta, W, both):
    """BASELINE configs[0] / configs[1] end to end through the hot path against tables recorded from the
    unmodified reference (float32 arm bit-exact, FP64 arm 1e-9)."""
    g = np.load(os.path.join(_cases.GOLDEN, name + ".npz"))
    info = gpu.ingest_fasta(os.path.join(DATA, fasta))
    n_seq = info["n_seq"]
    ltot = gpu.count(W, int(both))
    counts = gpu.get_counts()
    ref = np.zeros(4 ** W, np.uint64)
    ref[g["counts_nz_ids"]] = g["counts_nz_vals"]
    assert np.array_equal(counts, ref) and sha(counts) == str(g["counts_sha256"])
    v32 = gpu.background_model(2, precision=pkg.F32REF)
    v64 = gpu.background_model(2, precision=pkg.F64)
    assert np.array_equal(v32, g["v_f32"]) and np.array_equal(v64, g["v_f64"])
    gpu.stats(v32, v64, K=2, k=2, max_k=2)
    for k in range(3):
        assert sha(gpu.get_table(pkg.TABLE_BGPROB, pkg.F32REF, k)) == str(g[f"bgprob_k{k}_f32_sha256"])
    assert sha(gpu.get_table(pkg.TABLE_EXPECTED, pkg.F32REF)) == str(g["expected_f32_sha256"])
    assert sha(gpu.get_table(pkg.TABLE_ZSCORE, pkg.F32REF)) == str(g["z_f32_sha256"])
    # FP64 arm against the FP64 build of the reference
    codes, offs, _, _ = O.read_fasta(os.path.join(DATA, fasta))
    bg64 = O.f64.bg_probabilities(v64, W, 2)
    if both:
        bg64 = O.f64.aggregate_double_strand(bg64, W)
    e64, z64, lp64 = O.f64.stats(counts, bg64, ltot, W)
    assert sha(e64) == str(g["expected_f64_sha256"])  # the oracle itself is pinned to the reference here
    assert rel_ok(gpu.get_table(pkg.TABLE_BGPROB, pkg.F64, 2), bg64)
    assert rel_ok(gpu.get_table(pkg.TABLE_EXPECTED, pkg.F64), e64)
    assert rel_ok(gpu.get_table(pkg.TABLE_ZSCORE, pkg.F64), z64)
    assert rel_ok(gpu.get_table(pkg.TABLE_LOGP, pkg.F64), lp64)
    # float32 log p-values: device log() vs glibc log(): equal at float32 resolution, nearly all bit-equal
    bg32 = O.f32.bg_probabilities(v32, W, 2)
    if both:
        bg32 = O.f32.aggregate_double_strand(bg32, W)
    e32, z32, lp32 = O.f32.stats(counts, bg32, ltot, W)
    lp_gpu = gpu.get_table(pkg.TABLE_LOGP, pkg.F32REF)
    assert rel_ok(lp_gpu, lp32, 3e-7)
    assert int((lp_gpu != lp32).sum()) <= max(4, lp32.size // 100000)
    # seed list (std::sort tie order replayed on the host)
    assert np.array_equal(gpu.select_base_patterns(10.0, 3), g["seeds_f32"])
    # IUPAC sums of the patterns the reference's hill climb ended on
    ids = g["iupac_ids_f32"]
    sites, bgp, ex = gpu.iupac_aggregate(ids, pkg.F32REF)
    assert np.array_equal(sites, g["iupac_sites_f32"])
    assert np.array_equal(bgp, g["iupac_attrs_f32"][:, 0]) and np.array_equal(ex, g["iupac_attrs_f32"][:, 1])
    sites64, bgp64, ex64 = gpu.iupac_aggregate(ids, pkg.F64)
    assert np.array_equal(sites64, g["iupac_sites_f64"])
    assert rel_ok(bgp64, g["iupac_attrs_f64"][:, 0]) and rel_ok(ex64, g["iupac_attrs_f64"][:, 1])
    # EM: float32 arm bit-identical to the reference's PWMs, FP64 arm within 1e-9 of the FP64 build
    out, its = gpu.em(g["adv_pwm_f32"], pkg.F32REF)
    assert np.array_equal(out, g["em_pwm_f32"])
    out64, its64 = gpu.em(g["adv_pwm_f64"], pkg.F64)
    assert rel_ok(out64, g["em_pwm_f64"])
    assert np.array_equal(its, its64)


# ---------------------------------------------------------------------------------------------------------
# (c) tables on random inputs
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("W,both", [(4, True), (6, False), (8, True), (10, False)])
def test_tables_random(gpu, pkg, W, both):
    rng = np.random.RandomState(W)
    seqs = ["".join(rng.choice(list("ACGT"), rng.randint(50, 400), p=[.35, .15, .2, .3])) for _ in range(200)]
    codes, offs = O.encode(seqs)
    gpu.ingest_codes(codes, offs)
    ltot = gpu.count(W, int(both))
    counts = gpu.get_counts()
    v32, n32 = gpu.background_model(2, precision=pkg.F32REF, return_counts=True)
    rv32, rn = O.f32.background_model(codes, offs)
    assert np.array_equal(n32, rn) and np.array_equal(v32, rv32)
    v64 = gpu.background_model(2, precision=pkg.F64)
    assert np.array_equal(v64, O.f64.background_model(codes, offs)[0])
    kk = min(W - 1, 2)
    gpu.stats(v32, v64, K=2, k=kk, max_k=kk)
    for T, prec in ((O.f32, pkg.F32REF), (O.f64, pkg.F64)):
        v = v32 if prec == pkg.F32REF else v64
        for k in range(kk + 1):
            bg = T.bg_probabilities(v, W, k)
            if both:
                bg = T.aggregate_double_strand(bg, W)
            got = gpu.get_table(pkg.TABLE_BGPROB, prec, k)
            assert np.array_equal(got, bg) if prec == pkg.F32REF else rel_ok(got, bg)
        e, z, lp = T.stats(counts, bg, ltot, W)
        ge, gz = gpu.get_table(pkg.TABLE_EXPECTED, prec), gpu.get_table(pkg.TABLE_ZSCORE, prec)
        if prec == pkg.F32REF:
            assert np.array_equal(ge, e) and np.array_equal(gz, z)
        else:
            assert rel_ok(ge, e) and rel_ok(gz, z) and rel_ok(gpu.get_table(pkg.TABLE_LOGP, prec), lp)
    e32, z32, _ = O.f32.stats(counts, gpu.get_table(pkg.TABLE_BGPROB, pkg.F32REF, kk), ltot, W)
    for zt, ct, fn in ((3.0, 1, True), (1.0, 2, False)):
        assert np.array_equal(gpu.select_base_patterns(zt, ct, None, fn), O.f32.select(z32, counts, W, zt, ct, not both, fn))


# ---------------------------------------------------------------------------------------------------------
# (d) IUPAC sums
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("W,both", [(6, True), (6, False), (8, True), (8, False), (10, True)])
def test_iupac_sums_random(gpu, pkg, W, both):
    rng = np.random.RandomState(10 + W)
    seqs = ["".join(rng.choice(list("ACGT"), 300)) for _ in range(300)]
    codes, offs = O.encode(seqs)
    gpu.ingest_codes(codes, offs)
    ltot = gpu.count(W, int(both))
    counts = gpu.get_counts()
    v32 = gpu.background_model(2, precision=pkg.F32REF)
    v64 = gpu.background_model(2, precision=pkg.F64)
    gpu.stats(v32, v64, K=2)
    bg32, e32 = gpu.get_table(pkg.TABLE_BGPROB, pkg.F32REF, 2), gpu.get_table(pkg.TABLE_EXPECTED, pkg.F32REF)
    bg64, e64 = gpu.get_table(pkg.TABLE_BGPROB, pkg.F64, 2), gpu.get_table(pkg.TABLE_EXPECTED, pkg.F64)
    ids = _cases.random_iupac_ids(W, W, 64)
    ids += [pkg.iupac_id("N" * W), pkg.iupac_id("A" * W), pkg.iupac_id(("ACGT" * 4)[:W])]
    ids += [pkg.iupac_id(("SW" * 8)[:W]), pkg.iupac_id(("NNNNNNAC" * 2)[:W])]
    sites, bgp, ex = gpu.iupac_aggregate(ids, pkg.F32REF)
    s64, b64, x64 = gpu.iupac_aggregate(ids, pkg.F64)
    cc = gpu.iupac_count(ids)
    for i, iid in enumerate(ids):
        a = O.f32.iupac_aggregate(iid, W, both, counts, bg32, e32, len(seqs))
        assert sites[i] == a["n_sites"] and bgp[i] == a["bg_p"] and ex[i] == a["expected"], pkg.iupac_str(iid, W)
        b = O.f64.iupac_aggregate(iid, W, both, counts, bg64, e64, len(seqs))
        assert s64[i] == b["n_sites"] and rel_ok(b64[i], b["bg_p"]) and rel_ok(x64[i], b["expected"])
        assert cc[i] == O.count_combined(iid, W, both, counts)


# ---------------------------------------------------------------------------------------------------------
# (e) EM
# ---------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("W,both", [(4, True), (6, True), (8, False), (10, True)])
def test_em_random_pwms(gpu, pkg, W, both):
    rng = np.random.RandomState(20 + W)
    motif = "TGCTGAGTCAGCA"[:W]
    seqs = []
    for _ in range(400):
        s = list(rng.choice(list("ACGT"), 120))
        if rng.rand() < 0.5:
            p = rng.randint(0, 120 - W)
            s[p:p + W] = motif
        seqs.append("".join(s))
    codes, offs = O.encode(seqs)
    gpu.ingest_codes(codes, offs)
    gpu.count(W, int(both))
    counts = gpu.get_counts()
    v32 = gpu.background_model(2, precision=pkg.F32REF)
    v64 = gpu.background_model(2, precision=pkg.F64)
    kk = min(W - 1, 2)
    gpu.stats(v32, v64, K=2, k=kk, max_k=kk)
    bg32, bg64 = gpu.get_table(pkg.TABLE_BGPROB, pkg.F32REF, kk), gpu.get_table(pkg.TABLE_BGPROB, pkg.F64, kk)
    pwms = []
    for i in range(5):
        pw = rng.dirichlet(np.ones(4) * 0.5, W) * 0.2
        if i < 3:
            for p, ch in enumerate(motif):
                pw[p, "ACGT".index(ch)] += 0.8
        pw /= pw.sum(1, keepdims=True)
        pwms.append(pw)
    pwms = np.array(pwms)
    for thr, mx in ((0.08, 10), (1e-4, 3), (100.0, 10)):
        out, its = gpu.em(pwms.astype(np.float32), pkg.F32REF, 1e4, thr, mx, kk)
        out64, its64 = gpu.em(pwms, pkg.F64, 1e4, thr, mx, kk)
        for i in range(len(pwms)):
            r, it = O.f32.em(pwms[i].astype(np.float32), W, counts, bg32, 1e4, thr, mx)
            assert it == its[i] and np.array_equal(out[i], r), (W, thr, i)
            r64, it64 = O.f64.em(pwms[i], W, counts, bg64, 1e4, thr, mx)
            assert it64 == its64[i] and rel_ok(out64[i], r64), (W, thr, i)


def test_long_kmer_tables_iupac_em_W12(gpu, pkg):
    """The long-k-mer end of the path (BASELINE configs[4] asks for the EM table at large W): every stage at W = 12
    (16.8 M bins) against the oracle — tables of both arms, IUPAC sums, EM."""
    W, both = 12, False
    rng = np.random.RandomState(1212)
    motif = "TGCTGAGTCAGC"
    seqs = []
    for _ in range(300):
        s = list(rng.choice(list("ACGT"), 150))
        if rng.rand() < 0.6:
            p = rng.randint(0, 150 - W)
            s[p:p + W] = motif
        seqs.append("".join(s))
    codes, offs = O.encode(seqs)
    gpu.ingest_codes(codes, offs)
    ltot = gpu.count(W, int(both))
    counts = gpu.get_counts()
    ref_counts, ref_ltot = O.count(codes, offs, W, both)
    assert ltot == ref_ltot and np.array_equal(counts, ref_counts)
    v32 = gpu.background_model(2, precision=pkg.F32REF)
    v64 = gpu.background_model(2, precision=pkg.F64)
    gpu.stats(v32, v64, K=2, k=2, max_k=2)
    bg32 = O.f32.bg_probabilities(v32, W, 2)
    bg64 = O.f64.bg_probabilities(v64, W, 2)
    assert np.array_equal(gpu.get_table(pkg.TABLE_BGPROB, pkg.F32REF, 2), bg32)
    assert rel_ok(gpu.get_table(pkg.TABLE_BGPROB, pkg.F64, 2), bg64)
    e32, z32, _ = O.f32.stats(counts, bg32, ltot, W)
    assert np.array_equal(gpu.get_table(pkg.TABLE_EXPECTED, pkg.F32REF), e32)
    assert np.array_equal(gpu.get_table(pkg.TABLE_ZSCORE, pkg.F32REF), z32)
    assert np.array_equal(gpu.select_base_patterns(10.0, 3, None, True), O.f32.select(z32, counts, W, 10.0, 3, True, True))
    ids = [pkg.iupac_id("TGCTGASTCAGC"), pkg.iupac_id("TGCTGANNCAGC"), pkg.iupac_id("WSCTGAGTCAGY")]
    sites, bgp, ex = gpu.iupac_aggregate(ids, pkg.F32REF)
    for i, iid in enumerate(ids):
        a = O.f32.iupac_aggregate(iid, W, both, counts, bg32, e32, len(seqs))
        assert sites[i] == a["n_sites"] and bgp[i] == a["bg_p"] and ex[i] == a["expected"]
    pwms = []
    for i in range(2):
        pw = rng.dirichlet(np.ones(4) * 0.5, W) * 0.2
        for p, ch in enumerate(motif):
            pw[p, "ACGT".index(ch)] += 0.8
        pw /= pw.sum(1, keepdims=True)
        pwms.append(pw)
    pwms = np.array(pwms)
    out, its = gpu.em(pwms.astype(np.float32), pkg.F32REF, 1e4, 0.08, 3, 2)
    out64, its64 = gpu.em(pwms, pkg.F64, 1e4, 0.08, 3, 2)
    for i in range(len(pwms)):
        r, it = O.f32.em(pwms[i].astype(np.float32), W, counts, bg32, 1e4, 0.08, 3)
        assert it == its[i] and np.array_equal(out[i], r)
        r64, it64 = O.f64.em(pwms[i], W, counts, bg64, 1e4, 0.08, 3)
        assert it64 == its64[i] and rel_ok(out64[i], r64)


def test_em_sharded_protocol_single_gpu(gpu, pkg):
    """The multi-GPU EM protocol (index-sliced accumulation + sum of the accumulators) with both slices run
    on one GPU one after the other: must agree with the unsliced result to 1e-9."""
    g = np.load(os.path.join(_cases.GOLDEN, "cfg1_W8_BOTH.npz"))
    gpu.ingest_fasta(os.path.join(DATA, "MafK_100seqs.fasta"))
    gpu.count(8, pkg.STRAND_BOTH)
    gpu.stats(g["v_f32"], g["v_f64"], K=2)
    pw = g["adv_pwm_f64"]
    whole, its = gpu.em(pw, pkg.F64, threshold=0.0)  # never converges early: exactly 10 iterations
    import torch

    n_parts = 2
    # emulate the ranks sequentially: each accumulates its slice, the slices are summed, every "rank" updates
    state = pw.copy()
    for it in range(10):
        acc_sum = None
        for part in range(n_parts):
            acc = gpu.em_begin(state, max_iterations=1, part=part, n_parts=n_parts)
            gpu.em_accumulate()
            gpu.synchronize()  # the library runs on its own stream; torch reads on the default one
            t = torch.as_tensor(acc, device="cuda").clone()
            torch.cuda.synchronize()
            acc_sum = t if acc_sum is None else acc_sum + t
            gpu.em_end()
        rows = acc_sum.cpu().numpy().reshape(-1, 8, 4)
        state = rows / rows.sum(2, keepdims=True)
    state = state / state.sum(2, keepdims=True)
    assert np.all(its == 10)
    assert rel_ok(state, whole, 1e-9)


# ---------------------------------------------------------------------------------------------------------
# full-size properties (BASELINE configs[2]: 1M x 200 bp synthetic, W = 10)
# ---------------------------------------------------------------------------------------------------------
def test_full_size_properties(gpu, pkg):
    n_seq, L, W = 1_000_000, 200, 10
    gpu.ingest_synthetic(n_seq, L, seed=1234, motif="TGCTGASTCAGCA", plant_fraction=0.3)
    ltot = gpu.count(W, pkg.STRAND_BOTH)
    assert ltot == n_seq * (L - W + 1)  # no N: every window is looked at
    counts = gpu.get_counts()
    ids = np.arange(4 ** W, dtype=np.uint64)
    rc = np.zeros_like(ids)
    for p in range(W):
        rc |= (3 - ((ids >> np.uint64(2 * p)) & np.uint64(3))) << np.uint64(2 * (W - 1 - p))
    assert np.array_equal(counts, counts[rc])  # mirror symmetry
    canon = ids <= rc
    total = int(counts[canon].sum())
    assert total <= ltot and total > 0.999 * ltot  # only self-overlapping repeats are suppressed
    ltot_plus = gpu.count(W, pkg.STRAND_PLUS)
    plus = gpu.get_counts()
    assert ltot_plus == ltot and int(plus.sum()) <= ltot
    # both-strand pair count never exceeds the sum of the two single-strand counts
    assert np.all(counts <= plus + plus[rc])
    # the planted motif's core is massively enriched
    core = pkg.base_id("TGCTGACTCA")
    assert counts[core] > 50 * np.median(counts)
    # sharding property (what the multi-GPU path relies on): shard histograms add up exactly
    acc = np.zeros(4 ** W, np.uint64)
    lt = 0
    for r in range(4):
        first, n = pkg.shard_synthetic(n_seq, r, 4)
        gpu.ingest_synthetic(n, L, seed=1234, first_seq=first, motif="TGCTGASTCAGCA", plant_fraction=0.3)
        gpu.count_local(W, pkg.STRAND_BOTH)
        gpu.synchronize()
        import torch

        buf = torch.as_tensor(gpu.count_buffer(), device="cuda").cpu().numpy().view(np.uint64)
        acc += buf[:-1]
        lt += int(buf[-1])
    assert lt == ltot and np.array_equal(np.where(canon, acc, acc[rc]), counts)
    # a bounded sample of the same workload against the oracle (50k sequences)
    gpu.ingest_synthetic(50_000, L, seed=1234, motif="TGCTGASTCAGCA", plant_fraction=0.3)
    codes, offs = gpu.export_codes()
    lt = gpu.count(W, pkg.STRAND_BOTH)
    ref, rlt = O.count(codes, offs, W, True)
    assert lt == rlt and np.array_equal(gpu.get_counts(), ref)


def test_dirty_synthetic_against_oracle(gpu, pkg):
    """1 % N: exercises the resume rules at scale (200k x 150 bp)."""
    gpu.ingest_synthetic(200_000, 150, seed=99, motif="TGCTGASTCAGCA", plant_fraction=0.3, n_fraction=0.01)
    codes, offs = gpu.export_codes()
    assert 0.005 < (codes == 0).mean() < 0.015
    for W, both in ((10, True), (8, False), (6, True)):
        lt = gpu.count(W, int(both))
        ref, rlt = O.count(codes, offs, W, both)
        assert lt == rlt and np.array_equal(gpu.get_counts(), ref)
    v, n = gpu.background_model(2, return_counts=True)
    rv, rn = O.f32.background_model(codes, offs)
    assert np.array_equal(n, rn) and np.array_equal(v, rv)
