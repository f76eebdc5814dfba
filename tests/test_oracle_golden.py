"""The oracle (oracle/peng_oracle.c) against the golden vectors recorded from the UNMODIFIED reference
(tests/golden/make_golden.py) and against the identities the reference's own unit tests pin
(test/test_base_pattern.cpp:38-131).  CPU only."""
import os

import numpy as np
import pytest

import _cases
import _oracle as O

GOLDEN = _cases.GOLDEN
DATA = os.path.join(GOLDEN, "data")


# ---- the reference's own unit-test vectors (test/test_base_pattern.cpp) ---------------------------------
def test_reference_unit_vectors_encoding():
    CT, CTG, CTGA = 1 + 3 * 4, 1 + 3 * 4 + 2 * 16, 1 + 3 * 4 + 2 * 16 + 0 * 64
    assert O.base_id("CT") == CT and O.base_id("CTG") == CTG and O.base_id("CTGA") == CTGA  # :38-50
    TCAG = 3 + 1 * 4 + 0 * 16 + 2 * 64
    assert O.revcomp(CTGA, 4) == TCAG  # :52-68
    assert O.bg_id(CTGA, 4, 2) == 3 * 16 + 2 * 4 + 0  # :70-79
    assert O.base_str(1 + 3 * 4, 4)[0] == "C" and O.base_str(1 + 3 * 4, 4)[3] == "A"  # :120-131


def test_reference_unit_vectors_iupac_expansion():
    CTRA = 1 + 3 * 11 + 6 * 121 + 0
    r = O.iupac_expand(CTRA, 4, both=False)
    assert list(r) == [1 + 3 * 4 + 0 * 16, 1 + 3 * 4 + 2 * 16]  # CTAA, CTGA in this order (:81-95)
    NTRN = 10 + 3 * 11 + 6 * 121 + 10 * 1331
    r = O.iupac_expand(NTRN, 4, both=False)
    assert len(r) == 32 and len(set(r.tolist())) == 32  # (:97-105)
    YCCT = 7 + 1 * 11 + 1 * 121 + 3 * 1331
    assert sorted(O.iupac_expand(YCCT, 4, both=False).tolist()) == sorted([1 + 4 + 16 + 3 * 64, 3 + 4 + 16 + 3 * 64])


# ---- adversarial + random small cases recorded from the reference --------------------------------------
@pytest.mark.parametrize("case", _cases.golden_cases(), ids=lambda c: c["tag"])
def test_small_cases(case):
    path = _cases.write_fasta(case["fasta"])
    try:
        if case["exit"] != 0:
            with pytest.raises(O.FastaError):
                O.read_fasta(path)
            return
        codes, offs, mn, mx = O.read_fasta(path)
    finally:
        os.unlink(path)
    W, both = case["W"], case["strand"] == "BOTH"
    assert (len(offs) - 1, mn, mx) == (case["n_seq"], case["min_len"], case["max_len"])
    counts, ltot = O.count(codes, offs, W, both)
    assert ltot == case["ltot"]
    assert np.array_equal(counts, _cases.case_counts(case))
    v, _ = O.f32.background_model(codes, offs)
    assert [int(x) for x in v.view(np.uint32)] == case["v_bits"]
    k = min(W - 1, 2)
    bg = O.f32.bg_probabilities(v, W, k)
    if both:
        bg = O.f32.aggregate_double_strand(bg, W)
    e, z, lp = O.f32.stats(counts, bg, ltot, W)
    from tests_hash import sha

    assert sha(bg) == case["bgprob_sha256"] and sha(e) == case["expected_sha256"]
    assert sha(z) == case["z_sha256"] and sha(lp) == case["logp_sha256"]
    seeds = O.f32.select(z, counts, W, 10.0, 3, not both, True)
    assert seeds.tolist() == case["seeds"]


# ---- BASELINE configs[0] and configs[1] ------------------------------------------------------------------
@pytest.mark.parametrize("name,fasta,W,both", [("cfg1_W8_BOTH", "MafK_100seqs.fasta", 8, True),
                                               ("cfg1_W8_PLUS", "MafK_100seqs.fasta", 8, False),
                                               ("cfg2_W10_BOTH", "MafK.fasta", 10, True)])
@pytest.mark.parametrize("suf", ["f32", "f64"])
def test_config_tables(name, fasta, W, both, suf):
    from tests_hash import sha

    g = np.load(os.path.join(GOLDEN, name + ".npz"))
    T = O.f32 if suf == "f32" else O.f64
    codes, offs, mn, mx = O.read_fasta(os.path.join(DATA, fasta))
    counts, ltot = O.count(codes, offs, W, both)
    assert sha(counts) == str(g["counts_sha256"])
    ref = np.zeros(4 ** W, np.uint64)
    ref[g["counts_nz_ids"]] = g["counts_nz_vals"]
    assert np.array_equal(counts, ref)
    v, _ = T.background_model(codes, offs)
    assert np.array_equal(v, g["v_" + suf])
    bgs = []
    for k in range(3):
        bg = T.bg_probabilities(v, W, k)
        if both:
            bg = T.aggregate_double_strand(bg, W)
        assert sha(bg) == str(g[f"bgprob_k{k}_{suf}_sha256"])
        bgs.append(bg)
    e, z, lp = T.stats(counts, bgs[2], ltot, W)
    assert sha(e) == str(g[f"expected_{suf}_sha256"])
    assert sha(z) == str(g[f"z_{suf}_sha256"])
    assert sha(lp) == str(g[f"logp_{suf}_sha256"])
    seeds = T.select(z, counts, W, 10.0, 3, not both, True)
    assert np.array_equal(seeds, g["seeds_" + suf])
    n_seq = len(offs) - 1
    for i, iid in enumerate(g["iupac_ids_" + suf]):
        a = T.iupac_aggregate(int(iid), W, both, counts, bgs[2], e, n_seq)
        assert a["n_sites"] == g["iupac_sites_" + suf][i]
        assert (a["bg_p"], a["expected"], a["z"], a["logp"]) == tuple(g["iupac_attrs_" + suf][i])
        pwm = T.adv_pwm(int(iid), W, both, counts, v[:4])
        assert np.array_equal(pwm, g["adv_pwm_" + suf][i])
        if W <= 8 or i < 3:  # the full 17-PWM EM at W = 10 takes ~13 s of CPU; three pin it
            out, it = T.em(pwm, W, counts, bgs[2])
            assert np.array_equal(out, g["em_pwm_" + suf][i])
