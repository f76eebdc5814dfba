"""BASELINE configs[4]: the long-k-mer end of the path.  Tables, IUPAC sums and EM at W = 14 against the FP64 oracle
(4^14 = 2.7e8 bins: 2-4 GB of host tables, the slow tests of the suite), counting and id-sliced tables at the formal
maximum W = 16 (src/peng.cpp:325-330 of the reference) against sparse restatements, and the multi-GPU protocol for
sliced tables (reduce-scatter of the histogram, partial IUPAC sums, EM accumulators summed over the slices) emulated
slice by slice on one GPU.  Counts bit-exact, FP64 quantities <= 1e-9 relative (north_star)."""
import numpy as np
import pytest

import _oracle as O

pytestmark = pytest.mark.gpu
RTOL64 = 1e-9
MOTIF = "TGCTGASTCAGCA"


def rel_ok(a, b, rtol=RTOL64):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    both_inf = np.isinf(a) & np.isinf(b) & (np.sign(a) == np.sign(b))
    with np.errstate(invalid="ignore"):
        close = np.abs(a - b) <= rtol * np.maximum(np.abs(a), np.abs(b))
    return bool(np.all(close | both_inf))


def motif_pwms(rng, W, n):
    pw = rng.dirichlet(np.ones(4) * 0.5, (n, W)) * 0.2
    for m in range(n):
        for p in range(W):
            pw[m, p, "ACGT".index("TGCTGACTCAGCATT"[(m + p) % 15])] += 0.8
    return pw / pw.sum(2, keepdims=True)


@pytest.mark.parametrize("both", [False, True], ids=["PLUS", "BOTH"])
def test_w14_tables_iupac_em_against_fp64_oracle(gpu, pkg, both):
    """Every stage behind the count at W = 14 (configs[4]: --strand PLUS; BOTH as well): count table (deep partition
    path, bit-exact, whole table), bgprob / expected / z / logp (FP64, whole tables), three IUPAC sums, one PWM x 2 EM
    iterations — against oracle/peng_oracle.c in double."""
    W = 14
    gpu.set_count_path(0)
    gpu.ingest_synthetic(150_000, 300, 77, 0, MOTIF, 0.3, 0.01)  # 4.5e7 positions: the automatic deep path (>= 2^24)
    codes, offs = gpu.export_codes()
    ltot = gpu.count(W, int(both))
    assert gpu.last_count_path == 2
    ref_counts, ref_ltot = O.count(codes, offs, W, both)
    assert ltot == ref_ltot
    counts = gpu.get_counts()
    assert np.array_equal(counts, ref_counts)
    del counts
    v64 = gpu.background_model(2, precision=pkg.F64)
    rv64, _ = O.f64.background_model(codes, offs)
    assert rel_ok(v64, rv64)
    gpu.stats(None, v64, K=2, k=2, max_k=2)
    bg = O.f64.bg_probabilities(v64, W, 2)
    if both:
        bg = O.f64.aggregate_double_strand(bg, W)
    assert rel_ok(gpu.get_table(pkg.TABLE_BGPROB, pkg.F64, 2), bg)
    e, z, lp = O.f64.stats(ref_counts, bg, ltot, W)
    assert rel_ok(gpu.get_table(pkg.TABLE_EXPECTED, pkg.F64), e)
    gz = gpu.get_table(pkg.TABLE_ZSCORE, pkg.F64)
    # z = (count - e) / sqrt(e): where count ~ e the difference cancels, so the bound is relative to the operands
    assert np.all(np.abs(gz - z) <= RTOL64 * np.maximum(np.abs(z), (ref_counts + e) / np.sqrt(e)))
    del gz
    assert rel_ok(gpu.get_table(pkg.TABLE_LOGP, pkg.F64), lp, 1e-8)  # n log(mu / n) + n - mu: cancellation as well
    del z, lp
    ids = [pkg.iupac_id("TGCTGASTCAGCAT"), pkg.iupac_id("NNCTGAGTCAGCNN"), pkg.iupac_id("WSYTGAGTCAGCAK")]
    sites, bgp, ex = gpu.iupac_aggregate(ids, pkg.F64)
    for i, iid in enumerate(ids):
        a = O.f64.iupac_aggregate(iid, W, both, ref_counts, bg, e, len(offs) - 1)
        assert sites[i] == a["n_sites"] and rel_ok(bgp[i], a["bg_p"]) and rel_ok(ex[i], a["expected"])
    assert np.array_equal(gpu.iupac_count(ids), sites)
    pw = motif_pwms(np.random.RandomState(14), W, 1)
    out, its = gpu.em(pw, pkg.F64, 1e4, 0.0, 2, 2)
    r, it = O.f64.em(pw[0], W, ref_counts, bg, 1e4, 0.0, 2)
    assert its[0] == it == 2 and rel_ok(out[0], r)


def py_bg_prob(v, W, order, i):
    """calculate_bg_probability (src/base_pattern.cpp:300-325 of the reference) for one id, in Python floats (doubles)."""
    off = lambda k: (4 ** (k + 1) - 4) // 3
    p, y = 1.0, 0
    for j in range(W):
        y = ((y << 2) | ((i >> (2 * j)) & 3)) & (4 ** (order + 1) - 1)
        k = min(j, order)
        p *= float(v[off(k) + y])
    return p


def test_w16_count_and_sliced_tables_one_gpu(gpu, pkg):
    """The formal maximum pattern length: the whole-table FP64 arm does not fit one GPU at W = 16 (6 tables of 32 GiB),
    the id-sliced tables do.  Counting against the dict restatement; then, slice by slice (8 slices, as 8 ranks would
    hold them): bgprob / expected / z at sampled ids against a per-id Python restatement, partial IUPAC sums that add
    up to the whole-table sum, and the EM protocol over the slices against the count-weighted restatement."""
    import _cases

    W, both, n_parts = 16, False, 8
    rng = np.random.RandomState(16)
    seqs = []
    for _ in range(400):
        s = list(rng.choice(list("ACGT"), 120))
        if rng.rand() < 0.6:
            p = rng.randint(0, 120 - W)
            s[p:p + W] = "TGCTGAGTCAGCATTG"
        seqs.append("".join(s))
    codes, offs = O.encode(seqs)
    gpu.set_table_slice(0, 1)
    gpu.ingest_codes(codes, offs)
    ltot = gpu.count(W, int(both))
    ref, lt = _cases.py_count_sparse(seqs, W, both)
    assert ltot == lt
    keys = np.array(sorted(ref), np.uint64)
    vals = np.array([ref[int(k)] for k in keys], np.uint64)
    assert np.array_equal(gpu.gather_counts(keys), vals) and gpu.counts_total() == int(vals.sum())
    v64 = gpu.background_model(2, precision=pkg.F64)
    iid = pkg.iupac_id("TGCTGASTCAGCATNN")
    exp_ids = O.iupac_expand(iid, W, both)
    pw = motif_pwms(rng, W, 2)
    # restatement of one EM iteration over the occupied bins only (empty bins contribute 0)
    kd = np.array([[(int(k) >> (2 * p)) & 3 for p in range(W)] for k in keys])
    bgk = np.array([py_bg_prob(v64, W, 2, int(k)) for k in keys])
    ref_pw = []
    for m in range(len(pw)):
        prob = np.prod(pw[m][np.arange(W)[None, :], kd], axis=1)
        term = vals * 1e4 / (1.0 + 1e4 / (prob / bgk))
        new = np.zeros((W, 4))
        for p in range(W):
            for a in range(4):
                new[p, a] = term[kd[:, p] == a].sum()
        ref_pw.append(new / new.sum(1, keepdims=True))
    sites_tot, bg_tot, ex_tot = 0, 0.0, 0.0
    acc_sum = None
    try:
        for part in range(n_parts):
            gpu.set_table_slice(part, n_parts)
            lo, hi = gpu.table_slice()
            assert hi - lo == 4 ** W // n_parts
            gpu.stats(None, v64, K=2, k=2, max_k=2)
            mine = keys[(keys >= lo) & (keys < hi)]
            sample = np.concatenate([mine[:64], rng.randint(lo, hi, 64, dtype=np.int64).astype(np.uint64)])
            gb = gpu.gather_table(pkg.TABLE_BGPROB, sample, pkg.F64, 2)
            ge = gpu.gather_table(pkg.TABLE_EXPECTED, sample, pkg.F64)
            gz = gpu.gather_table(pkg.TABLE_ZSCORE, sample, pkg.F64)
            cnt = gpu.gather_counts(sample).astype(np.float64)
            rb = np.array([py_bg_prob(v64, W, 2, int(i)) for i in sample])
            assert rel_ok(gb, rb) and rel_ok(ge, rb * ltot)
            assert np.all(np.abs(gz - (cnt - rb * ltot) / np.sqrt(rb * ltot)) <= RTOL64 * (cnt + rb * ltot) / np.sqrt(rb * ltot))
            with pytest.raises(pkg.PengCudaError):
                gpu.gather_table(pkg.TABLE_ZSCORE, np.array([(hi + 5) % 4 ** W], np.uint64), pkg.F64)  # not this rank's id
            s, b, e = gpu.iupac_aggregate([iid], pkg.F64)  # partial sums over the slice
            sites_tot += int(s[0]); bg_tot += float(b[0]); ex_tot += float(e[0])
            acc = gpu.em_begin(pw, 1e4, 0.0, 1, 2, part, n_parts)
            gpu.em_accumulate()
            gpu.synchronize()
            import torch

            a = torch.as_tensor(acc, device="cuda").cpu().numpy().copy()
            acc_sum = a if acc_sum is None else acc_sum + a
            gpu.em_end()
        rb_all = np.array([py_bg_prob(v64, W, 2, int(i)) for i in exp_ids])
        assert sites_tot == sum(ref.get(int(i), 0) for i in exp_ids)
        assert rel_ok(bg_tot, rb_all.sum()) and rel_ok(ex_tot, (rb_all * ltot).sum())
        new = acc_sum.reshape(len(pw), W, 4)
        new = new / new.sum(2, keepdims=True)
        assert rel_ok(new, np.array(ref_pw))
    finally:
        gpu.set_table_slice(0, 1)


def test_sliced_protocol_equals_whole_tables_w12(gpu, pkg):
    """The complete sliced multi-GPU protocol at a size where the whole tables exist beside it (W = 12, PLUS, 4 ranks
    emulated on one GPU): count_local32 per shard -> reduce-scatter (emulated) -> count_finish32_sliced -> stats on the
    slice -> IUPAC partial sums / EM accumulators added over the ranks == the single-GPU tables, sums and PWMs."""
    import torch

    W, n_parts, n = 12, 4, 40_000
    gpu.set_table_slice(0, 1)
    gpu.ingest_synthetic(n, 200, 5, 0, MOTIF, 0.3, 0.01)
    ltot = gpu.count(W, pkg.STRAND_PLUS)
    whole_counts = gpu.get_counts()
    v64 = gpu.background_model(2, precision=pkg.F64)
    gpu.stats(None, v64, K=2, k=2, max_k=2)
    whole_z = gpu.get_table(pkg.TABLE_ZSCORE, pkg.F64)
    ids = [pkg.iupac_id("TGCTGASTCAGC"), pkg.iupac_id("NNCTGAGTCANN")]
    w_sites, w_bg, w_ex = gpu.iupac_aggregate(ids, pkg.F64)
    pw = motif_pwms(np.random.RandomState(3), W, 3)
    w_pwm, w_its = gpu.em(pw, pkg.F64, 1e4, 0.0, 3, 2)
    # the shards' unfolded uint32 tables, summed as the collective would
    acc = None
    for r in range(n_parts):
        gpu.ingest_synthetic(n // n_parts, 200, 5, r * (n // n_parts), MOTIF, 0.3, 0.01)
        gpu.count_local32(W, pkg.STRAND_PLUS)
        gpu.synchronize()
        buf = torch.as_tensor(gpu.count_buffer32(), device="cuda")
        acc = buf.clone() if acc is None else acc + buf
        torch.cuda.synchronize()  # torch's stream has read the table before the handle's stream reuses it
    sites = np.zeros(len(ids), np.uint64); bgs = np.zeros(len(ids)); exs = np.zeros(len(ids))
    em_acc, state = None, []
    try:
        for part in range(n_parts):
            gpu.set_table_slice(part, n_parts)
            gpu.count_local32(W, pkg.STRAND_PLUS)  # the rank's own count (its result is replaced by the emulated collective)
            gpu.synchronize()
            lo, hi = gpu.table_slice()
            buf = torch.as_tensor(gpu.count_buffer32(), device="cuda")
            buf[lo:hi] = acc[lo:hi]   # reduce-scatter in place: only the rank's slice holds the sum
            buf[-1] = acc[-1]         # all-reduce of the ltot word
            torch.cuda.synchronize()
            assert gpu.count_finish32_sliced() == ltot
            assert np.array_equal(gpu.get_counts()[: hi - lo], whole_counts[lo:hi])
            gpu.stats(None, v64, K=2, k=2, max_k=2)
            assert np.array_equal(gpu.get_table(pkg.TABLE_ZSCORE, pkg.F64), whole_z[lo:hi])
            s, b, e = gpu.iupac_aggregate(ids, pkg.F64)
            sites += s; bgs += b; exs += e
            with pytest.raises(pkg.PengCudaError):
                gpu.select_base_patterns()
            with pytest.raises(pkg.PengCudaError):
                gpu.em(pw, pkg.F64)
            a = gpu.em_begin(pw, 1e4, 0.0, 3, 2, part, n_parts)
            gpu.em_accumulate()
            gpu.synchronize()
            t = torch.as_tensor(a, device="cuda").cpu().numpy().copy()
            em_acc = t if em_acc is None else em_acc + t
            gpu.em_end()
        assert np.array_equal(sites, w_sites) and rel_ok(bgs, w_bg) and rel_ok(exs, w_ex)
        new = em_acc.reshape(len(pw), W, 4)
        new = new / new.sum(2, keepdims=True)
        # the first EM iteration of the whole-table run, recomputed: one iteration from the same start
        gpu.set_table_slice(0, 1)
        gpu.ingest_synthetic(n, 200, 5, 0, MOTIF, 0.3, 0.01)
        gpu.count(W, pkg.STRAND_PLUS)
        gpu.stats(None, v64, K=2, k=2, max_k=2)
        one, _ = gpu.em(pw, pkg.F64, 1e4, 0.0, 1, 2)
        assert rel_ok(new, one)
    finally:
        gpu.set_table_slice(0, 1)
