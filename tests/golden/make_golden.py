#!/usr/bin/env python
"""Generate the committed golden fixtures by RUNNING THE UNMODIFIED REFERENCE.

Run in the build container only (needs /root/reference and `make -C oracle ref`):

    python tests/golden/make_golden.py

It drives oracle/_ref/ref_dump (float32 as-is = "Oracle A") and oracle/_ref/ref_dump64
(FP64 variant = "Oracle B") plus the reference CLI oracle/_ref/peng_motif_ref and writes

  tests/golden/data/*.fa[sta]          input fixtures (the reference's shipped data + its unit-test FASTA)
  tests/golden/cfg1_W8_BOTH.npz        full tables for BASELINE configs[0]
  tests/golden/cfg1_W8_PLUS.npz        same input, --strand PLUS
  tests/golden/cfg2_W10_BOTH.npz       configs[1]: sparse counts + digests of the 4^10 tables + seeds/IUPAC/PWMs
  tests/golden/cfg*_stdout.txt, *.meme the reference's console report and MEME file
  tests/golden/cases.json              small adversarial + random FASTA cases with the reference's answers

Nothing here is read from /root/reference at test time; tests only read the files written here.
"""
import hashlib
import json
import os
import random
import shutil
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("PENG_REFERENCE", "/root/reference")
BIN = os.path.join(ROOT, "oracle", "_ref")


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def dump(fasta, W, strand, binary="ref_dump", extra=()):
    d = tempfile.mkdtemp(prefix="golden_")
    r = subprocess.run([os.path.join(BIN, binary), fasta, str(W), strand, d, *extra], capture_output=True, text=True)
    return d, r


def load_dump(d, dt):
    meta = dict(l.split() for l in open(os.path.join(d, "meta.txt")))
    out = {"meta": meta}
    for name, t in [("counts.u64", np.uint64), ("v.bin", dt), ("expected.bin", dt), ("z.bin", dt), ("logp.bin", dt),
                    ("seeds.u64", np.uint64), ("adv_pwm.bin", dt), ("em_pwm.bin", dt)]:
        p = os.path.join(d, name)
        if os.path.exists(p):
            out[name.split(".")[0]] = np.fromfile(p, t)
    k = 0
    while os.path.exists(os.path.join(d, f"bgprob_k{k}.bin")):
        out[f"bgprob_k{k}"] = np.fromfile(os.path.join(d, f"bgprob_k{k}.bin"), dt)
        k += 1
    p = os.path.join(d, "iupac.txt")
    if os.path.exists(p):
        out["iupac_rows"] = [l.split() for l in open(p)]
    return out


def config_fixture(name, fasta, W, strand, full):
    arrays = {}
    for suf, binary, dt in [("f32", "ref_dump", np.float32), ("f64", "ref_dump64", np.float64)]:
        d, r = dump(fasta, W, strand, binary)
        assert r.returncode == 0, r.stderr
        t = load_dump(d, dt)
        m = t["meta"]
        if suf == "f32":
            arrays["meta"] = np.array(json.dumps({k: m[k] for k in ("n_seq", "min_len", "max_len", "W", "strand", "ltot",
                                                                   "n_patterns", "k", "max_k")}))
            c = t["counts"]
            nz = np.nonzero(c)[0]
            arrays["counts_nz_ids"] = nz.astype(np.uint32)
            arrays["counts_nz_vals"] = c[nz].astype(np.uint32)
            arrays["counts_sha256"] = np.array(sha(c))
            arrays["seeds"] = t["seeds"]
        arrays[f"v_{suf}"] = t["v"]
        for key in ["bgprob_k0", "bgprob_k1", "bgprob_k2", "expected", "z", "logp"]:
            if key in t:
                arrays[f"{key}_{suf}_sha256"] = np.array(sha(t[key]))
                if full:
                    arrays[f"{key}_{suf}"] = t[key]
                else:  # keep the values at the seeds and at a fixed pseudo-random sample of ids
                    rs = np.random.RandomState(1234)
                    idx = np.unique(np.concatenate([t["seeds"].astype(np.int64),
                                                    rs.randint(0, len(t[key]), 4096)]))
                    arrays["sample_ids"] = idx.astype(np.uint32)
                    arrays[f"{key}_{suf}_sample"] = t[key][idx]
        arrays[f"seeds_{suf}"] = t["seeds"]
        rows = t.get("iupac_rows", [])
        arrays[f"iupac_ids_{suf}"] = np.array([int(r[0]) for r in rows], np.uint64)
        arrays[f"iupac_sites_{suf}"] = np.array([int(r[2]) for r in rows], np.uint64)
        arrays[f"iupac_attrs_{suf}"] = np.array([[float(x) for x in r[3:7]] for r in rows], dt).reshape(len(rows), 4)
        arrays[f"adv_pwm_{suf}"] = t.get("adv_pwm", np.zeros(0, dt)).reshape(-1, W, 4)
        arrays[f"em_pwm_{suf}"] = t.get("em_pwm", np.zeros(0, dt)).reshape(-1, W, 4)
        shutil.rmtree(d)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **arrays)
    # console report + MEME + JSON of the reference CLI (float32 arm)
    args = [os.path.join(BIN, "peng_motif_ref"), fasta, "-w", str(W), "--strand", strand]
    meme = os.path.join(HERE, name + ".meme")
    r = subprocess.run(args + ["-o", meme], capture_output=True, text=True)
    assert r.returncode == 0
    open(os.path.join(HERE, name + "_stdout.txt"), "w").write(r.stdout)
    js = os.path.join(HERE, name + ".json")
    r = subprocess.run(args + ["-j", js], capture_output=True, text=True)
    assert r.returncode == 0
    print("wrote", name)


def case_from_text(text, W, strand, tag):
    """Run the reference on a literal FASTA text and record what it produced."""
    with tempfile.NamedTemporaryFile("w", suffix=".fa", delete=False) as f:
        f.write(text)
        path = f.name
    d, r = dump(path, W, strand, "ref_dump", ("--tables-only",))
    case = {"tag": tag, "fasta": text, "W": W, "strand": strand, "exit": r.returncode}
    if r.returncode == 0 and os.path.exists(os.path.join(d, "z.bin")):
        t = load_dump(d, np.float32)
        m = t["meta"]
        c = t["counts"]
        nz = np.nonzero(c)[0]
        case.update(n_seq=int(m["n_seq"]), min_len=int(m["min_len"]), max_len=int(m["max_len"]), ltot=int(m["ltot"]),
                    counts={int(i): int(c[i]) for i in nz},
                    v_bits=[int(x) for x in t["v"].view(np.uint32)],
                    seeds=[int(x) for x in t["seeds"]],
                    z_sha256=sha(t["z"]), expected_sha256=sha(t["expected"]), logp_sha256=sha(t["logp"]),
                    bgprob_sha256=sha(t[f"bgprob_k{m['k']}"]))
    os.unlink(path)
    shutil.rmtree(d)
    return case


def random_fasta(rng):
    def rand_seq():
        L = rng.choice([0, 1, 2, 3, 5, 8, 9, 10, 11, 12, 20, 33, 64, 100, 300])
        at_rich = rng.random() < 0.2
        s = []
        while len(s) < L:
            r = rng.random()
            if r < 0.05:
                s.append("N")
            elif r < 0.08:
                s += ["N"] * rng.randint(1, 3)
            elif r < 0.12:
                s += [rng.choice("ACGT")] * rng.randint(3, 30)
            elif r < 0.16:
                u = "".join(rng.choice("ACGT") for _ in range(rng.randint(1, 4)))
                s += list(u * rng.randint(2, 10))
            elif r < 0.18:
                s.append(rng.choice("acgtnRYx"))
            else:
                s.append(rng.choice("AT" if at_rich else "ACGT"))
        return "".join(s[:L])

    lines = []
    for i in range(rng.randint(1, 12)):
        lines.append(">" if rng.random() < 0.1 else f">s{i}")
        s = rand_seq()
        w = rng.choice([7, 60, 1000])
        lines += [s[k:k + w] for k in range(0, len(s), w)]
        if rng.random() < 0.1:
            lines.append("")
    return "\n".join(lines) + ("\n" if rng.random() < 0.85 else "")


def main():
    assert os.path.exists(os.path.join(BIN, "ref_dump")), "run `make -C oracle ref` first"
    os.makedirs(os.path.join(HERE, "data"), exist_ok=True)
    for src in ["data/MafK.fasta", "data/MafK_100seqs.fasta", "test/test_data/default_sequence_set.fa"]:
        shutil.copy(os.path.join(REF, src), os.path.join(HERE, "data", os.path.basename(src)))
    d = os.path.join(HERE, "data")
    config_fixture("cfg1_W8_BOTH", os.path.join(d, "MafK_100seqs.fasta"), 8, "BOTH", full=True)
    config_fixture("cfg1_W8_PLUS", os.path.join(d, "MafK_100seqs.fasta"), 8, "PLUS", full=True)
    config_fixture("cfg2_W10_BOTH", os.path.join(d, "MafK.fasta"), 10, "BOTH", full=False)

    cases = []
    hand = [
        # SURVEY.md §8c probes (greedy non-overlap rule, the "resume at n+2" rule, palindromes)
        ("homopolymer20", ">a\n" + "A" * 20 + "\n", 4, "PLUS"),
        ("homopolymer20_both", ">a\n" + "A" * 20 + "\n", 4, "BOTH"),
        ("n_skip", ">a\nACGTACGTNACGTACGT\n", 4, "PLUS"),
        ("n_skip_both", ">a\nACGTACGTNACGTACGT\n", 4, "BOTH"),
        ("palin_gap", ">a\nCCGGTTNNCCGGTT\n", 4, "BOTH"),
        ("bg_n_quirk", ">a\n" + "C" * 12 + "N" + "C" * 12 + "\n", 4, "BOTH"),
        ("no_trailing_newline", ">a\nACGTACGTAC\n>b\nTTTTGGGGCCCCAAAA", 4, "BOTH"),
        ("exact_W_chain", ">a\nACGTNACGTNACGTNACGTNACGTNACGTAA\n", 4, "PLUS"),
        ("exact_W_chain2", ">a\nACGTANACGTNACGTNACGTNACGTNACGTAA\n", 4, "PLUS"),
        ("short_seqs", ">a\nACG\n>b\nA\n>c\nACGT\n>d\nACGTA\n", 4, "BOTH"),
        ("lower_and_other", ">a\nacgtacgtRYacgtnnacgtacgt\n", 4, "BOTH"),
        ("crlf", ">a\r\nACGTACGTACGT\r\nACGTACGTACGT\r\n", 4, "BOTH"),
        ("empty_record", ">a\n>b\nACGTACGTAC\n>c\n", 4, "BOTH"),
        ("header_only_index", ">\nACGTTGCAACGT\n>\nGGGGCCCCAAAATTTT\n", 6, "BOTH"),
        ("tandem_both", ">a\n" + "ACGT" * 12 + "\n>b\n" + "AT" * 20 + "\n>c\n" + "GC" * 20 + "\n", 6, "BOTH"),
        ("tandem_plus", ">a\n" + "ACGT" * 12 + "\n>b\n" + "AT" * 20 + "\n>c\n" + "GC" * 20 + "\n", 6, "PLUS"),
        ("unit_test_set_W4", open(os.path.join(d, "default_sequence_set.fa")).read(), 4, "BOTH"),
        ("space_in_sequence", ">a\nACGT ACGT\n", 4, "BOTH"),
        ("sequence_before_header", "ACGT\n>a\nACGT\n", 4, "BOTH"),
    ]
    for tag, text, W, strand in hand:
        cases.append(case_from_text(text, W, strand, tag))
    rng = random.Random(20260101)
    for i in range(80):
        W = rng.choice([2, 4, 6, 8])
        strand = rng.choice(["BOTH", "PLUS"])
        c = case_from_text(random_fasta(rng), W, strand, f"random{i}")
        if c.get("n_seq", 0) > 0 or c["exit"] != 0:
            cases.append(c)
    json.dump(cases, open(os.path.join(HERE, "cases.json"), "w"), indent=0)
    print("wrote", len(cases), "cases")


if __name__ == "__main__":
    sys.exit(main())
