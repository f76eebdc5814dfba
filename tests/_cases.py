"""Shared test inputs: golden cases + seeded random sequence sets (used by oracle, sim and GPU tests alike)."""
import json
import os
import tempfile

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_cases():
    return json.load(open(os.path.join(GOLDEN, "cases.json")))


def case_counts(c):
    ref = np.zeros(4 ** c["W"], np.uint64)
    for k, v in c["counts"].items():
        ref[int(k)] = v
    return ref


def write_fasta(text):
    f = tempfile.NamedTemporaryFile("w", suffix=".fa", delete=False, newline="")
    f.write(text)
    f.close()
    return f.name


def dirty_sequences(seed, n_max=6, len_max=400):
    """Random sequences with N runs, homopolymers and tandem repeats: the inputs on which the greedy
    non-overlap rule and the resume-at-n+2 rule actually bite."""
    rng = np.random.RandomState(seed)
    seqs = []
    for _ in range(rng.randint(1, n_max + 1)):
        L = rng.randint(1, len_max)
        s = np.array(list("ACGT"))[rng.randint(0, 4, L)]
        for _ in range(rng.randint(0, 4)):
            p = rng.randint(0, L)
            s[p:p + rng.randint(1, 3)] = "N"
        for _ in range(rng.randint(0, 3)):
            p = rng.randint(0, L)
            s[p:p + rng.randint(5, 80)] = rng.choice(list("ACGT"))
        for _ in range(rng.randint(0, 3)):
            p = rng.randint(0, L)
            u = rng.randint(2, 6)
            rep = np.tile(s[p:p + u].copy(), rng.randint(2, 20))[: max(0, L - p)]
            s[p:p + len(rep)] = rep
        seqs.append("".join(s))
    return seqs


def random_iupac_ids(seed, W, n):
    rng = np.random.RandomState(seed)
    out = []
    for _ in range(n):
        letters = rng.choice(11, W, p=[.16, .16, .16, .16, .06, .06, .06, .06, .05, .04, .03])
        out.append(sum(int(c) * 11 ** p for p, c in enumerate(letters)))
    return out


def py_count_sparse(seqs, W, both):
    """Independent dict-based restatement of the counting rule for pattern lengths whose 4^W table does
    not fit a test box (W = 14, 16).  src/base_pattern.cpp:331-441 of the reference."""
    from collections import Counter

    comp = {"A": "T", "C": "G", "G": "C", "T": "A"}

    def bid(s):
        return sum("ACGT".index(c) << (2 * p) for p, c in enumerate(s))

    cnt, ltot = Counter(), 0
    for s in seqs:
        L, i, last = len(s), 0, {}
        while i < L:
            p = 0
            while p < W and i < L and s[i] in "ACGT":
                p += 1
                i += 1
            if p < W:
                i += 1
                continue
            while True:
                k = s[i - W:i]
                key = bid(k)
                if both:
                    key = min(key, bid("".join(comp[c] for c in reversed(k))))
                if key not in last or last[key] + W <= i:
                    cnt[key] += 1
                    last[key] = i
                ltot += 1
                if i >= L or s[i] not in "ACGT":
                    break
                i += 1
            i += 2
    return cnt, ltot
