"""ctypes binding of oracle/libpeng_oracle.so — TEST INFRASTRUCTURE ONLY.

The oracle is the CPU restatement of the reference (oracle/peng_oracle.c); only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline legs may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ORACLE_DIR = os.path.join(os.path.dirname(_HERE), "oracle")
_LIB_PATH = os.path.join(_ORACLE_DIR, "libpeng_oracle.so")


def build():
    subprocess.check_call(["make", "-s", "-C", _ORACLE_DIR, "port"])


def _load():
    src_mtime = max(os.path.getmtime(os.path.join(_ORACLE_DIR, f))
                    for f in ("peng_oracle.c", "peng_oracle.h", "peng_oracle_select.cpp"))
    if not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < src_mtime:
        build()
    return C.CDLL(_LIB_PATH)


lib = _load()
u8p, u64p, i64p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint64), C.POINTER(C.c_int64)
f32p, f64p = C.POINTER(C.c_float), C.POINTER(C.c_double)


def _p(a, t):
    return a.ctypes.data_as(t)


lib.po_revcomp.restype = C.c_uint64
lib.po_revcomp.argtypes = [C.c_uint64, C.c_int]
lib.po_bg_id.restype = C.c_uint64
lib.po_bg_id.argtypes = [C.c_uint64, C.c_int, C.c_int]
lib.po_iupac_expansion_size.restype = C.c_uint64
lib.po_iupac_expansion_size.argtypes = [C.c_uint64, C.c_int]
lib.po_iupac_expand.restype = C.c_uint64
lib.po_iupac_expand.argtypes = [C.c_uint64, C.c_int, C.c_int, u64p, C.c_uint64]
lib.po_count_combined.restype = C.c_uint64
lib.po_count_combined.argtypes = [C.c_uint64, C.c_int, C.c_int, u64p]
lib.po_count.restype = None
lib.po_count.argtypes = [u8p, u64p, C.c_uint64, C.c_int, C.c_int, u64p, u64p]
lib.po_fasta_scan.argtypes = [C.c_char_p, u64p, u64p, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
lib.po_fasta_read.argtypes = [C.c_char_p, u8p, u64p]
lib.po_similar_letters.argtypes = [C.c_int, C.POINTER(C.c_int)]

IUPAC = "ACGTSWRYMKN"


def revcomp(i, W):
    return int(lib.po_revcomp(int(i), W))


def bg_id(pattern, length, k):
    return int(lib.po_bg_id(int(pattern), length, k))


def base_id(s):
    return sum("ACGT".index(c) << (2 * p) for p, c in enumerate(s))


def base_str(i, W):
    return "".join("ACGT"[(int(i) >> (2 * p)) & 3] for p in range(W))


def iupac_id(s):
    return sum(IUPAC.index(c) * 11 ** p for p, c in enumerate(s))


def iupac_str(i, W):
    return "".join(IUPAC[(int(i) // 11 ** p) % 11] for p in range(W))


def similar_letters(c):
    out = (C.c_int * 10)()
    n = lib.po_similar_letters(c, out)
    return list(out[:n])


class FastaError(Exception):
    pass


def read_fasta(path):
    """-> (codes u8[total], offsets u64[n+1], min_len, max_len); codes 0=other 1..4=ACGT."""
    n, tot = C.c_uint64(), C.c_uint64()
    mn, mx = C.c_uint32(), C.c_uint32()
    if lib.po_fasta_scan(path.encode(), C.byref(n), C.byref(tot), C.byref(mn), C.byref(mx)):
        raise FastaError(path)
    codes = np.zeros(max(tot.value, 1), np.uint8)
    offs = np.zeros(n.value + 1, np.uint64)
    if lib.po_fasta_read(path.encode(), _p(codes, u8p), _p(offs, u64p)):
        raise FastaError(path)
    return codes[: tot.value], offs, mn.value, mx.value


def encode(seqs):
    """list of str -> (codes, offsets) with the reference's letter codes."""
    tab = np.zeros(256, np.uint8)
    for i, ch in enumerate("ACGT"):
        tab[ord(ch)] = tab[ord(ch.lower())] = i + 1
    offs = np.zeros(len(seqs) + 1, np.uint64)
    parts = []
    for i, s in enumerate(seqs):
        a = np.frombuffer(s.encode("latin-1"), np.uint8)
        parts.append(tab[a])
        offs[i + 1] = offs[i] + np.uint64(len(a))
    codes = np.concatenate(parts) if parts else np.zeros(0, np.uint8)
    return np.ascontiguousarray(codes), offs


def count(codes, offsets, W, both=True):
    counts = np.zeros(4 ** W, np.uint64)
    ltot = C.c_uint64()
    codes = np.ascontiguousarray(codes, np.uint8)
    if codes.size == 0:
        codes = np.zeros(1, np.uint8)
    lib.po_count(_p(codes, u8p), _p(offsets, u64p), len(offsets) - 1, W, int(both), _p(counts, u64p),
                 C.byref(ltot))
    return counts, ltot.value


def iupac_expand(iid, W, both=True):
    n = lib.po_iupac_expansion_size(int(iid), W)
    out = np.zeros(n, np.uint64)
    lib.po_iupac_expand(int(iid), W, int(both), _p(out, u64p), n)
    return out


def count_combined(iid, W, both, counts):
    return int(lib.po_count_combined(int(iid), W, int(both), _p(counts, u64p)))


class _Typed:
    """The numeric half of the oracle in one precision ('f32' follows the reference bit for bit,
    'f64' is the same code with doubles)."""

    def __init__(self, suf):
        self.suf = suf
        self.dt = np.float32 if suf == "f32" else np.float64
        self.ct = C.c_float if suf == "f32" else C.c_double
        self.pt = f32p if suf == "f32" else f64p
        g = lambda name: getattr(lib, f"{name}_{suf}")
        pt, ct = self.pt, self.ct
        self._bgm = g("po_background_model"); self._bgm.restype = None
        self._bgm.argtypes = [u8p, u64p, C.c_uint64, C.c_int, pt, C.c_int, pt, i64p]
        self._bgp = g("po_bg_probabilities"); self._bgp.restype = None
        self._bgp.argtypes = [pt, C.c_int, C.c_int, pt]
        self._agg = g("po_aggregate_double_strand"); self._agg.restype = None
        self._agg.argtypes = [pt, C.c_int]
        self._stats = g("po_stats"); self._stats.restype = None
        self._stats.argtypes = [u64p, pt, C.c_uint64, C.c_int, pt, pt, pt]
        self._sel = g("po_select"); self._sel.restype = C.c_uint64
        self._sel.argtypes = [pt, u64p, C.c_int, ct, C.c_uint64, C.c_int, C.c_int, u64p, C.c_uint64]
        self._bmi = g("po_base_mi_score"); self._bmi.restype = ct
        self._bmi.argtypes = [C.c_uint64, ct, C.c_uint64]
        self._iagg = g("po_iupac_aggregate"); self._iagg.restype = None
        self._iagg.argtypes = [C.c_uint64, C.c_int, C.c_int, u64p, pt, pt, C.c_uint64, u64p, pt]
        self._adv = g("po_adv_pwm"); self._adv.restype = None
        self._adv.argtypes = [C.c_uint64, C.c_int, C.c_int, u64p, C.c_int, pt, pt]
        self._odds = g("po_prob_odds"); self._odds.restype = None
        self._odds.argtypes = [pt, C.c_int, pt, pt]
        self._em = g("po_em"); self._em.restype = C.c_int
        self._em.argtypes = [pt, C.c_int, u64p, pt, ct, ct, C.c_int, pt]
        self._emit = g("po_em_iteration"); self._emit.restype = None
        self._emit.argtypes = [pt, C.c_int, u64p, pt, ct, pt, pt]

    def background_model(self, codes, offsets, K=2, alpha=None, interpolate=True):
        alpha = np.ones(K + 1, self.dt) if alpha is None else np.asarray(alpha, self.dt)
        nv = (4 ** (K + 2) - 4) // 3
        v = np.zeros(nv, self.dt)
        n = np.zeros(nv, np.int64)
        codes = np.ascontiguousarray(codes, np.uint8)
        if codes.size == 0:
            codes = np.zeros(1, np.uint8)
        self._bgm(_p(codes, u8p), _p(offsets, u64p), len(offsets) - 1, K, _p(alpha, self.pt), int(interpolate),
                  _p(v, self.pt), _p(n, i64p))
        return v, n

    def bg_probabilities(self, v, W, order):
        out = np.zeros(4 ** W, self.dt)
        v = np.ascontiguousarray(v, self.dt)
        self._bgp(_p(v, self.pt), W, order, _p(out, self.pt))
        return out

    def aggregate_double_strand(self, bg, W):
        bg = np.array(bg, self.dt, copy=True)
        self._agg(_p(bg, self.pt), W)
        return bg

    def stats(self, counts, bgprob_k, ltot, W):
        n = 4 ** W
        e, z, lp = (np.zeros(n, self.dt) for _ in range(3))
        self._stats(_p(counts, u64p), _p(bgprob_k, self.pt), int(ltot), W, _p(e, self.pt), _p(z, self.pt),
                    _p(lp, self.pt))
        return e, z, lp

    def select(self, z, counts, W, z_thr=10.0, count_thr=3, single_stranded=False, filter_neighbors=True):
        cap = 4 ** W
        out = np.zeros(cap, np.uint64)
        n = self._sel(_p(z, self.pt), _p(counts, u64p), W, z_thr, count_thr, int(single_stranded),
                      int(filter_neighbors), _p(out, u64p), cap)
        return out[:n].copy()

    def base_mi_score(self, count, expected, n_seq):
        return self.dt(self._bmi(int(count), float(expected), int(n_seq)))

    def iupac_aggregate(self, iid, W, both, counts, bgprob_k, expected, n_seq):
        ns = C.c_uint64()
        out = np.zeros(5, self.dt)
        self._iagg(int(iid), W, int(both), _p(counts, u64p), _p(bgprob_k, self.pt), _p(expected, self.pt),
                   int(n_seq), C.byref(ns), _p(out, self.pt))
        return dict(n_sites=ns.value, bg_p=out[0], expected=out[1], z=out[2], logp=out[3], mi=out[4])

    def adv_pwm(self, iid, W, both, counts, v0, pseudo=10):
        pwm = np.zeros((W, 4), self.dt)
        v0 = np.ascontiguousarray(v0, self.dt)
        self._adv(int(iid), W, int(both), _p(counts, u64p), pseudo, _p(v0, self.pt), _p(pwm, self.pt))
        return pwm

    def prob_odds(self, pwm, W, bgprob):
        pwm = np.ascontiguousarray(pwm, self.dt)
        odds = np.zeros(4 ** W, self.dt)
        self._odds(_p(pwm, self.pt), W, _p(bgprob, self.pt), _p(odds, self.pt))
        return odds

    def em(self, pwm, W, counts, bgprob, saturation=1e4, thr=0.08, max_iter=10):
        pwm = np.ascontiguousarray(pwm, self.dt)
        out = np.zeros((W, 4), self.dt)
        it = self._em(_p(pwm, self.pt), W, _p(counts, u64p), _p(bgprob, self.pt), saturation, thr, max_iter,
                      _p(out, self.pt))
        return out, it

    def em_iteration(self, pwm, W, counts, bgprob, saturation=1e4):
        pwm = np.ascontiguousarray(pwm, self.dt)
        out = np.zeros((W, 4), self.dt)
        odds = np.zeros(4 ** W, self.dt)
        self._emit(_p(pwm, self.pt), W, _p(counts, u64p), _p(bgprob, self.pt), saturation, _p(out, self.pt),
                   _p(odds, self.pt))
        return out


f32 = _Typed("f32")
f64 = _Typed("f64")
