"""ctypes binding of tests/sim/sim_kernels.cpp — TEST INFRASTRUCTURE ONLY.

The per-thread CUDA kernel bodies compiled as host code and driven thread by thread (see the header of
sim_kernels.cpp).  Used by `-m "not gpu"` tests to check kernel logic without a GPU.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "sim", "sim_kernels.cpp")
_LIB = os.path.join(_HERE, "sim", "libpeng_sim.so")
_CSRC = os.path.join(os.path.dirname(_HERE), "peng-motif_b200", "csrc")
FRONT_PAD = 64


def _build():
    deps = [_SRC, os.path.join(_CSRC, "core_seq.cuh"), os.path.join(_CSRC, "core_tab.cuh"), os.path.join(_CSRC, "host_ingest.h"),
            os.path.join(_CSRC, "host_pack.cpp"), os.path.join(_CSRC, "host_pack.h")]
    if os.path.exists(_LIB) and os.path.getmtime(_LIB) >= max(os.path.getmtime(d) for d in deps):
        return
    subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-pthread", "-x", "c++",
                           _SRC, "-o", _LIB])


_build()
lib = C.CDLL(_LIB)
u8p, u32p, u64p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64)
f32p, f64p = C.POINTER(C.c_float), C.POINTER(C.c_double)
lib.sim_n_words32.restype = C.c_uint64
lib.sim_n_words32.argtypes = [C.c_uint64, C.c_uint64]
lib.sim_mark.restype = C.c_uint64
lib.sim_iupac_list.restype = C.c_uint64
lib.sim_revcomp.restype = C.c_uint32
lib.sim_count_keys.restype = C.c_uint64


def _p(a, t):
    return a.ctypes.data_as(t)


class Stream:
    def __init__(self, n_seq, total):
        self.n_seq, self.total = n_seq, total
        self.nw = int(lib.sim_n_words32(C.c_uint64(n_seq), C.c_uint64(total)))
        self.codes = np.zeros(self.nw * 2 + 8, np.uint32)
        self.valid = np.zeros(self.nw + 8, np.uint32)
        self.sstart = np.zeros(self.nw + 8, np.uint32)
        self.kstart = np.zeros(self.nw + 8, np.uint32)
        self.stream_end = FRONT_PAD + total + 2 * n_seq


def pack(src, offsets, ascii=False):
    n_seq = len(offsets) - 1
    total = int(offsets[-1]) if n_seq else 0
    st = Stream(n_seq, total)
    src = np.ascontiguousarray(src, np.uint8)
    if src.size == 0:
        src = np.zeros(1, np.uint8)
    lib.sim_pack(_p(src, u8p), _p(offsets, u64p), C.c_uint64(n_seq), int(ascii), _p(st.codes, u32p), _p(st.valid, u32p),
                 _p(st.sstart, u32p))
    return st


def synth(n_seq, seq_len, seed, first_seq=0, motif=None, plant_fraction=0.0, n_fraction=0.0):
    st = Stream(n_seq, n_seq * seq_len)
    m = np.ascontiguousarray(motif if motif is not None else [], np.uint8)
    lib.sim_synth(C.c_uint64(n_seq), C.c_uint32(seq_len), C.c_uint64(seed), C.c_uint64(first_seq),
                  _p(m, u8p) if motif is not None else None, C.c_uint32(len(m)), C.c_double(plant_fraction),
                  C.c_double(n_fraction), _p(st.codes, u32p), _p(st.valid, u32p), _p(st.sstart, u32p))
    return st


def unpack(st, offsets):
    out = np.zeros(max(st.total, 1), np.uint8)
    lib.sim_unpack(_p(st.codes, u32p), _p(st.valid, u32p), _p(offsets, u64p), C.c_uint64(st.n_seq), _p(out, u8p))
    return out[: st.total]


def count(st, W, both=True, wpc=1, mirror=True):
    ltot = int(lib.sim_mark(_p(st.valid, u32p), C.c_uint64(st.nw), W, _p(st.kstart, u32p)))
    hist = np.zeros(4 ** W, np.uint64)
    lib.sim_count(_p(st.codes, u32p), _p(st.kstart, u32p), C.c_uint64(st.stream_end), W, int(both), C.c_uint32(wpc),
                  _p(hist, u64p))
    if both and mirror:
        lib.sim_mirror(_p(hist, u64p), W)
    return hist, ltot


def count_keys(st, W, both=True, wpc=1):
    """W in (12, 14, 16): the multiset of counted keys without allocating a 4^W table."""
    ltot = int(lib.sim_mark(_p(st.valid, u32p), C.c_uint64(st.nw), W, _p(st.kstart, u32p)))
    cap = st.stream_end
    out = np.zeros(cap, np.uint32)
    n = lib.sim_count_keys(_p(st.codes, u32p), _p(st.kstart, u32p), C.c_uint64(st.stream_end), W, int(both),
                           C.c_uint32(wpc), _p(out, u32p), C.c_uint64(cap))
    return out[:n], ltot


def bg_counts(st, K=2):
    out = np.zeros((4 ** (K + 2) - 4) // 3, np.uint64)
    lib.sim_bg_counts(_p(st.codes, u32p), _p(st.valid, u32p), _p(st.sstart, u32p), C.c_uint64(st.stream_end), K,
                      _p(out, u64p))
    return out


def bgprob(v, W, order, both, dt=np.float32):
    out = np.zeros(4 ** W, dt)
    v = np.ascontiguousarray(v, dt)
    if dt == np.float32:
        lib.sim_bgprob_f32(_p(v, f32p), W, order, int(both), _p(out, f32p))
    else:
        lib.sim_bgprob_f64(_p(v, f64p), W, order, int(both), _p(out, f64p))
    return out


def stats(counts, bg, ltot, W):
    dt = bg.dtype
    e, z, lp = (np.zeros(4 ** W, dt) for _ in range(3))
    if dt == np.float32:
        lib.sim_stats_f32(_p(counts, u64p), _p(bg, f32p), C.c_uint64(ltot), W, _p(e, f32p), _p(z, f32p), _p(lp, f32p))
    else:
        lib.sim_stats_f64(_p(counts, u64p), _p(bg, f64p), C.c_uint64(ltot), W, _p(e, f64p), _p(z, f64p), _p(lp, f64p))
    return e, z, lp


SETS = [0x1, 0x2, 0x4, 0x8, 0x6, 0x9, 0x5, 0xA, 0x3, 0xC, 0xF]


def iupac_mask(iid, W):
    m = 0
    for p in range(W):
        m |= SETS[iid % 11] << (4 * p)
        iid //= 11
    return m


def iupac_list(iid, W, mode):
    mask = iupac_mask(int(iid), W)
    n = 1
    for p in range(W):
        n *= bin((mask >> (4 * p)) & 15).count("1")
    out = np.zeros(n, np.uint32)
    k = lib.sim_iupac_list(C.c_uint64(mask), W, mode, _p(out, u32p))
    return out[:k]


def em_acc_f32(pwm, W, counts, bg, s=1e4):
    pwm = np.ascontiguousarray(pwm, np.float32)
    acc = np.zeros((W, 4), np.float32)
    lib.sim_em_acc_f32(_p(pwm, f32p), W, _p(counts, u64p), _p(bg, f32p), C.c_float(s), _p(acc, f32p))
    return acc


def count_part(st, W, both=True, mirror=True):
    """The partition path's fast/slow split (k_count_part + k_count_slow + k_fold_canon), thread by thread.
    -> (counts, ltot, stats) with stats = dict(missed, needless_slow, fast, slow_units)."""
    hist = np.zeros(4 ** W, np.uint64)
    stats = np.zeros(5, np.uint64)
    lib.sim_count_part(_p(st.codes, u32p), _p(st.valid, u32p), C.c_uint64(st.nw), W, int(both), _p(hist, u64p),
                       _p(stats, u64p))
    if both:
        lib.sim_fold_canon(_p(hist, u64p), W)
        if mirror:
            lib.sim_mirror(_p(hist, u64p), W)
    return hist, int(stats[0]), dict(missed=int(stats[1]), needless_slow=int(stats[2]), fast=int(stats[3]),
                                     slow_units=int(stats[4]))


def count_part_keys(st, W, both=True):
    """W in (12, 14, 16): the partition path's split without a 4^W table -> (sorted canonical keys, ltot, stats)."""
    stats = np.zeros(5, np.uint64)
    cap = st.stream_end
    out = np.zeros(cap, np.uint32)
    lib.sim_count_part_keys.restype = C.c_uint64
    n = lib.sim_count_part_keys(_p(st.codes, u32p), _p(st.valid, u32p), C.c_uint64(st.nw), W, int(both), _p(out, u32p),
                                C.c_uint64(cap), _p(stats, u64p))
    return np.sort(out[:n]), int(stats[0]), dict(missed=int(stats[1]), needless_slow=int(stats[2]), fast=int(stats[3]),
                                                 slow_units=int(stats[4]))


def cand_planes_mismatch(st, W, both=True):
    """Bit-plane conflict scan against the 2-bit one over every unit of the stream -> (units that differ, flags set)."""
    n_flagged = C.c_uint64()
    lib.sim_cand_planes_mismatch.restype = C.c_uint64
    bad = lib.sim_cand_planes_mismatch(_p(st.codes, u32p), C.c_uint64(st.nw), W, int(both), C.byref(n_flagged))
    return int(bad), int(n_flagged.value)


def host_pack(src, offsets, ascii=False, n_threads=4, max_blocks=1 << 14):
    """host_pack.cpp: the packed ingest's host half, all phases -> (codes words, valid words, phases) in the device layout."""
    src = np.ascontiguousarray(src, np.uint8)
    offsets = np.ascontiguousarray(offsets, np.uint64)
    n_seq = len(offsets) - 1
    nw = int(lib.sim_n_words32(C.c_uint64(n_seq), C.c_uint64(int(offsets[-1]))))
    codes = np.full(2 * nw + 8, 0xDEADBEEF, np.uint32)
    valid = np.full(nw + 8, 0xDEADBEEF, np.uint32)
    n_exc = C.c_uint64()
    ph = lib.sim_host_pack(_p(src, u8p), _p(offsets, u64p), C.c_uint64(n_seq), int(ascii), int(n_threads), C.c_uint64(max_blocks),
                           C.c_uint64(nw), _p(codes, u32p), _p(valid, u32p), C.byref(n_exc))
    host_pack.last_exceptions = int(n_exc.value)  # validity words sent explicitly (the rest is implied by the offsets)
    return codes, valid, nw, ph


def host_pack_concurrent(src, offsets, n_callers, n_threads, rounds, max_blocks, want_codes):
    """n_callers threads pack the set through ONE pool at the same time (two phases in flight each) -> callers that went wrong"""
    src = np.ascontiguousarray(src, np.uint8)
    offsets = np.ascontiguousarray(offsets, np.uint64)
    n_seq = len(offsets) - 1
    nw = int(lib.sim_n_words32(C.c_uint64(n_seq), C.c_uint64(int(offsets[-1]))))
    want = np.ascontiguousarray(want_codes[: 2 * nw], np.uint32)
    return int(lib.sim_host_pack_concurrent(_p(src, u8p), _p(offsets, u64p), C.c_uint64(n_seq), int(n_callers), int(n_threads), int(rounds),
                                            C.c_uint64(max_blocks), C.c_uint64(nw), _p(want, u32p)))


def scan_fasta(path):
    """host_ingest.h: scan_fasta on a mapped file -> (codes in the reference's encoding, offsets) or raises ValueError."""
    n_seq, n_let = C.c_uint64(), C.c_uint64()
    rc = lib.sim_scan_fasta(os.fsencode(path), C.byref(n_seq), C.byref(n_let), None, None)
    if rc:
        raise ValueError(f"scan_fasta rc={rc}")
    letters = np.zeros(max(n_let.value, 1), np.uint8)
    offsets = np.zeros(n_seq.value + 1, np.uint64)
    lib.sim_scan_fasta(os.fsencode(path), C.byref(n_seq), C.byref(n_let), letters.ctypes.data_as(C.c_char_p), _p(offsets, u64p))
    letters = letters[: n_let.value]
    lut = np.zeros(256, np.uint8)
    for i, ch in enumerate("ACGT"):
        lut[ord(ch)] = lut[ord(ch.lower())] = i + 1
    return lut[letters], offsets


def scan_fasta_chunked(path, n_chunks, part=0, n_parts=1):
    """host_ingest.h: ChunkedScan (the parallel ingest's scan, chunk by chunk) -> (raw letters, offsets, warnings, last record reported)"""
    n_seq, n_let, n_warn, last = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_int()
    args = (os.fsencode(path), int(n_chunks), int(part), int(n_parts), C.byref(n_seq), C.byref(n_let))
    rc = lib.sim_scan_fasta_chunked(*args, None, None, C.byref(n_warn), C.byref(last))
    if rc:
        raise ValueError(f"scan_fasta rc={rc}")
    letters = np.zeros(max(n_let.value, 1), np.uint8)
    offsets = np.zeros(n_seq.value + 1, np.uint64)
    lib.sim_scan_fasta_chunked(*args, letters.ctypes.data_as(C.c_char_p), _p(offsets, u64p), C.byref(n_warn), C.byref(last))
    return letters[: n_let.value], offsets, int(n_warn.value), bool(last.value)


def select(z, counts, W, z_threshold=10.0, count_threshold=3, single_stranded=False, filter_neighbors=True):
    """host_ingest.h: select_base_patterns (pruned introsort) -> selected ids"""
    z = np.ascontiguousarray(z, np.float32)
    counts = np.ascontiguousarray(counts, np.uint64)
    out = np.zeros(len(z), np.uint64)
    lib.sim_select.restype = C.c_uint64
    n = lib.sim_select(_p(z, f32p), _p(counts, u64p), C.c_int(W), C.c_float(z_threshold), C.c_uint64(count_threshold),
                       C.c_int(int(single_stranded)), C.c_int(int(filter_neighbors)), _p(out, u64p), C.c_uint64(len(out)))
    return out[:n].copy()


def sort_leading(z, threshold):
    """-> (guaranteed length, order left by the pruned sort, order left by std::sort)"""
    z = np.ascontiguousarray(z, np.float32)
    a, b = np.zeros(len(z), np.uint64), np.zeros(len(z), np.uint64)
    lib.sim_sort_leading.restype = C.c_uint64
    lead = lib.sim_sort_leading(_p(z, f32p), C.c_uint64(len(z)), C.c_float(threshold), _p(a, u64p), _p(b, u64p))
    return int(lead), a, b


def partition_level_agrees(z):
    """host_ingest.h: partition_level_lists (the device's list formulation of a partition pass) == libstdc++'s pass"""
    z = np.ascontiguousarray(z, np.float32)
    lib.sim_partition_level_agrees.restype = C.c_int
    return bool(lib.sim_partition_level_agrees(_p(z, f32p), C.c_uint64(len(z))))


def select_levels(z, counts, W, z_threshold=10.0, count_threshold=3, single_stranded=False, filter_neighbors=True, stop_below=64):
    """host_ingest.h: sort_leading_by_levels + select_walk with a host engine -> (selected ids, partition levels run)"""
    z = np.ascontiguousarray(z, np.float32)
    counts = np.ascontiguousarray(counts, np.uint64)
    out = np.zeros(len(z), np.uint64)
    lv = C.c_uint64()
    lib.sim_select_levels.restype = C.c_uint64
    n = lib.sim_select_levels(_p(z, f32p), _p(counts, u64p), C.c_int(W), C.c_float(z_threshold), C.c_uint64(count_threshold),
                              C.c_int(int(single_stranded)), C.c_int(int(filter_neighbors)), C.c_uint64(stop_below), _p(out, u64p),
                              C.c_uint64(len(out)), C.byref(lv))
    return out[:n].copy(), lv.value


def pack_nibbles(src, offsets, src_base, n_threads=3):
    """nibble_pack (host) + pack_word<2> (device body) for the words behind src_base -> (Stream, first word written)"""
    n_seq = len(offsets) - 1
    st = Stream(n_seq, int(offsets[-1]))
    src = np.ascontiguousarray(src, np.uint8)
    fw = C.c_uint64()
    lib.sim_pack_nibbles(_p(src, u8p), _p(offsets, u64p), C.c_uint64(n_seq), C.c_uint64(src_base), C.c_int(n_threads), _p(st.codes, u32p),
                         _p(st.valid, u32p), _p(st.sstart, u32p), C.byref(fw))
    return st, fw.value
