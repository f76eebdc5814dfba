"""peng_motif_b200 — Python plumbing around libpeng_cuda.so (the sm_100a hot path of PEnG-motif).

The product is the C-ABI library (include/peng_cuda.h, csrc/) plus the C++ host driver (host/); this module
only binds the C ABI with ctypes for tests and bench.py and mirrors the reference's class seam
(`BasePattern` ctor + getters, src/base_pattern.h:36-80 of the reference) closely enough that the parity
tests read like the reference's own.  There is NO CPU fallback: if the library is missing or no CUDA device
is present every entry point raises.

The directory is called `peng-motif_b200` (not importable by name); load it with `_pkg.load()` at the repo
root, which registers it as module `peng_motif_b200`.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpeng_cuda.so")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "peng_cuda.h")

STRAND_PLUS, STRAND_BOTH = 0, 1
F32REF, F64 = 0, 1
TABLE_BGPROB, TABLE_EXPECTED, TABLE_ZSCORE, TABLE_LOGP = 0, 1, 2, 3


class PengCudaError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libpeng_cuda error {code}: {msg}")
        self.code = code


def build(verbose=False):
    """Compile libpeng_cuda.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-C", _HERE], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("building libpeng_cuda.so failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stdout)
    return LIB_PATH


_lib = None
u8p, u32p, u64p, i64p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_int64)
f32p, f64p, ip = C.POINTER(C.c_float), C.POINTER(C.c_double), C.POINTER(C.c_int)
H = C.c_void_p

_PROTOS = {
    "peng_cuda_device_count": (C.c_int, []),
    "peng_cuda_create": (C.c_int, [C.c_int, C.POINTER(H)]),
    "peng_cuda_destroy": (None, [H]),
    "peng_cuda_last_error": (C.c_char_p, [H]),
    "peng_cuda_launch_count": (C.c_uint64, [H]),
    "peng_cuda_set_stream": (C.c_int, [H, C.c_void_p]),
    "peng_cuda_synchronize": (C.c_int, [H]),
    "peng_cuda_enable_timing": (C.c_int, [H, C.c_int]),
    "peng_cuda_get_timing": (C.c_int, [H, C.c_int, f32p]),
    "peng_cuda_set_ingest_mode": (C.c_int, [H, C.c_int, C.c_double]),
    "peng_cuda_last_ingest_host_packed": (C.c_uint64, [H]),
    "peng_cuda_last_ingest_h2d_bytes": (C.c_uint64, [H]),
    "peng_cuda_ingest_fasta": (C.c_int, [H, C.c_char_p, u64p, u64p, u32p, u32p]),
    "peng_cuda_ingest_fasta_shard": (C.c_int, [H, C.c_char_p, C.c_int, C.c_int, u64p, u64p, u32p, u32p]),
    "peng_cuda_background_from_counts": (C.c_int, [C.c_int, f32p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "peng_cuda_ingest_codes": (C.c_int, [H, C.c_void_p, u64p, C.c_uint64]),
    "peng_cuda_ingest_codes_device": (C.c_int, [H, C.c_void_p, u64p, C.c_uint64]),
    "peng_cuda_ingest_sequences": (C.c_int, [H, C.c_void_p, C.c_void_p, C.c_uint64]),
    "peng_cuda_ingest_synthetic": (C.c_int, [H, C.c_uint64, C.c_uint32, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint32,
                                             C.c_double, C.c_double]),
    "peng_cuda_sequences_info": (C.c_int, [H, u64p, u64p, u32p, u32p]),
    "peng_cuda_export_codes": (C.c_int, [H, C.c_void_p, C.c_void_p]),
    "peng_cuda_background_model": (C.c_int, [H, C.c_int, f32p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "peng_cuda_background_counts": (C.c_int, [H, C.c_int]),
    "peng_cuda_background_finish": (C.c_int, [H, C.c_int, f32p, C.c_int, C.c_int, C.c_void_p]),
    "peng_cuda_count": (C.c_int, [H, C.c_int, C.c_int, u64p]),
    "peng_cuda_get_counts": (C.c_int, [H, C.c_void_p]),
    "peng_cuda_get_counts32": (C.c_int, [H, C.c_void_p]),
    "peng_cuda_set_count_prefetch": (C.c_int, [H, C.c_int, C.c_int]),
    "peng_cuda_count_local": (C.c_int, [H, C.c_int, C.c_int]),
    "peng_cuda_set_count_path": (C.c_int, [H, C.c_int]),
    "peng_cuda_last_count_path": (C.c_int, [H]),
    "peng_cuda_count_buffer": (C.c_int, [H, C.POINTER(C.c_void_p), u64p]),
    "peng_cuda_count_finish": (C.c_int, [H, u64p]),
    "peng_cuda_count_local32": (C.c_int, [H, C.c_int, C.c_int]),
    "peng_cuda_count_buffer32": (C.c_int, [H, C.POINTER(C.c_void_p), u64p]),
    "peng_cuda_count_finish32": (C.c_int, [H, u64p]),
    "peng_cuda_peer_export": (C.c_int, [H, C.c_int, C.c_void_p]),
    "peng_cuda_peer_attach": (C.c_int, [H, C.c_int, C.c_int, C.c_void_p]),
    "peng_cuda_peer_detach": (C.c_int, [H]),
    "peng_cuda_count_exchange32": (C.c_int, [H, u64p]),
    "peng_cuda_peer_trace": (C.c_int, [H, f64p]),
    "peng_cuda_set_table_slice": (C.c_int, [H, C.c_int, C.c_int]),
    "peng_cuda_table_slice": (C.c_int, [H, u64p, u64p]),
    "peng_cuda_count_finish32_sliced": (C.c_int, [H, u64p]),
    "peng_cuda_gather_counts": (C.c_int, [H, C.c_void_p, C.c_uint64, C.c_void_p]),
    "peng_cuda_counts_total": (C.c_int, [H, u64p]),
    "peng_cuda_set_counts": (C.c_int, [H, C.c_int, C.c_int, C.c_void_p, C.c_uint64]),
    "peng_cuda_stats": (C.c_int, [H, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    "peng_cuda_get_table": (C.c_int, [H, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "peng_cuda_gather_table": (C.c_int, [H, C.c_int, C.c_int, C.c_int, u64p, C.c_uint64, C.c_void_p]),
    "peng_cuda_select_base_patterns": (C.c_int, [H, C.c_float, C.c_uint64, C.c_int, C.c_int, C.c_void_p, C.c_uint64, u64p]),
    "peng_cuda_iupac_aggregate": (C.c_int, [H, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]),
    "peng_cuda_iupac_count": (C.c_int, [H, C.c_void_p, C.c_int, C.c_void_p]),
    "peng_cuda_em": (C.c_int, [H, C.c_int, C.c_void_p, C.c_int, C.c_double, C.c_double, C.c_int, C.c_int, C.c_void_p,
                               C.c_void_p]),
    "peng_cuda_em_begin": (C.c_int, [H, C.c_int, C.c_void_p, C.c_double, C.c_double, C.c_int, C.c_int, C.c_int, C.c_int,
                                     C.POINTER(C.c_void_p)]),
    "peng_cuda_em_accumulate": (C.c_int, [H]),
    "peng_cuda_em_update": (C.c_int, [H, ip]),
    "peng_cuda_em_end": (C.c_int, [H, C.c_void_p, C.c_void_p]),
}


def load_library():
    """dlopen the in-tree libpeng_cuda.so and declare every prototype of include/peng_cuda.h."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PengCudaError(-2, f"{LIB_PATH} is missing: run __graft_entry__.build() (there is no CPU fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def exported_symbols():
    return sorted(_PROTOS)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


# --------------------------------------------------------------------------------------------------------
# host-side helpers (encodings of the reference, sharding)
# --------------------------------------------------------------------------------------------------------
IUPAC_LETTERS = "ACGTSWRYMKN"


def base_id(s):
    """'ATGC' -> id with position 0 least significant (src/base_pattern.h:24-29 of the reference)."""
    return sum("ACGT".index(c) << (2 * p) for p, c in enumerate(s))


def base_str(i, W):
    return "".join("ACGT"[(int(i) >> (2 * p)) & 3] for p in range(W))


def iupac_id(s):
    return sum(IUPAC_LETTERS.index(c) * 11 ** p for p, c in enumerate(s))


def iupac_str(i, W):
    return "".join(IUPAC_LETTERS[(int(i) // 11 ** p) % 11] for p in range(W))


def shard_sequences(offsets, rank, world):
    """Contiguous block of sequences for `rank`, balanced by bases: returns (first, last) with
    sequences [first, last) owned by the rank.  Every sequence is an independent unit for counting, ltot and
    the background (k+1)-mer counts, so the per-rank tables simply add up."""
    offsets = np.asarray(offsets, np.uint64)
    n = len(offsets) - 1
    total = int(offsets[-1]) if n else 0
    cuts = [0]
    for r in range(1, world):
        target = total * r // world
        cuts.append(int(np.searchsorted(offsets, np.uint64(target), side="left")))
    cuts.append(n)
    cuts = [min(max(c, 0), n) for c in cuts]
    for i in range(1, len(cuts)):
        cuts[i] = max(cuts[i], cuts[i - 1])
    return cuts[rank], cuts[rank + 1]


def shard_synthetic(n_seq, rank, world):
    """Equal split of a fixed-length synthetic set: (first_seq, n_local)."""
    first = n_seq * rank // world
    last = n_seq * (rank + 1) // world
    return first, last - first


class DeviceArray:
    """A raw device pointer exposed through __cuda_array_interface__ so that torch can alias it
    (torch.as_tensor(DeviceArray(...), device='cuda')) for NCCL collectives — no copy."""

    def __init__(self, ptr, n, typestr="<u8"):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr, "data": (int(ptr), False), "version": 2}


# --------------------------------------------------------------------------------------------------------
# handle wrapper
# --------------------------------------------------------------------------------------------------------
class PengCuda:
    """One handle = one GPU.  Mirrors the seams of the reference: ingest (SequenceSet), background model,
    BasePattern tables, IUPAC sums, EM."""

    def __init__(self, device=0):
        self.lib = load_library()
        h = H()
        rc = self.lib.peng_cuda_create(device, C.byref(h))
        if rc != 0:
            raise PengCudaError(rc, self.lib.peng_cuda_last_error(None).decode())
        self.h = h
        self.W = None
        self.strand = None
        self.n_seq = 0

    def close(self):
        if getattr(self, "h", None):
            self.lib.peng_cuda_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise PengCudaError(rc, self.lib.peng_cuda_last_error(self.h).decode())

    @property
    def launches(self):
        return int(self.lib.peng_cuda_launch_count(self.h))

    def set_stream(self, cuda_stream_ptr):
        self._ck(self.lib.peng_cuda_set_stream(self.h, cuda_stream_ptr))

    def synchronize(self):
        self._ck(self.lib.peng_cuda_synchronize(self.h))

    def enable_timing(self, on=True):
        self._ck(self.lib.peng_cuda_enable_timing(self.h, int(on)))

    def get_timing(self, phase):
        ms = C.c_float()
        self._ck(self.lib.peng_cuda_get_timing(self.h, phase, C.byref(ms)))
        return ms.value

    # ---- (a) ingest ----
    def ingest_fasta(self, path):
        n, tot, mn, mx = C.c_uint64(), C.c_uint64(), C.c_uint32(), C.c_uint32()
        self._ck(self.lib.peng_cuda_ingest_fasta(self.h, os.fsencode(path), C.byref(n), C.byref(tot), C.byref(mn), C.byref(mx)))
        self.n_seq = n.value
        return dict(n_seq=n.value, total_bases=tot.value, min_len=mn.value, max_len=mx.value)

    def ingest_fasta_shard(self, path, part, n_parts):
        n, tot, mn, mx = C.c_uint64(), C.c_uint64(), C.c_uint32(), C.c_uint32()
        self._ck(self.lib.peng_cuda_ingest_fasta_shard(self.h, os.fsencode(path), part, n_parts, C.byref(n), C.byref(tot), C.byref(mn),
                                                       C.byref(mx)))
        self.n_seq = n.value
        return dict(n_seq=n.value, total_bases=tot.value, min_len=mn.value, max_len=mx.value)

    def ingest_codes(self, codes, offsets):
        codes = np.ascontiguousarray(codes, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        self._ck(self.lib.peng_cuda_ingest_codes(self.h, _ptr(codes), offsets.ctypes.data_as(u64p), len(offsets) - 1))
        self.n_seq = len(offsets) - 1

    def ingest_sequences(self, seq_ptrs, lengths):
        """seq_ptrs: uint64 array of host addresses (one per sequence, reference letter codes), lengths: uint32 array."""
        seq_ptrs = np.ascontiguousarray(seq_ptrs, np.uint64)
        lengths = np.ascontiguousarray(lengths, np.uint32)
        self._ck(self.lib.peng_cuda_ingest_sequences(self.h, _ptr(seq_ptrs), _ptr(lengths), len(lengths)))
        self.n_seq = len(lengths)

    def set_ingest_mode(self, mode=0, host_fraction=-1.0):
        """0 automatic (hybrid host / device packing), 1 raw bytes, 2 host packing only."""
        self._ck(self.lib.peng_cuda_set_ingest_mode(self.h, int(mode), float(host_fraction)))

    @property
    def last_ingest_host_packed(self):
        return int(self.lib.peng_cuda_last_ingest_host_packed(self.h))

    @property
    def last_ingest_h2d_bytes(self):
        return int(self.lib.peng_cuda_last_ingest_h2d_bytes(self.h))

    def ingest_codes_device(self, device_ptr, offsets):
        offsets = np.ascontiguousarray(offsets, np.uint64)
        self._ck(self.lib.peng_cuda_ingest_codes_device(self.h, device_ptr, offsets.ctypes.data_as(u64p), len(offsets) - 1))
        self.n_seq = len(offsets) - 1

    def ingest_synthetic(self, n_seq, seq_len, seed=1234, first_seq=0, motif=None, plant_fraction=0.0, n_fraction=0.0):
        """motif: string over ACGT plus at most one 'S' (C or G with p = 1/2), e.g. 'TGCTGASTCAGCA'."""
        m = None
        if motif:
            m = np.array(["ACGTS".index(c) for c in motif], np.uint8)
        self._ck(self.lib.peng_cuda_ingest_synthetic(self.h, n_seq, seq_len, seed, first_seq, _ptr(m), len(m) if m is not None else 0,
                                                     plant_fraction, n_fraction))
        self.n_seq = n_seq

    def sequences_info(self):
        n, tot, mn, mx = C.c_uint64(), C.c_uint64(), C.c_uint32(), C.c_uint32()
        self._ck(self.lib.peng_cuda_sequences_info(self.h, C.byref(n), C.byref(tot), C.byref(mn), C.byref(mx)))
        return dict(n_seq=n.value, total_bases=tot.value, min_len=mn.value, max_len=mx.value)

    def export_codes(self):
        info = self.sequences_info()
        codes = np.zeros(max(info["total_bases"], 1), np.uint8)
        offsets = np.zeros(info["n_seq"] + 1, np.uint64)
        self._ck(self.lib.peng_cuda_export_codes(self.h, _ptr(codes), _ptr(offsets)))
        return codes[: info["total_bases"]], offsets

    # ---- background model ----
    def background_model(self, K=2, alpha=None, interpolate=True, precision=F32REF, return_counts=False):
        alpha = np.ones(K + 1, np.float32) if alpha is None else np.ascontiguousarray(alpha, np.float32)
        nv = (4 ** (K + 2) - 4) // 3
        v = np.zeros(nv, np.float32 if precision == F32REF else np.float64)
        n = np.zeros(nv, np.int64)
        self._ck(self.lib.peng_cuda_background_model(self.h, K, alpha.ctypes.data_as(f32p), int(interpolate), precision,
                                                     _ptr(v), _ptr(n)))
        return (v, n) if return_counts else v

    # ---- (b) counting ----
    def count(self, W, strand=STRAND_BOTH, want_ltot=True):
        """want_ltot=False: nothing is read back, nothing waits (count_finish() fetches ltot later)."""
        ltot = C.c_uint64()
        self._ck(self.lib.peng_cuda_count(self.h, W, strand, C.byref(ltot) if want_ltot else None))
        self.W, self.strand = W, strand
        return ltot.value if want_ltot else None

    # peer group: fused cross-GPU sum + fold over NVLink peer memory
    PEER_BLOB_BYTES = 256

    def peer_export(self, W):
        blob = np.zeros(self.PEER_BLOB_BYTES, np.uint8)
        self._ck(self.lib.peng_cuda_peer_export(self.h, W, _ptr(blob)))
        self.W = W
        return blob

    def peer_attach(self, rank, world, blobs):
        blobs = np.ascontiguousarray(blobs, np.uint8).reshape(world, self.PEER_BLOB_BYTES)
        self._ck(self.lib.peng_cuda_peer_attach(self.h, rank, world, _ptr(blobs)))

    def peer_detach(self):
        self._ck(self.lib.peng_cuda_peer_detach(self.h))

    def peer_trace(self):
        out = np.zeros(5, np.float64)
        self._ck(self.lib.peng_cuda_peer_trace(self.h, out.ctypes.data_as(f64p)))
        return out

    def count_exchange32(self, want_ltot=True):
        ltot = C.c_uint64()
        self._ck(self.lib.peng_cuda_count_exchange32(self.h, C.byref(ltot) if want_ltot else None))
        return ltot.value if want_ltot else None

    def set_count_path(self, path):
        """0 auto, 1 L2-atomic kernel, 2 partition path (W >= 10), 3 partition path with L2-RED second stage (W >= 12)."""
        self._ck(self.lib.peng_cuda_set_count_path(self.h, int(path)))

    @property
    def last_count_path(self):
        return int(self.lib.peng_cuda_last_count_path(self.h))

    def count_local(self, W, strand=STRAND_BOTH):
        self._ck(self.lib.peng_cuda_count_local(self.h, W, strand))
        self.W, self.strand = W, strand

    def count_buffer(self):
        p, n = C.c_void_p(), C.c_uint64()
        self._ck(self.lib.peng_cuda_count_buffer(self.h, C.byref(p), C.byref(n)))
        return DeviceArray(p.value, n.value, "<i8")  # uint64 counts viewed as int64 (torch / NCCL sum is the same)

    def count_finish(self):
        ltot = C.c_uint64()
        self._ck(self.lib.peng_cuda_count_finish(self.h, C.byref(ltot)))
        return ltot.value

    # 32-bit exchange: sets of < 2^32 positions in total over all ranks, <= 2^30 per rank
    def count_local32(self, W, strand=STRAND_BOTH):
        self._ck(self.lib.peng_cuda_count_local32(self.h, W, strand))
        self.W, self.strand = W, strand

    def count_buffer32(self):
        p, n = C.c_void_p(), C.c_uint64()
        self._ck(self.lib.peng_cuda_count_buffer32(self.h, C.byref(p), C.byref(n)))
        return DeviceArray(p.value, n.value, "<i4")  # uint32 bins viewed as int32 (two's-complement sum is the same)

    def count_finish32(self):
        ltot = C.c_uint64()
        self._ck(self.lib.peng_cuda_count_finish32(self.h, C.byref(ltot)))
        return ltot.value

    # id-sliced tables (long pattern lengths on several GPUs)
    def set_table_slice(self, part=0, n_parts=1):
        self._ck(self.lib.peng_cuda_set_table_slice(self.h, int(part), int(n_parts)))
        self._sliced = n_parts > 1

    def table_slice(self):
        lo, hi = C.c_uint64(), C.c_uint64()
        self._ck(self.lib.peng_cuda_table_slice(self.h, C.byref(lo), C.byref(hi)))
        return lo.value, hi.value

    def count_finish32_sliced(self):
        ltot = C.c_uint64()
        self._ck(self.lib.peng_cuda_count_finish32_sliced(self.h, C.byref(ltot)))
        return ltot.value

    def _table_len(self):
        if getattr(self, "_sliced", False):
            lo, hi = self.table_slice()
            return hi - lo
        return 4 ** self.W

    def get_counts(self):
        out = np.zeros(4 ** self.W, np.uint64)
        self._ck(self.lib.peng_cuda_get_counts(self.h, _ptr(out)))
        return out

    def get_counts32(self):
        out = np.zeros(4 ** self.W, np.uint32)
        self._ck(self.lib.peng_cuda_get_counts32(self.h, _ptr(out)))
        return out

    def set_count_prefetch(self, W=0, strand=STRAND_BOTH):
        """Arm ingest_codes to run stage P of the count (W = 10 partition path) while the data is still arriving."""
        self._ck(self.lib.peng_cuda_set_count_prefetch(self.h, int(W), int(strand)))

    def gather_counts(self, ids):
        ids = np.ascontiguousarray(ids, np.uint64)
        out = np.zeros(len(ids), np.uint64)
        self._ck(self.lib.peng_cuda_gather_counts(self.h, _ptr(ids), len(ids), _ptr(out)))
        return out

    def counts_total(self):
        t = C.c_uint64()
        self._ck(self.lib.peng_cuda_counts_total(self.h, C.byref(t)))
        return t.value

    def set_counts(self, W, strand, counts, ltot):
        counts = np.ascontiguousarray(counts, np.uint64)
        assert counts.size == 4 ** W
        self._ck(self.lib.peng_cuda_set_counts(self.h, W, strand, _ptr(counts), int(ltot)))
        self.W, self.strand = W, strand

    # ---- (c) tables ----
    def stats(self, v32=None, v64=None, K=2, k=None, max_k=None):
        k = min(self.W - 1, K) if k is None else k
        max_k = k if max_k is None else max_k
        a32 = np.ascontiguousarray(v32, np.float32) if v32 is not None else None
        a64 = np.ascontiguousarray(v64, np.float64) if v64 is not None else None
        self._ck(self.lib.peng_cuda_stats(self.h, K, _ptr(a32), _ptr(a64), k, max_k))

    def get_table(self, table, precision=F32REF, order=0):
        out = np.zeros(self._table_len(), np.float32 if precision == F32REF else np.float64)
        self._ck(self.lib.peng_cuda_get_table(self.h, table, order, precision, _ptr(out)))
        return out

    def gather_table(self, table, ids, precision=F32REF, order=0):
        ids = np.ascontiguousarray(ids, np.uint64)
        out = np.zeros(len(ids), np.float32 if precision == F32REF else np.float64)
        self._ck(self.lib.peng_cuda_gather_table(self.h, table, order, precision, ids.ctypes.data_as(u64p), len(ids), _ptr(out)))
        return out

    def select_base_patterns(self, z_threshold=10.0, count_threshold=3, single_stranded=None, filter_neighbors=True):
        if single_stranded is None:
            single_stranded = self.strand == STRAND_PLUS
        cap = 4 ** self.W
        out = np.zeros(cap, np.uint64)
        n = C.c_uint64()
        self._ck(self.lib.peng_cuda_select_base_patterns(self.h, z_threshold, count_threshold, int(single_stranded),
                                                         int(filter_neighbors), _ptr(out), cap, C.byref(n)))
        return out[: n.value].copy()

    # ---- (d) IUPAC ----
    def iupac_aggregate(self, iupac_ids, precision=F32REF):
        ids = np.ascontiguousarray(iupac_ids, np.uint64)
        n = len(ids)
        dt = np.float32 if precision == F32REF else np.float64
        sites, bg, ex = np.zeros(n, np.uint64), np.zeros(n, dt), np.zeros(n, dt)
        self._ck(self.lib.peng_cuda_iupac_aggregate(self.h, _ptr(ids), n, precision, _ptr(sites), _ptr(bg), _ptr(ex)))
        return sites, bg, ex

    def iupac_count(self, iupac_ids):
        ids = np.ascontiguousarray(iupac_ids, np.uint64)
        out = np.zeros(len(ids), np.uint64)
        self._ck(self.lib.peng_cuda_iupac_count(self.h, _ptr(ids), len(ids), _ptr(out)))
        return out

    # ---- (e) EM ----
    def em(self, pwms, precision=F32REF, saturation=1e4, threshold=0.08, max_iterations=10, bg_order=None):
        dt = np.float32 if precision == F32REF else np.float64
        pw = np.ascontiguousarray(pwms, dt).reshape(-1, self.W, 4)
        out = np.zeros_like(pw)
        its = np.zeros(len(pw), np.int32)
        bg_order = min(self.W - 1, 2) if bg_order is None else bg_order
        self._ck(self.lib.peng_cuda_em(self.h, len(pw), _ptr(pw), precision, saturation, threshold, max_iterations, bg_order,
                                       _ptr(out), _ptr(its)))
        return out, its

    def em_begin(self, pwms, saturation=1e4, threshold=0.08, max_iterations=10, bg_order=None, part=0, n_parts=1):
        pw = np.ascontiguousarray(pwms, np.float64).reshape(-1, self.W, 4)
        bg_order = min(self.W - 1, 2) if bg_order is None else bg_order
        p = C.c_void_p()
        self._ck(self.lib.peng_cuda_em_begin(self.h, len(pw), _ptr(pw), saturation, threshold, max_iterations, bg_order, part,
                                             n_parts, C.byref(p)))
        self._em_n = len(pw)
        return DeviceArray(p.value, len(pw) * self.W * 4, "<f8")

    def em_accumulate(self):
        self._ck(self.lib.peng_cuda_em_accumulate(self.h))

    def em_update(self, want_active=True):
        n = C.c_int(-1)
        self._ck(self.lib.peng_cuda_em_update(self.h, C.byref(n) if want_active else None))
        return n.value

    def em_end(self):
        out = np.zeros((self._em_n, self.W, 4), np.float64)
        its = np.zeros(self._em_n, np.int32)
        self._ck(self.lib.peng_cuda_em_end(self.h, _ptr(out), _ptr(its)))
        return out, its
