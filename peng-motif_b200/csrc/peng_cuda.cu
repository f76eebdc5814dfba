// peng_cuda.cu — libpeng_cuda.so: __global__ wrappers + the C ABI of include/peng_cuda.h.
//
// sm_100a only.  No tensor cores: nothing on this path is a dense contraction (integer histogramming,
// element-wise table passes, small reductions).  The per-thread bodies live in core_seq.cuh / core_tab.cuh.
#include <cuda_runtime.h>

#include <unistd.h>

#include <algorithm>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <type_traits>
#include <memory>
#include <vector>

#include "../../include/peng_cuda.h"
#include "core_seq.cuh"
#include "core_tab.cuh"
#include "host_ingest.h"
#include "host_pack.h"

using namespace peng;

// =====================================================================================================
// context
// =====================================================================================================
#define PENG_MAX_ORDER 5
#define PENG_MAX_W 16

struct peng_cuda_ctx {
  int device = 0;
  cudaStream_t own_stream = nullptr, stream = nullptr;
  std::string err;
  uint64_t launches = 0;
  int sm_count = 148;

  // ---- sequences ----
  uint64_t n_seq = 0, total_bases = 0;
  uint32_t min_len = 0, max_len = 0;
  uint32_t fixed_len = 0;          // > 0: all sequences have this length (synthetic), offsets implicit
  uint64_t stream_end = 0;         // first position behind the last separator
  uint64_t n_words32 = 0;          // 32-position words allocated (incl. pads)
  uint32_t *d_codes = nullptr, *d_valid = nullptr, *d_sstart = nullptr, *d_kstart = nullptr;
  uint64_t* d_offsets = nullptr;   // n_seq+1 (base coordinates), unused when fixed_len
  size_t cap_words32 = 0, cap_offsets = 0;  // allocated capacities: re-ingesting a set of similar size reuses the buffers
  cudaStream_t copy_stream = nullptr;       // H2D of ingest chunks, overlapped with packing on `stream`
  cudaEvent_t copy_ev[2] = {nullptr, nullptr};
  uint8_t* pin[2] = {nullptr, nullptr};    // pinned staging buffers of the streaming FASTA ingest
  size_t pin_cap = 0;                       // their size: as large as the input needs, at most PENG_PIN_BYTES (pinning 2 x 16 MiB
                                            // costs ~10 ms: most of a 1 MB file's ingest)
  cudaEvent_t pin_ev[2] = {nullptr, nullptr};
  std::vector<uint64_t> h_offsets;
  bool offsets_in_pin = false;     // the host's copy of the offsets is pin_offs (hybrid ingest), not h_offsets
  uint64_t* pin_offs = nullptr;    // pinned copy of the offsets for the H2D copy (hybrid ingest), grow-only
  size_t pin_offs_cap = 0;

  // ---- count table ----
  int W = 0, strand = PENG_STRAND_BOTH;
  uint64_t n_patterns = 0;
  uint32_t* d_hist32 = nullptr;    // per-launch accumulation target of the atomics
  uint64_t* d_counts = nullptr;    // [n_patterns] + [1] ltot slot (summed over ranks together)
  uint64_t ltot = 0;
  bool counts_final = false;       // mirror step applied, ltot published
  cudaStream_t side_stream = nullptr;  // the slow units run beside stage H (W = 10) / stage P2 (deep path)
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  bool fuse_mirror = false;        // peng_cuda_count (single GPU): the final fold may write the mirror copy as well
  bool mirrored = false;           // ... and did
  bool defer_fold = false;         // peng_cuda_count_local32: leave the uint32 table unfolded for a 32-bit exchange
  bool fold_deferred = false;      // ... and it was left

  // ---- background ----
  int K = -1;
  uint64_t* d_bgcounts = nullptr;  // v-layout, 64-bit
  float* d_v32 = nullptr;
  double* d_v64 = nullptr;
  int k_used = -1, max_k = -1;
  float* d_bg32[PENG_MAX_ORDER + 1] = {};
  double* d_bg64[PENG_MAX_ORDER + 1] = {};
  float *d_exp32 = nullptr, *d_z32 = nullptr, *d_logp32 = nullptr;
  double *d_exp64 = nullptr, *d_z64 = nullptr, *d_logp64 = nullptr;
  bool stats_ready = false;
  // ---- id-sliced tables (multi-GPU long k-mers, peng_cuda_set_table_slice): this rank keeps [slice_lo, slice_hi) of every
  // float table and — after peng_cuda_count_finish32_sliced — of the 64-bit count table.  A sliced buffer is addressed
  // through a pointer moved back by slice_lo elements, so that kernels keep indexing by the absolute id.
  int slice_part = 0, slice_parts = 1;
  uint64_t slice_lo = 0, slice_hi = 0;  // valid after a count (slice_hi = 0: not yet derived)
  bool counts_sliced = false;
  uint64_t tab_lo = 0, tab_hi = 0;      // what the float tables hold (set by peng_cuda_stats)

  // ---- partition path (W = 10) ----
  int count_path = 0;              // 0 auto, 1 L2-atomic path, 2 partition path
  bool used_partition = false;     // what the last count ran
  uint2* d_planes = nullptr;       // bit planes of d_codes, {bit 1, bit 0} per 32 positions (ensure_planes)
  size_t cap_planes = 0;
  bool planes_valid = false;
  uint64_t planes_from = 0;        // words [planes_from, n_words32) were written by the packing kernel of the ingest, the rest is missing
  uint16_t* d_part_buckets = nullptr;
  size_t part_bucket_elems = 0;
  uint32_t* d_part_fill = nullptr;
  size_t part_fill_elems = 0;
  int ingest_mode = 0;             // 0 auto (hybrid where it pays), 1 raw bytes + device packing, 2 host packing
  double ingest_fraction = -1;     // share of the bases packed on the host (< 0: from the thread count)
  uint64_t last_ingest_packed = 0; // sequences the host packed in the last ingest
  uint64_t last_ingest_h2d = 0;    // bytes the last ingest from host buffers sent over PCIe
  std::vector<uint64_t> seq_offs;  // peng_cuda_ingest_sequences: offsets built from the lengths (kept between calls)
  int part_cfg = 0;                // stage-P geometry (tuning knob, PENG_PART_CFG)
  bool fold_untiled = false;       // PENG_FOLD_UNTILED: one-thread-per-id fold / mirror at every W (A/B)
  uint2* d_slow_list = nullptr;
  size_t slow_list_elems = 0;
  unsigned int* d_slow_count = nullptr;
  // ---- deep partition path (W >= 12): bucket levels 1 (32-bit ids), 2a (32-bit, W = 16 only), 2 (16-bit leaf entries) ----
  void* d_deep_buckets[3] = {nullptr, nullptr, nullptr};
  size_t deep_bucket_bytes[3] = {0, 0, 0};
  uint32_t* d_deep_fill[3] = {nullptr, nullptr, nullptr};
  size_t deep_fill_elems[3] = {0, 0, 0};
  int deep_mode = 0;               // 0: stage P -> P2 -> leaf histograms, 1: stage P -> L2 REDs (reference variant, tests)

  // ---- phase timing (bench) ----
  bool timing = false;
  cudaEvent_t ev[11] = {};  // 9, 10: around the peer exchange kernel

  // ---- scratch ----
  void* d_scratch = nullptr;
  size_t scratch_bytes = 0;
  uint32_t* d_sel = nullptr;       // seed selection: order | list A | list B | block counts | state
  size_t sel_words = 0;
  bool select_on_host = false;     // PENG_SELECT_HOST: whole-table host replay (A/B, tests)
  bool use_nibble = false;         // PENG_NIBBLE: nibble staging of the raw part of the hybrid ingest

  // ---- peer group: fused cross-GPU sum + fold over NVLink peer memory (peng_cuda_peer_*, peng_cuda_count_exchange32) ----
  struct {
    bool attached = false;
    int rank = 0, world = 1, W = 0;
    uint32_t* hist[8] = {};      // every rank's uint32 launch table (own entry: d_hist32)
    uint64_t* counts[8] = {};    // every rank's 64-bit table
    uint32_t* flags[8] = {};     // every rank's flag words
    void* opened[3][8] = {};     // IPC mappings to close on detach
    uint32_t* d_flags = nullptr; // own: [0 .. 7] arrival epoch per rank, [8] CTAs done, [9] error
    uint32_t epoch = 0;
    double timeout_ms = 20000.0; // PENG_PEER_TIMEOUT_MS
    uint32_t* d_units = nullptr; // BOTH strands: the exchange units this rank owns (see k_peer_exchange32)
    uint32_t n_owned = 0;
    int units_W = 0, units_rank = -1, units_world = 0;
    int colocated = 1;           // ranks of the group on THIS device (tests; several handles of one process): their exchange
                                 // kernels wait for each other and must be resident together
  } peer;
  bool count_pending = false;    // a count has run and its ltot has not been fetched yet (lazy finish)
  // count prefetch (peng_cuda_set_count_prefetch): stage P of the W = 10 partition path runs inside peng_cuda_ingest_codes,
  // unit range by unit range as the packed blocks land, under the DMA of the later ones
  struct {
    int W = 0, strand = 0;       // armed for this count (W = 0: off)
    bool done = false;           // the current ingest ran stage P for (W, strand): the next matching count skips it
  } pre;
  uint64_t table_gen = 1;        // generation of {count table, float tables}: consumers that cache derived data compare it
  bool hist_clean = false;       // every bin of the uint32 launch table is 0 (each complete fold leaves it so): no clearing pass

  // ---- EM session (multi-GPU protocol) ----
  struct {
    bool open = false;
    int n = 0, max_it = 0, order = 0, part = 0, n_parts = 1, nbx = 0;
    double sat = 0, thr = 0;
    double *d_old = nullptr, *d_acc = nullptr, *d_part = nullptr, *d_change = nullptr;
    int *d_iter = nullptr, *d_active = nullptr, *d_nactive = nullptr;
    double2* d_pairs = nullptr;  // {(double) count, bgprob} of the ids of this rank's slice (k_em_pairs)
    // the buffers are kept between sessions (cudaMalloc / cudaFree cost more than an EM iteration at W = 10) and the pair
    // table is reused while the tables it was made from stand (table_gen: bumped by every count / stats)
    size_t cap_small = 0, cap_part = 0, cap_pairs = 0;
    uint64_t pairs_gen = 0;
    int pairs_order = -1;
    uint32_t pairs_ub = 0, pairs_ue = 0;
    uint32_t ub = 0, ue = 0;     // the slice in units of 256 ids
  } em;
};

static thread_local std::string g_create_error;

static int fail(peng_handle_t h, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  if (h) h->err = buf; else g_create_error = buf;
  return code;
}

#define CU(h, call)                                                                                   \
  do {                                                                                                \
    cudaError_t e_ = (call);                                                                          \
    if (e_ != cudaSuccess)                                                                            \
      return fail(h, PENG_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

#define LAUNCH_CHECK(h)                \
  do {                                 \
    (h)->launches++;                   \
    CU(h, cudaGetLastError());         \
  } while (0)

template <typename T>
static int dev_alloc(peng_handle_t h, T** p, size_t n) {
  if (*p) { cudaFree(*p); *p = nullptr; }
  if (n == 0) n = 1;
  cudaError_t e = cudaMalloc((void**)p, n * sizeof(T));
  if (e != cudaSuccess) {
    *p = nullptr;
    return fail(h, PENG_E_NOMEM, "cudaMalloc(%zu bytes) failed: %s", n * sizeof(T), cudaGetErrorString(e));
  }
  return PENG_OK;
}
template <typename T>
static void dev_free(T** p) {
  if (*p) cudaFree(*p);
  *p = nullptr;
}
static int need_scratch(peng_handle_t h, size_t bytes) {
  if (h->scratch_bytes >= bytes) return PENG_OK;
  if (h->d_scratch) cudaFree(h->d_scratch);
  h->d_scratch = nullptr;
  h->scratch_bytes = 0;
  cudaError_t e = cudaMalloc(&h->d_scratch, bytes);
  if (e != cudaSuccess) return fail(h, PENG_E_NOMEM, "scratch cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
  h->scratch_bytes = bytes;
  return PENG_OK;
}
#define TRY(expr)            \
  do {                       \
    int rc_ = (expr);        \
    if (rc_ != PENG_OK) return rc_; \
  } while (0)

static inline unsigned grid_for(uint64_t n, unsigned block) { return (unsigned)((n + block - 1) / block); }
static bool use_partition(peng_handle_t h, int W);
static int ensure_count_tables(peng_handle_t h, int W);
static int ensure_planes(peng_handle_t h);
static int make_planes(peng_handle_t h, uint64_t begin, uint64_t end);
struct PrefetchRun;
static int prefetch_begin(peng_handle_t h, PrefetchRun* pr);
static int prefetch_range(peng_handle_t h, PrefetchRun* pr, uint64_t ub, uint64_t ue);
struct PrefetchRun {
  bool on = false, first = true;
  unsigned grid_p = 0;
  uint32_t region_cap = 0;
  size_t smem_h = 0;
  uint64_t done_u = 0;      // units [0, done_u) of the host-packed front have been through stage P
  uint64_t planes_done = 0; // words [0, planes_done) of the host-packed front have their bit planes
};

// id-sliced tables: a buffer that holds [lo, hi) is addressed through a pointer moved back by lo, so that kernels index by
// the absolute id (never dereferenced outside [lo, hi))
template <typename T>
static inline T* vbase(T* p, uint64_t lo) { return p ? p - lo : nullptr; }
// this rank's slice of the id range: whole tiles of 1024 ids (the EM slices of peng_cuda_em_begin use the same cut)
static void slice_range(peng_handle_t h, uint64_t* lo, uint64_t* hi) {
  const uint64_t n = h->n_patterns;
  if (h->slice_parts <= 1 || n < 1024) { *lo = 0; *hi = n; return; }
  const uint64_t n_tiles = n >> 10;
  *lo = (n_tiles * (uint64_t)h->slice_part / (uint64_t)h->slice_parts) << 10;
  *hi = (n_tiles * (uint64_t)(h->slice_part + 1) / (uint64_t)h->slice_parts) << 10;
}
// the ids the count table is valid for
static inline uint64_t counts_lo(peng_handle_t h) { return h->counts_sliced ? h->slice_lo : 0; }
static inline uint64_t counts_hi(peng_handle_t h) { return h->counts_sliced ? h->slice_hi : h->n_patterns; }
static inline bool tables_whole(peng_handle_t h) { return !h->counts_sliced && h->tab_lo == 0 && h->tab_hi == h->n_patterns; }

// =====================================================================================================
// kernels: ingest
// =====================================================================================================
// The packing kernels also write the bit planes of the words they produce (uint2 {bit 1, bit 0} per 32 positions, read by
// stage P's conflict scan): the codes are in registers anyway, so the planes cost 8 more bytes written per 32 positions
// and no pass of their own (k_make_planes, 26 us + 100 MB of traffic on configs[2], only runs for words the HOST packed).
template <int MODE>
__global__ void k_pack(const uint8_t* __restrict__ src, const uint64_t* __restrict__ offsets, uint64_t n_seq,
                       uint32_t* __restrict__ codes, uint32_t* __restrict__ valid, uint32_t* __restrict__ sstart,
                       uint2* __restrict__ planes, uint64_t t_begin, uint64_t t_end, uint64_t src_base) {
  const uint64_t t = t_begin + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < t_end) {
    uint32_t c[2];
    pack_word<MODE>(t, src, offsets, n_seq, codes, valid, sstart, c, src_base);
    uint2 o;
    planes_of_unit(c[0], c[1], &o.x, &o.y);
    planes[t] = o;
  }
}

__global__ void k_synth(SynthParams P, uint32_t* __restrict__ codes, uint32_t* __restrict__ valid,
                        uint32_t* __restrict__ sstart, uint2* __restrict__ planes, uint64_t n_words32) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n_words32) {
    uint32_t c[2];
    synth_word(t, P, codes, valid, sstart, c);
    uint2 o;
    planes_of_unit(c[0], c[1], &o.x, &o.y);
    planes[t] = o;
  }
}

// one thread per stream position: write the reference letter code of real bases to out[base index]
__global__ void k_export(const uint32_t* __restrict__ codes, const uint32_t* __restrict__ valid,
                         const uint64_t* __restrict__ offsets, uint64_t n_seq, uint32_t fixed_len,
                         uint64_t stream_end, uint8_t* __restrict__ out) {
  const uint64_t q = PENG_FRONT_PAD + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= stream_end) return;
  uint64_t s, i, len, base0;
  if (fixed_len) {
    const uint64_t stride = (uint64_t)fixed_len + 2;
    s = (q - PENG_FRONT_PAD) / stride;
    i = (q - PENG_FRONT_PAD) - s * stride;
    len = fixed_len;
    base0 = s * fixed_len;
  } else {
    uint64_t lo = 0, hi = n_seq - 1;
    while (lo < hi) {
      const uint64_t mid = (lo + hi + 1) >> 1;
      if (PENG_FRONT_PAD + offsets[mid] + 2 * mid <= q) lo = mid; else hi = mid - 1;
    }
    s = lo;
    i = q - (PENG_FRONT_PAD + offsets[s] + 2 * s);
    len = offsets[s + 1] - offsets[s];
    base0 = offsets[s];
  }
  if (i < len) out[base0 + i] = unpack_code(codes, valid, q);
}

// =====================================================================================================
// kernels: counting
// =====================================================================================================
__global__ void k_mark_kstart(const uint32_t* __restrict__ valid, uint32_t* __restrict__ kstart, uint64_t n_words32,
                              int W, unsigned long long* __restrict__ ltot) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int n = 0;
  if (t < n_words32) n = mark_kstart_word(t, n_words32, valid, W, kstart);
  // block reduction: warp shuffles, then one atomic per warp
  for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
  if ((threadIdx.x & 31) == 0 && n) atomicAdd(ltot, (unsigned long long)n);
}

struct EmitGlobal {
  uint32_t* hist;
  __device__ __forceinline__ void operator()(uint32_t key, uint32_t) const { atomicAdd(hist + key, 1u); }  // RED.ADD, no return
};

// Generic path (any W): one L2 atomic per counted occurrence into the 4^W uint32 table (L2-resident up to
// W = 12).  A thread owns `words_per_chunk` code words = 16*words_per_chunk k-mer end positions.
template <int W, bool BOTH>
__global__ void __launch_bounds__(128) k_count_global(const uint32_t* __restrict__ codes,
                                                      const uint32_t* __restrict__ kstart,
                                                      uint32_t* __restrict__ hist, uint64_t first_chunk,
                                                      uint64_t n_chunks, uint32_t words_per_chunk) {
  const uint64_t c = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_chunks) return;
  const uint64_t e0 = PENG_FRONT_PAD + (first_chunk + c) * 16ull * words_per_chunk;
  count_chunk<W, BOTH>(codes, kstart, e0, words_per_chunk, EmitGlobal{hist});
}

// Shared-memory privatised path for 4^W * 4 B <= smem (W <= 6 as uint32 bins; W = 8 as 2 x 16-bit packed
// bins is handled by k_count_smem16): every CTA builds a private sub-histogram with shared-memory atomics
// and merges it into the global table once at the end.
struct EmitShared {
  uint32_t* sh;
  __device__ __forceinline__ void operator()(uint32_t key, uint32_t) const { atomicAdd(sh + key, 1u); }
};

template <int W, bool BOTH>
__global__ void __launch_bounds__(256) k_count_smem(const uint32_t* __restrict__ codes,
                                                    const uint32_t* __restrict__ kstart,
                                                    uint32_t* __restrict__ hist, uint64_t first_chunk, uint64_t n_chunks,
                                                    uint32_t words_per_chunk) {
  extern __shared__ uint32_t sh[];
  constexpr uint32_t NB = 1u << (2 * W);
  for (uint32_t i = threadIdx.x; i < NB; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  for (uint64_t c = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; c < n_chunks;
       c += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t e0 = PENG_FRONT_PAD + (first_chunk + c) * 16ull * words_per_chunk;
    count_chunk<W, BOTH>(codes, kstart, e0, words_per_chunk, EmitShared{sh});
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < NB; i += blockDim.x) {
    const uint32_t v = sh[i];
    if (v) atomicAdd(hist + i, v);
  }
}

// W = 8: 65536 bins do not fit shared memory as uint32 (256 KB) but do as 16-bit halves (128 KB).  A
// private half-word can overflow, so every increment uses the returning form of the atomic: the thread
// that moves a half from 0x7FFF to 0x8000 owns the carry, takes 0x8000 back out of the half and adds it
// to the global bin.  A half can therefore never reach 0x10000 (that would need 32768 further hits on the
// same bin inside the few cycles between the two atomics of the carrying thread).
struct EmitShared16 {
  uint32_t* sh;
  uint32_t* hist;
  __device__ __forceinline__ void operator()(uint32_t key, uint32_t) const {
    const uint32_t sft = (key & 1u) * 16u;
    const uint32_t old = atomicAdd(sh + (key >> 1), 1u << sft);
    if (((old >> sft) & 0xFFFFu) == 0x7FFFu) {
      atomicSub(sh + (key >> 1), 0x8000u << sft);
      atomicAdd(hist + key, 0x8000u);
    }
  }
};

template <int W, bool BOTH>
__global__ void __launch_bounds__(512) k_count_smem16(const uint32_t* __restrict__ codes,
                                                      const uint32_t* __restrict__ kstart,
                                                      uint32_t* __restrict__ hist, uint64_t first_chunk, uint64_t n_chunks,
                                                      uint32_t words_per_chunk) {
  extern __shared__ uint32_t sh[];
  constexpr uint32_t NW = 1u << (2 * W - 1);  // packed words
  for (uint32_t i = threadIdx.x; i < NW; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  for (uint64_t c = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; c < n_chunks;
       c += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t e0 = PENG_FRONT_PAD + (first_chunk + c) * 16ull * words_per_chunk;
    count_chunk<W, BOTH>(codes, kstart, e0, words_per_chunk, EmitShared16{sh, hist});
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < NW; i += blockDim.x) {
    const uint32_t v = sh[i];
    if (v & 0xFFFFu) atomicAdd(hist + 2 * i, v & 0xFFFFu);
    if (v >> 16) atomicAdd(hist + 2 * i + 1, v >> 16);
  }
}

// -----------------------------------------------------------------------------------------------------
// Partition path (W = 10): beyond the L2-atomic ceiling.
//
// One global atomic per counted W-mer caps k_count_global at the ~190 G/s random-RED rate of the L2 (DESIGN.md
// §3.1); shared-memory atomics are ~12x faster but 4^10 bins do not fit one SM.  So the table is cut into
// PART_NCLASS contiguous id ranges ("classes", 28 340 bins = 111 KB of uint32 each) and counting runs in stages:
//   P  k_count_part : streams the packed sequence once; per 32 positions it derives the k-mer start mask, the
//                     bit-parallel conflict candidates (core_seq.cuh: conflict_candidates) and pushes the FORWARD
//                     id of every "fast" W-mer (looked at, no equal key within W-1 positions: counted for sure)
//                     into the CTA's shared-memory queue of its class; whole 128-entry blocks of 16-bit keys go to
//                     the CTA's private region of the class bucket in global memory (no global atomics).
//                     Units holding a "slow" W-mer are appended to a list.
//   S  k_count_slow : exact ring logic of count_chunk for the listed units only (tandem repeats, palindromes).
//   H  k_hist_part  : 4 CTAs per class build the class's bins in shared memory with non-returning shared atomics
//                     and merge them into the uint32 table.
// k_fold_canon then adds the forward-id bins of a reverse-complement pair into the canonical bin (BOTH).
// Skewed input cannot break exactness: a class queue that would overflow in a round is rolled back and the
// round's keys of that class go straight to the global table; a full bucket region does the same per block.
#define PART_NCLASS 37
#define PART_SLOWCAP 512u
#define PART_CPC 4
#define PART_HT 1024

__host__ __device__ __forceinline__ uint32_t part_class_start(uint32_t c) {  // first id of class c (4^10 ids)
  return (uint32_t)((((uint64_t)c << 20) + PART_NCLASS - 1) / PART_NCLASS);
}

// Class geometry of stage P.
//   W = 10 : 37 id ranges of 28 340 bins (one shared-memory histogram each in stage H), 16-bit entries (the low half of
//            the id; a class spans < 2^16 ids).
//   W >= 12: "deep" partition path (see k_part2 / k_hist_leaf below): 32 classes = the top 5 bits of the 2W-bit id, 32-bit
//            entries (the whole id); stage P2 splits every class again until a leaf of 2^15 bins fits shared memory.
template <int W>
struct PartGeom {
  static constexpr uint32_t NCLASS = 32;
  static constexpr bool RAW_KEYS = false;
  typedef uint32_t entry_t;
  typedef uint4 vec_t;  // 4 entries
  __device__ __forceinline__ static uint32_t cls(uint32_t key, uint32_t link = 0) { return (key ^ link) >> (2 * W - 5); }
  __device__ __forceinline__ static uint32_t id_of(uint32_t e, uint32_t) { return e; }
};
template <>
struct PartGeom<10> {
  static constexpr uint32_t NCLASS = PART_NCLASS;
  static constexpr bool RAW_KEYS = true;  // stage P pushes words whose bits above 2 W are not cleared
  typedef uint16_t entry_t;
  typedef uint2 vec_t;
  // floor(id * NCLASS / 2^20) of the low 20 bits: the shift clears what lies above them (multiply + high multiply, FMA pipe)
  __device__ __forceinline__ static uint32_t cls(uint32_t key, uint32_t link = 0) { return __umulhi(key * 4096u + link, (uint32_t)PART_NCLASS); }
  __device__ __forceinline__ static uint32_t id_of(uint32_t e, uint32_t c) {
    const uint32_t start = part_class_start(c);
    return start + ((e - start) & 0xFFFFu);
  }
};

// One push: class c (W = 10: floor(key * NCLASS / 2^20) as a single high multiply), slot = cnt[c]++, store if the slot
// exists.  W-mers that are not pushed go to a scratch queue: ptxas turns a PREDICATED returning shared atomic into a
// branch (BSSY / BRA / BSYNC per push), an unconditional one on a dummy counter is cheaper.  Stage P is bound by the
// half-rate ALU pipe (ncu: pipe_alu 75 %), so the address arithmetic is written as multiply-adds (FMA pipe).
template <int W, uint32_t CAP>
__device__ __forceinline__ uint32_t part_push(uint32_t key, uint32_t push, uint32_t cnt_s, uint32_t q_s, uint32_t link) {
  typedef PartGeom<W> G;
  uint32_t c = G::cls(key, link);  // link: 0 at run time (k_count_part), a dependency for the scheduler only
  c = push ? c : G::NCLASS;
  const uint32_t cnt_addr = c * 4u + cnt_s;
  const uint32_t base = c * (CAP * (uint32_t)sizeof(typename G::entry_t)) + q_s;
  uint32_t slot;
  if constexpr (sizeof(typename G::entry_t) == 2) {
    asm volatile(
        "{\n\t.reg .pred q;\n\t.reg .u32 a;\n\t"
        "atom.shared.add.u32 %0, [%1], 1;\n\t"
        "setp.lt.u32 q, %0, %2;\n\t"
        "mad.lo.u32 a, %0, 2, %3;\n\t"
        "@q st.shared.u16 [a], %4;\n\t}"
        : "=r"(slot)
        : "r"(cnt_addr), "r"(CAP), "r"(base), "h"((uint16_t)key)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred q;\n\t.reg .u32 a;\n\t"
        "atom.shared.add.u32 %0, [%1], 1;\n\t"
        "setp.lt.u32 q, %0, %2;\n\t"
        "mad.lo.u32 a, %0, 4, %3;\n\t"
        "@q st.shared.u32 [a], %4;\n\t}"
        : "=r"(slot)
        : "r"(cnt_addr), "r"(CAP), "r"(base), "r"(key)
        : "memory");
  }
  return slot;
}

template <int W, bool BOTH, int PART_NT, uint32_t PART_CAP, int MINB, bool LINK, bool PL>
__global__ void __launch_bounds__(PART_NT, MINB) k_count_part(
    const uint32_t* __restrict__ codes, const uint2* __restrict__ planes, const uint32_t* __restrict__ valid, uint32_t* __restrict__ kstart,
    uint64_t n_words32, uint64_t unit_begin, uint64_t unit_end, typename PartGeom<W>::entry_t* __restrict__ buckets, uint32_t region_cap,
    uint32_t* __restrict__ fill_out, uint32_t* __restrict__ hist, unsigned long long* __restrict__ ltot,
    uint2* __restrict__ slow_list, unsigned int* __restrict__ slow_count, uint32_t zero, uint32_t carry_fill) {
  static_assert(W >= 10 && W <= 16, "class geometries exist for 4^10 ids (one level) and 4^12 .. 4^16 ids (deep path)");
  typedef PartGeom<W> G;
  typedef typename G::entry_t entry_t;
  typedef typename G::vec_t vec_t;
  constexpr uint32_t PART_NCLASS_ = G::NCLASS;
  constexpr uint32_t KEYMASK = W == 16 ? 0xFFFFFFFFu : (1u << (2 * (W & 15))) - 1u;
  extern __shared__ __align__(16) unsigned char q_raw[];
  entry_t* q = reinterpret_cast<entry_t*>(q_raw);  // [NCLASS + 1][PART_CAP]
  __shared__ uint32_t cnt[PART_NCLASS_ + 1], prev[PART_NCLASS_], fill[PART_NCLASS_], ovf[PART_NCLASS_];
  __shared__ uint32_t any_ovf, n_slow, slow_base;
  __shared__ uint2 slow_q[PART_SLOWCAP];
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // carry_fill: an earlier launch of the same count (another unit range: the count prefetch of the ingest) left keys in
  // this CTA's bucket regions — go on behind them
  if (threadIdx.x < PART_NCLASS_) {
    cnt[threadIdx.x] = 0; prev[threadIdx.x] = 0; ovf[threadIdx.x] = 0;
    fill[threadIdx.x] = carry_fill ? fill_out[threadIdx.x * gridDim.x + blockIdx.x] : 0u;
  }
  if (threadIdx.x == 0) { cnt[PART_NCLASS_] = 0; any_ovf = 0; n_slow = 0; }
  __syncthreads();
  const uint32_t cnt_s = (uint32_t)__cvta_generic_to_shared(cnt), q_s = (uint32_t)__cvta_generic_to_shared(q);
  const uint64_t n_units = unit_end - unit_begin;
  const uint64_t n_rounds = (n_units + PART_NT - 1) / PART_NT;
  uint32_t my_ltot = 0;
  // What one thread needs of a unit (32 k-mer starts): the three code words its keys are cut from, the mask of
  // fast W-mers and the mask of slow ones.  Scanning a unit is ALU work only, pushing is shared-memory work only:
  // the loop scans round r+1 in the shadow of the pushes of round r, one scan step behind every other push
  // (ptxas keeps the source order of independent work, it does not interleave the two streams by itself).
  // PL: the scan runs on the bit planes of units u-1, u, u+1 (sw = H[3], L[3]), else on the code words 2u-1 .. 2u+3.
  struct Unit { uint32_t w1, w2, w3, fast, slow, u; };
  typedef typename std::conditional<PL, CandStatePl, CandState>::type State;
  typedef typename std::conditional<PL, CandStepsPl<W, BOTH>, CandSteps<W, BOTH>>::type Steps;
  uint32_t sw[6], kw[3], ks;
  State cs;
  auto scan_begin = [&](uint64_t r) -> uint32_t {  // loads + k-mer starts of the thread's unit in round r
    const uint64_t u = unit_begin + r * PART_NT + threadIdx.x;
    const bool live = r < n_rounds && u < unit_end;
    const uint64_t uc = live ? u : unit_begin;  // clamp: dead threads read a legal address and drop the result
    ks = kstart_word(uc, n_words32, valid, W);
    ks = live ? ks : 0u;
    if (live) kstart[u] = ks;
    my_ltot += __popc(ks);
    const uint2 a = __ldg(reinterpret_cast<const uint2*>(codes + 2 * uc));
    if constexpr (PL) {
      const uint2 p0 = uc ? __ldg(planes + uc - 1) : make_uint2(0u, 0u);
      const uint2 p1 = __ldg(planes + uc), p2 = __ldg(planes + uc + 1);
      sw[0] = p0.x; sw[1] = p1.x; sw[2] = p2.x; sw[3] = p0.y; sw[4] = p1.y; sw[5] = p2.y;
      kw[0] = a.x; kw[1] = a.y; kw[2] = __ldg(codes + 2 * uc + 2);
      cand_pl_init(cs);
    } else {
      sw[0] = uc ? __ldg(codes + 2 * uc - 1) : 0u;
      const uint2 b = __ldg(reinterpret_cast<const uint2*>(codes + 2 * uc + 2));
      sw[1] = a.x; sw[2] = a.y; sw[3] = b.x; sw[4] = b.y;
      kw[0] = a.x; kw[1] = a.y; kw[2] = b.x;
      cand_init(cs);
    }
    return (uint32_t)u;
  };
  auto scan_end = [&](uint32_t u) -> Unit {
    uint32_t cand;
    if constexpr (PL) cand = cs.c;
    else cand = cand_result(cs);
    return Unit{kw[0], kw[1], kw[2], ks & ~cand, ks & cand, u};
  };
  auto link_in = [&](uint32_t v) { sw[PL ? 1 : 2] ^= v; };  // the unit's own word
  auto link_out = [&]() -> uint32_t {
    if constexpr (PL) return cs.c | cs.p[0];
    else return cs.c1 | cs.p[0];
  };
  const uint64_t region_stride = (uint64_t)gridDim.x * region_cap, region0 = (uint64_t)blockIdx.x * region_cap;  // bucket region of (class c, this CTA)
  constexpr bool STEP_PER_PUSH = Steps::N > 16;  // W >= 12: one scan step behind every push instead of every other
  static_assert(Steps::N <= 32, "at most one scan step per push");
  uint32_t un = scan_begin(blockIdx.x);
#pragma unroll
  for (int I = 0; I < Steps::N; I++) Steps::run(I, sw, cs);
  Unit cur = scan_end(un);
  for (uint64_t r = blockIdx.x; r < n_rounds; r += gridDim.x) {
    const uint32_t fast = cur.fast, w1 = cur.w1, w2 = cur.w2, w3 = cur.w3;
    un = scan_begin(r + gridDim.x);
    // `zero` is 0 at run time but opaque to the compiler: the two data dependencies it fakes (slot of push i -> inputs
    // of scan step i/2 -> key of push i+1) are what keeps ptxas from sorting the two instruction streams apart again.
    uint32_t link = 0;
#pragma unroll
    for (int i = 0; i < 32; i++) {
      // W = 10: the letters behind the W-mer stay in the word — cls_raw() shifts them out, the 16-bit store drops them
      uint32_t key = i < 16 ? __funnelshift_r(w1, w2, 2 * i) : __funnelshift_r(w2, w3, 2 * (i - 16));
      if (!G::RAW_KEYS) key &= KEYMASK;
      const uint32_t slot = part_push<W, PART_CAP>(key, fast & (1u << i), cnt_s, q_s, link);
      const int step = STEP_PER_PUSH ? i : (i >> 1);
      if ((STEP_PER_PUSH || (i & 1)) && step < Steps::N) {
        if (LINK) link_in(slot & zero);
        Steps::run(step, sw, cs);
        if (LINK) link = link_out() & zero;
      }
    }
    const Unit nxt = scan_end(un);
    if (cur.slow) {
      const uint32_t i = atomicAdd(&n_slow, 1u);
      if (i < PART_SLOWCAP) slow_q[i] = make_uint2(cur.u, cur.slow);
      else slow_list[atomicAdd(slow_count, 1u)] = make_uint2(cur.u, cur.slow);
    }
    __syncthreads();
    // ---- flush: warp w owns classes w, w+16, w+32 ----
    for (uint32_t c = warp; c < PART_NCLASS_; c += PART_NT / 32) {
      const uint32_t n = cnt[c];
      if (n > PART_CAP) {  // this round did not fit: roll the queue back, the round's keys of c are redone below
        if (lane == 0) { ovf[c] = 1; any_ovf = 1; cnt[c] = prev[c]; }
        continue;
      }
      const uint32_t nfl = n & ~127u;
      if (nfl) {
        const uint32_t base = fill[c];
        const entry_t* src = q + c * PART_CAP;
        if (base + nfl <= region_cap) {
          // whole 128-entry blocks, 16 bytes per lane and step, at most CAP entries: unrolled, 32-bit offsets (the rolled
          // loop with its 64-bit address arithmetic was a quarter of stage P's instructions: profiles/r2_count_part source page)
          unsigned char* d8 = reinterpret_cast<unsigned char*>(buckets + (c * region_stride + region0 + base));
          const unsigned char* s8 = reinterpret_cast<const unsigned char*>(src);
          const uint32_t bytes = nfl * (uint32_t)sizeof(entry_t);
#pragma unroll
          for (uint32_t k = 0; k < PART_CAP * sizeof(entry_t) / 512; k++) {
            const uint32_t off = lane * 16u + k * 512u;
            if (off < bytes) *reinterpret_cast<uint4*>(d8 + off) = *reinterpret_cast<const uint4*>(s8 + off);
          }
          if (lane == 0) fill[c] = base + nfl;
        } else {  // private bucket region full: straight to the global table
          for (uint32_t i = lane; i < nfl; i += 32) atomicAdd(hist + G::id_of((uint32_t)src[i], c), 1u);
        }
        const uint32_t rem = n - nfl;
        vec_t v = vec_t();
        if (lane * 4 < rem) v = *(const vec_t*)(src + nfl + lane * 4);
        __syncwarp();
        if (lane * 4 < rem) *(vec_t*)(q + c * PART_CAP + lane * 4) = v;
        if (lane == 0) { cnt[c] = rem; prev[c] = rem; }
      } else if (lane == 0) prev[c] = n;
    }
    if (threadIdx.x == 0) cnt[PART_NCLASS_] = 0;
    if (warp == PART_NT / 32 - 1) {  // slow units of this round -> global list (one global atomic per CTA round)
      const uint32_t ns = min(n_slow, PART_SLOWCAP);
      if (ns) {
        if (lane == 0) slow_base = atomicAdd(slow_count, ns);
        __syncwarp();
        const uint32_t sb = slow_base;
        for (uint32_t i = lane; i < ns; i += 32) slow_list[sb + i] = slow_q[i];
        __syncwarp();
        if (lane == 0) n_slow = 0;
      }
    }
    __syncthreads();
    if (any_ovf) {
#pragma unroll 1
      for (int i = 0; i < 32; i++) {
        if (fast & (1u << i)) {
          const uint32_t key = (i < 16 ? __funnelshift_r(w1, w2, 2 * i) : __funnelshift_r(w2, w3, 2 * (i - 16))) & KEYMASK;
          if (ovf[G::cls(key)]) atomicAdd(hist + key, 1u);
        }
      }
      __syncthreads();
      if (threadIdx.x < PART_NCLASS_) ovf[threadIdx.x] = 0;
      if (threadIdx.x == 0) any_ovf = 0;
      __syncthreads();
    }
    cur = nxt;
  }
  // leftovers (< 128 per class and CTA): straight to the global table
  for (uint32_t c = warp; c < PART_NCLASS_; c += PART_NT / 32) {
    if (lane == 0) fill_out[c * gridDim.x + blockIdx.x] = fill[c];
    const uint32_t n = cnt[c];
    for (uint32_t i = lane; i < n; i += 32) atomicAdd(hist + G::id_of((uint32_t)q[c * PART_CAP + i], c), 1u);
  }
  for (int o = 16; o > 0; o >>= 1) my_ltot += __shfl_xor_sync(0xffffffffu, my_ltot, o);
  if (lane == 0 && my_ltot) atomicAdd(ltot, (unsigned long long)my_ltot);
}

// exact resolution of the units that hold a W-mer with an equal key close in front of it
struct EmitSlow {
  uint32_t* hist;
  uint32_t mask;
  int w1;  // W - 1
  __device__ __forceinline__ void operator()(uint32_t key, uint32_t i) const {
    const int b = (int)i - w1;  // start position relative to the unit
    if (b >= 0 && b < 32 && ((mask >> b) & 1u)) atomicAdd(hist + key, 1u);
  }
};
template <int W, bool BOTH>
__global__ void __launch_bounds__(128) k_count_slow(const uint32_t* __restrict__ codes, const uint32_t* __restrict__ kstart,
                                                    const uint2* __restrict__ slow_list,
                                                    const unsigned int* __restrict__ slow_count, uint32_t* __restrict__ hist) {
  const uint32_t n = *slow_count;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const uint2 e = slow_list[i];
    count_chunk<W, BOTH>(codes, kstart, (uint64_t)e.x * 32u, 3, EmitSlow{hist, e.y, W - 1});
  }
}

// stage H: CTA (class c, part p) walks the regions p, p + PART_CPC, ... of the class, one region per warp at a time
__global__ void __launch_bounds__(PART_HT) k_hist_part(const uint16_t* __restrict__ buckets, uint32_t region_cap,
                                                       uint32_t n_regions, const uint32_t* __restrict__ fill,
                                                       uint32_t* __restrict__ hist) {
  extern __shared__ uint32_t bins[];
  const uint32_t c = blockIdx.x / PART_CPC, part = blockIdx.x % PART_CPC;
  const uint32_t start = part_class_start(c), nb = part_class_start(c + 1) - start;
  for (uint32_t i = threadIdx.x; i < nb; i += PART_HT) bins[i] = 0;
  __syncthreads();
  const uint32_t s16 = start & 0xFFFFu;
  const uint32_t bins_s = (uint32_t)__cvta_generic_to_shared(bins);
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (uint32_t reg = part + PART_CPC * warp; reg < n_regions; reg += PART_CPC * (PART_HT / 32)) {
    const uint32_t n8 = fill[c * n_regions + reg] >> 3;  // 16-byte groups (fill is a multiple of 128 keys)
    const uint4* src = reinterpret_cast<const uint4*>(buckets + ((uint64_t)c * n_regions + reg) * region_cap);
#pragma unroll 2
    for (uint32_t g = lane; g < n8; g += 32) {
      const uint4 v = __ldg(src + g);
      const uint32_t ws[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const uint32_t i0 = ((ws[k] & 0xFFFFu) - s16) & 0xFFFFu, i1 = ((ws[k] >> 16) - s16) & 0xFFFFu;
        asm volatile("red.shared.add.u32 [%0], 1;" :: "r"(bins_s + i0 * 4u) : "memory");
        asm volatile("red.shared.add.u32 [%0], 1;" :: "r"(bins_s + i1 * 4u) : "memory");
      }
    }
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < nb; i += PART_HT) {
    const uint32_t v = bins[i];
    if (v) atomicAdd(hist + start + i, v);
  }
}

// -----------------------------------------------------------------------------------------------------
// Deep partition path (W = 12, 14, 16): 4^W bins neither fit one SM (W = 10: 37 classes did) nor, from W = 14 on, the
// L2 — one RED per counted W-mer then pays a DRAM read-modify-write of a 32-byte sector (W = 14: 24 Gbases/s).
// Stage P (k_count_part with PartGeom<W>) splits the ids by their top 5 bits into 32-bit entries; stage P2 (k_part2)
// splits every class again by the next NC2 bits — into 32-bit entries while more than 15 + 8 bits are left (W = 16),
// finally into 16-bit entries holding the low 15 bits — and stage H (k_hist_leaf) builds each leaf of 2^15 bins in
// shared memory and adds it to the table with coalesced plain read-modify-writes (a leaf has one owner CTA).
// Every stage writes CTA-private regions (no global atomics), keeps exactness under skew the way stage P does (queue
// overflow: roll the round back, its keys of that class go straight to the table; region full: same per block).
#define DEEP_LEAF_BITS 15
#define DEEP_P2_NT 256
#ifndef DEEP_P2_KPT
#define DEEP_P2_KPT 16  // keys per thread and round (16-byte loads)
#endif
#ifndef DEEP_P2_MINB
#define DEEP_P2_MINB 4  // CTAs per SM
#endif
#define DEEP_P2_CAP (DEEP_P2_KPT == 8 ? 88 : DEEP_P2_KPT == 16 ? 104 : 136)  // 63 left over + one round's share + 5 sigma, multiple of 8
#define DEEP_P2_NQ 256  // shared-memory queues per CTA: one per lane of its 8 warps in the flush

// Stage P2.  in : regions [c1][r] of in_cap 32-bit ids, in_fill[c1 * n_in + r] of them valid (multiple of 4);
// CTA (c1, g) owns the input regions r = g, g + G, ... of class c1.  A key goes to queue
//   qi = (class field << SUB_BITS) | (lane & (2^SUB_BITS - 1)),   class field = (id >> shift) & (2^NC2_BITS - 1):
// few classes (W = 12: 16) are spread over sub-queues so that the slot atomics of a warp do not pile up on a handful
// of counters.  Queue qi of CTA g flushes whole 64-entry blocks to its private region
//   [(leaf * G + g) << SUB_BITS | sub],  leaf = c1 << NC2_BITS | class field,
// so a leaf owns G << SUB_BITS regions.  Per round a lane owns one queue in the flush (fill / rollback state in registers).
template <int NC2_BITS, int SUB_BITS, uint32_t CAP, typename OUT_T>
__global__ void __launch_bounds__(DEEP_P2_NT, DEEP_P2_MINB) k_part2(const uint32_t* __restrict__ in, uint32_t in_cap, uint32_t n_in,
                                                         const uint32_t* __restrict__ in_fill, uint32_t G, uint32_t shift,
                                                         OUT_T* __restrict__ out, uint32_t out_cap,
                                                         uint32_t* __restrict__ out_fill, uint32_t* __restrict__ hist) {
  static_assert(NC2_BITS + SUB_BITS == 8 && DEEP_P2_NQ == 256 && DEEP_P2_NT == 256, "one queue per lane of the 8 warps");
  constexpr uint32_t GRAN = 64;  // entries per flushed block
  constexpr uint32_t NQ = DEEP_P2_NQ, CMASK = (1u << NC2_BITS) - 1u, SMASK = (1u << SUB_BITS) - 1u;
  extern __shared__ __align__(16) unsigned char q_raw[];
  OUT_T* q = reinterpret_cast<OUT_T*>(q_raw);  // [NQ + 1][CAP]
  __shared__ uint32_t cnt[NQ + 1];
  __shared__ uint8_t ovf[NQ];
  __shared__ uint32_t any_ovf;
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  cnt[threadIdx.x] = 0;
  ovf[threadIdx.x] = 0;
  if (threadIdx.x == 0) { cnt[NQ] = 0; any_ovf = 0; }
  __syncthreads();
  const uint32_t c1 = blockIdx.x / G, g = blockIdx.x % G;
  const uint32_t cnt_s = (uint32_t)__cvta_generic_to_shared(cnt), q_s = (uint32_t)__cvta_generic_to_shared(q);
  const uint32_t sub = lane & SMASK;
  // flush ownership: this thread owns queue myq = threadIdx.x
  const uint32_t myq = threadIdx.x;
  const uint32_t my_leaf = (c1 << NC2_BITS) | (myq >> SUB_BITS);
  const uint64_t my_region = ((uint64_t)(my_leaf * G + g) << SUB_BITS) | (myq & SMASK);
  uint32_t my_fill = 0, my_prev = 0;
  constexpr uint32_t ROUND = DEEP_P2_NT * DEEP_P2_KPT;
  // The CTA walks (region r, offset base) round by round; the keys of round i + 1 are loaded right behind the pushes of
  // round i (into the same registers), so that their latency hides behind the barrier and the flush.
  uint32_t r = g, base = 0, n = 0;
  while (r < n_in && (n = in_fill[c1 * n_in + r]) == 0) r += G;
  uint32_t key[DEEP_P2_KPT];
  uint32_t live = 0;
  auto load_round = [&](uint32_t rr, uint32_t bb, uint32_t nn) {
    const uint4* src = reinterpret_cast<const uint4*>(in + ((uint64_t)c1 * n_in + rr) * in_cap);
    live = 0;
#pragma unroll
    for (int j = 0; j < DEEP_P2_KPT / 4; j++) {
      const uint32_t i = bb + (j * DEEP_P2_NT + threadIdx.x) * 4;  // nn is a multiple of 4: a 16-byte group is whole
      const bool ok = i < nn;
      const uint4 v = ok ? __ldg(src + (i >> 2)) : make_uint4(0, 0, 0, 0);
      key[4 * j] = v.x; key[4 * j + 1] = v.y; key[4 * j + 2] = v.z; key[4 * j + 3] = v.w;
      live |= ok ? (0xFu << (4 * j)) : 0u;
    }
  };
  if (r < n_in) load_round(r, base, n);
  while (r < n_in) {
    {
      const uint32_t cur_r = r, cur_base = base, cur_n = n;
      // ---- pushes: slot atomics in batches of 8 (their results are not awaited one by one), then the stores ----
#pragma unroll
      for (int k0 = 0; k0 < DEEP_P2_KPT; k0 += 8) {
        uint32_t slot[8], qb[8];
#pragma unroll
        for (int k = 0; k < 8; k++) {
          uint32_t c = (((key[k0 + k] >> shift) & CMASK) << SUB_BITS) | sub;
          c = (live >> (k0 + k)) & 1u ? c : NQ;
          qb[k] = c * (CAP * (uint32_t)sizeof(OUT_T)) + q_s;
          asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(slot[k]) : "r"(c * 4u + cnt_s) : "memory");
        }
#pragma unroll
        for (int k = 0; k < 8; k++) {
          if constexpr (sizeof(OUT_T) == 2) {
            asm volatile(
                "{\n\t.reg .pred q;\n\t.reg .u32 a;\n\t"
                "setp.lt.u32 q, %0, %1;\n\t"
                "mad.lo.u32 a, %0, 2, %2;\n\t"
                "@q st.shared.u16 [a], %3;\n\t}"
                :: "r"(slot[k]), "r"(CAP), "r"(qb[k]), "h"((uint16_t)(key[k0 + k] & ((1u << DEEP_LEAF_BITS) - 1u)))
                : "memory");
          } else {
            asm volatile(
                "{\n\t.reg .pred q;\n\t.reg .u32 a;\n\t"
                "setp.lt.u32 q, %0, %1;\n\t"
                "mad.lo.u32 a, %0, 4, %2;\n\t"
                "@q st.shared.u32 [a], %3;\n\t}"
                :: "r"(slot[k]), "r"(CAP), "r"(qb[k]), "r"(key[k0 + k])
                : "memory");
          }
        }
      }
      // ---- next round's keys ----
      base += ROUND;
      while (r < n_in && base >= n) {
        r += G;
        base = 0;
        n = r < n_in ? in_fill[c1 * n_in + r] : 0;
      }
      if (r < n_in) load_round(r, base, n);
      __syncthreads();
      // ---- flush: lane l of warp w owns queue 32 w + l.  The four 8-lane groups of the warp work on four queues at a
      // time (one 64-entry block = 8 lanes x 16 bytes of 16-bit entries): group s walks the full queues among 8 s .. 8 s + 7.
      {
        uint32_t nq = cnt[myq];
        const bool over = nq > CAP;  // this round did not fit: roll the queue back, its keys of the round are redone below
        if (over) { ovf[myq] = 1; any_ovf = 1; cnt[myq] = my_prev; nq = my_prev; }
        const bool full = nq >= GRAN;  // my_prev < GRAN: a rolled-back queue has nothing to flush
        if (!full) my_prev = nq;
        uint32_t m = __ballot_sync(0xffffffffu, full);
        const uint32_t grp = lane >> 3, gl = lane & 7;
        while (m) {
          const uint32_t mg = (m >> (8 * grp)) & 0xFFu;  // this group's full queues
          const bool act = mg != 0;
          const uint32_t l = act ? 8 * grp + (uint32_t)__ffs(mg) - 1u : lane;
          m &= ~((mg & (0u - mg)) << (8 * grp));
          m = __reduce_and_sync(0xffffffffu, m);  // every group cleared its own bit
          const uint32_t qi = warp * 32 + l;
          const uint32_t n_l = __shfl_sync(0xffffffffu, nq, l), fb = __shfl_sync(0xffffffffu, my_fill, l);
          const uint32_t nfl_l = n_l & ~(GRAN - 1);
          OUT_T* qs = q + qi * CAP;
          const uint64_t region = ((uint64_t)(((c1 << NC2_BITS) | (qi >> SUB_BITS)) * G + g) << SUB_BITS) | (qi & SMASK);
          const bool fits = fb + nfl_l <= out_cap;
          if (act) {
            if (fits) {
              OUT_T* dst = out + region * out_cap + fb;
              for (uint32_t i = gl * 8; i < nfl_l; i += GRAN) {  // one trip unless a queue took >= 2 blocks in a round
#pragma unroll
                for (uint32_t v = 0; v < sizeof(OUT_T) / 2; v++)
                  *(uint4*)((unsigned char*)(dst + i) + 16 * v) = *(const uint4*)((const unsigned char*)(qs + i) + 16 * v);
              }
            } else {  // region full: straight to the table
              const uint32_t leaf = (c1 << NC2_BITS) | (qi >> SUB_BITS);
              for (uint32_t i = gl; i < nfl_l; i += 8) {
                const uint32_t id = sizeof(OUT_T) == 2 ? ((leaf << DEEP_LEAF_BITS) | (uint32_t)qs[i]) : (uint32_t)qs[i];
                atomicAdd(hist + id, 1u);
              }
            }
          }
          // the < 64 entries behind the flushed blocks move to the front (whole 8-entry pieces: what lies behind the
          // last valid entry is never read)
          const uint32_t rem = n_l - nfl_l;
          uint4 mv[sizeof(OUT_T) / 2];
          const bool mover = act && gl * 8 < rem;
          if (mover) {
#pragma unroll
            for (uint32_t v = 0; v < sizeof(OUT_T) / 2; v++) mv[v] = *(const uint4*)((const unsigned char*)(qs + nfl_l + gl * 8) + 16 * v);
          }
          __syncwarp();
          if (mover) {
#pragma unroll
            for (uint32_t v = 0; v < sizeof(OUT_T) / 2; v++) *(uint4*)((unsigned char*)(qs + gl * 8) + 16 * v) = mv[v];
          }
          if (act && lane == l) {
            if (fits) my_fill = fb + nfl_l;
            cnt[myq] = rem;
            my_prev = rem;
          }
        }
      }
      if (threadIdx.x == 0) cnt[NQ] = 0;
      __syncthreads();
      if (any_ovf) {  // rare: the round's keys are read again (the registers already hold the next round)
        const uint4* src = reinterpret_cast<const uint4*>(in + ((uint64_t)c1 * n_in + cur_r) * in_cap);
#pragma unroll 1
        for (int j = 0; j < DEEP_P2_KPT / 4; j++) {
          const uint32_t i = cur_base + (j * DEEP_P2_NT + threadIdx.x) * 4;
          if (i < cur_n) {
            const uint4 v = __ldg(src + (i >> 2));
            const uint32_t ks[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int t = 0; t < 4; t++)
              if (ovf[(((ks[t] >> shift) & CMASK) << SUB_BITS) | sub]) atomicAdd(hist + ks[t], 1u);
          }
        }
        __syncthreads();
        ovf[threadIdx.x] = 0;
        if (threadIdx.x == 0) any_ovf = 0;
        __syncthreads();
      }
    }
  }
  // leftovers (< GRAN per queue): straight to the table
  out_fill[my_region] = my_fill;
  __syncthreads();
  for (uint32_t qi = warp; qi < NQ; qi += DEEP_P2_NT / 32) {
    const uint32_t leaf = (c1 << NC2_BITS) | (qi >> SUB_BITS);
    const uint32_t nq = cnt[qi];
    for (uint32_t i = lane; i < nq; i += 32) {
      const uint32_t e = (uint32_t)q[qi * CAP + i];
      atomicAdd(hist + (sizeof(OUT_T) == 2 ? ((leaf << DEEP_LEAF_BITS) | e) : e), 1u);
    }
  }
}

// stage H of the deep path: CTA = leaf; regions [leaf][0 .. n_reg) of 16-bit entries (low 15 bits of the id), one region
// per warp at a time; the leaf's 2^15 bins are built in shared memory and added to the table with 16-byte plain
// read-modify-writes (this CTA is the only writer of the leaf's bins in this launch).
__global__ void __launch_bounds__(1024, 1) k_hist_leaf(const uint16_t* __restrict__ buckets, uint32_t region_cap, uint32_t n_reg,
                                                       const uint32_t* __restrict__ fill, uint32_t* __restrict__ hist) {
  extern __shared__ __align__(16) uint32_t bins[];
  constexpr uint32_t NB = 1u << DEEP_LEAF_BITS;
  __shared__ uint32_t any;
  const uint32_t leaf = blockIdx.x;
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) any = 0;
  __syncthreads();
  uint32_t mine = 0;
  for (uint32_t r = threadIdx.x; r < n_reg; r += 1024) mine |= fill[(uint64_t)leaf * n_reg + r];
  if (mine) any = 1;
  uint4* b4 = reinterpret_cast<uint4*>(bins);
#pragma unroll
  for (uint32_t i = threadIdx.x; i < NB / 4; i += 1024) b4[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  if (!any) return;
  const uint32_t bins_s = (uint32_t)__cvta_generic_to_shared(bins);
  for (uint32_t r = warp; r < n_reg; r += 32) {
    const uint32_t n8 = fill[(uint64_t)leaf * n_reg + r] >> 3;  // 16-byte groups (fill is a multiple of 64 entries)
    const uint4* src = reinterpret_cast<const uint4*>(buckets + ((uint64_t)leaf * n_reg + r) * region_cap);
#pragma unroll 2
    for (uint32_t gi = lane; gi < n8; gi += 32) {
      const uint4 v = __ldg(src + gi);
      const uint32_t ws[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; k++) {
        asm volatile("red.shared.add.u32 [%0], 1;" :: "r"(bins_s + (ws[k] & 0xFFFFu) * 4u) : "memory");
        asm volatile("red.shared.add.u32 [%0], 1;" :: "r"(bins_s + (ws[k] >> 16) * 4u) : "memory");
      }
    }
  }
  __syncthreads();
  uint4* dst = reinterpret_cast<uint4*>(hist + ((uint64_t)leaf << DEEP_LEAF_BITS));
  uint4 d[NB / 4 / 1024];
#pragma unroll
  for (uint32_t k = 0; k < NB / 4 / 1024; k++) d[k] = dst[threadIdx.x + k * 1024];
#pragma unroll
  for (uint32_t k = 0; k < NB / 4 / 1024; k++) {
    const uint4 v = b4[threadIdx.x + k * 1024];
    d[k].x += v.x; d[k].y += v.y; d[k].z += v.z; d[k].w += v.w;
    if (v.x | v.y | v.z | v.w) dst[threadIdx.x + k * 1024] = d[k];
  }
}

// Reference variant of the deep path (PENG_DEEP_MODE=1, tests): the 32-bit ids of stage P go to the table with one
// RED each, class after class, so that the bins in flight (1/32 of the table) stay L2-resident up to W = 14.
__global__ void __launch_bounds__(512) k_hist_l2(const uint32_t* __restrict__ buckets, uint32_t region_cap, uint32_t n_regions,
                                                 uint32_t n_class, const uint32_t* __restrict__ fill, uint32_t* __restrict__ hist) {
  for (uint32_t c = 0; c < n_class; c++) {
    for (uint32_t r = blockIdx.x; r < n_regions; r += gridDim.x) {
      const uint32_t n = fill[c * n_regions + r];
      const uint32_t* src = buckets + ((uint64_t)c * n_regions + r) * region_cap;
      for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) atomicAdd(hist + __ldg(src + i), 1u);
    }
  }
}

// BOTH on the partition path: fast W-mers sit in their forward-id bin, slow ones in the canonical bin; the pair
// count is their sum, kept in the canonical bin (k_mirror copies it to the other id afterwards).
__global__ void k_fold_canon(uint32_t* __restrict__ hist, uint64_t* __restrict__ acc, int W, uint64_t n) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const uint32_t rc = revcomp_id((uint32_t)i, W);
    if ((uint32_t)i <= rc) {
      uint64_t v = hist[i];
      if ((uint32_t)i < rc) v += hist[rc];
      if (v) {
        acc[i] += v;
        hist[i] = 0;
        hist[rc] = 0;
      }
    }
  }
}
// The only fold of a count (sets below 2^30 positions): the 64-bit table is written, not added to, so it needs no
// clearing beforehand; MIRROR also writes the pair count at the reverse-complement id (single-GPU peng_cuda_count: no
// separate k_mirror pass).  Without MIRROR the non-canonical entries are written as 0.
template <bool MIRROR>
__global__ void k_fold_canon_first(uint32_t* __restrict__ hist, uint64_t* __restrict__ acc, int W, uint64_t n) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const uint32_t rc = revcomp_id((uint32_t)i, W);
    if ((uint32_t)i <= rc) {
      uint64_t v = hist[i];
      if ((uint32_t)i < rc) v += hist[rc];
      acc[i] = v;
      if ((uint32_t)i < rc) acc[rc] = MIRROR ? v : 0ull;
      if (v) {
        hist[i] = 0;
        hist[rc] = 0;
      }
    }
  }
}

// fold the uint32 launch histogram into the 64-bit table and clear it ("global merge")
__global__ void k_fold(uint32_t* __restrict__ hist, uint64_t* __restrict__ acc, uint64_t n) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const uint32_t v = hist[i];
    if (v) {
      acc[i] += v;
      hist[i] = 0;
    }
  }
}
__global__ void k_fold_first(uint32_t* __restrict__ hist, uint64_t* __restrict__ acc, uint64_t n) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const uint32_t v = hist[i];
    acc[i] = v;
    if (v) hist[i] = 0;
  }
}

// ---- tiled fold / mirror for tables beyond the L2 (W >= 12) -----------------------------------------------------------
// One thread per id reads its partner at revcomp(id): neighbouring ids have partners 4^(W-1) apart, every partner read
// (and mirror write) moves a whole 32-byte sector for 4 or 8 bytes (W = 14, BOTH: fold 7.9 ms + mirror 3.9 ms).  Written
// as id = (a, m, b) with a = top 3 letters, b = low 3 letters, m = the W-6 letters between them,
//     revcomp(a, m, b) = (revcomp3(b), revcomp(m), revcomp3(a)):
// the 64 x 64 ids of a middle m ("tile": 64 rows of 64 consecutive ids) have their partners exactly in the tile of
// revcomp(m).  A CTA loads the two tiles row by row, pairs them through shared memory and writes rows: every global
// access is a 256- or 512-byte run.
#define FOLD_TS 64  // ids per tile row = 4^3
__global__ void k_mirror(uint64_t* __restrict__ counts, int W, uint64_t n);
__device__ __forceinline__ uint32_t revcomp3(uint32_t x) { return revcomp_id(x, 3); }
// FIRST: the 64-bit table is written (= pair sum at canonical ids, at the others too with MIRROR, else 0); otherwise the
// pair sums are added at the canonical ids.  The uint32 table is cleared either way.
template <bool FIRST, bool MIRROR>
__global__ void __launch_bounds__(256) k_fold_canon_tiled(uint32_t* __restrict__ hist, uint64_t* __restrict__ acc, int W) {
  __shared__ uint32_t tA[FOLD_TS * FOLD_TS], tB[FOLD_TS * FOLD_TS];
  const uint32_t m = blockIdx.x, mp = revcomp_id(m, W - 6);
  if (m > mp) return;  // the pair of tiles belongs to the CTA of the smaller middle
  const bool self = m == mp;
  const int sh_a = 2 * W - 6;
  for (uint32_t e = threadIdx.x; e < FOLD_TS * FOLD_TS; e += 256) {
    const uint32_t a = e >> 6, b = e & 63u;
    tA[e] = hist[((uint64_t)a << sh_a) | ((uint64_t)m << 6) | b];
    if (!self) tB[e] = hist[((uint64_t)a << sh_a) | ((uint64_t)mp << 6) | b];
  }
  __syncthreads();
  for (uint32_t e = threadIdx.x; e < FOLD_TS * FOLD_TS; e += 256) {
    const uint32_t a = e >> 6, b = e & 63u;
    const uint32_t pe = revcomp3(b) * FOLD_TS + revcomp3(a);  // where the partner sits in the other tile
#pragma unroll
    for (int t = 0; t < 2; t++) {
      if (t == 1 && self) break;
      const uint32_t mid = t ? mp : m, pmid = t ? m : mp;
      const uint64_t j = ((uint64_t)a << sh_a) | ((uint64_t)mid << 6) | b;
      const uint64_t rj = ((uint64_t)(pe >> 6) << sh_a) | ((uint64_t)pmid << 6) | (pe & 63u);
      const uint32_t own = t ? tB[e] : tA[e];
      const uint32_t other = (t || self) ? tA[pe] : tB[pe];
      const uint64_t v = j == rj ? (uint64_t)own : (uint64_t)own + other;
      if (FIRST) acc[j] = (j <= rj || MIRROR) ? v : 0ull;
      else if (j <= rj && v) acc[j] += v;
      if (own) hist[j] = 0;
    }
  }
}
// the pair count of every canonical id copied to its reverse complement (k_mirror), tile by tile
__global__ void __launch_bounds__(256) k_mirror_tiled(uint64_t* __restrict__ counts, int W) {
  extern __shared__ uint64_t tiles[];  // two tiles of 64 x 64
  uint64_t* tA = tiles;
  uint64_t* tB = tiles + FOLD_TS * FOLD_TS;
  const uint32_t m = blockIdx.x, mp = revcomp_id(m, W - 6);
  if (m > mp) return;
  const bool self = m == mp;
  const int sh_a = 2 * W - 6;
  for (uint32_t e = threadIdx.x; e < FOLD_TS * FOLD_TS; e += 256) {
    const uint32_t a = e >> 6, b = e & 63u;
    tA[e] = counts[((uint64_t)a << sh_a) | ((uint64_t)m << 6) | b];
    if (!self) tB[e] = counts[((uint64_t)a << sh_a) | ((uint64_t)mp << 6) | b];
  }
  __syncthreads();
  for (uint32_t e = threadIdx.x; e < FOLD_TS * FOLD_TS; e += 256) {
    const uint32_t a = e >> 6, b = e & 63u;
    const uint32_t pe = revcomp3(b) * FOLD_TS + revcomp3(a);
#pragma unroll
    for (int t = 0; t < 2; t++) {
      if (t == 1 && self) break;
      const uint32_t mid = t ? mp : m, pmid = t ? m : mp;
      const uint64_t j = ((uint64_t)a << sh_a) | ((uint64_t)mid << 6) | b;
      const uint64_t rj = ((uint64_t)(pe >> 6) << sh_a) | ((uint64_t)pmid << 6) | (pe & 63u);
      if (j > rj) counts[j] = (t || self) ? tA[pe] : tB[pe];
    }
  }
}
// tiled from W = 8 on (2 letters of middle): at W = 10 the one-thread-per-id fold took 23 us of a 0.48 ms step although the
// table is L2-resident (scattered 4-byte partner reads), the tiled one 256- / 512-byte runs only.  PENG_FOLD_TILED_MIN_W for A/B.
static int fold_tiled_min_w() {
  static const int v = getenv("PENG_FOLD_TILED_MIN_W") ? atoi(getenv("PENG_FOLD_TILED_MIN_W")) : 8;
  return std::max(8, v);
}
#define FOLD_TILED_MIN_W fold_tiled_min_w()
static int launch_fold_canon(peng_handle_t h, int W, bool first, bool mirror) {
  const uint64_t n = h->n_patterns;
  if (W >= FOLD_TILED_MIN_W && !h->fold_untiled) {
    const unsigned grid = 1u << (2 * (W - 6));
    if (first && mirror) k_fold_canon_tiled<true, true><<<grid, 256, 0, h->stream>>>(h->d_hist32, h->d_counts, W);
    else if (first) k_fold_canon_tiled<true, false><<<grid, 256, 0, h->stream>>>(h->d_hist32, h->d_counts, W);
    else k_fold_canon_tiled<false, false><<<grid, 256, 0, h->stream>>>(h->d_hist32, h->d_counts, W);
  } else {
    if (first && mirror) k_fold_canon_first<true><<<grid_for(n, 256), 256, 0, h->stream>>>(h->d_hist32, h->d_counts, W, n);
    else if (first) k_fold_canon_first<false><<<grid_for(n, 256), 256, 0, h->stream>>>(h->d_hist32, h->d_counts, W, n);
    else k_fold_canon<<<grid_for(n, 256), 256, 0, h->stream>>>(h->d_hist32, h->d_counts, W, n);
  }
  LAUNCH_CHECK(h);
  return PENG_OK;
}
static int launch_mirror(peng_handle_t h) {
  if (h->W >= FOLD_TILED_MIN_W && !h->fold_untiled) {
    const size_t smem = 2 * FOLD_TS * FOLD_TS * sizeof(uint64_t);
    CU(h, cudaFuncSetAttribute(k_mirror_tiled, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_mirror_tiled<<<1u << (2 * (h->W - 6)), 256, smem, h->stream>>>(h->d_counts, h->W);
  } else {
    k_mirror<<<grid_for(h->n_patterns, 256), 256, 0, h->stream>>>(h->d_counts, h->W, h->n_patterns);
  }
  LAUNCH_CHECK(h);
  return PENG_OK;
}

// 32-bit exchange (peng_cuda_count_local32 / _finish32): ltot rides in word n of the uint32 table
__global__ void k_ltot_to32(const uint64_t* __restrict__ ltot64, uint32_t* __restrict__ word) { *word = (uint32_t)*ltot64; }
__global__ void k_ltot_from32(uint32_t* __restrict__ word, uint64_t* __restrict__ ltot64) {
  *ltot64 = *word;
  *word = 0;
}

// -----------------------------------------------------------------------------------------------------
// Fused cross-GPU sum + fold + mirror over NVLink peer memory (multi-GPU counting step).
//
// What it replaces: k_ltot_to32 -> ncclAllReduce(4^W uint32) on NCCL's stream -> k_ltot_from32 -> k_fold_canon_first ->
// read-back — a flat +0.14 ms on a 0.5 ms step, the same at 2 and at 8 GPUs (latency, not bandwidth).  Here ONE kernel
// per rank does the whole exchange on the handle's stream, two-shot:
//   arrive  : every rank tells every peer "my launch table is complete" (flag word in the peer's memory, release / acquire
//             at system scope) and waits for all of them;
//   reduce  : the rank that owns a unit reads that unit's uint32 bins from ALL ranks' launch tables (16-byte loads over
//             NVLink), adds them, folds reverse-complement pairs through shared memory, and
//   publish : writes the finished 64-bit bins — canonical AND mirrored copy — straight into EVERY rank's 64-bit table
//             (32-byte peer stores): reduce-scatter, fold, mirror and all-gather in one pass;
//   depart  : the last CTA of a rank tells the peers "my stores have landed" and waits for theirs, so that whatever
//             follows on the stream sees a complete table and the launch table may be cleared again.
// BOTH strands: a unit is a quarter of a pair of 64 x 64 tiles (k_fold_canon_tiled's geometry: id = (a, m, b), the partner
// of (a, m, b) is (revcomp3(b), revcomp(m), revcomp3(a))): the 16 rows a with a & 3 == q of tile m and the 16-column band
// b >> 4 == 3 - q of tile revcomp(m) hold each other's partners, every access is a 256- or 64-byte run.  Units of a middle
// m belong to rank m % world.  PLUS: a unit is 1024 consecutive ids.
// Nothing here can deadlock on a peer that never arrives: the waits give up after PENG_PEER_TIMEOUT_MS (default 20 s:
// ranks may legitimately arrive far apart — uneven shards, a first-touch allocation) and raise the error word.
struct PeerArgs {
  uint32_t* hist[8];
  uint64_t* counts[8];
  uint32_t* flags[8];
  int rank, world, W;
  uint32_t epoch;
  uint64_t n;
  long long timeout_cycles;  // how long a rank waits for the others before it gives up and raises the error word
  const uint32_t* units;     // BOTH: the units (m << 2 | q) this rank owns, n_owned of them (built by the host: even over the CTAs)
  uint32_t n_owned;
};
__device__ __forceinline__ void st_release_sys(uint32_t* a, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(a), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* a) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(a) : "memory");
  return v;
}
// Table reads from a peer: ordinary (weak) 16-byte loads.  They are ordered behind the acquire of the peer's arrival flag
// (ld.acquire.sys, which also invalidates this SM's L1: CCTL.IVALL in the SASS), so they see the peer's finished table.
// `ld.relaxed.sys` here made every lane's 16 bytes a request of its own (LDG.STRONG.SYS is not coalesced): ~150 GB/s over
// the link and 20 us per unit at N = 2 (profiles/r2_peer_exchange_trace.txt).
__device__ __forceinline__ uint4 ld_peer_u4(const uint32_t* p) {
  uint4 v;
  asm volatile("ld.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void peer_wait_all(const uint32_t* my_flags, int world, uint32_t epoch, uint32_t* err, uint32_t phase,
                                              long long timeout_cycles) {
  if ((int)threadIdx.x < world) {
    const long long t0 = clock64();
    while ((int32_t)(ld_acquire_sys(my_flags + threadIdx.x) - epoch) < 0) {
      if (clock64() - t0 > timeout_cycles) { *err = 0x100u | (phase << 4) | threadIdx.x; break; }  // a peer never arrived
      __nanosleep(100);
    }
  }
  __syncthreads();
}
__device__ __forceinline__ void st_peer_4u64(uint64_t* p, uint64_t a, uint64_t b, uint64_t c, uint64_t d) {
  asm volatile("st.global.v2.u64 [%0], {%1, %2};" :: "l"(p), "l"(a), "l"(b) : "memory");
  asm volatile("st.global.v2.u64 [%0], {%1, %2};" :: "l"(p + 2), "l"(c), "l"(d) : "memory");
}

__device__ __forceinline__ void st_peer_u4(uint32_t* p, uint4 v) {
  asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" :: "l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// What crosses the link (second version; the first one published finished 64-bit bins, canonical and mirrored copy: 12 B
// per bin and rank, and the all-to-all of 16-byte accesses only reaches ~300 GB/s per GPU — 86 us at N = 8): the owner of
// a unit reads the unit's uint32 bins from all ranks and writes the pair sums back IN PLACE, as uint32, at the unit's HOME
// positions of every rank's launch table (nobody else reads or writes those positions during the exchange): 4 B read +
// 2 B written per bin and rank.  Home = the rows a & 3 == q of tile m for m < revcomp(m) (the band of tile revcomp(m) is
// read, never written), the whole tile for a self-paired middle (one CTA loads it completely before it stores).  The
// mirror copy and the widening to 64 bits are local work afterwards: k_expand_home_tiled (k_fold_first for PLUS), which
// also leaves the launch table clean.
template <bool BOTH>
__global__ void __launch_bounds__(256) k_peer_exchange32(PeerArgs A) {
  __shared__ __align__(16) uint32_t tS[4096];  // normal unit: tX = tS[0 .. 1023], tY = tS[1024 .. 2047]; self-paired middle: the whole tile
  __shared__ uint32_t is_last;
  uint32_t* tX = tS;
  uint32_t* tY = tS + 1024;
  uint32_t* my_flags = A.flags[A.rank];
  uint32_t* err = my_flags + 9;
  const uint32_t t = threadIdx.x;
  // trace (words 16 .. of the flag buffer, nanoseconds): CTA 0 stamps start / arrived / units done, the last CTA its end
  unsigned long long* stamp = reinterpret_cast<unsigned long long*>(my_flags + 16);
  auto now_ns = [] { unsigned long long v; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(v)); return v; };
  if (blockIdx.x == 0 && t == 0) stamp[0] = now_ns();
  // ---- arrive ----
  if (blockIdx.x == 0 && (int)t < A.world) {
    __threadfence_system();
    st_release_sys(A.flags[t] + A.rank, A.epoch);
  }
  peer_wait_all(my_flags, A.world, A.epoch, err, 0u, A.timeout_cycles);
  if (blockIdx.x == 0 && t == 0) stamp[1] = now_ns();
  if (blockIdx.x == 0 && t == 0) {  // ltot rides in word 4^W of every launch table
    uint64_t lt = 0;
    for (int p = 0; p < A.world; p++) {
      uint32_t v;
      asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(A.hist[p] + A.n) : "memory");
      lt += v;
    }
    A.counts[A.rank][A.n] = lt;
  }
  if (BOTH) {
    const int W = A.W, sh_a = 2 * W - 6;
    for (uint32_t ju = blockIdx.x; ju < A.n_owned; ju += gridDim.x) {
      const uint32_t u = A.units[ju];
      const uint32_t m = u >> 3, q = u & 7u, mp = revcomp_id(m, W - 6);
      if (q == 4u) {  // self-paired middle: the 64 x 64 tile holds its own partners
#pragma unroll 1
        for (uint32_t k = 0; k < 4; k++) {
          const uint32_t e0 = 4u * (t + 256u * k);  // elements e0 .. e0 + 3 = (a, b0 .. b0 + 3)
          const uint64_t j = ((uint64_t)(e0 >> 6) << sh_a) | ((uint64_t)m << 6) | (e0 & 63u);
          uint4 v[8];
#pragma unroll
          for (int p = 0; p < 8; p++)
            if (p < A.world) v[p] = ld_peer_u4(A.hist[p] + j);
          uint4 sv = make_uint4(0, 0, 0, 0);
#pragma unroll
          for (int p = 0; p < 8; p++)
            if (p < A.world) { sv.x += v[p].x; sv.y += v[p].y; sv.z += v[p].z; sv.w += v[p].w; }
          reinterpret_cast<uint4*>(tS)[t + 256u * k] = sv;
        }
        __syncthreads();
#pragma unroll 1
        for (uint32_t k = 0; k < 4; k++) {
          const uint32_t e0 = 4u * (t + 256u * k), a = e0 >> 6;
          const uint64_t j = ((uint64_t)a << sh_a) | ((uint64_t)m << 6) | (e0 & 63u);
          uint32_t o[4];
#pragma unroll
          for (int i = 0; i < 4; i++) {
            const uint32_t e = e0 + i, pe = revcomp3(e & 63u) * 64u + revcomp3(a);
            o[i] = e == pe ? tS[e] : tS[e] + tS[pe];
          }
          const uint4 ov = make_uint4(o[0], o[1], o[2], o[3]);
#pragma unroll
          for (int p = 0; p < 8; p++)
            if (p < A.world) st_peer_u4(A.hist[p] + j, ov);
        }
        __syncthreads();
        continue;
      }
      // X (home): rows a = 4 rx + q of tile m, thread t -> row rx = t >> 4, ids b0 .. b0 + 3
      const uint32_t rx = t >> 4, b0 = (t & 15u) * 4u;
      const uint64_t jX = ((uint64_t)(rx * 4u + q) << sh_a) | ((uint64_t)m << 6) | b0;
      // Y (partners): the 16-column band 3 - q of tile mp, thread t -> row ay = t >> 2, band columns yl0 .. yl0 + 3
      const uint32_t ay = t >> 2, yl0 = (t & 3u) * 4u, band = 16u * (3u - q);
      const uint64_t jY = ((uint64_t)ay << sh_a) | ((uint64_t)mp << 6) | (band + yl0);
      uint4 vx[8], vy[8];
#pragma unroll
      for (int p = 0; p < 8; p++) {
        if (p < A.world) {
          vx[p] = ld_peer_u4(A.hist[p] + jX);
          vy[p] = ld_peer_u4(A.hist[p] + jY);
        }
      }
      uint4 sx = make_uint4(0, 0, 0, 0), sy = make_uint4(0, 0, 0, 0);
#pragma unroll
      for (int p = 0; p < 8; p++) {
        if (p < A.world) {
          sx.x += vx[p].x; sx.y += vx[p].y; sx.z += vx[p].z; sx.w += vx[p].w;
          sy.x += vy[p].x; sy.y += vy[p].y; sy.z += vy[p].z; sy.w += vy[p].w;
        }
      }
      reinterpret_cast<uint4*>(tY)[t] = sy;  // tY[ay * 16 + yl]
      __syncthreads();
      if (blockIdx.x == 0 && t == 0 && ju == 0) stamp[6] = now_ns();
      // the partner of (a, b) sits at Y row revcomp3(b), band column (3 - a2) + 4 (3 - a1), a1 = rx & 3, a2 = rx >> 2
      const uint32_t ylp = (3u - (rx >> 2)) + 4u * (3u - (rx & 3u));
      uint4 ov;
      ov.x = sx.x + tY[revcomp3(b0) * 16 + ylp];
      ov.y = sx.y + tY[revcomp3(b0 + 1) * 16 + ylp];
      ov.z = sx.z + tY[revcomp3(b0 + 2) * 16 + ylp];
      ov.w = sx.w + tY[revcomp3(b0 + 3) * 16 + ylp];
#pragma unroll
      for (int p = 0; p < 8; p++)
        if (p < A.world) st_peer_u4(A.hist[p] + jX, ov);
      __syncthreads();
      if (blockIdx.x == 0 && t == 0 && ju == 0) stamp[7] = now_ns();
    }
    (void)tX;
  } else {
    const uint64_t n_mine = ((A.n >> 10) - (uint64_t)A.rank + (uint64_t)A.world - 1) / (uint64_t)A.world;  // units rank, rank + world, ...
    for (uint64_t ju = blockIdx.x; ju < n_mine; ju += gridDim.x) {
      const uint64_t u = (uint64_t)A.rank + ju * (uint64_t)A.world;
      const uint64_t j = (u << 10) + t * 4u;
      uint4 v[8];
#pragma unroll
      for (int p = 0; p < 8; p++)
        if (p < A.world) v[p] = ld_peer_u4(A.hist[p] + j);
      uint4 sv = make_uint4(0, 0, 0, 0);
#pragma unroll
      for (int p = 0; p < 8; p++)
        if (p < A.world) { sv.x += v[p].x; sv.y += v[p].y; sv.z += v[p].z; sv.w += v[p].w; }
#pragma unroll
      for (int p = 0; p < 8; p++)
        if (p < A.world) st_peer_u4(A.hist[p] + j, sv);
    }
  }
  // ---- depart ----
  if (blockIdx.x == 0 && t == 0) stamp[2] = now_ns();
  __threadfence_system();
  __syncthreads();
  if (blockIdx.x == 0 && t == 0) stamp[3] = now_ns();
  if (t == 0) is_last = atomicAdd(my_flags + 8, 1u) == gridDim.x - 1 ? 1u : 0u;
  __syncthreads();
  if (is_last) {
    __threadfence();
    if (t == 0) stamp[4] = now_ns();
    if (t == 0) my_flags[8] = 0u;
    if ((int)t < A.world) {
      __threadfence_system();
      st_release_sys(A.flags[t] + A.rank, A.epoch + 1u);
    }
    peer_wait_all(my_flags, A.world, A.epoch + 1u, err, 1u, A.timeout_cycles);
    if (t == 0) A.hist[A.rank][A.n] = 0u;  // every rank has read this rank's ltot word
    if (t == 0) stamp[5] = now_ns();
  }
}

// Local half of the exchange (BOTH): the launch table holds the pair sums at the HOME positions (tile m of every pair of
// middles m < revcomp(m), all of a self-paired tile).  One CTA per pair writes tile m as it is and tile revcomp(m) as its
// transposed copy — (a, b) <-> (revcomp3(b), revcomp3(a)) — into the 64-bit table, and clears both tiles of the launch table.
__global__ void __launch_bounds__(256) k_expand_home_tiled(uint32_t* __restrict__ hist, uint64_t* __restrict__ acc, int W) {
  __shared__ __align__(16) uint32_t tA[FOLD_TS * FOLD_TS];
  const uint32_t m = blockIdx.x, mp = revcomp_id(m, W - 6);
  if (m > mp) return;
  const int sh_a = 2 * W - 6;
  const uint32_t t = threadIdx.x;
  const uint4 zero = make_uint4(0, 0, 0, 0);
#pragma unroll
  for (uint32_t k = 0; k < 4; k++) {
    const uint32_t e0 = 4u * (t + 256u * k);
    uint4* src = reinterpret_cast<uint4*>(hist + (((uint64_t)(e0 >> 6) << sh_a) | ((uint64_t)m << 6) | (e0 & 63u)));
    const uint4 v = *src;
    reinterpret_cast<uint4*>(tA)[t + 256u * k] = v;
    *src = zero;
    ulonglong2* dst = reinterpret_cast<ulonglong2*>(acc + (((uint64_t)(e0 >> 6) << sh_a) | ((uint64_t)m << 6) | (e0 & 63u)));
    dst[0] = make_ulonglong2(v.x, v.y);
    dst[1] = make_ulonglong2(v.z, v.w);
  }
  if (m == mp) return;
  __syncthreads();
#pragma unroll
  for (uint32_t k = 0; k < 4; k++) {
    const uint32_t e0 = 4u * (t + 256u * k), ap = e0 >> 6, bp0 = e0 & 63u;  // elements (ap, bp0 .. bp0 + 3) of tile mp
    const uint64_t j = ((uint64_t)ap << sh_a) | ((uint64_t)mp << 6) | bp0;
    const uint32_t col = revcomp3(ap);  // source column in tile m; source rows revcomp3(bp)
    ulonglong2* dst = reinterpret_cast<ulonglong2*>(acc + j);
    dst[0] = make_ulonglong2(tA[revcomp3(bp0) * 64 + col], tA[revcomp3(bp0 + 1) * 64 + col]);
    dst[1] = make_ulonglong2(tA[revcomp3(bp0 + 2) * 64 + col], tA[revcomp3(bp0 + 3) * 64 + col]);
    *reinterpret_cast<uint4*>(hist + j) = zero;
  }
}

// second copy of the pair count at the reverse-complement id (src/base_pattern.cpp:387-392). Only
// non-canonical entries are written and only canonical entries are read, so in place is race-free.
__global__ void k_mirror(uint64_t* __restrict__ counts, int W, uint64_t n) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const uint32_t rc = revcomp_id((uint32_t)i, W);
    if ((uint32_t)i > rc) counts[i] = counts[rc];
  }
}

__global__ void k_gather_u64(const uint64_t* __restrict__ table, const uint64_t* __restrict__ ids, uint64_t n,
                             uint64_t* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = table[ids[i]];
}

__global__ void k_sum_u64(const uint64_t* __restrict__ table, uint64_t n, unsigned long long* __restrict__ total) {
  unsigned long long s = 0;
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) s += table[i];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0 && s) atomicAdd(total, s);
}

// =====================================================================================================
// kernels: background model counts
// =====================================================================================================
__global__ void __launch_bounds__(256) k_bg_counts(const uint32_t* __restrict__ codes,
                                                   const uint32_t* __restrict__ valid,
                                                   const uint32_t* __restrict__ sstart, uint64_t stream_end, int K,
                                                   unsigned long long* __restrict__ out, uint32_t n_bins) {
  extern __shared__ uint32_t shb[];  // n_bins private bins
  for (uint32_t i = threadIdx.x; i < n_bins; i += blockDim.x) shb[i] = 0;
  __syncthreads();
  const uint64_t n = stream_end - PENG_FRONT_PAD;
  for (uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (uint64_t)gridDim.x * blockDim.x) {
    const uint64_t q = PENG_FRONT_PAD + t;
    // separators: the two positions in front of a sequence start / the last two of the stream
    const bool sep = q + 2 >= stream_end || bit_at(sstart, q + 1) || bit_at(sstart, q + 2);
    if (sep) continue;
    uint32_t ids[PENG_MAX_ORDER + 1];
    const uint32_t m = background_ids(codes, valid, sstart, q, K, ids);
    for (int k = 0; k <= K; k++)
      if ((m >> k) & 1u) atomicAdd(shb + v_offset(k) + ids[k], 1u);
  }
  __syncthreads();
  for (uint32_t i = threadIdx.x; i < n_bins; i += blockDim.x)
    if (shb[i]) atomicAdd(out + i, (unsigned long long)shb[i]);
}

// =====================================================================================================
// kernels: tables
// =====================================================================================================
// ids [lo, lo + n); every table pointer is indexed by the absolute id (sliced tables: moved back by the slice start)
template <typename R>
__global__ void k_bgprob(const R* __restrict__ v, int W, int order, int both, R* __restrict__ out, uint64_t lo, uint64_t n) {
  extern __shared__ unsigned char shv_raw[];
  R* shv = reinterpret_cast<R*>(shv_raw);
  const uint32_t nv = v_offset(order + 1);
  for (uint32_t i = threadIdx.x; i < nv; i += blockDim.x) shv[i] = v[i];
  __syncthreads();
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[lo + i] = bg_prob_entry<R>(shv, W, order, (uint32_t)(lo + i), both != 0);
}

template <typename R>
__global__ void k_stats(const uint64_t* __restrict__ counts, const R* __restrict__ bg, uint64_t ltot,
                        R* __restrict__ expected, R* __restrict__ z, R* __restrict__ logp, uint64_t lo, uint64_t n) {
  const uint64_t i = lo + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < lo + n) stats_of<R>(counts[i], bg[i], ltot, expected + i, z + i, logp + i);
}

// =====================================================================================================
// kernels: IUPAC sums
// =====================================================================================================
__device__ __forceinline__ double warp_sum(double v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ unsigned long long warp_sum(unsigned long long v) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Order-free sums (PENG_F64, and the integer count sums of every mode): block per pattern.
template <typename R, bool BOTH, bool WITH_FLOATS>
__global__ void __launch_bounds__(256) k_iupac_any(const uint64_t* __restrict__ masks, int W,
                                                   const uint64_t* __restrict__ counts, const R* __restrict__ bg,
                                                   const R* __restrict__ expected,
                                                   unsigned long long* __restrict__ out_n, double* __restrict__ out_bg,
                                                   double* __restrict__ out_exp, uint64_t id_lo, uint64_t id_hi) {
  const uint64_t mask = masks[blockIdx.x];
  const IupacShape shape = iupac_shape(mask, W);
  const uint64_t N = 1ull << shape.total_bits;
  unsigned long long sn = 0;
  double sb = 0, se = 0;
  for (uint64_t t = threadIdx.x; t < N; t += blockDim.x) {
    const uint32_t e = iupac_expansion_any(mask, W, shape, t);
    if (e < id_lo || e >= id_hi) continue;  // id-sliced tables: this rank's partial sum
    if (BOTH && !iupac_is_representative(mask, W, e)) continue;
    sn += counts[e];  // counts / bgprob / expected hold the pair value at both ids of a pair
    if (WITH_FLOATS) {
      sb += (double)bg[e];
      se += (double)expected[e];
    }
  }
  __shared__ unsigned long long shn[8];
  __shared__ double shbg[8], shex[8];
  sn = warp_sum(sn);
  if (WITH_FLOATS) { sb = warp_sum(sb); se = warp_sum(se); }
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) { shn[w] = sn; shbg[w] = sb; shex[w] = se; }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long tn = 0;
    double tb = 0, te = 0;
    for (int i = 0; i < (int)(blockDim.x >> 5); i++) { tn += shn[i]; tb += shbg[i]; te += shex[i]; }
    out_n[blockIdx.x] = tn;
    if (WITH_FLOATS) { out_bg[blockIdx.x] = tb; out_exp[blockIdx.x] = te; }
  }
}

// PENG_F32REF, PLUS strand: float sums accumulated sequentially in the reference's expansion order
// (src/iupac_pattern.cpp:428-441).  The block gathers 1024 values at a time, thread 0 adds them in order.
__global__ void __launch_bounds__(256) k_iupac_f32_plus(const uint64_t* __restrict__ masks, int W,
                                                        const float* __restrict__ bg,
                                                        const float* __restrict__ expected,
                                                        float* __restrict__ out_bg, float* __restrict__ out_exp) {
  __shared__ float sbg[1024], sex[1024];
  const uint64_t mask = masks[blockIdx.x];
  const IupacShape shape = iupac_shape(mask, W);
  const uint64_t N = 1ull << shape.total_bits;
  float acc_b = 0.f, acc_e = 0.f;
  for (uint64_t base = 0; base < N; base += 1024) {
    const uint32_t m = (uint32_t)min((uint64_t)1024, N - base);
    for (uint32_t i = threadIdx.x; i < m; i += blockDim.x) {
      const uint32_t e = iupac_expansion_dfs(mask, W, shape, base + i);
      sbg[i] = bg[e];
      sex[i] = expected[e];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (uint32_t i = 0; i < m; i++) {
        acc_b = __fadd_rn(acc_b, sbg[i]);
        acc_e = __fadd_rn(acc_e, sex[i]);
      }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) { out_bg[blockIdx.x] = acc_b; out_exp[blockIdx.x] = acc_e; }
}

// PENG_F32REF, BOTH strands: the reference adds in ascending canonical-id order over distinct ids
// (sorted list, adjacent duplicates skipped).  The block marks the canonical ids in a 4^W-bit bitmap in
// global scratch (L2-resident) and then walks the bitmap in order: ascending + distinct for free.
#define IUPAC_TILE_WORDS 128
__global__ void __launch_bounds__(IUPAC_TILE_WORDS) k_iupac_f32_both(const uint64_t* __restrict__ masks, int W,
                                                                     const float* __restrict__ bg,
                                                                     const float* __restrict__ expected,
                                                                     uint32_t* __restrict__ bitmaps,
                                                                     float* __restrict__ out_bg,
                                                                     float* __restrict__ out_exp) {
  __shared__ float sbg[IUPAC_TILE_WORDS * 32], sex[IUPAC_TILE_WORDS * 32];
  __shared__ uint32_t warp_tot[IUPAC_TILE_WORDS / 32];
  __shared__ uint32_t tile_any[(1u << 20) / 32 / IUPAC_TILE_WORDS / 32 + 1];  // up to W = 10 tracked exactly
  const uint64_t mask = masks[blockIdx.x];
  const IupacShape shape = iupac_shape(mask, W);
  const uint64_t N = 1ull << shape.total_bits;
  const uint32_t n_words = W >= 3 ? (1u << (2 * W - 5)) : 1u;
  uint32_t* bm = bitmaps + (size_t)blockIdx.x * n_words;
  const uint32_t n_tiles = (n_words + IUPAC_TILE_WORDS - 1) / IUPAC_TILE_WORDS;
  const bool track = n_tiles <= 32u * (sizeof(tile_any) / 4);
  for (uint32_t i = threadIdx.x; i < n_words; i += blockDim.x) bm[i] = 0;
  for (uint32_t i = threadIdx.x; i < sizeof(tile_any) / 4; i += blockDim.x) tile_any[i] = 0;
  __syncthreads();
  for (uint64_t t = threadIdx.x; t < N; t += blockDim.x) {
    const uint32_t e = iupac_expansion_any(mask, W, shape, t);
    const uint32_t rc = revcomp_id(e, W);
    const uint32_t c = e < rc ? e : rc;
    atomicOr(bm + (c >> 5), 1u << (c & 31));
    if (track) {
      const uint32_t tile = (c >> 5) / IUPAC_TILE_WORDS;
      atomicOr(tile_any + (tile >> 5), 1u << (tile & 31));
    }
  }
  __syncthreads();
  float acc_b = 0.f, acc_e = 0.f;
  for (uint32_t tile = 0; tile < n_tiles; tile++) {
    if (track && !((tile_any[tile >> 5] >> (tile & 31)) & 1u)) continue;  // block-uniform
    const uint32_t wi = tile * IUPAC_TILE_WORDS + threadIdx.x;
    const uint32_t w = wi < n_words ? bm[wi] : 0u;
    const uint32_t pc = __popc(w);
    // exclusive scan of pc over the block: warp scan + warp totals
    uint32_t incl = pc;
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t y = __shfl_up_sync(0xffffffffu, incl, o);
      if ((threadIdx.x & 31) >= (unsigned)o) incl += y;
    }
    if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint32_t off = incl - pc, total = 0;
    for (int i = 0; i < IUPAC_TILE_WORDS / 32; i++) {
      if (i < (int)(threadIdx.x >> 5)) off += warp_tot[i];
      total += warp_tot[i];
    }
    uint32_t ww = w;
    while (ww) {
      const uint32_t b = __ffs(ww) - 1;
      ww &= ww - 1;
      const uint32_t id = (wi << 5) | b;
      sbg[off] = bg[id];
      sex[off] = expected[id];
      off++;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (uint32_t i = 0; i < total; i++) {
        acc_b = __fadd_rn(acc_b, sbg[i]);
        acc_e = __fadd_rn(acc_e, sex[i]);
      }
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) { out_bg[blockIdx.x] = acc_b; out_exp[blockIdx.x] = acc_e; }
}

// =====================================================================================================
// kernels: EM, FP64 (fast path)
// =====================================================================================================
// new_pwm[p][a] = sum over ids with digit_p = a of  cnt*s / (1 + s/odds),  odds = prod_p pwm[p][digit_p] / bg
//              = s * sum cnt*prob' / (prob' + bg),   prob' = prob / s
// (the common factor s cancels in the row normalisation that follows and is dropped from the accumulators).
//
// ONE pass over the table per iteration for ALL PWMs (16 B per bin and iteration, whatever the number of
// PWMs).  The work is laid out PWM-per-thread, not id-per-thread:
//   * k_em_pairs (once per EM call) writes {(double) count, bg} per id — the two operands every (id, PWM) term needs,
//     one 16-byte load;
//   * a CTA of 256 threads hosts R = 256 / P "slots"; thread (r, m) = (t / P, t % P) evaluates PWM m on the ids of slot
//     r.  A slot is a contiguous range of units of 256 ids (the low 4 letters), handed out evenly over all slots of the
//     grid; every 16-id block of every slot of the CTA arrives in shared memory through a 4-stage cp.async pipeline
//     (each table byte crosses HBM -> SM once) and is read by the P threads of its slot as a broadcast;
//   * because a thread owns a PWM for a whole id range, NOTHING is reduced across lanes inside the loop: the sums for
//     positions 0 and 1 sit in 8 statically indexed registers (the 16 ids of a block are unrolled), those of the other
//     positions in a private shared-memory column that takes one add per block (positions 2, 3) or per unit (>= 4);
//     the prefix product over the positions >= 2 is formed once per block, so a term costs 1 multiply for its
//     probability, 1 add for the denominator, the reciprocal (1 MUFU seed + 2 Newton steps = 4 FMA, 1-2 ulp; the IEEE
//     division is 9+ instructions and a slow path), 2 multiplies and ~1.6 adds: ~10.6 FP64 instructions, against 23
//     (+ 108 shuffles per tile and PWM) in the id-per-lane kernel this replaces;
//   * blocks / groups of 4 ids whose counts are all zero are skipped (sparse long-k-mer tables);
//   * at the end the R slots of a PWM are added in fixed order: one partial row per CTA and PWM, reduced in fixed order
//     by k_em_reduce* -> deterministic.
#define EM_NT 256
#define EM_STAGES 4
#define EM_MAX_R 32
#define EM_MAX_GROUP 64  // PWMs per launch (larger batches run group after group)
#define EM_UNIT_BITS 8   // 256 ids per unit

__device__ __forceinline__ double em_rcp(double d) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));  // ~2^-20 relative (the seed looks at the upper 32 bits)
  double e = fma(-d, r, 1.0);
  r = fma(r, e, r);  // 2^-40
  e = fma(-d, r, 1.0);
  r = fma(r, e, r);  // below the rounding error of a double
  return r;
}

__global__ void k_em_pairs(const uint64_t* __restrict__ counts, const double* __restrict__ bg, uint64_t id_begin, uint64_t n,
                           double2* __restrict__ pairs) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) pairs[i] = make_double2((double)counts[id_begin + i], bg[id_begin + i]);
}

// rows of the private shared-memory columns: positions 2, 3 (one add per 16-id block) and positions >= 6 (one add per
// 256-id unit); positions 0, 1 (per id) and 4, 5 (per unit, predicated adds) live in registers.  The same rows are reused
// at the end to hand the 16 register sums to the cross-slot reduction.
#define EM_ROWS(W) ((4 * (W) - 16) > 16 ? (4 * (W) - 16) : 16)

template <int W>
__global__ void __launch_bounds__(EM_NT, 2) k_em_accumulate_all(
    const double2* __restrict__ pairs /* indexed by id - (unit_begin << 8) */, const double* __restrict__ pwms /* [P][W*4] */,
    const int* __restrict__ active, int P, int R, double inv_saturation, uint32_t unit_begin, uint32_t unit_end,
    double* __restrict__ partials /* [P][nbx][W*4] */, int nbx) {
  static_assert(W >= 6, "positions 0 .. 5 are handled in registers / per block");
  constexpr int WA = W * 4;
  extern __shared__ __align__(16) unsigned char em_smem[];
  double2* stage = reinterpret_cast<double2*>(em_smem);      // [EM_STAGES][R][16]
  double* spw = reinterpret_cast<double*>(stage + EM_STAGES * R * 16);  // [WA][P]: entry-major, the PWMs of a warp side by side
  double* sacc = spw + WA * P;                                // [EM_ROWS(W)][EM_NT]: one private column per thread
  const int t = threadIdx.x;
  const int r = t / P, m = t - r * P;
  const bool worker = r < R;
  const bool live = worker && active[m] != 0;
  if (!__syncthreads_or(live)) return;  // every PWM of this launch has converged
  for (int i = t; i < WA * P; i += EM_NT) {
    const int mm = i / WA, e = i - mm * WA;
    spw[e * P + mm] = pwms[i];
  }
#pragma unroll 4
  for (int e = 0; e < EM_ROWS(W); e++) sacc[e * EM_NT + t] = 0.0;
  // ---- slots: slot s of S = gridDim.x * R owns the units [lo(s), lo(s + 1)) ----
  // the first `rem` slots take q + 1 units, the others q (no 64-bit divisions per slot: they were microseconds of the
  // 40 us a whole W = 10 pass takes)
  const uint32_t S = gridDim.x * (uint32_t)R, n_units = unit_end - unit_begin;
  const uint32_t q_units = n_units / S, rem_units = n_units - q_units * S;
  auto slot_lo = [&](uint32_t s) { return s * q_units + min(s, rem_units); };  // relative to unit_begin
  const uint32_t s0 = blockIdx.x * (uint32_t)R;
  uint32_t n_steps = 0;
  for (int rr = 0; rr < R; rr++) n_steps = max(n_steps, (slot_lo(s0 + rr + 1) - slot_lo(s0 + rr)) << 4);
  // loader role: thread t copies 16-byte piece (t & 15) of the blocks of slots (t >> 4) and (t >> 4) + 16
  uint32_t ld_lo[2], ld_n[2];
#pragma unroll
  for (int j = 0; j < 2; j++) {
    const int lr = (t >> 4) + 16 * j;
    const uint32_t lo = lr < R ? slot_lo(s0 + lr) : 0u, hi = lr < R ? slot_lo(s0 + lr + 1) : 0u;
    ld_lo[j] = lo;
    ld_n[j] = (hi - lo) << 4;
  }
  const uint32_t stage_s = (uint32_t)__cvta_generic_to_shared(stage);
  auto issue = [&](uint32_t k) {
#pragma unroll
    for (int j = 0; j < 2; j++) {
      if (k < ld_n[j]) {
        const int lr = (t >> 4) + 16 * j;
        const double2* src = pairs + (((uint64_t)ld_lo[j] << EM_UNIT_BITS) + ((uint64_t)k << 4) + (t & 15));
        const uint32_t dst = stage_s + (uint32_t)((((k % EM_STAGES) * R + lr) * 16 + (t & 15)) * sizeof(double2));
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dst), "l"(src) : "memory");
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  __syncthreads();  // spw complete
  // compute role
  const uint32_t my_lo = worker ? slot_lo(s0 + r) : 0u;
  const uint32_t my_n = live ? ((slot_lo(s0 + r + 1) - my_lo) << 4) : 0u;
  double pw0[4], pw1[4], acc0[4] = {0, 0, 0, 0}, acc1[4] = {0, 0, 0, 0}, acc4[4] = {0, 0, 0, 0}, acc5[4] = {0, 0, 0, 0};
#pragma unroll
  for (int a = 0; a < 4; a++) {
    pw0[a] = live ? spw[a * P + m] : 0.0;
    pw1[a] = live ? spw[(4 + a) * P + m] : 0.0;
  }
  double pre4 = 0.0, unit_total = 0.0;
#pragma unroll 1
  for (uint32_t k = 0; k < EM_STAGES - 1; k++) issue(k);
#pragma unroll 1
  for (uint32_t k = 0; k < n_steps; k++) {
    asm volatile("cp.async.wait_group %0;" :: "n"(EM_STAGES - 2) : "memory");
    __syncthreads();  // block k of every slot has landed; everybody is done with block k - 1, whose buffer is refilled now
    issue(k + EM_STAGES - 1);
    if (k < my_n) {
      const uint32_t b = k & 15u, u = unit_begin + my_lo + (k >> 4);
      if (b == 0) {
        pre4 = inv_saturation;
#pragma unroll
        for (int q = 4; q < W; q++) pre4 *= spw[(q * 4 + ((u >> (2 * (q - 4))) & 3u)) * P + m];
        unit_total = 0.0;
      }
      const uint32_t d2 = b & 3u, d3 = b >> 2;
      const double2* src = stage + ((k % EM_STAGES) * R + r) * 16;
      const double pre2 = pre4 * spw[(12 + d3) * P + m] * spw[(8 + d2) * P + m];
      double blk = 0.0;
#pragma unroll
      for (int half = 0; half < 2; half++) {  // 8 ids at a time: 32 registers of table data in flight
        double2 v[8];
        unsigned long long any = 0;
#pragma unroll
        for (int i = 0; i < 8; i++) {
          v[i] = src[8 * half + i];
          any |= (unsigned long long)__double_as_longlong(v[i].x);  // counts are >= 0: the bit pattern is zero exactly for an empty bin
        }
        if (any) {
#pragma unroll
          for (int dd = 0; dd < 2; dd++) {
            const int d1 = 2 * half + dd;
            const unsigned long long any4 = (unsigned long long)__double_as_longlong(v[4 * dd].x) | (unsigned long long)__double_as_longlong(v[4 * dd + 1].x) |
                                            (unsigned long long)__double_as_longlong(v[4 * dd + 2].x) | (unsigned long long)__double_as_longlong(v[4 * dd + 3].x);
            if (any4) {
              const double p1 = pre2 * pw1[d1];
              double tt[4];
#pragma unroll
              for (int d0 = 0; d0 < 4; d0++) {
                const double prob = p1 * pw0[d0];
                const double2 cb = v[4 * dd + d0];
                tt[d0] = (cb.x * prob) * em_rcp(prob + cb.y);
                acc0[d0] += tt[d0];
              }
              const double g = (tt[0] + tt[1]) + (tt[2] + tt[3]);
              acc1[d1] += g;
              blk += g;
            }
          }
        }
      }
      sacc[d2 * EM_NT + t] += blk;
      sacc[(4 + d3) * EM_NT + t] += blk;
      unit_total += blk;
      if (b == 15u) {
        const uint32_t d4 = u & 3u, d5 = (u >> 2) & 3u;
#pragma unroll
        for (int a = 0; a < 4; a++) {
          acc4[a] += d4 == (uint32_t)a ? unit_total : 0.0;
          acc5[a] += d5 == (uint32_t)a ? unit_total : 0.0;
        }
#pragma unroll
        for (int q = 6; q < W; q++) sacc[(8 + (q - 6) * 4 + ((u >> (2 * (q - 4))) & 3u)) * EM_NT + t] += unit_total;
      }
    }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncthreads();
  // ---- the R slots of a PWM are added in fixed order: first the shared-memory rows (positions 2, 3, >= 6) ... ----
  for (int i = t; i < P * (WA - 16); i += EM_NT) {
    const int mm = i / (WA - 16), row = i - mm * (WA - 16);
    const int e = row < 8 ? 8 + row : 16 + row;  // row -> entry p * 4 + a: rows 0 .. 7 are positions 2, 3, rows 8 .. are positions 6 ..
    double sum = 0.0;
    for (int rr = 0; rr < R; rr++) sum += sacc[row * EM_NT + rr * P + mm];
    partials[((size_t)mm * nbx + blockIdx.x) * WA + e] = sum;
  }
  __syncthreads();
  // ... then the register sums (positions 0, 1, 4, 5) through the same rows
#pragma unroll
  for (int a = 0; a < 4; a++) {
    sacc[a * EM_NT + t] = acc0[a];
    sacc[(4 + a) * EM_NT + t] = acc1[a];
    sacc[(8 + a) * EM_NT + t] = acc4[a];
    sacc[(12 + a) * EM_NT + t] = acc5[a];
  }
  __syncthreads();
  for (int i = t; i < P * 16; i += EM_NT) {
    const int mm = i >> 4, row = i & 15;
    const int e = row < 8 ? row : 8 + row;  // rows 8 .. 15 are positions 4, 5
    double sum = 0.0;
    for (int rr = 0; rr < R; rr++) sum += sacc[row * EM_NT + rr * P + mm];
    partials[((size_t)mm * nbx + blockIdx.x) * WA + e] = sum;
  }
}

// small-W fallback (W = 2, 4): one thread per accumulator walks the table
__global__ void k_em_accumulate_small(const uint64_t* __restrict__ counts, const double* __restrict__ bg,
                                      const double* __restrict__ pwms, const int* __restrict__ active, int W,
                                      double saturation, uint32_t id_begin, uint32_t id_end,
                                      double* __restrict__ partials) {
  const int m = blockIdx.x;
  if (!active[m]) return;
  const int c = threadIdx.x;
  if (c >= W * 4) return;
  const int p = c >> 2, a = c & 3;
  const double* pw = pwms + (size_t)m * W * 4;
  double s = 0;
  for (uint32_t id = id_begin; id < id_end; id++) {
    if (((id >> (2 * p)) & 3u) != (uint32_t)a) continue;
    const uint64_t cnt = counts[id];
    if (!cnt) continue;
    double prob = 1.0;
    for (int q = 0; q < W; q++) prob *= pw[q * 4 + ((id >> (2 * q)) & 3u)];
    s += ((double)cnt * saturation) * prob / (prob + saturation * bg[id]);
  }
  partials[(size_t)m * (W * 4) + c] = s;
}

// acc[m][i] = sum_b partials[m][b][i] in fixed order (16 strided part sums per accumulator, then those in order)
__global__ void __launch_bounds__(PENG_MAX_W * 4 * 16) k_em_reduce(const double* __restrict__ partials, const int* __restrict__ active,
                                                                   int nbx, int WA, double* __restrict__ acc) {
  const int m = blockIdx.x;
  if (!active[m]) return;
  __shared__ double part[16][PENG_MAX_W * 4];
  const int i = threadIdx.x % (PENG_MAX_W * 4), r = threadIdx.x / (PENG_MAX_W * 4);
  if (i < WA) {
    double s = 0;
    for (int b = r; b < nbx; b += 16) s += partials[((size_t)m * nbx + b) * WA + i];
    part[r][i] = s;
  }
  __syncthreads();
  if (r == 0 && i < WA) {
    double s = 0;
#pragma unroll
    for (int k = 0; k < 16; k++) s += part[k][i];
    acc[(size_t)m * WA + i] = s;
  }
}

// normalise rows, L1 change, iteration bookkeeping (src/peng.cpp:104-144); one warp per PWM
__global__ void k_em_update_f64(const double* __restrict__ acc, double* __restrict__ old_pwm,
                                double* __restrict__ change, int* __restrict__ iters, int* __restrict__ active, int W,
                                double threshold, int max_iterations, int* __restrict__ n_active) {
  const int m = blockIdx.x;
  if (threadIdx.x != 0) return;
  if (active[m]) {
    double ch = 0;
    for (int p = 0; p < W; p++) {
      double s = 0;
      for (int a = 0; a < 4; a++) s += acc[(size_t)m * W * 4 + p * 4 + a];
      for (int a = 0; a < 4; a++) {
        const double nv = acc[(size_t)m * W * 4 + p * 4 + a] / s;
        ch += fabs(nv - old_pwm[(size_t)m * W * 4 + p * 4 + a]);
        old_pwm[(size_t)m * W * 4 + p * 4 + a] = nv;
      }
    }
    change[m] = ch;
    iters[m] += 1;
    const int still = !(ch <= threshold || iters[m] >= max_iterations);
    active[m] = still;
    if (still) atomicAdd(n_active, 1);
  }
}

// single-GPU iteration tail: k_em_reduce + k_em_update_f64 in one launch (one block per PWM).  16 threads per accumulator
// take every 16th partial row each, their sums are added in fixed order: deterministic, and 296 rows cost 19 dependent
// loads instead of 296 (the one-thread-per-accumulator version took 37 us per iteration — as long as the table pass at W = 10).
#define EM_RED_PARTS 16
__global__ void __launch_bounds__(PENG_MAX_W * 4 * EM_RED_PARTS) k_em_reduce_update(
    const double* __restrict__ partials, int nbx, double* __restrict__ acc, double* __restrict__ old_pwm, double* __restrict__ change,
    int* __restrict__ iters, int* __restrict__ active, int W, double threshold, int max_iterations, int* __restrict__ n_active) {
  const int m = blockIdx.x, WA = W * 4;
  if (!active[m]) return;
  __shared__ double part[EM_RED_PARTS][PENG_MAX_W * 4];
  __shared__ double sh[PENG_MAX_W * 4];
  const int i = threadIdx.x % (PENG_MAX_W * 4), r = threadIdx.x / (PENG_MAX_W * 4);
  if (i < WA) {
    double s = 0;
    for (int b = r; b < nbx; b += EM_RED_PARTS) s += partials[((size_t)m * nbx + b) * WA + i];
    part[r][i] = s;
  }
  __syncthreads();
  if (r == 0 && i < WA) {
    double s = 0;
#pragma unroll
    for (int k = 0; k < EM_RED_PARTS; k++) s += part[k][i];
    sh[i] = s;
    acc[(size_t)m * WA + i] = s;
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  double ch = 0;
  for (int p = 0; p < W; p++) {
    double s = 0;
    for (int a = 0; a < 4; a++) s += sh[p * 4 + a];
    for (int a = 0; a < 4; a++) {
      const double nv = sh[p * 4 + a] / s;
      ch += fabs(nv - old_pwm[(size_t)m * WA + p * 4 + a]);
      old_pwm[(size_t)m * WA + p * 4 + a] = nv;
    }
  }
  change[m] = ch;
  iters[m] += 1;
  const int still = !(ch <= threshold || iters[m] >= max_iterations);
  active[m] = still;
  if (still) atomicAdd(n_active, 1);
}

__global__ void k_em_init(double* __restrict__ change, int* __restrict__ iters, int* __restrict__ active, int n, int W,
                          double threshold, int max_iterations) {
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m < n) {
    change[m] = (double)W;  // (src/peng.cpp:101)
    iters[m] = 0;
    active[m] = !((double)W <= threshold || 0 >= max_iterations);
  }
}

// final normalisation of IUPACPattern(ori, pwm) (src/iupac_pattern.cpp:61)
template <typename R>
__global__ void k_em_final_normalize(R* __restrict__ pwm, int n, int W) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;  // row index over n*W
  if (r < n * W) {
    R s = (R)0;
    for (int a = 0; a < 4; a++) s = Ar<R>::add(s, pwm[(size_t)r * 4 + a]);
    for (int a = 0; a < 4; a++) pwm[(size_t)r * 4 + a] = Ar<R>::div(pwm[(size_t)r * 4 + a], s);
  }
}

// =====================================================================================================
// kernels: EM, float32 reference-faithful path
// =====================================================================================================
// (1) every term in the reference's float32 arithmetic, element-wise and therefore order-free
__global__ void k_em_terms_f32(const uint64_t* __restrict__ counts, const float* __restrict__ bg,
                               const float* __restrict__ pwms, const int* __restrict__ active, int W, float saturation,
                               uint64_t n, float* __restrict__ terms) {
  const int m = blockIdx.y;
  if (!active[m]) return;
  __shared__ float spwm[PENG_MAX_W * 4];
  for (int i = threadIdx.x; i < W * 4; i += blockDim.x) spwm[i] = pwms[(size_t)m * W * 4 + i];
  __syncthreads();
  const uint64_t id = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (id < n) terms[(size_t)m * n + id] = em_term_f32(spwm, W, (uint32_t)id, counts[id], bg[id], saturation);
}

// (2) the reference adds the terms to new_pwm[p][a] one at a time in ascending id order
// (src/peng.cpp:121-127): 4W independent sequential float32 chains per PWM.  One warp per chain: the lanes
// gather the next 128 chain elements (coalesced where the chain is contiguous), lane 0 adds them in order
// out of shared memory while the loads of the following batch are already in flight.
#define CHAIN_BATCH 128
__global__ void __launch_bounds__(128) k_em_chains_f32(const float* __restrict__ terms, const int* __restrict__ active,
                                                       int W, uint64_t n, float* __restrict__ acc) {
  const int m = blockIdx.y;
  if (!active[m]) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int p = blockIdx.x;          // position
  const uint32_t a = warp;           // letter: 4 warps per block
  __shared__ __align__(16) float buf[4][CHAIN_BATCH];
  const float* t = terms + (size_t)m * n;
  const uint32_t len = (uint32_t)(n >> 2);  // elements per chain
  float s = 0.f;
  float nxt[CHAIN_BATCH / 32];
#pragma unroll
  for (int k = 0; k < CHAIN_BATCH / 32; k++) {
    const uint32_t idx = k * 32 + lane;
    nxt[k] = idx < len ? t[em_chain_id(p, a, idx)] : 0.f;
  }
  for (uint32_t base = 0; base < len; base += CHAIN_BATCH) {
#pragma unroll
    for (int k = 0; k < CHAIN_BATCH / 32; k++) buf[warp][k * 32 + lane] = nxt[k];
    __syncwarp();
    const uint32_t nb = base + CHAIN_BATCH;
#pragma unroll
    for (int k = 0; k < CHAIN_BATCH / 32; k++) {
      const uint32_t idx = nb + k * 32 + lane;
      nxt[k] = idx < len ? t[em_chain_id(p, a, idx)] : 0.f;
    }
    if (lane == 0) {
      const uint32_t cnt = min((uint32_t)CHAIN_BATCH, len - base);
      if (cnt == CHAIN_BATCH) {
        const float4* b4 = reinterpret_cast<const float4*>(buf[warp]);
#pragma unroll 8
        for (int i = 0; i < CHAIN_BATCH / 4; i++) {
          const float4 x = b4[i];
          s = __fadd_rn(s, x.x); s = __fadd_rn(s, x.y); s = __fadd_rn(s, x.z); s = __fadd_rn(s, x.w);
        }
      } else {
        for (uint32_t i = 0; i < cnt; i++) s = __fadd_rn(s, buf[warp][i]);
      }
    }
    __syncwarp();
  }
  if (lane == 0) acc[(size_t)m * W * 4 + p * 4 + a] = s;
}

// (3) normalize_pwm + change + swap (src/iupac_pattern.cpp:291-303, src/peng.cpp:129-143), float32, in order
__global__ void k_em_update_f32(const float* __restrict__ acc, float* __restrict__ old_pwm, int* __restrict__ iters,
                                int* __restrict__ active, int W, float threshold, int max_iterations,
                                int* __restrict__ n_active) {
  const int m = blockIdx.x;
  if (threadIdx.x != 0 || !active[m]) return;
  float nw[PENG_MAX_W * 4];
  for (int p = 0; p < W; p++) {
    float s = 0.f;
    for (int a = 0; a < 4; a++) s = __fadd_rn(s, acc[(size_t)m * W * 4 + p * 4 + a]);
    for (int a = 0; a < 4; a++) nw[p * 4 + a] = __fdiv_rn(acc[(size_t)m * W * 4 + p * 4 + a], s);
  }
  float ch = 0.f;
  for (int i = 0; i < W * 4; i++) {
    ch = __fadd_rn(ch, fabsf(__fsub_rn(nw[i], old_pwm[(size_t)m * W * 4 + i])));
    old_pwm[(size_t)m * W * 4 + i] = nw[i];
  }
  iters[m] += 1;
  const int still = !(ch <= threshold || iters[m] >= max_iterations);
  active[m] = still;
  if (still) atomicAdd(n_active, 1);
}

__global__ void k_em_init_f32(int* __restrict__ iters, int* __restrict__ active, int n, int W, float threshold,
                              int max_iterations) {
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m < n) {
    iters[m] = 0;
    active[m] = !((float)W <= threshold || 0 >= max_iterations);
  }
}

// =====================================================================================================
// C ABI: lifetime
// =====================================================================================================
extern "C" int peng_cuda_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

extern "C" int peng_cuda_create(int device, peng_handle_t* out) {
  if (!out) return fail(nullptr, PENG_E_INVALID, "out is NULL");
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0)
    return fail(nullptr, PENG_E_CUDA, "no CUDA device (%s): libpeng_cuda has no CPU fallback",
                e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
  if (device < 0 || device >= n) return fail(nullptr, PENG_E_INVALID, "device %d out of range [0,%d)", device, n);
  e = cudaSetDevice(device);
  if (e != cudaSuccess) return fail(nullptr, PENG_E_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
  peng_cuda_ctx* h = new peng_cuda_ctx();
  h->device = device;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) h->sm_count = prop.multiProcessorCount;
  e = cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) {
    delete h;
    return fail(nullptr, PENG_E_CUDA, "cudaStreamCreate: %s", cudaGetErrorString(e));
  }
  h->stream = h->own_stream;
  if (cudaStreamCreateWithFlags(&h->side_stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming) != cudaSuccess) {
    g_create_error = "peng_cuda_create: side stream / events";
    delete h;
    return PENG_E_CUDA;
  }
  if (const char* e = getenv("PENG_PART_CFG")) h->part_cfg = atoi(e);
  h->fold_untiled = getenv("PENG_FOLD_UNTILED") != nullptr;
  if (const char* e = getenv("PENG_INGEST_MODE")) h->ingest_mode = !strcmp(e, "raw") ? 1 : !strcmp(e, "packed") ? 2 : 0;
  if (const char* e = getenv("PENG_HOST_PACK_FRACTION")) h->ingest_fraction = atof(e);
  if (const char* e = getenv("PENG_DEEP_MODE")) h->deep_mode = atoi(e);
  h->select_on_host = getenv("PENG_SELECT_HOST") != nullptr;
  h->use_nibble = getenv("PENG_NIBBLE") != nullptr;
  if (const char* e = getenv("PENG_PEER_TIMEOUT_MS")) h->peer.timeout_ms = std::max(1.0, atof(e));
  *out = h;
  return PENG_OK;
}

static void free_tables(peng_handle_t h) {
  for (int k = 0; k <= PENG_MAX_ORDER; k++) { dev_free(&h->d_bg32[k]); dev_free(&h->d_bg64[k]); }
  dev_free(&h->d_exp32); dev_free(&h->d_z32); dev_free(&h->d_logp32);
  dev_free(&h->d_exp64); dev_free(&h->d_z64); dev_free(&h->d_logp64);
  h->stats_ready = false;
  h->table_gen++;
  dev_free(&h->em.d_pairs);  // as large as two tables: not kept across table changes
  h->em.cap_pairs = 0;
}
static void free_em_buffers(peng_handle_t h) {
  dev_free(&h->em.d_old); dev_free(&h->em.d_acc); dev_free(&h->em.d_part); dev_free(&h->em.d_change);
  dev_free(&h->em.d_iter); dev_free(&h->em.d_active); dev_free(&h->em.d_nactive);
  dev_free(&h->em.d_pairs);
  h->em.cap_small = h->em.cap_part = h->em.cap_pairs = 0;
  h->em.pairs_gen = 0;
}
static void free_em(peng_handle_t h) {  // closes the session; the buffers stay for the next one
  h->em.open = false;
}

extern "C" void peng_cuda_destroy(peng_handle_t h) {
  if (!h) return;
  cudaSetDevice(h->device);
  cudaStreamSynchronize(h->stream);
  peng_cuda_peer_detach(h);
  dev_free(&h->peer.d_flags);
  dev_free(&h->peer.d_units);
  dev_free(&h->d_codes); dev_free(&h->d_valid); dev_free(&h->d_sstart); dev_free(&h->d_kstart);
  dev_free(&h->d_offsets); dev_free(&h->d_hist32); dev_free(&h->d_counts); dev_free(&h->d_bgcounts);
  dev_free(&h->d_v32); dev_free(&h->d_v64);
  free_tables(h);
  free_em_buffers(h);
  if (h->d_scratch) cudaFree(h->d_scratch);
  dev_free(&h->d_sel);
  if (h->copy_stream) cudaStreamDestroy(h->copy_stream);
  for (int i = 0; i < 2; i++) if (h->copy_ev[i]) cudaEventDestroy(h->copy_ev[i]);
  for (int i = 0; i < 2; i++) if (h->pin[i]) cudaFreeHost(h->pin[i]);
  if (h->pin_offs) cudaFreeHost(h->pin_offs);
  for (int i = 0; i < 2; i++) if (h->pin_ev[i]) cudaEventDestroy(h->pin_ev[i]);
  dev_free(&h->d_planes);
  dev_free(&h->d_part_buckets); dev_free(&h->d_part_fill); dev_free(&h->d_slow_list); dev_free(&h->d_slow_count);
  for (int i = 0; i < 3; i++) { dev_free(&h->d_deep_buckets[i]); dev_free(&h->d_deep_fill[i]); }
  for (int i = 0; i < 11; i++) if (h->ev[i]) cudaEventDestroy(h->ev[i]);
  if (h->own_stream) cudaStreamDestroy(h->own_stream);
  if (h->side_stream) cudaStreamDestroy(h->side_stream);
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  if (h->ev_join) cudaEventDestroy(h->ev_join);
  delete h;
}

extern "C" const char* peng_cuda_last_error(peng_handle_t h) { return h ? h->err.c_str() : g_create_error.c_str(); }
extern "C" uint64_t peng_cuda_launch_count(peng_handle_t h) { return h ? h->launches : 0; }

extern "C" int peng_cuda_set_stream(peng_handle_t h, void* s) {
  if (!h) return PENG_E_INVALID;
  h->stream = s ? (cudaStream_t)s : h->own_stream;
  return PENG_OK;
}
extern "C" int peng_cuda_enable_timing(peng_handle_t h, int on) {
  if (!h) return PENG_E_INVALID;
  CU(h, cudaSetDevice(h->device));
  if (on && !h->ev[0])
    for (int i = 0; i < 11; i++) CU(h, cudaEventCreate(&h->ev[i]));
  h->timing = on != 0;
  return PENG_OK;
}
extern "C" int peng_cuda_get_timing(peng_handle_t h, int phase, float* ms) {
  if (!h || !h->ev[0] || phase < 0 || phase > 7 || !ms) return fail(h, PENG_E_INVALID, "get_timing: timing not enabled");
  CU(h, cudaSetDevice(h->device));
  if (phase == 7) {  // the last peer exchange kernel (waits for the other ranks included)
    CU(h, cudaEventSynchronize(h->ev[10]));
    CU(h, cudaEventElapsedTime(ms, h->ev[9], h->ev[10]));
    return PENG_OK;
  }
  CU(h, cudaEventSynchronize(h->ev[5]));
  // events: 0 start | 1 k-mer starts marked | 4 counting done | 5 folded; of the first chunk of the partition path:
  // 6 stage P launched | 2 stage P done | 3 slow units done (side stream, beside stage H / P2) | 8 stage P2 done (deep
  // path, else = 2) | 7 stage H done
  static const int from[7] = {0, 1, 4, 6, 2, 8, 2}, to[7] = {1, 4, 5, 2, 3, 7, 8};
  CU(h, cudaEventElapsedTime(ms, h->ev[from[phase]], h->ev[to[phase]]));
  return PENG_OK;
}
#define MARK(h, i) do { if ((h)->timing) CU(h, cudaEventRecord((h)->ev[i], (h)->stream)); } while (0)

extern "C" int peng_cuda_synchronize(peng_handle_t h) {
  if (!h) return PENG_E_INVALID;
  CU(h, cudaSetDevice(h->device));
  CU(h, cudaStreamSynchronize(h->stream));
  return PENG_OK;
}

// =====================================================================================================
// C ABI: ingest
// =====================================================================================================
static int alloc_stream(peng_handle_t h, uint64_t n_seq, uint64_t total_bases) {
  h->n_seq = n_seq;
  h->total_bases = total_bases;
  h->stream_end = PENG_FRONT_PAD + total_bases + 2 * n_seq;
  h->n_words32 = (h->stream_end + PENG_TAIL_PAD + 31) / 32 + 8;
  if (h->cap_words32 < h->n_words32) {
    TRY(dev_alloc(h, &h->d_codes, h->n_words32 * 2 + 8));
    TRY(dev_alloc(h, &h->d_valid, h->n_words32 + 8));
    TRY(dev_alloc(h, &h->d_sstart, h->n_words32 + 8));
    TRY(dev_alloc(h, &h->d_kstart, h->n_words32 + 8));
    TRY(dev_alloc(h, &h->d_planes, h->n_words32 + 8));
    h->cap_planes = h->n_words32 + 8;
    h->cap_words32 = h->n_words32;
  }
  // the few guard words behind n_words32 are read by funnel shifts at the very end of the stream
  CU(h, cudaMemsetAsync(h->d_codes + h->n_words32 * 2, 0, 8 * 4, h->stream));
  CU(h, cudaMemsetAsync(h->d_valid + h->n_words32, 0, 8 * 4, h->stream));
  CU(h, cudaMemsetAsync(h->d_sstart + h->n_words32, 0, 8 * 4, h->stream));
  CU(h, cudaMemsetAsync(h->d_kstart + h->n_words32, 0, 8 * 4, h->stream));  // the rest is written by every count
  CU(h, cudaMemsetAsync(h->d_planes + h->n_words32, 0, 8 * sizeof(uint2), h->stream));
  h->planes_from = h->n_words32;  // no word has its planes yet
  h->pre.done = false;
  h->counts_final = false;
  h->stats_ready = false;
  h->planes_valid = false;
  return PENG_OK;
}

static void seq_lengths(peng_handle_t h, const uint64_t* offsets, uint64_t n_seq) {
  uint32_t mn = 0x7fffffffu, mx = 0;
  for (uint64_t s = 0; s < n_seq; s++) {
    const uint64_t L = offsets[s + 1] - offsets[s];
    if (L > mx) mx = (uint32_t)L;
    if (L < mn) mn = (uint32_t)L;
  }
  h->min_len = mn;
  h->max_len = mx;
}

// host copy of the offsets + shortest / longest sequence, on the pool's threads (1.4 ms single-threaded for 10^6 records)
struct BookTask {
  const uint64_t* offsets;
  uint64_t* dst;
  uint64_t* dst2;  // optional second copy (pinned, for the DMA engine)
  uint64_t n_seq;
  int n_tasks;
  uint32_t mn[64], mx[64];
};
static void book_task(int i, void* arg) {
  BookTask& b = *(BookTask*)arg;
  const uint64_t s0 = b.n_seq * (uint64_t)i / b.n_tasks, s1 = b.n_seq * (uint64_t)(i + 1) / b.n_tasks;
  uint32_t mn = 0x7fffffffu, mx = 0;
  for (uint64_t s = s0; s < s1; s++) {
    const uint64_t L = b.offsets[s + 1] - b.offsets[s];
    if (L > mx) mx = (uint32_t)L;
    if (L < mn) mn = (uint32_t)L;
  }
  b.mn[i] = mn;
  b.mx[i] = mx;
  if (b.dst) memcpy(b.dst + s0, b.offsets + s0, (s1 - s0 + (i == b.n_tasks - 1 ? 1 : 0)) * sizeof(uint64_t));
  if (b.dst2) memcpy(b.dst2 + s0, b.offsets + s0, (s1 - s0 + (i == b.n_tasks - 1 ? 1 : 0)) * sizeof(uint64_t));
}
static void host_bookkeeping_parallel(peng_handle_t h, const uint64_t* offsets, uint64_t n_seq, uint64_t* pinned_copy = nullptr) {
  // with a pinned copy at hand that copy is also the host's record of the offsets (peng_cuda_export_codes): one copy, not two
  h->offsets_in_pin = pinned_copy != nullptr;
  if (!pinned_copy && h->h_offsets.size() != n_seq + 1) h->h_offsets.resize(n_seq + 1);
  peng_hostpack::Pool& pool = peng_hostpack::global_pool();
  BookTask b;
  b.offsets = offsets;
  b.dst = pinned_copy ? nullptr : h->h_offsets.data();
  b.dst2 = pinned_copy;
  b.n_seq = n_seq;
  b.n_tasks = std::min(pool.size(), 64);
  pool.run(b.n_tasks, book_task, &b);
  uint32_t mn = 0x7fffffffu, mx = 0;
  for (int i = 0; i < b.n_tasks; i++) { mn = std::min(mn, b.mn[i]); mx = std::max(mx, b.mx[i]); }
  h->min_len = mn;
  h->max_len = mx;
}

// mode 0 / 1 / 2: letter codes / FASTA letters / nibbles relative to base src_base (pack_word)
static int launch_pack(peng_handle_t h, const uint8_t* d_src, int mode, uint64_t t_begin, uint64_t t_end, uint64_t src_base = 0) {
  if (t_end <= t_begin) return PENG_OK;
  const unsigned block = 256, grid = grid_for(t_end - t_begin, block);
  if (mode == 1)
    k_pack<1><<<grid, block, 0, h->stream>>>(d_src, h->d_offsets, h->n_seq, h->d_codes, h->d_valid, h->d_sstart, h->d_planes, t_begin, t_end, 0);
  else if (mode == 2)
    k_pack<2><<<grid, block, 0, h->stream>>>(d_src, h->d_offsets, h->n_seq, h->d_codes, h->d_valid, h->d_sstart, h->d_planes, t_begin, t_end, src_base);
  else
    k_pack<0><<<grid, block, 0, h->stream>>>(d_src, h->d_offsets, h->n_seq, h->d_codes, h->d_valid, h->d_sstart, h->d_planes, t_begin, t_end, 0);
  LAUNCH_CHECK(h);
  return PENG_OK;
}

static int prepare_ingest(peng_handle_t h, const uint64_t* offsets, uint64_t n_seq) {
  const uint64_t total = n_seq ? offsets[n_seq] : 0;
  h->fixed_len = 0;
  TRY(alloc_stream(h, n_seq, total));
  if (h->cap_offsets < n_seq + 1) {
    TRY(dev_alloc(h, &h->d_offsets, n_seq + 1));
    h->cap_offsets = n_seq + 1;
  }
  return PENG_OK;
}

// src = concatenated sequences on the DEVICE (letter codes or ASCII), offsets on the host
static int ingest_device_src(peng_handle_t h, const uint8_t* d_src, const uint64_t* offsets, uint64_t n_seq, bool ascii) {
  TRY(prepare_ingest(h, offsets, n_seq));
  CU(h, cudaMemcpyAsync(h->d_offsets, offsets, (n_seq + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, h->stream));
  TRY(launch_pack(h, d_src, ascii ? 1 : 0, 0, h->n_words32));
  h->planes_valid = true;
  h->offsets_in_pin = false;
  h->h_offsets.assign(offsets, offsets + n_seq + 1);  // host bookkeeping in the shadow of the device work
  seq_lengths(h, offsets, n_seq);
  return PENG_OK;
}

#define PENG_PIN_BYTES (16u << 20)
#ifndef PENG_PHASE_MB_DEFAULT
#define PENG_PHASE_MB_DEFAULT 16
#endif
static int need_pins(peng_handle_t h, size_t bytes) {
  bytes = std::min<size_t>(PENG_PIN_BYTES, std::max<size_t>(bytes, (size_t)64 << 10));
  if (h->pin_cap >= bytes) return PENG_OK;
  for (int i = 0; i < 2; i++) {
    if (h->pin[i]) { CU(h, cudaStreamSynchronize(h->copy_stream ? h->copy_stream : h->stream)); cudaFreeHost(h->pin[i]); h->pin[i] = nullptr; }
    CU(h, cudaHostAlloc((void**)&h->pin[i], bytes, cudaHostAllocDefault));
    if (!h->pin_ev[i]) CU(h, cudaEventCreateWithFlags(&h->pin_ev[i], cudaEventDisableTiming));
  }
  h->pin_cap = bytes;
  return PENG_OK;
}
static int need_copy_stream_and_pins(peng_handle_t h, bool pins) {
  if (!h->copy_stream) {
    CU(h, cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; i++) CU(h, cudaEventCreateWithFlags(&h->copy_ev[i], cudaEventDisableTiming));
  }
  if (pins) TRY(need_pins(h, PENG_PIN_BYTES));
  return PENG_OK;
}

// sequence-start bits of the sequences whose first base lies in front of word t_end (the words are cleared beforehand)
// Packed ingest, phases without letters other than ACGT: the validity words follow from the offsets (core_seq.cuh:
// implied_valid_word); the few blocks where they do not (host_pack.h: ValidExc) are listed and written over them.
__global__ void k_valid_implied(const uint64_t* __restrict__ offsets, uint64_t n_seq, uint64_t t0, uint64_t t1, uint32_t* __restrict__ valid) {
  const uint64_t t = t0 + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < t1) valid[t] = implied_valid_word(t, offsets, n_seq);
}
__global__ void k_valid_exceptions(const uint64_t* __restrict__ exc, uint32_t n, uint32_t* __restrict__ valid) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t blk = exc[2 * i], v = exc[2 * i + 1];
  valid[2 * blk] = (uint32_t)v;
  valid[2 * blk + 1] = (uint32_t)(v >> 32);
}

__global__ void k_sstart_scatter(const uint64_t* __restrict__ offsets, uint64_t n_seq, uint64_t t_end, uint32_t* __restrict__ sstart) {
  const uint64_t s = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_seq) return;
  const uint64_t q = PENG_FRONT_PAD + offsets[s] + 2 * s;
  if ((q >> 5) < t_end && offsets[s + 1] > offsets[s]) atomicOr(sstart + (q >> 5), 1u << (q & 31));
}

// src on the HOST, hybrid ingest.  One byte per base over PCIe (~46 GB/s) is the bound of the plain path below, so the
// host threads pack the FIRST part of the set into the device layout themselves (host_pack.h: 0.375 B/base cross the bus)
// while the DMA engine moves the raw bytes of the REST, which k_pack packs on the device as before.  The split
// f = Rh / (Rp + 0.625 Rh) equalises the two (Rh = host packing rate, Rp = PCIe rate in bases/s).  The 64-position block
// that holds the split is written whole by k_pack (the raw copy starts 128 bases early), the host's part of it is dropped.
// returns 1 when the set does not qualify (few threads, no AVX-512, a sequence longer than a phase): the caller runs the
// plain path.
static int ingest_host_hybrid(peng_handle_t h, const uint8_t* src, const uint64_t* offsets, uint64_t n_seq, bool ascii,
                              const uint8_t* const* seq_ptrs = nullptr) {
  using namespace peng_hostpack;
  const double t_entry = std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
  double t_marks[4] = {0, 0, 0, 0};
  const uint64_t total = n_seq ? offsets[n_seq] : 0;
  if (h->ingest_mode == 1) return 1;
  // per-sequence arrays (seq_ptrs): there is no contiguous byte range for the DMA engine to take, the host packs everything
  const bool force = h->ingest_mode == 2 || seq_ptrs != nullptr;
  Pool& pool = global_pool();
  // automatic choice only where it pays; an explicit mode or fraction is taken as given (tests, tuning)
  if (!force && h->ingest_fraction < 0 && (!have_avx512() || pool.size() < 4 || total < (1ull << 22))) return 1;
  h->last_ingest_packed = 0;
  // PENG_NIBBLE=1 (off by default): the part the host does not pack completely travels as NIBBLES (host_pack.h: nibble_pack,
  // 0.5 instead of 1 B/base over PCIe).  Measured on the B200 box (16 host threads, tools/ingest_bench.py,
  // profiles/r2_ingest_nibble_vs_raw.txt): best split 2.94 ms against 2.63 ms for raw bytes straight out of the caller's
  // buffer — the host threads are the co-bottleneck of the hybrid ingest, and every byte they touch a second time costs
  // more than the PCIe bytes it saves.  Kept for hosts with more cores per GPU than PCIe bandwidth.
  const bool nib = !ascii && !seq_ptrs && h->use_nibble;
  double f = 1.0;
  if (!force && nib) {
    // host: f / rh + (1 - f) / rn of packing, bus: (0.375 f + 0.5 (1 - f)) / rp — equal at
    //   f = (0.5 / rp - 1 / rn) / (1 / rh - 1 / rn + 0.125 / rp)       (rn = nibble rate: ~3 x the full packing rate)
    static const double rn_factor = getenv("PENG_NIBBLE_RATE_FACTOR") ? atof(getenv("PENG_NIBBLE_RATE_FACTOR")) : 3.0;
    const double rh = 0.8 * (5.7 * std::min(pool.size(), 8) + 2.2 * std::max(0, pool.size() - 8)), rn = rn_factor * rh, rp = 46.0;
    f = (0.5 / rp - 1.0 / rn) / (1.0 / rh - 1.0 / rn + 0.125 / rp);
    if (h->ingest_fraction >= 0) f = h->ingest_fraction;
    f = std::min(1.0, std::max(0.0, f));
  } else if (!force) {
    // Gbases/s: measured packing rate (tools/probe/host_pack_bench.cpp on the B200 box: 11 / 42 / 73 / 115 with 1 / 4 / 8 / 16
    // threads — the host's memory system from 8 threads on), PCIe rate; x 0.85: the packing runs beside the DMA traffic
    // (measured inside the ingest: 104 at 16 threads; optimum of tools/ingest_bench.py's sweep: 0.9)
    const double rh = 0.85 * (9.5 * std::min(pool.size(), 8) + 5.0 * std::max(0, pool.size() - 8)), rp = 46.0;
    // packed bytes per base: 0.375, or 0.25 where the validity words are implied (no letters other than ACGT: 256 probes say)
    bool clean = !ascii && !seq_ptrs && lists_exceptions() && getenv("PENG_NO_IMPLIED_VALID") == nullptr;
    for (uint64_t i = 0; clean && i < 256; i++) clean = (uint8_t)(src[(total - 1) / 256 * i] - 1u) < 4u;
    f = rh / (rp + (clean ? 0.75 : 0.625) * rh);
    if (h->ingest_fraction >= 0) f = h->ingest_fraction;
    f = std::min(1.0, std::max(0.0, f));
  } else if (h->ingest_fraction >= 0 && !seq_ptrs) {
    f = std::min(1.0, std::max(0.0, h->ingest_fraction));
  }
  uint64_t s_split = n_seq;
  if (f < 1.0) s_split = (uint64_t)(std::upper_bound(offsets, offsets + n_seq + 1, (uint64_t)(f * (double)total)) - offsets) - 1;
  if (s_split == 0) return 1;
  // 16 B of codes + 8 B of validity per block of 64 positions.  A phase fills at most PENG_PHASE_MB of its staging
  // buffer: the copy of the LAST phase cannot hide behind packing, so phases are kept shorter than the buffers allow
  static const uint64_t phase_bytes = [] {
    const char* e = getenv("PENG_PHASE_MB");
    const uint64_t mb = e ? (uint64_t)atoi(e) : PENG_PHASE_MB_DEFAULT;
    return std::min<uint64_t>(PENG_PIN_BYTES, std::max<uint64_t>(1, mb) << 20);
  }();
  const uint64_t max_blocks = phase_bytes / 24;
  // every phase must hold at least one sequence
  {
    uint64_t s = 0;
    bool first = true;
    while (s < s_split) {
      const uint64_t e = std::min(s_split, plan_phase(offsets, n_seq, s, PENG_FRONT_PAD, first, max_blocks - 4));
      if (e == s) return 1;
      s = e;
      first = false;
    }
  }
  auto mark = [&](int i) { t_marks[i] = std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count() - t_entry; };
  mark(0);
  CU(h, cudaSetDevice(h->device));
  TRY(prepare_ingest(h, offsets, n_seq));
  TRY(need_copy_stream_and_pins(h, true));
  mark(1);
  const uint64_t q_total = ((h->n_words32 * 32 + 63) / 64) * 64;
  const uint64_t n_blocks = q_total / 64;
  const uint64_t q_split = PENG_FRONT_PAD + offsets[s_split] + 2 * s_split;
  const uint64_t blk_split = s_split < n_seq ? (q_split >> 6) : n_blocks;
  const uint64_t b_lo = s_split < n_seq ? (offsets[s_split] - std::min<uint64_t>(offsets[s_split], 128)) & ~(uint64_t)1 : total;
  // implied validity (see k_valid_implied): a phase lists at most 1/16 of its blocks (+ 64), else it sends its validity words
  static const bool implied_on = getenv("PENG_NO_IMPLIED_VALID") == nullptr;
  const bool implied = implied_on && lists_exceptions();
  const size_t raw_bytes = s_split < n_seq ? (((size_t)(total - b_lo) + 16 + 255) & ~(size_t)255) : 0;
  const uint64_t exc_cap = implied ? blk_split / 16 + 64 * (blk_split / ((2u << 20) / 24) + 8) : 0;  // entries, all phases together
  if (raw_bytes + exc_cap) TRY(need_scratch(h, raw_bytes + (size_t)exc_cap * 16));
  uint64_t* const d_exc = (uint64_t*)((uint8_t*)h->d_scratch + raw_bytes);
  uint64_t exc_used = 0, h2d = 0;
  std::vector<ValidExc> exc;
  bool exc_complete = false;
  if (h->pin_offs_cap < n_seq + 1) {
    if (h->pin_offs) cudaFreeHost(h->pin_offs);
    h->pin_offs = nullptr;
    h->pin_offs_cap = 0;
    CU(h, cudaHostAlloc((void**)&h->pin_offs, (n_seq + 1 + (n_seq >> 2)) * sizeof(uint64_t), cudaHostAllocDefault));
    h->pin_offs_cap = n_seq + 1 + (n_seq >> 2);
  }
  // count prefetch: stage P of the armed count follows the data as it lands (see PrefetchRun / prefetch_range)
  PrefetchRun pr;
  TRY(prefetch_begin(h, &pr));
  CU(h, cudaEventRecord(h->copy_ev[0], h->stream));
  CU(h, cudaStreamWaitEvent(h->copy_stream, h->copy_ev[0], 0));
  mark(2);
  // ---- the raw rest: queued first, the DMA engine works on it while the host packs ----
  int buf = 0;
  if (s_split < n_seq && nib) {
    // nibbles through the two pinned staging buffers: the pool packs chunk i while the DMA engine moves chunk i - 1
    const uint64_t chunk = 2 * (uint64_t)h->pin_cap;  // bases per staging buffer
    for (uint64_t b0 = b_lo; b0 < total; b0 += chunk) {
      const uint64_t b1 = std::min(total, b0 + chunk);
      CU(h, cudaEventSynchronize(h->pin_ev[buf]));  // the previous copy out of this staging buffer is done
      nibble_pack(pool, src + b0, b1 - b0, h->pin[buf]);
      CU(h, cudaMemcpyAsync((uint8_t*)h->d_scratch + ((b0 - b_lo) >> 1), h->pin[buf], (size_t)((b1 - b0 + 1) >> 1), cudaMemcpyHostToDevice, h->copy_stream));
      h2d += (b1 - b0 + 1) >> 1;
      CU(h, cudaEventRecord(h->pin_ev[buf], h->copy_stream));
      buf ^= 1;
    }
  } else if (s_split < n_seq) {
    const uint64_t chunk = 32ull << 20;
    for (uint64_t b0 = b_lo; b0 < total; b0 += chunk) {
      const uint64_t b1 = std::min(total, b0 + chunk);
      CU(h, cudaMemcpyAsync((uint8_t*)h->d_scratch + (b0 - b_lo), src + b0, b1 - b0, cudaMemcpyHostToDevice, h->copy_stream));
      h2d += b1 - b0;
    }
  }
  // ---- the offsets: a cudaMemcpyAsync out of the caller's pageable array would block this thread for the whole copy
  // (0.38 ms per 10^6 sequences), so the pool's threads copy them into a pinned buffer — in the same pass that keeps the
  // host copy and finds the shortest / longest sequence — and the DMA engine takes them from there
  host_bookkeeping_parallel(h, offsets, n_seq, h->pin_offs);
  CU(h, cudaMemcpyAsync(h->d_offsets, h->pin_offs, (n_seq + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, h->copy_stream));
  h2d += (n_seq + 1) * sizeof(uint64_t);
  if (s_split < n_seq) {  // raw bytes and offsets have arrived: the device packs the rest in one go
    const uint8_t* d_src = (const uint8_t*)h->d_scratch - b_lo;  // indexed by absolute base offset, never read below b_lo
    CU(h, cudaEventRecord(h->copy_ev[0], h->copy_stream));
    CU(h, cudaStreamWaitEvent(h->stream, h->copy_ev[0], 0));
    if (nib) TRY(launch_pack(h, (const uint8_t*)h->d_scratch, 2, blk_split * 2, h->n_words32, b_lo));
    else TRY(launch_pack(h, d_src, ascii ? 1 : 0, blk_split * 2, h->n_words32));
    // the units behind the one that holds the first position of sequence s_split look at k_pack's words only (every
    // look-back of the k-mer start rule stops at that sequence's start): their stage P runs under the DMA of the phases
    if (pr.on) TRY(prefetch_range(h, &pr, (q_split >> 5) + 1, h->n_words32));
  }
  // ---- the packed part, phase by phase through the two pinned staging buffers ----
  bool has_carry = false;
  Partial carry{};
  static const bool dbg = getenv("PENG_INGEST_DEBUG") != nullptr;
  static const bool taper = getenv("PENG_NO_TAPER") == nullptr;
  double t_pack = 0, t_wait = 0, t_issue = 0;
  auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  const double t_start = now();
  mark(3);
  // Two phases are in flight: while the pool packs phase i + 1 into one staging buffer, this thread queues the copies and
  // kernels of phase i out of the other (the packing is the bottleneck of the ingest; nothing else should sit on its path).
  struct Slot {
    uint64_t s, e, blk0, blk1;
    bool first, last;
    int buf;
    uint64_t *codes64, *valid64;
    Phase ph;
    PhaseJob* job = nullptr;
    std::vector<ValidExc> exc;
    bool exc_complete = false;
  } slots[2];
  auto begin_phase = [&](Slot& c, uint64_t s, bool first, int buf) -> int {
    // the copy of the LAST phase cannot hide behind packing: towards the end every phase takes half of what is left
    // (down to 2 MiB of blocks), so that the tail of the DMA queue is short
    uint64_t cap = max_blocks - 4;
    if (taper) {
      const uint64_t blk_here = first ? 0 : (PENG_FRONT_PAD + offsets[s] + 2 * s) >> 6;
      const uint64_t remaining = blk_split > blk_here ? blk_split - blk_here : 0;
      if (remaining < 2 * cap) cap = std::min(cap, std::max<uint64_t>((2u << 20) / 24, remaining / 2));
    }
    uint64_t e = std::min(s_split, plan_phase(offsets, n_seq, s, PENG_FRONT_PAD, first, cap));
    if (e == s) e = std::min(s_split, plan_phase(offsets, n_seq, s, PENG_FRONT_PAD, first, max_blocks - 4));  // a long sequence
    // the last phase of a fully packed set also writes the tail pad; if that does not fit, the tail gets a phase of its own
    bool last = false;
    if (s_split == n_seq && e == n_seq) {
      const uint64_t q0 = first ? 0 : PENG_FRONT_PAD + offsets[s] + 2 * s;
      last = n_blocks - (q0 >> 6) <= max_blocks - 4;
    }
    double t0 = now();
    CU(h, cudaEventSynchronize(h->pin_ev[buf]));  // the previous copy out of this staging buffer is done
    t_wait += now() - t0;
    c.s = s;
    c.e = e;
    c.first = first;
    c.last = last;
    c.buf = buf;
    c.codes64 = (uint64_t*)h->pin[buf];
    c.valid64 = c.codes64 + 2 * max_blocks;
    c.ph = Phase{src, offsets, s, e, ascii, PENG_FRONT_PAD, first, last, q_total, c.codes64, c.valid64, seq_ptrs};
    if (implied) {
      c.ph.exc = &c.exc;
      c.ph.exc_complete = &c.exc_complete;
    }
    c.exc_complete = false;
    phase_blocks(c.ph, &c.blk0, &c.blk1);
    c.job = pack_phase_begin(pool, c.ph);
    return PENG_OK;
  };
  auto issue_phase = [&](Slot& c) -> int {
    const uint64_t blk0 = c.blk0, blk1 = std::min(c.blk1, blk_split);  // the block of the split belongs to k_pack
    const int buf = c.buf;
    std::vector<ValidExc>& exc = c.exc;
    uint32_t n_exc = 0;
    bool phase_implied = false;
    if (blk1 > blk0) {
      CU(h, cudaMemcpyAsync(h->d_codes + blk0 * 4, c.codes64, (blk1 - blk0) * 16, cudaMemcpyHostToDevice, h->copy_stream));
      h2d += (blk1 - blk0) * 16;
      if (implied && c.exc_complete) {
        while (!exc.empty() && exc.back().blk >= blk1) exc.pop_back();  // sorted by block; the block of the split is k_pack's
        phase_implied = exc.size() <= (blk1 - blk0) / 16 + 64 && exc_used + exc.size() <= exc_cap;
      }
      if (phase_implied) {  // the listed blocks travel through the staging area of the validity words that stay here
        n_exc = (uint32_t)exc.size();
        if (n_exc) {
          memcpy(c.valid64, exc.data(), (size_t)n_exc * 16);
          CU(h, cudaMemcpyAsync(d_exc + 2 * exc_used, c.valid64, (size_t)n_exc * 16, cudaMemcpyHostToDevice, h->copy_stream));
          h2d += (uint64_t)n_exc * 16;
        }
      } else {
        CU(h, cudaMemcpyAsync(h->d_valid + blk0 * 2, c.valid64, (blk1 - blk0) * 8, cudaMemcpyHostToDevice, h->copy_stream));
        h2d += (blk1 - blk0) * 8;
      }
    }
    CU(h, cudaEventRecord(h->pin_ev[buf], h->copy_stream));
    if (phase_implied) {  // (the offsets were queued in front of every phase)
      CU(h, cudaStreamWaitEvent(h->stream, h->pin_ev[buf], 0));
      k_valid_implied<<<grid_for((blk1 - blk0) * 2, 256), 256, 0, h->stream>>>(h->d_offsets, n_seq, blk0 * 2, blk1 * 2, h->d_valid);
      LAUNCH_CHECK(h);
      if (n_exc) {
        k_valid_exceptions<<<grid_for(n_exc, 256), 256, 0, h->stream>>>(d_exc + 2 * exc_used, n_exc, h->d_valid);
        LAUNCH_CHECK(h);
      }
      exc_used += n_exc;
    }
    if (pr.on && blk1 > blk0) {  // the blocks of this phase once they have landed: bit planes, then stage P up to two units short of the frontier
      CU(h, cudaStreamWaitEvent(h->stream, h->pin_ev[buf], 0));
      TRY(make_planes(h, pr.planes_done, blk1 * 2));
      pr.planes_done = blk1 * 2;
      if (blk1 * 2 >= 2) {
        TRY(prefetch_range(h, &pr, pr.done_u, blk1 * 2 - 2));
        pr.done_u = std::max(pr.done_u, blk1 * 2 - 2);
      }
    }
    return PENG_OK;
  };
  {
    int rc = begin_phase(slots[0], 0, true, buf);
    for (int cur = 0; rc == PENG_OK;) {
      Slot& c = slots[cur];
      double t0 = now();
      pack_phase_end(pool, c.job, &has_carry, &carry);
      c.job = nullptr;
      t_pack += now() - t0;
      const bool more = !(c.last || (c.e >= s_split && s_split < n_seq));
      // the next phase starts before this one's copies are queued — unless its staging buffer is still being read by the
      // DMA engine: then the bus is the bottleneck, and this phase's copy goes out first
      bool next_free = false;
      if (more) {
        const cudaError_t q = cudaEventQuery(h->pin_ev[c.buf ^ 1]);
        next_free = q == cudaSuccess;
        if (q == cudaErrorNotReady) (void)cudaGetLastError();  // "not ready" is an answer, not an error: do not leave it for the next launch check
      }
      if (next_free) rc = begin_phase(slots[cur ^ 1], c.e, false, c.buf ^ 1);
      t0 = now();
      if (rc == PENG_OK) rc = issue_phase(c);
      t_issue += now() - t0;
      if (more && !next_free && rc == PENG_OK) rc = begin_phase(slots[cur ^ 1], c.e, false, c.buf ^ 1);
      if (!more) break;
      cur ^= 1;
    }
    if (rc != PENG_OK) {  // a job that was started must be joined before its Phase goes away
      for (Slot& c : slots)
        if (c.job) {
          bool hc = false;
          Partial pc{};
          pack_phase_end(pool, c.job, &hc, &pc);
        }
      return rc;
    }
  }
  // sequence starts of the packed part (k_pack writes those of its own words)
  const uint64_t t_split = std::min<uint64_t>(blk_split * 2, h->n_words32);
  CU(h, cudaMemsetAsync(h->d_sstart, 0, t_split * 4, h->stream));
  CU(h, cudaEventRecord(h->copy_ev[1], h->copy_stream));
  CU(h, cudaStreamWaitEvent(h->stream, h->copy_ev[1], 0));  // the offsets and every staged block have arrived
  k_sstart_scatter<<<grid_for(std::max<uint64_t>(n_seq, 1), 256), 256, 0, h->stream>>>(h->d_offsets, n_seq, t_split, h->d_sstart);
  LAUNCH_CHECK(h);
  // k_pack wrote the planes of its words; those of the host-packed words follow here (what the prefetch has not made yet)
  TRY(make_planes(h, pr.planes_done, t_split));
  h->planes_valid = true;
  if (pr.on) {  // what is left of stage P: the front's last units and the units around the split
    TRY(prefetch_range(h, &pr, pr.done_u, s_split < n_seq ? std::min<uint64_t>((q_split >> 5) + 1, h->n_words32) : h->n_words32));
    h->pre.done = true;
  }
  const double t_loop = now();
  const double t_book = now();
  CU(h, cudaStreamSynchronize(h->stream));
  if (dbg)
    fprintf(stderr, "[ingest] before the phases: plan %.2f, prepare %.2f, offsets copy issued %.2f, raw rest issued %.2f ms\n",
            1e3 * t_marks[0], 1e3 * (t_marks[1] - t_marks[0]), 1e3 * (t_marks[2] - t_marks[1]), 1e3 * (t_marks[3] - t_marks[2]));
  if (dbg)
    fprintf(stderr, "[ingest] threads %d split %.2f: phases %.2f ms (pack %.2f, wait %.2f, issue %.2f), bookkeeping %.2f, final sync %.2f\n",
            pool.size(), (double)offsets[s_split] / (double)std::max<uint64_t>(total, 1), 1e3 * (t_loop - t_start), 1e3 * t_pack,
            1e3 * t_wait, 1e3 * t_issue, 1e3 * (t_book - t_loop), 1e3 * (now() - t_book));
  h->last_ingest_packed = s_split;
  h->last_ingest_h2d = h2d;
  return PENG_OK;
}

// src on the HOST: the copy is cut into chunks on a second stream, the packing kernel follows chunk by chunk on the
// handle's stream (a word is packed once every base of its 32 positions has arrived), so that packing, and the host's
// own bookkeeping, hide behind the PCIe transfer.
static int ingest_host_src(peng_handle_t h, const uint8_t* src, const uint64_t* offsets, uint64_t n_seq, bool ascii) {
  h->last_ingest_packed = 0;
  h->last_ingest_h2d = (n_seq ? offsets[n_seq] : 0) + (n_seq + 1) * sizeof(uint64_t);  // the plain path; the hybrid one counts its own
  {
    const int rc = ingest_host_hybrid(h, src, offsets, n_seq, ascii);
    if (rc != 1) return rc;  // done (or failed); 1 = not applicable, plain path
  }
  CU(h, cudaSetDevice(h->device));
  const uint64_t total = n_seq ? offsets[n_seq] : 0;
  TRY(need_scratch(h, (size_t)total + 16));
  TRY(prepare_ingest(h, offsets, n_seq));
  if (!h->copy_stream) {
    CU(h, cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; i++) CU(h, cudaEventCreateWithFlags(&h->copy_ev[i], cudaEventDisableTiming));
  }
  uint8_t* d_src = (uint8_t*)h->d_scratch;
  // the copy stream must not run ahead of earlier work on the handle's stream that still reads the scratch buffer
  CU(h, cudaEventRecord(h->copy_ev[0], h->stream));
  CU(h, cudaStreamWaitEvent(h->copy_stream, h->copy_ev[0], 0));
  const uint64_t chunk = 32ull << 20;
  uint64_t t_done = 0, s_hint = 0;
  int flip = 0;
  for (uint64_t b0 = 0; b0 < total || b0 == 0; b0 += chunk) {
    const uint64_t b1 = std::min(total, b0 + chunk);
    if (b1 > b0) CU(h, cudaMemcpyAsync(d_src + b0, src + b0, b1 - b0, cudaMemcpyHostToDevice, h->copy_stream));
    // the offsets follow the first chunk: out of the caller's pageable array the call blocks this thread for the whole
    // staging (0.38 ms per 10^6 sequences) — with a chunk already queued the DMA engine is busy meanwhile
    if (b0 == 0) CU(h, cudaMemcpyAsync(h->d_offsets, offsets, (n_seq + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, h->copy_stream));
    CU(h, cudaEventRecord(h->copy_ev[flip], h->copy_stream));
    CU(h, cudaStreamWaitEvent(h->stream, h->copy_ev[flip], 0));
    flip ^= 1;
    uint64_t t_end = h->n_words32;
    if (b1 < total) {  // every position in front of base b1 is complete: q1 = FRONT + b1 + 2 * (sequence holding b1)
      const uint64_t* it = std::upper_bound(offsets + s_hint, offsets + n_seq + 1, b1);
      s_hint = (uint64_t)(it - offsets) - 1;
      t_end = (PENG_FRONT_PAD + b1 + 2 * s_hint) / 32;
    }
    TRY(launch_pack(h, d_src, ascii ? 1 : 0, t_done, t_end));
    t_done = std::max(t_done, t_end);
    if (total == 0) break;
  }
  h->planes_valid = true;
  h->offsets_in_pin = false;
  h->h_offsets.assign(offsets, offsets + n_seq + 1);  // host bookkeeping while the DMA engine works
  seq_lengths(h, offsets, n_seq);
  CU(h, cudaStreamSynchronize(h->stream));  // the caller may free / reuse src
  return PENG_OK;
}

extern "C" int peng_cuda_ingest_codes(peng_handle_t h, const uint8_t* codes, const uint64_t* offsets, uint64_t n_seq) {
  if (!h || !offsets || (!codes && n_seq && offsets[n_seq])) return fail(h, PENG_E_INVALID, "ingest_codes: NULL argument");
  return ingest_host_src(h, codes, offsets, n_seq, false);
}

// Sequences as the reference holds them: one array per Sequence object (src/shared/Sequence.h:18-19).  The host threads pack
// them straight from where they lie (pageable memory: nothing is pinned or concatenated); without AVX-512 / threads the
// arrays are gathered into one buffer for the plain path.
struct GatherTask {
  const uint8_t* const* seqs;
  const uint64_t* offs;
  uint8_t* dst;
  uint64_t n_seq;
  int n_tasks;
};
static void gather_task(int i, void* arg) {
  GatherTask& g = *(GatherTask*)arg;
  const uint64_t s0 = g.n_seq * (uint64_t)i / g.n_tasks, s1 = g.n_seq * (uint64_t)(i + 1) / g.n_tasks;
  for (uint64_t s = s0; s < s1; s++) memcpy(g.dst + g.offs[s], g.seqs[s], g.offs[s + 1] - g.offs[s]);
}
struct LengthScan {
  const uint32_t* lengths;
  uint64_t* offs;
  uint64_t n_seq;
  int n_tasks;
  std::vector<uint64_t> sums;  // [i + 1] = sum of chunk i, then (after the serial pass) [i] = total in front of chunk i
};
static void length_sum_task(int i, void* arg) {
  LengthScan& sc = *(LengthScan*)arg;
  const uint64_t a = sc.n_seq * (uint64_t)i / (uint64_t)sc.n_tasks, b = sc.n_seq * (uint64_t)(i + 1) / (uint64_t)sc.n_tasks;
  uint64_t t = 0;
  for (uint64_t s = a; s < b; s++) t += sc.lengths[s];
  sc.sums[(size_t)i + 1] = t;
}
static void length_scan_task(int i, void* arg) {
  LengthScan& sc = *(LengthScan*)arg;
  const uint64_t a = sc.n_seq * (uint64_t)i / (uint64_t)sc.n_tasks, b = sc.n_seq * (uint64_t)(i + 1) / (uint64_t)sc.n_tasks;
  uint64_t t = sc.sums[(size_t)i];
  for (uint64_t s = a; s < b; s++) {
    sc.offs[s] = t;
    t += sc.lengths[s];
  }
}

// h->seq_offs = prefix sums of the lengths: two passes on the pool (chunk sums, then the chunks' offsets), into a buffer the
// handle keeps (a fresh 8 MB vector per call costs more in page faults than the sums)
static int offsets_from_lengths(peng_handle_t h, const uint32_t* lengths, uint64_t n_seq) {
  peng_hostpack::Pool& pool = peng_hostpack::global_pool();
  std::vector<uint64_t>& offs = h->seq_offs;
  if (offs.size() < n_seq + 1) offs.resize(n_seq + 1 + (n_seq >> 3));
  LengthScan sc{lengths, offs.data(), n_seq, std::max(1, std::min<int>(pool.size() * 2, (int)(n_seq / 65536 + 1))), {}};
  sc.sums.assign((size_t)sc.n_tasks + 1, 0);
  pool.run(sc.n_tasks, length_sum_task, &sc);
  for (int i = 0; i < sc.n_tasks; i++) sc.sums[(size_t)i + 1] += sc.sums[(size_t)i];  // sums[i]: total in front of chunk i
  pool.run(sc.n_tasks, length_scan_task, &sc);
  offs[n_seq] = sc.sums[(size_t)sc.n_tasks];
  return PENG_OK;
}

extern "C" int peng_cuda_ingest_sequences(peng_handle_t h, const uint8_t* const* seqs, const uint32_t* lengths, uint64_t n_seq) {
  if (!h || (n_seq && (!seqs || !lengths))) return fail(h, PENG_E_INVALID, "ingest_sequences: NULL argument");
  peng_hostpack::Pool& pool = peng_hostpack::global_pool();
  std::vector<uint64_t>& offs = h->seq_offs;
  TRY(offsets_from_lengths(h, lengths, n_seq));
  h->last_ingest_packed = 0;
  const int rc = ingest_host_hybrid(h, nullptr, offs.data(), n_seq, false, seqs);
  if (rc != 1) return rc;
  std::vector<uint8_t> buf(offs[n_seq] + 1);
  GatherTask g{seqs, offs.data(), buf.data(), n_seq, std::max(1, std::min(pool.size() * 4, 256))};
  pool.run(g.n_tasks, gather_task, &g);
  return ingest_host_src(h, buf.data(), offs.data(), n_seq, false);
}

extern "C" int peng_cuda_set_ingest_mode(peng_handle_t h, int mode, double host_fraction) {
  if (!h || mode < 0 || mode > 2 || host_fraction > 1.0) return fail(h, PENG_E_INVALID, "set_ingest_mode: mode 0..2, fraction <= 1 (< 0: automatic)");
  h->ingest_mode = mode;
  h->ingest_fraction = host_fraction;
  return PENG_OK;
}
extern "C" uint64_t peng_cuda_last_ingest_host_packed(peng_handle_t h) { return h ? h->last_ingest_packed : 0; }
extern "C" uint64_t peng_cuda_last_ingest_h2d_bytes(peng_handle_t h) { return h ? h->last_ingest_h2d : 0; }

extern "C" int peng_cuda_ingest_codes_device(peng_handle_t h, const void* d_codes, const uint64_t* offsets,
                                             uint64_t n_seq) {
  if (!h || !offsets) return fail(h, PENG_E_INVALID, "ingest_codes_device: NULL argument");
  CU(h, cudaSetDevice(h->device));
  return ingest_device_src(h, (const uint8_t*)d_codes, offsets, n_seq, false);
}

// Streaming FASTA ingest: the file is mapped, one host pass splits it into records with the reference's line rules
// (host_ingest.h: scan_fasta) and copies the sequence bytes into two pinned staging buffers; every full buffer goes to
// the device on the copy stream while the scan fills the other one.  Letters are translated and packed on the device
// once the record table is known (k_pack).
struct StagingSink {
  peng_handle_t h;
  uint8_t* d_src;
  int cur = 0;
  size_t fill = 0;
  uint64_t sent = 0;  // bytes handed to the DMA engine so far
  int rc = PENG_OK;
  std::vector<uint64_t> offsets{0};
  std::vector<std::pair<const char*, size_t>> segments;  // lines of the record being read (for the warning below)
  const char* last_header = nullptr;
  size_t last_header_len = 0;
  std::vector<std::pair<const char*, size_t>> last_segments;
  // sharded ingest (peng_cuda_ingest_fasta_shard): only the records whose first sequence line starts in [keep_lo, keep_hi)
  // of the file are staged; every shard scans the whole file with the whole rule set, so the cut is the same for all
  const char* text = nullptr;
  size_t keep_lo = 0, keep_hi = (size_t)-1;
  bool cur_keep = true, last_mine = false;
  int flush() {
    if (fill == 0 || rc != PENG_OK) return rc;
    peng_handle_t hh = h;
    auto fail_cuda = [&](cudaError_t e) { rc = fail(hh, PENG_E_CUDA, "ingest_fasta: %s", cudaGetErrorString(e)); return rc; };
    cudaError_t e = cudaMemcpyAsync(d_src + sent, h->pin[cur], fill, cudaMemcpyHostToDevice, h->copy_stream);
    if (e != cudaSuccess) return fail_cuda(e);
    if ((e = cudaEventRecord(h->pin_ev[cur], h->copy_stream)) != cudaSuccess) return fail_cuda(e);
    sent += fill;
    fill = 0;
    cur ^= 1;
    if ((e = cudaEventSynchronize(h->pin_ev[cur])) != cudaSuccess) return fail_cuda(e);  // its previous copy has left
    return rc;
  }
  void letters(const char* p, size_t n) {
    if (segments.empty()) cur_keep = (size_t)(p - text) >= keep_lo && (size_t)(p - text) < keep_hi;
    segments.emplace_back(p, n);
    if (!cur_keep) return;
    while (n && rc == PENG_OK) {
      const size_t m = std::min<size_t>(n, h->pin_cap - fill);
      memcpy(h->pin[cur] + fill, p, m);
      fill += m;
      p += m;
      n -= m;
      if (fill == h->pin_cap) flush();
    }
  }
  void end_record(const char* header, size_t header_len) {
    last_mine = cur_keep;
    if (cur_keep) {
      offsets.push_back(sent + fill);
      last_header = header;
      last_header_len = header_len;
      last_segments.swap(segments);
    }
    segments.clear();
  }
  bool last_kept = false;
  void end_of_file(bool last_record_kept) { last_kept = last_record_kept; }
};

// Parallel FASTA ingest (the default; the streaming path above is the fallback).  The file is cut at header lines into
// chunks, the pool's threads scan them with the same line rules (scan_fasta), and instead of copying letters a RecordSink
// notes WHERE every record's sequence lies: single-line records (and those are the long ones of a ChIP-seq peak file) are
// read in place from the mapping, wrapped records are joined in a side arena.  The pointer table then goes through the
// packed ingest (ingest_host_hybrid with per-sequence pointers, ASCII): the host threads write the device layout straight
// from the page cache.
struct PtrGather {
  peng_host::ChunkedScan* f;
  std::vector<uint64_t> first;  // index of chunk i's first record
  const uint8_t** ptrs;
  uint32_t* lens;
};
static void ptr_gather_task(int i, void* arg) {
  PtrGather& g = *(PtrGather*)arg;
  const peng_host::RecordSink& k = g.f->sinks[(size_t)i];
  if (k.ptrs.empty()) return;
  memcpy(g.ptrs + g.first[(size_t)i], k.ptrs.data(), k.ptrs.size() * sizeof(const uint8_t*));
  memcpy(g.lens + g.first[(size_t)i], k.lens.data(), k.lens.size() * sizeof(uint32_t));
}

// returns 1 when the set does not qualify (the caller takes the streaming path), else a PENG_ code
static int ingest_fasta_parallel(peng_handle_t h, const char* path, int part, int n_parts) {
  using namespace peng_hostpack;
  if (getenv("PENG_FASTA_SERIAL") || !lists_exceptions()) return 1;  // (the masked loads that make reading in place safe are the AVX-512 path's)
  // Shards of a multi-GPU run come from one host thread per device at the same time.  The pool lets their jobs take turns,
  // but that combination has only run on the CPU so far: shards stay on the streaming reader unless
  // asked for.
  if (n_parts > 1 && !getenv("PENG_FASTA_SHARD_PARALLEL")) return 1;
  Pool& pool = global_pool();
  peng_host::MappedFile file;
  if (file.open_file(path, false) != 0) return fail(h, PENG_E_IO, "Error: Cannot open FASTA file: %s", path);
  if (file.size < (2u << 20)) return 1;  // nothing to gain below a few chunks: first-call costs dominate small files
  peng_host::ChunkedScan f;
  const int n = (int)std::max<size_t>(1, std::min<size_t>((size_t)pool.size() * 4, file.size >> 18));
  f.prepare(file.data, file.size, path, n, part, n_parts);
  pool.run(n, [](int i, void* arg) { ((peng_host::ChunkedScan*)arg)->scan_chunk(i); }, &f);
  // warnings and the first error in file order, as one pass over the file would have printed them
  for (int i = 0; i < n; i++) {
    if (!f.warns[(size_t)i].empty()) fputs(f.warns[(size_t)i].c_str(), stderr);
    if (f.rcs[(size_t)i] != 0) return fail(h, PENG_E_FORMAT, "%s", f.msgs[(size_t)i].c_str());
  }
  PtrGather g;
  g.f = &f;
  uint64_t ns = 0;
  int last_chunk = -1;  // the chunk that holds the end of the file
  for (int i = 0; i < n; i++) {
    g.first.push_back(ns);
    ns += f.sinks[(size_t)i].ptrs.size();
    if (f.sinks[(size_t)i].too_long) return 1;
    if (f.cut[(size_t)i + 1] > f.cut[(size_t)i]) last_chunk = i;
  }
  if (ns == 0) return 1;
  std::vector<const uint8_t*> ptrs(ns);
  std::vector<uint32_t> lens(ns);
  g.ptrs = ptrs.data();
  g.lens = lens.data();
  pool.run(n, ptr_gather_task, &g);
  // SequenceSet::readFASTA reports undefined letters for the LAST record only (src/shared/SequenceSet.cpp:395-405)
  const peng_host::RecordSink& lastk = f.sinks[(size_t)last_chunk];
  if (lastk.last_kept && lastk.last_mine) {
    const std::string name = lastk.last_header_len ? std::string(lastk.last_header, lastk.last_header_len) : std::to_string(ns);
    for (const auto& seg : lastk.last_segments)
      for (size_t i = 0; i < seg.second; i++) {
        const char c = seg.first[i], up = (char)(c & 0xDF);
        if (up != 'A' && up != 'C' && up != 'G' && up != 'T')
          fprintf(stderr, "Warning: The FASTA file contains an undefined base: %c at sequence %s\n", c, name.c_str());
      }
  }
  std::vector<uint64_t>& offs = h->seq_offs;
  TRY(offsets_from_lengths(h, lens.data(), ns));
  return ingest_host_hybrid(h, nullptr, offs.data(), ns, true, ptrs.data());
}

static int ingest_fasta_impl(peng_handle_t h, const char* path, int part, int n_parts, uint64_t* n_seq, uint64_t* total_bases,
                             uint32_t* min_len, uint32_t* max_len);
extern "C" int peng_cuda_ingest_fasta(peng_handle_t h, const char* path, uint64_t* n_seq, uint64_t* total_bases,
                                      uint32_t* min_len, uint32_t* max_len) {
  return ingest_fasta_impl(h, path, 0, 1, n_seq, total_bases, min_len, max_len);
}
// Shard `part` of `n_parts` of the file for a multi-GPU run: contiguous runs of records, cut by position in the file (i.e.
// balanced by bases), every record in exactly one shard.  A shard may be empty (n_seq = 0).
extern "C" int peng_cuda_ingest_fasta_shard(peng_handle_t h, const char* path, int part, int n_parts, uint64_t* n_seq,
                                            uint64_t* total_bases, uint32_t* min_len, uint32_t* max_len) {
  if (n_parts < 1 || part < 0 || part >= n_parts) return fail(h, PENG_E_INVALID, "ingest_fasta_shard: bad shard %d/%d", part, n_parts);
  return ingest_fasta_impl(h, path, part, n_parts, n_seq, total_bases, min_len, max_len);
}
static int ingest_fasta_impl(peng_handle_t h, const char* path, int part, int n_parts, uint64_t* n_seq, uint64_t* total_bases,
                             uint32_t* min_len, uint32_t* max_len) {
  if (!h || !path) return fail(h, PENG_E_INVALID, "ingest_fasta: NULL argument");
  CU(h, cudaSetDevice(h->device));
  {
    const int rc = ingest_fasta_parallel(h, path, part, n_parts);
    if (rc != 1) {  // done (or failed); 1 = not applicable, streaming path
      if (rc != PENG_OK) return rc;
      if (n_seq) *n_seq = h->n_seq;
      if (total_bases) *total_bases = h->total_bases;
      if (min_len) *min_len = h->min_len;
      if (max_len) *max_len = h->max_len;
      return PENG_OK;
    }
  }
  h->last_ingest_packed = 0;
  peng_host::MappedFile file;
  if (file.open_file(path) != 0) return fail(h, PENG_E_IO, "Error: Cannot open FASTA file: %s", path);
  TRY(need_scratch(h, file.size + 16));  // the sequence bytes are a subset of the file's
  if (!h->copy_stream) {
    CU(h, cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; i++) CU(h, cudaEventCreateWithFlags(&h->copy_ev[i], cudaEventDisableTiming));
  }
  TRY(need_pins(h, file.size + 1));
  // the copy stream must not run ahead of earlier work on the handle's stream that still reads the scratch buffer
  CU(h, cudaEventRecord(h->copy_ev[0], h->stream));
  CU(h, cudaStreamWaitEvent(h->copy_stream, h->copy_ev[0], 0));
  StagingSink sink;
  sink.h = h;
  sink.d_src = (uint8_t*)h->d_scratch;
  sink.text = file.data;
  if (n_parts > 1) {
    sink.keep_lo = (size_t)((unsigned __int128)file.size * (unsigned)part / (unsigned)n_parts);
    sink.keep_hi = part == n_parts - 1 ? (size_t)-1 : (size_t)((unsigned __int128)file.size * (unsigned)(part + 1) / (unsigned)n_parts);
  }
  std::string msg;
  const int rc = peng_host::scan_fasta(file.data, file.size, path, sink, &msg);
  if (rc != 0) {
    cudaStreamSynchronize(h->copy_stream);
    return fail(h, PENG_E_FORMAT, "%s", msg.c_str());
  }
  TRY(sink.flush());
  TRY(sink.rc);
  const uint64_t ns = sink.offsets.size() - 1;
  // SequenceSet::readFASTA reports undefined letters for the LAST record only (src/shared/SequenceSet.cpp:395-405)
  if (ns && sink.last_kept && sink.last_mine) {
    const std::string name = sink.last_header_len ? std::string(sink.last_header, sink.last_header_len) : std::to_string(ns);
    for (const auto& seg : sink.last_segments)
      for (size_t i = 0; i < seg.second; i++) {
        const char c = seg.first[i], up = (char)(c & 0xDF);
        if (up != 'A' && up != 'C' && up != 'G' && up != 'T')
          fprintf(stderr, "Warning: The FASTA file contains an undefined base: %c at sequence %s\n", c, name.c_str());
      }
  }
  TRY(prepare_ingest(h, sink.offsets.data(), ns));
  CU(h, cudaMemcpyAsync(h->d_offsets, sink.offsets.data(), (ns + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, h->copy_stream));
  CU(h, cudaEventRecord(h->copy_ev[1], h->copy_stream));
  CU(h, cudaStreamWaitEvent(h->stream, h->copy_ev[1], 0));
  TRY(launch_pack(h, sink.d_src, 1, 0, h->n_words32));
  h->planes_valid = true;
  seq_lengths(h, sink.offsets.data(), ns);
  h->offsets_in_pin = false;
  h->h_offsets.swap(sink.offsets);
  CU(h, cudaStreamSynchronize(h->stream));
  if (n_seq) *n_seq = h->n_seq;
  if (total_bases) *total_bases = h->total_bases;
  if (min_len) *min_len = h->min_len;
  if (max_len) *max_len = h->max_len;
  return PENG_OK;
}

extern "C" int peng_cuda_ingest_synthetic(peng_handle_t h, uint64_t n_seq, uint32_t seq_len, uint64_t seed,
                                          uint64_t first_seq, const uint8_t* motif, uint32_t motif_len,
                                          double plant_fraction, double n_fraction) {
  if (!h || seq_len == 0 || motif_len > 32) return fail(h, PENG_E_INVALID, "ingest_synthetic: bad argument");
  CU(h, cudaSetDevice(h->device));
  h->fixed_len = seq_len;
  h->offsets_in_pin = false;
  h->h_offsets.clear();
  dev_free(&h->d_offsets);
  h->cap_offsets = 0;
  h->min_len = h->max_len = seq_len;
  TRY(alloc_stream(h, n_seq, n_seq * (uint64_t)seq_len));
  SynthParams P;
  P.n_seq = n_seq;
  P.first_seq = first_seq;
  P.seed = seed;
  P.seq_len = seq_len;
  P.motif_len = motif ? motif_len : 0;
  P.plant_threshold = (uint32_t)(std::min(1.0, std::max(0.0, plant_fraction)) * 16777216.0);
  P.n_threshold = (uint32_t)(std::min(1.0, std::max(0.0, n_fraction)) * 16777216.0);
  P.motif_bits = 0;
  P.degenerate_pos = 0xFFFFFFFFu;
  for (uint32_t i = 0; i < P.motif_len; i++) {
    uint32_t c = motif[i];
    if (c > 3) { P.degenerate_pos = i; c = 1; }  // code 4 = 'S' (C or G)
    P.motif_bits |= (uint64_t)c << (2 * i);
  }
  const unsigned block = 256;
  k_synth<<<grid_for(h->n_words32, block), block, 0, h->stream>>>(P, h->d_codes, h->d_valid, h->d_sstart, h->d_planes, h->n_words32);
  LAUNCH_CHECK(h);
  h->planes_valid = true;
  return PENG_OK;
}

extern "C" int peng_cuda_sequences_info(peng_handle_t h, uint64_t* n_seq, uint64_t* total_bases, uint32_t* min_len,
                                        uint32_t* max_len) {
  if (!h || !h->d_codes) return fail(h, PENG_E_INVALID, "no sequences ingested");
  if (n_seq) *n_seq = h->n_seq;
  if (total_bases) *total_bases = h->total_bases;
  if (min_len) *min_len = h->min_len;
  if (max_len) *max_len = h->max_len;
  return PENG_OK;
}

extern "C" int peng_cuda_export_codes(peng_handle_t h, uint8_t* codes, uint64_t* offsets) {
  if (!h || !h->d_codes) return fail(h, PENG_E_INVALID, "no sequences ingested");
  CU(h, cudaSetDevice(h->device));
  if (offsets) {
    if (h->fixed_len) for (uint64_t s = 0; s <= h->n_seq; s++) offsets[s] = s * h->fixed_len;
    else if (h->offsets_in_pin) std::copy(h->pin_offs, h->pin_offs + h->n_seq + 1, offsets);
    else std::copy(h->h_offsets.begin(), h->h_offsets.end(), offsets);
  }
  if (codes && h->total_bases) {
    TRY(need_scratch(h, h->total_bases));
    const uint64_t n = h->stream_end - PENG_FRONT_PAD;
    k_export<<<grid_for(n, 256), 256, 0, h->stream>>>(h->d_codes, h->d_valid, h->d_offsets, h->n_seq, h->fixed_len,
                                                      h->stream_end, (uint8_t*)h->d_scratch);
    LAUNCH_CHECK(h);
    CU(h, cudaMemcpyAsync(codes, h->d_scratch, h->total_bases, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
  }
  return PENG_OK;
}

// =====================================================================================================
// C ABI: counting
// =====================================================================================================
// bit planes of the codes (core_seq.cuh: conflict_candidates_planes), once per ingest, by the first count that wants them
__global__ void __launch_bounds__(256) k_make_planes(const uint2* __restrict__ codes, uint2* __restrict__ planes, uint64_t begin, uint64_t end) {
  const uint64_t u = begin + (uint64_t)blockIdx.x * 256 + threadIdx.x;
  if (u >= end) return;
  const uint2 c = __ldg(codes + u);
  uint2 o;
  planes_of_unit(c.x, c.y, &o.x, &o.y);
  planes[u] = o;
}
// the planes of the words no packing kernel produced (the part of a hybrid ingest the host threads packed)
static int make_planes(peng_handle_t h, uint64_t begin, uint64_t end) {
  if (end <= begin) return PENG_OK;
  k_make_planes<<<grid_for(end - begin, 256), 256, 0, h->stream>>>(reinterpret_cast<const uint2*>(h->d_codes), h->d_planes, begin, end);
  LAUNCH_CHECK(h);
  return PENG_OK;
}
static int ensure_planes(peng_handle_t h) {
  if (h->planes_valid) return PENG_OK;
  TRY(make_planes(h, 0, std::min<uint64_t>(h->planes_from, h->n_words32)));
  h->planes_valid = true;
  return PENG_OK;
}

// Stage-P launch geometry: threads per CTA, queue entries per class, CTAs per SM.
struct PartArgs {
  uint64_t ub, ue;
  unsigned grid;
  uint32_t region_cap;
  void* buckets;
  uint32_t* fill;
  uint32_t carry_fill = 0;
};
template <int W, bool BOTH, int NT, uint32_t CAP, int PER_SM, bool LINK, bool PL = true>
static int launch_part_cfg(peng_handle_t h, const PartArgs& a) {
  if (PL) TRY(ensure_planes(h));
  typedef PartGeom<W> G;
  const size_t smem_p = (size_t)(G::NCLASS + 1) * CAP * sizeof(typename G::entry_t);
  auto kern = k_count_part<W, BOTH, NT, CAP, PER_SM, LINK, PL>;
  CU(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_p));
  kern<<<a.grid, NT, smem_p, h->stream>>>(h->d_codes, h->d_planes, h->d_valid, h->d_kstart, h->n_words32, a.ub, a.ue,
                                          (typename G::entry_t*)a.buckets, a.region_cap, a.fill, h->d_hist32,
                                          (unsigned long long*)(h->d_counts + h->n_patterns), h->d_slow_list, h->d_slow_count, 0u,
                                          a.carry_fill);
  LAUNCH_CHECK(h);
  return PENG_OK;
}
// PENG_PART_CFG: geometry 0..2, +4 = without the scan/push interleaving dependencies
static int part_cfg_per_sm(int cfg) { return (cfg & 3) == 1 ? 4 : (cfg & 3) == 2 ? 6 : 2; }
template <int W, bool BOTH>
static int launch_part(peng_handle_t h, int cfg, const PartArgs& a) {
  switch (cfg) {
    case 1: return launch_part_cfg<W, BOTH, 256, 512, 4, true>(h, a);
    case 2: return launch_part_cfg<W, BOTH, 128, 384, 6, true>(h, a);
    case 4: return launch_part_cfg<W, BOTH, 512, 1024, 2, false>(h, a);
    case 5: return launch_part_cfg<W, BOTH, 256, 512, 4, false>(h, a);
    case 6: return launch_part_cfg<W, BOTH, 128, 384, 6, false>(h, a);
    case 8: return launch_part_cfg<W, BOTH, 512, 1024, 2, true, false>(h, a);  // the 2-bit conflict scan (A/B)
    default: return launch_part_cfg<W, BOTH, 512, 1024, 2, true>(h, a);
  }
}

// The exact resolution of the slow units (few threads, latency-bound) runs on a side stream beside the next stage, which
// only touches the table through atomics; `ev_join` is awaited before anything that does not.
template <int W>
static int launch_slow_beside(peng_handle_t h, bool both, bool mark) {
  CU(h, cudaEventRecord(h->ev_fork, h->stream));
  CU(h, cudaStreamWaitEvent(h->side_stream, h->ev_fork, 0));
  if (both) k_count_slow<W, true><<<h->sm_count * 4, 128, 0, h->side_stream>>>(h->d_codes, h->d_kstart, h->d_slow_list, h->d_slow_count, h->d_hist32);
  else k_count_slow<W, false><<<h->sm_count * 4, 128, 0, h->side_stream>>>(h->d_codes, h->d_kstart, h->d_slow_list, h->d_slow_count, h->d_hist32);
  LAUNCH_CHECK(h);
  if (mark && h->timing) CU(h, cudaEventRecord(h->ev[3], h->side_stream));
  CU(h, cudaEventRecord(h->ev_join, h->side_stream));
  return PENG_OK;
}

// Partition path driver: stage P (+ fused k-mer start marking), slow units, stage H; chunks of <= 2^30 positions so
// that a uint32 bin cannot overflow between folds.
struct PartPlan {
  unsigned grid_p;
  uint32_t region_cap;
  size_t smem_h;
};
static int part_reserve(peng_handle_t h, uint64_t chunk_units, PartPlan* pl) {
  const unsigned grid_p = (unsigned)h->sm_count * part_cfg_per_sm(h->part_cfg);
  // private region of (class, stage-P CTA): 2x the expected share (composition skew of real genomes: a class of an AT-rich set takes ~1.6x its uniform share), at least 1024 keys, multiple of 128
  uint64_t region_cap = (chunk_units * 32 * 2) / ((uint64_t)grid_p * PART_NCLASS);
  region_cap = std::max<uint64_t>(1024, (region_cap + 127) & ~127ull);
  const size_t bucket_elems = (size_t)region_cap * grid_p * PART_NCLASS;
  if (h->part_bucket_elems < bucket_elems) {
    TRY(dev_alloc(h, &h->d_part_buckets, bucket_elems));
    h->part_bucket_elems = bucket_elems;
  }
  if (h->part_fill_elems < (size_t)grid_p * PART_NCLASS) {
    TRY(dev_alloc(h, &h->d_part_fill, (size_t)grid_p * PART_NCLASS));
    h->part_fill_elems = (size_t)grid_p * PART_NCLASS;
  }
  if (!h->d_slow_count) TRY(dev_alloc(h, &h->d_slow_count, (size_t)1));
  if (h->slow_list_elems < chunk_units) {
    TRY(dev_alloc(h, &h->d_slow_list, (size_t)chunk_units));
    h->slow_list_elems = chunk_units;
  }
  pl->grid_p = grid_p;
  pl->region_cap = (uint32_t)region_cap;
  pl->smem_h = (size_t)(part_class_start(1) + 1) * 4;
  CU(h, cudaFuncSetAttribute(k_hist_part, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl->smem_h));
  return PENG_OK;
}
// stage P on the units [ub, ue); carry: behind the keys an earlier launch of the same count left in the bucket regions
template <int W>
static int part_stage_p(peng_handle_t h, bool both, const PartPlan& pl, uint64_t ub, uint64_t ue, bool carry) {
  if (ue <= ub) return PENG_OK;
  PartArgs a{ub, ue, pl.grid_p, pl.region_cap, h->d_part_buckets, h->d_part_fill};
  a.carry_fill = carry ? 1u : 0u;
  return both ? launch_part<W, true>(h, h->part_cfg, a) : launch_part<W, false>(h, h->part_cfg, a);
}
// slow units (side stream) beside stage H, joined
template <int W>
static int part_stage_h(peng_handle_t h, bool both, const PartPlan& pl, bool mark) {
  TRY(launch_slow_beside<W>(h, both, mark));
  if (mark) MARK(h, 8);
  k_hist_part<<<PART_NCLASS * PART_CPC, PART_HT, pl.smem_h, h->stream>>>(h->d_part_buckets, pl.region_cap, pl.grid_p, h->d_part_fill,
                                                                        h->d_hist32);
  LAUNCH_CHECK(h);
  CU(h, cudaStreamWaitEvent(h->stream, h->ev_join, 0));
  if (mark) MARK(h, 7);
  return PENG_OK;
}

template <int W>
static int launch_count_partition(peng_handle_t h, bool both) {
  const uint64_t n_units = h->n_words32;
  const uint64_t units_per_chunk = 1ull << 25;
  PartPlan pl;
  TRY(part_reserve(h, std::min(n_units, units_per_chunk), &pl));
  if (h->pre.done) {  // stage P already ran, range by range, under the DMA of the ingest (peng_cuda_set_count_prefetch)
    MARK(h, 6); MARK(h, 2);
    TRY(part_stage_h<W>(h, both, pl, true));
    h->used_partition = true;
    return PENG_OK;
  }
  for (uint64_t ub = 0; ub < n_units; ub += units_per_chunk) {
    const uint64_t ue = std::min(n_units, ub + units_per_chunk);
    CU(h, cudaMemsetAsync(h->d_slow_count, 0, 4, h->stream));
    if (ub == 0) MARK(h, 6);  // phases 3..5 time the stages of the FIRST chunk (a full one whenever there are several)
    TRY(part_stage_p<W>(h, both, pl, ub, ue, false));
    if (ub == 0) MARK(h, 2);
    TRY(part_stage_h<W>(h, both, pl, ub == 0));
    if (ue < n_units) {
      if (both) TRY(launch_fold_canon(h, W, false, false));
      else k_fold<<<grid_for(h->n_patterns, 256), 256, 0, h->stream>>>(h->d_hist32, h->d_counts, h->n_patterns);
      LAUNCH_CHECK(h);
    }
  }
  h->used_partition = true;
  return PENG_OK;
}


// Count prefetch inside the hybrid ingest (peng_cuda_set_count_prefetch): tables and buckets of the armed count are set up
// before the first byte moves; prefetch_range runs stage P on a unit range whose words have landed.
static int prefetch_begin(peng_handle_t h, PrefetchRun* pr) {
  h->pre.done = false;
  pr->on = false;
  const int W = h->pre.W;
  if (W != 10 || h->n_words32 > (1ull << 25) || !use_partition(h, W)) return PENG_OK;  // the one-level partition path, one chunk
  const uint64_t n = 1ull << (2 * W);
  TRY(ensure_count_tables(h, W));
  h->counts_final = false;
  h->count_pending = false;
  h->stats_ready = false;
  if (!h->hist_clean) CU(h, cudaMemsetAsync(h->d_hist32, 0, (n + 4) * 4, h->stream));
  h->hist_clean = false;
  CU(h, cudaMemsetAsync(h->d_counts + n, 0, 8, h->stream));
  PartPlan pl;
  TRY(part_reserve(h, h->n_words32, &pl));
  CU(h, cudaMemsetAsync(h->d_slow_count, 0, 4, h->stream));
  pr->grid_p = pl.grid_p;
  pr->region_cap = pl.region_cap;
  pr->smem_h = pl.smem_h;
  pr->on = true;
  pr->first = true;
  return PENG_OK;
}
static int prefetch_range(peng_handle_t h, PrefetchRun* pr, uint64_t ub, uint64_t ue) {
  if (ue <= ub) return PENG_OK;
  const PartPlan pl{pr->grid_p, pr->region_cap, pr->smem_h};
  TRY(part_stage_p<10>(h, h->pre.strand == PENG_STRAND_BOTH, pl, ub, ue, !pr->first));
  pr->first = false;
  return PENG_OK;
}

extern "C" int peng_cuda_set_count_prefetch(peng_handle_t h, int W, int strand) {
  if (!h || (strand != PENG_STRAND_PLUS && strand != PENG_STRAND_BOTH)) return fail(h, PENG_E_INVALID, "set_count_prefetch: bad argument");
  h->pre.W = W == 10 ? 10 : 0;  // other pattern lengths: the hint is ignored, the count runs in one piece as before
  h->pre.strand = strand;
  h->pre.done = false;
  return PENG_OK;
}

// Deep partition path driver (W = 12, 14, 16), see k_part2.  Chunks of <= 2^30 positions bound the bucket memory; the
// uint32 table is folded into the 64-bit one only when a bin could overflow (2^32 positions since the last fold).
template <typename T>
static int deep_reserve(peng_handle_t h, int level, size_t bucket_elems, size_t fill_elems) {
  const size_t bytes = bucket_elems * sizeof(T);
  if (h->deep_bucket_bytes[level] < bytes) {
    unsigned char* p = (unsigned char*)h->d_deep_buckets[level];
    TRY(dev_alloc(h, &p, bytes));
    h->d_deep_buckets[level] = p;
    h->deep_bucket_bytes[level] = bytes;
  }
  if (h->deep_fill_elems[level] < fill_elems) {
    TRY(dev_alloc(h, &h->d_deep_fill[level], fill_elems));
    h->deep_fill_elems[level] = fill_elems;
  }
  return PENG_OK;
}
static uint32_t deep_region_cap(uint64_t keys, uint64_t n_regions, uint32_t min_cap) {
  uint64_t cap = keys * 2 / n_regions;
  cap = std::max<uint64_t>(min_cap, (cap + 127) & ~127ull);
  return (uint32_t)cap;
}
template <int NC2_BITS, int SUB_BITS, uint32_t CAP, typename OUT_T>
static int launch_part2(peng_handle_t h, const uint32_t* in, uint32_t in_cap, uint32_t n_in, const uint32_t* in_fill,
                        uint32_t n_c1, uint32_t G, uint32_t shift, OUT_T* out, uint32_t out_cap, uint32_t* out_fill) {
  const size_t smem = (size_t)(DEEP_P2_NQ + 1) * CAP * sizeof(OUT_T) + 256;  // the remainder move reads whole 16-byte pieces
  auto kern = k_part2<NC2_BITS, SUB_BITS, CAP, OUT_T>;
  CU(h, cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  CU(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<n_c1 * G, DEEP_P2_NT, smem, h->stream>>>(in, in_cap, n_in, in_fill, G, shift, out, out_cap, out_fill, h->d_hist32);
  LAUNCH_CHECK(h);
  return PENG_OK;
}

template <int W>
static int launch_count_deep(peng_handle_t h, bool both) {
  typedef PartGeom<W> G1;
  constexpr uint32_t NC1 = G1::NCLASS;
  const uint64_t n_units = h->n_words32;
  const uint64_t units_per_chunk = 1ull << 25;
  const unsigned grid_p = (unsigned)h->sm_count * 2;
  const uint64_t chunk_units = std::min(n_units, units_per_chunk);
  const uint64_t chunk_keys = chunk_units * 32;
  // stage P2: 4 CTAs of 256 threads per SM; G2 CTAs per stage-P class.  A leaf owns n_reg regions (see k_part2).
  const uint32_t G2 = std::max(1u, (unsigned)h->sm_count * DEEP_P2_MINB / NC1);
  const uint32_t cap1 = deep_region_cap(chunk_keys, (uint64_t)grid_p * NC1, 1024);
  TRY(deep_reserve<uint32_t>(h, 0, (size_t)cap1 * grid_p * NC1, (size_t)grid_p * NC1));
  constexpr uint32_t LEAVES = 1u << (2 * W - DEEP_LEAF_BITS);
  // W = 12: 5 + 4 class bits (16 sub-queues each); W = 14: 5 + 8; W = 16: 5 + 4 (32-bit entries, 16 sub-queues) + 8
  const uint32_t n_reg2a = G2 << 4;                      // W = 16: regions per middle class
  const uint32_t n_reg = W == 12 ? (G2 << 4) : W == 14 ? G2 : 1u;  // regions per leaf
  uint32_t cap2a = 0, cap2 = 0;
  if (h->deep_mode == 0) {
    if (W == 16) {
      cap2a = deep_region_cap(chunk_keys, (uint64_t)NC1 * 16 * n_reg2a, 256);
      TRY(deep_reserve<uint32_t>(h, 1, (size_t)cap2a * NC1 * 16 * n_reg2a, (size_t)NC1 * 16 * n_reg2a));
    }
    cap2 = deep_region_cap(chunk_keys, (uint64_t)LEAVES * n_reg, 256);
    TRY(deep_reserve<uint16_t>(h, 2, (size_t)cap2 * LEAVES * n_reg, (size_t)LEAVES * n_reg));
  }
  if (!h->d_slow_count) TRY(dev_alloc(h, &h->d_slow_count, (size_t)1));
  if (h->slow_list_elems < chunk_units) {
    TRY(dev_alloc(h, &h->d_slow_list, (size_t)chunk_units));
    h->slow_list_elems = chunk_units;
  }
  const size_t smem_leaf = (size_t)4 << DEEP_LEAF_BITS;
  CU(h, cudaFuncSetAttribute(k_hist_leaf, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_leaf));
  uint32_t* b1 = (uint32_t*)h->d_deep_buckets[0];
  uint32_t* f1 = h->d_deep_fill[0];
  uint64_t pos_since_fold = 0;
  for (uint64_t ub = 0; ub < n_units; ub += units_per_chunk) {
    const uint64_t ue = std::min(n_units, ub + units_per_chunk);
    if (pos_since_fold + (ue - ub) * 32 > 0xFFFFFFFFull) {  // a uint32 bin could overflow: fold first
      if (both) TRY(launch_fold_canon(h, W, false, false));
      else k_fold<<<grid_for(h->n_patterns, 256), 256, 0, h->stream>>>(h->d_hist32, h->d_counts, h->n_patterns);
      LAUNCH_CHECK(h);
      pos_since_fold = 0;
    }
    pos_since_fold += (ue - ub) * 32;
    CU(h, cudaMemsetAsync(h->d_slow_count, 0, 4, h->stream));
    const PartArgs a{ub, ue, grid_p, cap1, b1, f1};
    if (ub == 0) MARK(h, 6);
    const int rc_p = both ? launch_part_cfg<W, true, 512, 768, 2, false>(h, a) : launch_part_cfg<W, false, 512, 768, 2, false>(h, a);
    TRY(rc_p);
    if (ub == 0) MARK(h, 2);
    TRY(launch_slow_beside<W>(h, both, ub == 0));
    if (h->deep_mode == 1) {
      if (ub == 0) MARK(h, 8);
      k_hist_l2<<<grid_p, 512, 0, h->stream>>>(b1, cap1, grid_p, NC1, f1, h->d_hist32);
      LAUNCH_CHECK(h);
      CU(h, cudaStreamWaitEvent(h->stream, h->ev_join, 0));
    } else {
      uint16_t* b2 = (uint16_t*)h->d_deep_buckets[2];
      uint32_t* f2 = h->d_deep_fill[2];
      if (W == 12) {
        TRY((launch_part2<4, 4, DEEP_P2_CAP, uint16_t>(h, b1, cap1, grid_p, f1, NC1, G2, DEEP_LEAF_BITS, b2, cap2, f2)));
      } else if (W == 14) {
        TRY((launch_part2<8, 0, DEEP_P2_CAP, uint16_t>(h, b1, cap1, grid_p, f1, NC1, G2, DEEP_LEAF_BITS, b2, cap2, f2)));
      } else {
        uint32_t* b2a = (uint32_t*)h->d_deep_buckets[1];
        uint32_t* f2a = h->d_deep_fill[1];
        TRY((launch_part2<4, 4, DEEP_P2_CAP, uint32_t>(h, b1, cap1, grid_p, f1, NC1, G2, DEEP_LEAF_BITS + 8, b2a, cap2a, f2a)));
        TRY((launch_part2<8, 0, DEEP_P2_CAP, uint16_t>(h, b2a, cap2a, n_reg2a, f2a, NC1 * 16, 1, DEEP_LEAF_BITS, b2, cap2, f2)));
      }
      CU(h, cudaStreamWaitEvent(h->stream, h->ev_join, 0));  // the leaf merge is not atomic: the slow units must be in
      if (ub == 0) MARK(h, 8);
      k_hist_leaf<<<LEAVES, 1024, smem_leaf, h->stream>>>(b2, cap2, n_reg, f2, h->d_hist32);
      LAUNCH_CHECK(h);
    }
    if (ub == 0) MARK(h, 7);
  }
  h->used_partition = true;
  return PENG_OK;
}

// which inputs take a partition path (count_path 0 = automatic)
static bool use_partition(peng_handle_t h, int W) {
  if (W < 10 || h->count_path == 1) return false;
  if (h->count_path == 2) return true;
  // measured crossovers against the L2-atomic kernel (tools/count_bench.py, PENG_COUNT_PATH=1|2): the deep path pays fixed
  // costs per level (W = 16: 2^17 leaves), the L2 path pays per id once the table leaves the L2 (W >= 14)
  const int log2_min = W == 10 ? 22 : W == 12 ? 25 : W == 14 ? 24 : 28;
  return h->stream_end - PENG_FRONT_PAD >= (1ull << log2_min);
}

template <int W>
static int launch_count_W(peng_handle_t h, bool both) {
  // chunk geometry: 16 code words (256 end positions) per thread for large inputs, less when the input is
  // too small to fill the machine (the 48-position warm-up is then the larger share, but the run is short)
  const uint64_t n_words16 = (h->stream_end - PENG_FRONT_PAD + 15) / 16 + 1;
  uint32_t wpc = 16;
  while (wpc > 1 && n_words16 / wpc < (uint64_t)h->sm_count * 1024) wpc >>= 1;
  const uint64_t n_chunks = (n_words16 + wpc - 1) / wpc;
  constexpr size_t bins_bytes = (size_t)4 << (2 * W);
  // every launch covers at most 2^31 positions and is followed by a fold, so that no uint32 bin can wrap (few bins at
  // small W: one bin may take a large share of a multi-Gbp set)
  const uint64_t max_chunks_per_launch = (1ull << 31) / (16ull * wpc);
  if constexpr (W <= 8) {
    constexpr bool PACKED = W == 8;  // 128 KB of packed 16-bit bins
    const unsigned block = PACKED ? 512 : 256;
    const size_t smem = PACKED ? bins_bytes / 2 : bins_bytes;
    if constexpr (PACKED) {
      if (both) CU(h, cudaFuncSetAttribute(k_count_smem16<W, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      else CU(h, cudaFuncSetAttribute(k_count_smem16<W, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    for (uint64_t first = 0; first < n_chunks; first += max_chunks_per_launch) {
      const uint64_t nc = std::min(max_chunks_per_launch, n_chunks - first);
      const unsigned grid = (unsigned)std::min<uint64_t>(grid_for(nc, block), (uint64_t)h->sm_count * (PACKED ? 1 : 8));
      if constexpr (PACKED) {
        if (both) k_count_smem16<W, true><<<grid, block, smem, h->stream>>>(h->d_codes, h->d_kstart, h->d_hist32, first, nc, wpc);
        else k_count_smem16<W, false><<<grid, block, smem, h->stream>>>(h->d_codes, h->d_kstart, h->d_hist32, first, nc, wpc);
      } else {
        if (both) k_count_smem<W, true><<<grid, block, smem, h->stream>>>(h->d_codes, h->d_kstart, h->d_hist32, first, nc, wpc);
        else k_count_smem<W, false><<<grid, block, smem, h->stream>>>(h->d_codes, h->d_kstart, h->d_hist32, first, nc, wpc);
      }
      LAUNCH_CHECK(h);
      if (first + nc < n_chunks) {
        k_fold<<<grid_for(h->n_patterns, 256), 256, 0, h->stream>>>(h->d_hist32, h->d_counts, h->n_patterns);
        LAUNCH_CHECK(h);
      }
    }
    return PENG_OK;
  } else {
  if constexpr (W == 10) {
    if (use_partition(h, W)) return launch_count_partition<W>(h, both);
  } else {
    if (use_partition(h, W)) return launch_count_deep<W>(h, both);
  }
  // W >= 10: global atomics; a uint32 bin cannot overflow within 2^32 positions, so split and fold
  for (uint64_t first = 0; first < n_chunks; first += max_chunks_per_launch) {
    const uint64_t nc = std::min(max_chunks_per_launch, n_chunks - first);
    const unsigned block = 128;
    if (both) k_count_global<W, true><<<grid_for(nc, block), block, 0, h->stream>>>(h->d_codes, h->d_kstart, h->d_hist32, first, nc, wpc);
    else k_count_global<W, false><<<grid_for(nc, block), block, 0, h->stream>>>(h->d_codes, h->d_kstart, h->d_hist32, first, nc, wpc);
    LAUNCH_CHECK(h);
    if (first + nc < n_chunks) {
      k_fold<<<grid_for(h->n_patterns, 256), 256, 0, h->stream>>>(h->d_hist32, h->d_counts, h->n_patterns);
      LAUNCH_CHECK(h);
    }
  }
  return PENG_OK;
  }
}

// the uint32 launch table and the 64-bit table of pattern length W (kept while W stays the same)
static int ensure_count_tables(peng_handle_t h, int W) {
  const uint64_t n = 1ull << (2 * W);
  if (h->W != W || !h->d_counts) {
    if (h->peer.attached) return fail(h, PENG_E_INVALID, "the count tables of W = %d are shared with the peer group: peng_cuda_peer_detach first", h->peer.W);
    free_tables(h);
    TRY(dev_alloc(h, &h->d_hist32, n + 4));
    TRY(dev_alloc(h, &h->d_counts, n + 1));
    h->W = W;
    h->n_patterns = n;
    h->counts_final = false;
    h->count_pending = false;
    h->hist_clean = false;
  }
  return PENG_OK;
}

extern "C" int peng_cuda_count_local(peng_handle_t h, int W, int strand) {
  if (!h || !h->d_codes) return fail(h, PENG_E_INVALID, "count: no sequences ingested");
  if (W < 2 || W > PENG_MAX_W || (W & 1)) return fail(h, PENG_E_INVALID, "count: pattern length %d must be even and in [2,16]", W);
  if (strand != PENG_STRAND_PLUS && strand != PENG_STRAND_BOTH) return fail(h, PENG_E_INVALID, "count: bad strand");
  CU(h, cudaSetDevice(h->device));
  const uint64_t n = 1ull << (2 * W);
  TRY(ensure_count_tables(h, W));
  h->W = W;
  h->strand = strand;
  h->n_patterns = n;
  h->counts_final = false;
  h->count_pending = false;
  h->stats_ready = false;
  h->counts_sliced = false;
  // stage P of exactly this count already ran inside the ingest: the launch table and the ltot word hold its part
  const bool prefetched = h->pre.done && h->pre.W == W && h->pre.strand == strand && use_partition(h, W) && h->n_words32 <= (1ull << 25);
  h->pre.done = prefetched;
  if (!prefetched && !h->hist_clean) CU(h, cudaMemsetAsync(h->d_hist32, 0, (n + 4) * 4, h->stream));
  h->hist_clean = false;
  const bool both = strand == PENG_STRAND_BOTH;
  const bool partition = use_partition(h, W);
  // up to 2^30 stream positions (pads included) every path folds exactly once, at the end; the deep partition path
  // (W >= 12) folds between its chunks only when a uint32 bin could overflow, i.e. never below 2^32 positions.  A single
  // fold WRITES the whole 64-bit table: no clearing pass, no read of the old values, the mirror copy in the same kernel.
  const bool single_chunk = h->n_words32 <= (1ull << 25);
  const bool single_fold = single_chunk || (partition && W >= 12 && h->n_words32 * 32 <= 0xFFFFFFFFull);
  if (prefetched) {}
  else if (single_fold) CU(h, cudaMemsetAsync(h->d_counts + n, 0, 8, h->stream));
  else CU(h, cudaMemsetAsync(h->d_counts, 0, (n + 1) * 8, h->stream));
  h->mirrored = false;
  h->used_partition = false;
  MARK(h, 0);
  if (!partition) {  // the partition path derives the k-mer starts inside its first stage
    k_mark_kstart<<<grid_for(h->n_words32, 256), 256, 0, h->stream>>>(h->d_valid, h->d_kstart, h->n_words32, W,
                                                                     (unsigned long long*)(h->d_counts + n));
    LAUNCH_CHECK(h);
  }
  MARK(h, 1);
  if (!partition) { MARK(h, 6); MARK(h, 2); MARK(h, 3); MARK(h, 8); MARK(h, 7); }
  int rc = PENG_OK;
  switch (W) {
    case 2: rc = launch_count_W<2>(h, both); break;
    case 4: rc = launch_count_W<4>(h, both); break;
    case 6: rc = launch_count_W<6>(h, both); break;
    case 8: rc = launch_count_W<8>(h, both); break;
    case 10: rc = launch_count_W<10>(h, both); break;
    case 12: rc = launch_count_W<12>(h, both); break;
    case 14: rc = launch_count_W<14>(h, both); break;
    case 16: rc = launch_count_W<16>(h, both); break;
  }
  TRY(rc);
  MARK(h, 4);
  h->fold_deferred = false;
  if (h->defer_fold && single_chunk) {
    // 32-bit exchange: the caller sums the uint32 table (+ ltot in word n) over the ranks, peng_cuda_count_finish32 folds
    k_ltot_to32<<<1, 1, 0, h->stream>>>(h->d_counts + n, h->d_hist32 + n);
    LAUNCH_CHECK(h);
    h->fold_deferred = true;
  } else if (h->used_partition && both) {
    if (single_fold && h->fuse_mirror) {
      TRY(launch_fold_canon(h, W, true, true));
      h->mirrored = true;
    } else TRY(launch_fold_canon(h, W, single_fold, false));
  } else {
    if (single_fold) k_fold_first<<<grid_for(n, 256), 256, 0, h->stream>>>(h->d_hist32, h->d_counts, n);
    else k_fold<<<grid_for(n, 256), 256, 0, h->stream>>>(h->d_hist32, h->d_counts, n);
    LAUNCH_CHECK(h);
  }
  MARK(h, 5);
  h->pre.done = false;                // consumed: a second count of the same set starts from scratch
  h->hist_clean = !h->fold_deferred;  // every fold kernel zeroes the bins it reads
  return PENG_OK;
}

extern "C" int peng_cuda_set_count_path(peng_handle_t h, int path) {
  if (!h || path < 0 || path > 3)
    return fail(h, PENG_E_INVALID, "set_count_path: path must be 0 (auto), 1 (L2 atomics), 2 (partition) or 3 (partition, L2-RED second stage)");
  h->count_path = path == 3 ? 2 : path;
  h->deep_mode = path == 3 ? 1 : 0;
  return PENG_OK;
}
extern "C" int peng_cuda_last_count_path(peng_handle_t h) { return h ? (h->used_partition ? 2 : 1) : 0; }

extern "C" int peng_cuda_count_buffer(peng_handle_t h, void** device_ptr, uint64_t* n_u64) {
  if (!h || !h->d_counts) return fail(h, PENG_E_INVALID, "count_buffer: nothing counted");
  if (h->fold_deferred) return fail(h, PENG_E_INVALID, "count_buffer: the table is unfolded (peng_cuda_count_buffer32)");
  if (device_ptr) *device_ptr = h->d_counts;
  if (n_u64) *n_u64 = h->n_patterns + 1;
  return PENG_OK;
}

// 32-bit exchange for multi-GPU sets of fewer than 2^32 positions IN TOTAL (no uint32 bin, nor ltot, can overflow in the
// sum): half the bytes of the 64-bit exchange cross NVLink, and the fold runs once, after the sum, mirror included.
extern "C" int peng_cuda_count_local32(peng_handle_t h, int W, int strand) {
  if (!h) return PENG_E_INVALID;
  h->defer_fold = true;
  const int rc = peng_cuda_count_local(h, W, strand);
  h->defer_fold = false;
  TRY(rc);
  if (!h->fold_deferred) return fail(h, PENG_E_INVALID, "count_local32: the local set has more than 2^30 positions, use the 64-bit exchange");
  return PENG_OK;
}
extern "C" int peng_cuda_count_buffer32(peng_handle_t h, void** device_ptr, uint64_t* n_u32) {
  if (!h || !h->d_hist32 || !h->fold_deferred) return fail(h, PENG_E_INVALID, "count_buffer32: no unfolded table (peng_cuda_count_local32 first)");
  if (device_ptr) *device_ptr = h->d_hist32;
  if (n_u32) *n_u32 = h->n_patterns + 1;
  return PENG_OK;
}
extern "C" int peng_cuda_count_finish32(peng_handle_t h, uint64_t* ltot) {
  if (!h || !h->d_hist32 || !h->fold_deferred) return fail(h, PENG_E_INVALID, "count_finish32: no unfolded table (peng_cuda_count_local32 first)");
  CU(h, cudaSetDevice(h->device));
  const uint64_t n = h->n_patterns;
  k_ltot_from32<<<1, 1, 0, h->stream>>>(h->d_hist32 + n, h->d_counts + n);
  LAUNCH_CHECK(h);
  // The summed table may mix the two bin conventions (ranks choose their counting path from their own shard size: the
  // partition paths leave fast W-mers in their forward-id bin, the others count at the canonical id): the pair fold is
  // right for both — a canonical-only table has 0 at the partner bin — so BOTH strands always take it.
  if (h->strand == PENG_STRAND_BOTH) {
    TRY(launch_fold_canon(h, h->W, true, true));
    h->mirrored = true;
  } else {
    k_fold_first<<<grid_for(n, 256), 256, 0, h->stream>>>(h->d_hist32, h->d_counts, n);
    LAUNCH_CHECK(h);
  }
  h->fold_deferred = false;
  h->hist_clean = true;
  return peng_cuda_count_finish(h, ltot);
}

// Sliced finish of the 32-bit exchange (PLUS strand): the caller has reduce-scattered the uint32 table in place (this
// rank's slice [lo, hi) of it holds the sum over the ranks) and all-reduced the ltot word; only the slice is folded into
// the 64-bit table, and every later table pass (stats, IUPAC sums, EM) works on the slice.
extern "C" int peng_cuda_count_finish32_sliced(peng_handle_t h, uint64_t* ltot) {
  if (!h || !h->d_hist32 || !h->fold_deferred) return fail(h, PENG_E_INVALID, "count_finish32_sliced: no unfolded table (peng_cuda_count_local32 first)");
  if (h->slice_parts <= 1) return fail(h, PENG_E_INVALID, "count_finish32_sliced: no table slice set (peng_cuda_set_table_slice)");
  if (h->strand != PENG_STRAND_PLUS)
    return fail(h, PENG_E_INVALID, "count_finish32_sliced: both strands pair ids of different slices — sum the whole table (peng_cuda_count_finish32)");
  CU(h, cudaSetDevice(h->device));
  const uint64_t n = h->n_patterns;
  slice_range(h, &h->slice_lo, &h->slice_hi);
  k_ltot_from32<<<1, 1, 0, h->stream>>>(h->d_hist32 + n, h->d_counts + n);
  LAUNCH_CHECK(h);
  const uint64_t ns = h->slice_hi - h->slice_lo;
  k_fold_first<<<grid_for(ns, 256), 256, 0, h->stream>>>(h->d_hist32 + h->slice_lo, h->d_counts + h->slice_lo, ns);
  LAUNCH_CHECK(h);
  h->fold_deferred = false;
  h->counts_sliced = true;
  h->mirrored = true;  // nothing to mirror
  return peng_cuda_count_finish(h, ltot);
}

extern "C" int peng_cuda_set_table_slice(peng_handle_t h, int part, int n_parts) {
  if (!h || n_parts < 1 || part < 0 || part >= n_parts) return fail(h, PENG_E_INVALID, "set_table_slice: bad slice %d/%d", part, n_parts);
  h->slice_part = part;
  h->slice_parts = n_parts;
  h->stats_ready = false;
  return PENG_OK;
}
extern "C" int peng_cuda_table_slice(peng_handle_t h, uint64_t* lo, uint64_t* hi) {
  if (!h || !h->n_patterns) return fail(h, PENG_E_INVALID, "table_slice: nothing counted");
  uint64_t a, b;
  slice_range(h, &a, &b);
  if (lo) *lo = a;
  if (hi) *hi = b;
  return PENG_OK;
}

// device half of the finish: everything that is left to do on the stream (the mirror copy where no earlier kernel wrote it)
static int count_finish_device(peng_handle_t h) {
  if (h->strand == PENG_STRAND_BOTH && !h->mirrored) {
    TRY(launch_mirror(h));
    h->mirrored = true;
  }
  return PENG_OK;
}

extern "C" int peng_cuda_count_finish(peng_handle_t h, uint64_t* ltot) {
  if (!h || !h->d_counts) return fail(h, PENG_E_INVALID, "count_finish: nothing counted");
  if (h->fold_deferred) return fail(h, PENG_E_INVALID, "count_finish: the table is unfolded (peng_cuda_count_finish32)");
  CU(h, cudaSetDevice(h->device));
  if (!h->counts_final) {
    TRY(count_finish_device(h));
    uint32_t peer_err = 0;
    CU(h, cudaMemcpyAsync(&h->ltot, h->d_counts + h->n_patterns, 8, cudaMemcpyDeviceToHost, h->stream));
    if (h->peer.attached) CU(h, cudaMemcpyAsync(&peer_err, h->peer.d_flags + 9, 4, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    if (peer_err) {
      cudaMemsetAsync(h->peer.d_flags + 9, 0, 4, h->stream);
      return fail(h, PENG_E_CUDA, "peer exchange: rank %u of the peer group did not arrive within %.0f ms (%s of the exchange) — the count table is incomplete",
                  peer_err & 15u, h->peer.timeout_ms, (peer_err >> 4) & 1u ? "end" : "start");
    }
    h->counts_final = true;
    h->count_pending = false;
  }
  if (ltot) *ltot = h->ltot;
  return PENG_OK;
}
// consumers of the finished table: fetch ltot now if the count was left pending (ltot == NULL in peng_cuda_count / _count_exchange32)
static int ensure_final(peng_handle_t h) {
  if (h && !h->counts_final && h->count_pending && !h->fold_deferred) return peng_cuda_count_finish(h, nullptr);
  return PENG_OK;
}

extern "C" int peng_cuda_count(peng_handle_t h, int W, int strand, uint64_t* ltot) {
  if (h) h->fuse_mirror = true;
  const int rc = peng_cuda_count_local(h, W, strand);
  if (h) h->fuse_mirror = false;
  TRY(rc);
  if (!ltot) {  // no read-back, no synchronisation: the table is complete in stream order, ltot is fetched by whoever asks first
    TRY(count_finish_device(h));
    h->count_pending = true;
    return PENG_OK;
  }
  return peng_cuda_count_finish(h, ltot);
}

// ---- peer group ---------------------------------------------------------------------------------------------------------
struct PeerBlob {
  uint32_t magic;
  int32_t W, device, pid;
  uint64_t hist, counts, flags;  // raw device pointers (used when exporter and attacher are the same process)
  cudaIpcMemHandle_t ipc_hist, ipc_counts, ipc_flags;
};
static_assert(sizeof(PeerBlob) <= 256, "PENG_PEER_BLOB_BYTES");
#define PEER_MAGIC 0x50454E47u

extern "C" int peng_cuda_peer_export(peng_handle_t h, int W, void* blob) {
  if (!h || !blob) return fail(h, PENG_E_INVALID, "peer_export: NULL argument");
  if (W < 8 || W > PENG_MAX_W || (W & 1)) return fail(h, PENG_E_INVALID, "peer_export: the fused exchange covers W = 8 .. 16");
  CU(h, cudaSetDevice(h->device));
  if (h->peer.attached) return fail(h, PENG_E_INVALID, "peer_export: already attached (peng_cuda_peer_detach first)");
  TRY(ensure_count_tables(h, W));
  if (!h->peer.d_flags) {
    TRY(dev_alloc(h, &h->peer.d_flags, (size_t)64));
    CU(h, cudaMemsetAsync(h->peer.d_flags, 0, 64 * 4, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
  }
  PeerBlob b;
  memset(&b, 0, sizeof b);
  b.magic = PEER_MAGIC;
  b.W = W;
  b.device = h->device;
  b.pid = (int32_t)getpid();
  b.hist = (uint64_t)h->d_hist32;
  b.counts = (uint64_t)h->d_counts;
  b.flags = (uint64_t)h->peer.d_flags;
  CU(h, cudaIpcGetMemHandle(&b.ipc_hist, h->d_hist32));
  CU(h, cudaIpcGetMemHandle(&b.ipc_counts, h->d_counts));
  CU(h, cudaIpcGetMemHandle(&b.ipc_flags, h->peer.d_flags));
  memset(blob, 0, 256);
  memcpy(blob, &b, sizeof b);
  return PENG_OK;
}

extern "C" int peng_cuda_peer_detach(peng_handle_t h) {
  if (!h) return PENG_E_INVALID;
  cudaSetDevice(h->device);
  if (h->peer.attached) cudaStreamSynchronize(h->stream);
  for (int k = 0; k < 3; k++)
    for (int p = 0; p < 8; p++)
      if (h->peer.opened[k][p]) { cudaIpcCloseMemHandle(h->peer.opened[k][p]); h->peer.opened[k][p] = nullptr; }
  h->peer.attached = false;
  h->peer.world = 1;
  return PENG_OK;
}

// BOTH strands: the exchange units this rank owns — the quarter-tile pairs of the middles m <= revcomp(m) with
// m mod world == rank — listed so that the CTAs take them one by one (a strided walk over all m left half of the CTAs
// without work: 296 = 4 * 74 keeps m mod 2 fixed per CTA).  Built when the group is attached: nothing that allocates or
// synchronises is left to do once the first exchange kernel of the group spins on a device.
static int peer_build_units(peng_handle_t h) {
  if (h->peer.units_W == h->W && h->peer.units_rank == h->peer.rank && h->peer.units_world == h->peer.world) return PENG_OK;
  std::vector<uint32_t> units;
  const uint32_t n_mid = 1u << (2 * (h->W - 6));
  for (uint32_t m = (uint32_t)h->peer.rank; m < n_mid; m += (uint32_t)h->peer.world)
  {
    const uint32_t mp = (uint32_t)peng_host::revcomp_id64(m, h->W - 6);
    if (m < mp) for (uint32_t q = 0; q < 4; q++) units.push_back(m << 3 | q);  // four quarter units
    else if (m == mp) units.push_back(m << 3 | 4u);                              // self-paired middle: one whole-tile unit
  }
  TRY(dev_alloc(h, &h->peer.d_units, std::max<size_t>(units.size(), 1)));
  CU(h, cudaMemcpyAsync(h->peer.d_units, units.data(), units.size() * 4, cudaMemcpyHostToDevice, h->stream));
  CU(h, cudaStreamSynchronize(h->stream));
  h->peer.n_owned = (uint32_t)units.size();
  h->peer.units_W = h->W; h->peer.units_rank = h->peer.rank; h->peer.units_world = h->peer.world;
  return PENG_OK;
}

extern "C" int peng_cuda_peer_attach(peng_handle_t h, int rank, int world, const void* blobs) {
  if (!h || !blobs || world < 1 || world > 8 || rank < 0 || rank >= world) return fail(h, PENG_E_INVALID, "peer_attach: bad rank %d / world %d (1 .. 8)", rank, world);
  if (h->peer.attached) return fail(h, PENG_E_INVALID, "peer_attach: already attached");
  if (!h->peer.d_flags || !h->d_hist32) return fail(h, PENG_E_INVALID, "peer_attach: peng_cuda_peer_export first");
  CU(h, cudaSetDevice(h->device));
  const unsigned char* base = (const unsigned char*)blobs;
  for (int p = 0; p < world; p++) {
    PeerBlob b;
    memcpy(&b, base + (size_t)p * 256, sizeof b);
    if (b.magic != PEER_MAGIC || b.W != h->W) { peng_cuda_peer_detach(h); return fail(h, PENG_E_INVALID, "peer_attach: blob %d is not a peer export of W = %d", p, h->W); }
    if (p == rank) {
      h->peer.hist[p] = h->d_hist32; h->peer.counts[p] = h->d_counts; h->peer.flags[p] = h->peer.d_flags;
    } else if (b.pid == (int32_t)getpid()) {  // another handle of this process: its pointers are valid here
      if (b.device != h->device) {
        const cudaError_t e = cudaDeviceEnablePeerAccess(b.device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { peng_cuda_peer_detach(h); return fail(h, PENG_E_CUDA, "peer_attach: no peer access from device %d to %d: %s", h->device, b.device, cudaGetErrorString(e)); }
        cudaGetLastError();
      }
      h->peer.hist[p] = (uint32_t*)b.hist; h->peer.counts[p] = (uint64_t*)b.counts; h->peer.flags[p] = (uint32_t*)b.flags;
    } else {
      void* ptr[3] = {nullptr, nullptr, nullptr};
      const cudaIpcMemHandle_t* hs[3] = {&b.ipc_hist, &b.ipc_counts, &b.ipc_flags};
      for (int k = 0; k < 3; k++) {
        const cudaError_t e = cudaIpcOpenMemHandle(&ptr[k], *hs[k], cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) { peng_cuda_peer_detach(h); return fail(h, PENG_E_CUDA, "peer_attach: cudaIpcOpenMemHandle(rank %d): %s", p, cudaGetErrorString(e)); }
        h->peer.opened[k][p] = ptr[k];
      }
      h->peer.hist[p] = (uint32_t*)ptr[0]; h->peer.counts[p] = (uint64_t*)ptr[1]; h->peer.flags[p] = (uint32_t*)ptr[2];
    }
  }
  h->peer.colocated = 0;
  for (int p = 0; p < world; p++) {
    PeerBlob b;
    memcpy(&b, base + (size_t)p * 256, sizeof b);
    if (b.pid == (int32_t)getpid() && b.device == h->device) h->peer.colocated++;
  }
  h->peer.colocated = std::max(1, h->peer.colocated);
  h->peer.rank = rank;
  h->peer.world = world;
  h->peer.W = h->W;
  h->peer.epoch = 0;
  h->peer.attached = true;
  TRY(peer_build_units(h));
  // CUDA loads a kernel at its first launch, and that load can wait for the device to drain — forever, if it is requested
  // while another rank's exchange kernel spins on this device waiting for THIS rank (seen with several handles of one
  // process on one device).  Everything launched behind an exchange kernel is therefore loaded now.
  {
    cudaFuncAttributes fa;
    CU(h, cudaFuncGetAttributes(&fa, k_peer_exchange32<true>));
    CU(h, cudaFuncGetAttributes(&fa, k_peer_exchange32<false>));
    CU(h, cudaFuncGetAttributes(&fa, k_expand_home_tiled));
    CU(h, cudaFuncGetAttributes(&fa, k_fold_first));
  }
  return PENG_OK;
}

// trace of the last exchange kernel on this rank, nanoseconds relative to its start: arrived (all ranks' tables complete),
// units of CTA 0 done, CTA 0's stores visible system-wide, all CTAs done, departed (all ranks' stores have landed)
extern "C" int peng_cuda_peer_trace(peng_handle_t h, double* out5) {
  if (!h || !h->peer.d_flags || !out5) return fail(h, PENG_E_INVALID, "peer_trace: no peer group");
  CU(h, cudaSetDevice(h->device));
  unsigned long long st[8];
  CU(h, cudaMemcpyAsync(st, h->peer.d_flags + 16, sizeof st, cudaMemcpyDeviceToHost, h->stream));
  CU(h, cudaStreamSynchronize(h->stream));
  for (int i = 0; i < 5; i++) out5[i] = (double)(long long)(st[i + 1] - st[0]);
  if (getenv("PENG_PEER_TRACE")) fprintf(stderr, "[peer trace] first unit of CTA 0: loads landed %+.2f us, stores issued %+.2f us (since kernel start)\n",
                                         1e-3 * (double)(long long)(st[6] - st[0]), 1e-3 * (double)(long long)(st[7] - st[0]));
  return PENG_OK;
}

// The multi-GPU counting step behind peng_cuda_count_local32: sum over the ranks, fold, mirror and publication of the
// finished table on every rank in ONE kernel over peer memory (k_peer_exchange32).  Every rank of the group must call it.
// ltot == NULL: nothing is read back and nothing waits.
extern "C" int peng_cuda_count_exchange32(peng_handle_t h, uint64_t* ltot) {
  if (!h || !h->d_hist32 || !h->fold_deferred) return fail(h, PENG_E_INVALID, "count_exchange32: no unfolded table (peng_cuda_count_local32 first)");
  if (!h->peer.attached || h->peer.W != h->W) return fail(h, PENG_E_INVALID, "count_exchange32: no peer group for W = %d (peng_cuda_peer_export / _attach)", h->W);
  CU(h, cudaSetDevice(h->device));
  PeerArgs A;
  for (int p = 0; p < 8; p++) { A.hist[p] = h->peer.hist[p]; A.counts[p] = h->peer.counts[p]; A.flags[p] = h->peer.flags[p]; }
  A.rank = h->peer.rank;
  A.world = h->peer.world;
  A.W = h->W;
  A.n = h->n_patterns;
  A.timeout_cycles = (long long)(h->peer.timeout_ms * 1.9e6);  // SM clock ~1.9 GHz
  h->peer.epoch += 2;
  A.epoch = h->peer.epoch - 1;  // arrive = epoch, depart = epoch + 1
  const bool both = h->strand == PENG_STRAND_BOTH;
  if (both) TRY(peer_build_units(h));
  A.units = h->peer.d_units;
  A.n_owned = h->peer.n_owned;
  const uint64_t n_units = both ? (uint64_t)std::max<uint32_t>(h->peer.n_owned, 1) : ((h->n_patterns >> 10) + h->peer.world - 1) / h->peer.world;
  // One rank per device (the product): two CTAs per SM (128 registers x 256 threads).  Several ranks on ONE device (tests,
  // handles of one process): their exchange kernels spin until the slowest rank has finished counting, so together they
  // may hold at most half of the SMs — the counting kernels of the late ranks (1024-thread CTAs) need whole SMs to run.
  const uint64_t cta_budget = h->peer.colocated > 1 ? std::max<uint64_t>(1, (uint64_t)h->sm_count / (2 * (uint64_t)h->peer.colocated))
                                                    : (uint64_t)h->sm_count * 2;
  unsigned grid = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>(n_units, cta_budget));
  static const int grid_env = getenv("PENG_PEER_GRID") ? atoi(getenv("PENG_PEER_GRID")) : 0;  // tuning knob
  if (grid_env > 0) grid = (unsigned)std::min<uint64_t>(grid, (uint64_t)grid_env);
  MARK(h, 9);
  if (both) k_peer_exchange32<true><<<grid, 256, 0, h->stream>>>(A);
  else k_peer_exchange32<false><<<grid, 256, 0, h->stream>>>(A);
  LAUNCH_CHECK(h);
  // local half: widen to 64 bits (+ the mirror copy: transposed tiles) and clear the launch table
  if (both) k_expand_home_tiled<<<1u << (2 * (h->W - 6)), 256, 0, h->stream>>>(h->d_hist32, h->d_counts, h->W);
  else k_fold_first<<<grid_for(h->n_patterns, 256), 256, 0, h->stream>>>(h->d_hist32, h->d_counts, h->n_patterns);
  LAUNCH_CHECK(h);
  MARK(h, 10);
  h->fold_deferred = false;
  h->mirrored = true;
  h->hist_clean = true;
  if (!ltot) {
    h->count_pending = true;
    return PENG_OK;
  }
  return peng_cuda_count_finish(h, ltot);
}

extern "C" int peng_cuda_get_counts(peng_handle_t h, uint64_t* counts) {
  TRY(ensure_final(h));
  if (!h || !h->d_counts || !h->counts_final) return fail(h, PENG_E_INVALID, "get_counts: count table not final");
  CU(h, cudaSetDevice(h->device));
  // sliced count table: the slice only (out holds hi - lo elements, peng_cuda_table_slice)
  CU(h, cudaMemcpyAsync(counts, h->d_counts + counts_lo(h), (counts_hi(h) - counts_lo(h)) * 8, cudaMemcpyDeviceToHost, h->stream));
  CU(h, cudaStreamSynchronize(h->stream));
  return PENG_OK;
}

__global__ void k_narrow_u64(const uint64_t* __restrict__ in, uint32_t* __restrict__ out, uint64_t n) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (uint32_t)in[i];
}
extern "C" int peng_cuda_get_counts32(peng_handle_t h, uint32_t* counts) {
  TRY(ensure_final(h));
  if (!h || !h->d_counts || !h->counts_final || !counts) return fail(h, PENG_E_INVALID, "get_counts32: count table not final");
  if (h->counts_sliced) return fail(h, PENG_E_INVALID, "get_counts32: sliced count table (peng_cuda_get_counts)");
  if (h->ltot > 0xFFFFFFFFull) return fail(h, PENG_E_INVALID, "get_counts32: %llu W-mers were looked at — a bin may exceed 32 bits (peng_cuda_get_counts)", (unsigned long long)h->ltot);
  CU(h, cudaSetDevice(h->device));
  TRY(need_scratch(h, h->n_patterns * 4));
  k_narrow_u64<<<grid_for(h->n_patterns, 256), 256, 0, h->stream>>>(h->d_counts, (uint32_t*)h->d_scratch, h->n_patterns);
  LAUNCH_CHECK(h);
  CU(h, cudaMemcpyAsync(counts, h->d_scratch, h->n_patterns * 4, cudaMemcpyDeviceToHost, h->stream));
  CU(h, cudaStreamSynchronize(h->stream));
  return PENG_OK;
}

extern "C" int peng_cuda_gather_counts(peng_handle_t h, const uint64_t* ids, uint64_t n, uint64_t* out) {
  TRY(ensure_final(h));
  if (!h || !h->d_counts || !h->counts_final || (n && (!ids || !out))) return fail(h, PENG_E_INVALID, "gather_counts: count table not final");
  if (n == 0) return PENG_OK;
  for (uint64_t i = 0; i < n; i++)
    if (ids[i] < counts_lo(h) || ids[i] >= counts_hi(h)) return fail(h, PENG_E_INVALID, "gather_counts: id out of range (of this rank's slice)");
  CU(h, cudaSetDevice(h->device));
  TRY(need_scratch(h, n * 16));
  uint64_t* d_ids = (uint64_t*)h->d_scratch;
  uint64_t* d_out = d_ids + n;
  CU(h, cudaMemcpyAsync(d_ids, ids, n * 8, cudaMemcpyHostToDevice, h->stream));
  k_gather_u64<<<grid_for(n, 256), 256, 0, h->stream>>>(h->d_counts, d_ids, n, d_out);
  LAUNCH_CHECK(h);
  CU(h, cudaMemcpyAsync(out, d_out, n * 8, cudaMemcpyDeviceToHost, h->stream));
  CU(h, cudaStreamSynchronize(h->stream));
  return PENG_OK;
}

extern "C" int peng_cuda_counts_total(peng_handle_t h, uint64_t* total) {
  TRY(ensure_final(h));
  if (!h || !h->d_counts || !h->counts_final || !total) return fail(h, PENG_E_INVALID, "counts_total: count table not final");
  CU(h, cudaSetDevice(h->device));
  TRY(need_scratch(h, 8));
  CU(h, cudaMemsetAsync(h->d_scratch, 0, 8, h->stream));
  k_sum_u64<<<h->sm_count * 8, 256, 0, h->stream>>>(h->d_counts + counts_lo(h), counts_hi(h) - counts_lo(h), (unsigned long long*)h->d_scratch);
  LAUNCH_CHECK(h);
  CU(h, cudaMemcpyAsync(total, h->d_scratch, 8, cudaMemcpyDeviceToHost, h->stream));
  CU(h, cudaStreamSynchronize(h->stream));
  return PENG_OK;
}

extern "C" int peng_cuda_set_counts(peng_handle_t h, int W, int strand, const uint64_t* counts, uint64_t ltot) {
  if (!h || !counts) return fail(h, PENG_E_INVALID, "set_counts: NULL argument");
  if (W < 2 || W > PENG_MAX_W || (W & 1)) return fail(h, PENG_E_INVALID, "set_counts: bad pattern length %d", W);
  CU(h, cudaSetDevice(h->device));
  const uint64_t n = 1ull << (2 * W);
  TRY(ensure_count_tables(h, W));
  h->W = W;
  h->strand = strand;
  h->n_patterns = n;
  h->count_pending = false;
  CU(h, cudaMemcpyAsync(h->d_counts, counts, n * 8, cudaMemcpyHostToDevice, h->stream));
  CU(h, cudaMemcpyAsync(h->d_counts + n, &ltot, 8, cudaMemcpyHostToDevice, h->stream));
  CU(h, cudaStreamSynchronize(h->stream));
  h->ltot = ltot;
  h->counts_final = true;
  h->counts_sliced = false;
  h->stats_ready = false;
  return PENG_OK;
}

// =====================================================================================================
// C ABI: background model
// =====================================================================================================
extern "C" int peng_cuda_background_counts(peng_handle_t h, int K) {
  if (!h || !h->d_codes) return fail(h, PENG_E_INVALID, "background: no sequences ingested");
  if (K < 0 || K > PENG_MAX_ORDER) return fail(h, PENG_E_INVALID, "background: order %d not in [0,%d]", K, PENG_MAX_ORDER);
  CU(h, cudaSetDevice(h->device));
  const uint32_t n_bins = v_offset(K + 1);
  TRY(dev_alloc(h, &h->d_bgcounts, (size_t)n_bins));
  CU(h, cudaMemsetAsync(h->d_bgcounts, 0, n_bins * 8, h->stream));
  const uint64_t n = h->stream_end - PENG_FRONT_PAD;
  const unsigned block = 256;
  const unsigned grid = (unsigned)std::max<uint64_t>(1, std::min<uint64_t>(grid_for(n, block), (uint64_t)h->sm_count * 8));
  k_bg_counts<<<grid, block, n_bins * 4, h->stream>>>(h->d_codes, h->d_valid, h->d_sstart, h->stream_end, K,
                                                     (unsigned long long*)h->d_bgcounts, n_bins);
  LAUNCH_CHECK(h);
  h->K = K;
  return PENG_OK;
}

extern "C" int peng_cuda_background_finish(peng_handle_t h, int K, const float* alpha, int interpolate, int precision,
                                           void* v_out) {
  if (!h || !h->d_bgcounts || h->K != K || !alpha || !v_out) return fail(h, PENG_E_INVALID, "background_finish: bad call");
  CU(h, cudaSetDevice(h->device));
  const uint32_t n_bins = v_offset(K + 1);
  std::vector<uint64_t> n(n_bins);
  CU(h, cudaMemcpyAsync(n.data(), h->d_bgcounts, n_bins * 8, cudaMemcpyDeviceToHost, h->stream));
  CU(h, cudaStreamSynchronize(h->stream));
  if (precision == PENG_F32REF) peng_host::calculate_v<float>(n.data(), K, alpha, interpolate != 0, (float*)v_out);
  else peng_host::calculate_v<double>(n.data(), K, alpha, interpolate != 0, (double*)v_out);
  return PENG_OK;
}

// calculateV on (k+1)-mer counts the caller has summed over the shards of a multi-GPU run (pure host arithmetic)
extern "C" int peng_cuda_background_from_counts(int K, const float* alpha, int interpolate, int precision, const int64_t* kmer_counts,
                                                void* v_out) {
  if (K < 0 || K > PENG_MAX_ORDER || !alpha || !kmer_counts || !v_out) return PENG_E_INVALID;
  const uint32_t n_bins = v_offset(K + 1);
  std::vector<uint64_t> n(kmer_counts, kmer_counts + n_bins);
  if (precision == PENG_F32REF) peng_host::calculate_v<float>(n.data(), K, alpha, interpolate != 0, (float*)v_out);
  else peng_host::calculate_v<double>(n.data(), K, alpha, interpolate != 0, (double*)v_out);
  return PENG_OK;
}

extern "C" int peng_cuda_background_model(peng_handle_t h, int K, const float* alpha, int interpolate, int precision,
                                          void* v_out, int64_t* kmer_counts_out) {
  TRY(peng_cuda_background_counts(h, K));
  TRY(peng_cuda_background_finish(h, K, alpha, interpolate, precision, v_out));
  if (kmer_counts_out) {
    const uint32_t n_bins = v_offset(K + 1);
    CU(h, cudaMemcpyAsync(kmer_counts_out, h->d_bgcounts, n_bins * 8, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
  }
  return PENG_OK;
}

// =====================================================================================================
// C ABI: tables
// =====================================================================================================
extern "C" int peng_cuda_stats(peng_handle_t h, int K, const float* v32, const double* v64, int k, int max_k) {
  TRY(ensure_final(h));
  if (!h || !h->counts_final) return fail(h, PENG_E_INVALID, "stats: count table not final");
  if (K < 0 || K > PENG_MAX_ORDER || k < 0 || k > K || max_k < k || max_k > K || max_k > h->W - 1)
    return fail(h, PENG_E_INVALID, "stats: bad orders K=%d k=%d max_k=%d (W=%d)", K, k, max_k, h->W);
  if (!v32 && !v64) return fail(h, PENG_E_INVALID, "stats: no background model given");
  CU(h, cudaSetDevice(h->device));
  const uint32_t nv = v_offset(K + 1);
  const int both = h->strand == PENG_STRAND_BOTH;
  const unsigned block = 256;
  free_tables(h);
  // id-sliced tables (peng_cuda_set_table_slice): only this rank's ids are computed and kept
  uint64_t lo, hi;
  slice_range(h, &lo, &hi);
  if (h->counts_sliced && (lo != h->slice_lo || hi != h->slice_hi)) return fail(h, PENG_E_INVALID, "stats: the table slice changed since the sliced count");
  const uint64_t n = hi - lo;
  h->tab_lo = lo;
  h->tab_hi = hi;
  if (v32) {
    TRY(dev_alloc(h, &h->d_v32, (size_t)nv));
    CU(h, cudaMemcpyAsync(h->d_v32, v32, nv * 4, cudaMemcpyHostToDevice, h->stream));
    for (int o = 0; o <= max_k; o++) {
      TRY(dev_alloc(h, &h->d_bg32[o], n));
      k_bgprob<float><<<grid_for(n, block), block, v_offset(o + 1) * 4, h->stream>>>(h->d_v32, h->W, o, both, vbase(h->d_bg32[o], lo), lo, n);
      LAUNCH_CHECK(h);
    }
    TRY(dev_alloc(h, &h->d_exp32, n)); TRY(dev_alloc(h, &h->d_z32, n)); TRY(dev_alloc(h, &h->d_logp32, n));
    k_stats<float><<<grid_for(n, block), block, 0, h->stream>>>(h->d_counts, vbase(h->d_bg32[k], lo), h->ltot, vbase(h->d_exp32, lo),
                                                                vbase(h->d_z32, lo), vbase(h->d_logp32, lo), lo, n);
    LAUNCH_CHECK(h);
  }
  if (v64) {
    TRY(dev_alloc(h, &h->d_v64, (size_t)nv));
    CU(h, cudaMemcpyAsync(h->d_v64, v64, nv * 8, cudaMemcpyHostToDevice, h->stream));
    for (int o = 0; o <= max_k; o++) {
      TRY(dev_alloc(h, &h->d_bg64[o], n));
      k_bgprob<double><<<grid_for(n, block), block, v_offset(o + 1) * 8, h->stream>>>(h->d_v64, h->W, o, both, vbase(h->d_bg64[o], lo), lo, n);
      LAUNCH_CHECK(h);
    }
    TRY(dev_alloc(h, &h->d_exp64, n)); TRY(dev_alloc(h, &h->d_z64, n)); TRY(dev_alloc(h, &h->d_logp64, n));
    k_stats<double><<<grid_for(n, block), block, 0, h->stream>>>(h->d_counts, vbase(h->d_bg64[k], lo), h->ltot, vbase(h->d_exp64, lo),
                                                                 vbase(h->d_z64, lo), vbase(h->d_logp64, lo), lo, n);
    LAUNCH_CHECK(h);
  }
  CU(h, cudaStreamSynchronize(h->stream));  // v32 / v64 host buffers may go away
  h->K = K;
  h->k_used = k;
  h->max_k = max_k;
  h->stats_ready = true;
  return PENG_OK;
}

static const void* table_ptr(peng_handle_t h, int table, int order, int precision) {
  const bool f32 = precision == PENG_F32REF;
  switch (table) {
    case PENG_TABLE_BGPROB:
      if (order < 0 || order > h->max_k) return nullptr;
      return f32 ? (const void*)h->d_bg32[order] : (const void*)h->d_bg64[order];
    case PENG_TABLE_EXPECTED: return f32 ? (const void*)h->d_exp32 : (const void*)h->d_exp64;
    case PENG_TABLE_ZSCORE: return f32 ? (const void*)h->d_z32 : (const void*)h->d_z64;
    case PENG_TABLE_LOGP: return f32 ? (const void*)h->d_logp32 : (const void*)h->d_logp64;
  }
  return nullptr;
}

extern "C" int peng_cuda_get_table(peng_handle_t h, int table, int order, int precision, void* out) {
  if (!h || !h->stats_ready || !out) return fail(h, PENG_E_INVALID, "get_table: stats not computed");
  const void* p = table_ptr(h, table, order, precision);
  if (!p) return fail(h, PENG_E_INVALID, "get_table: table %d order %d precision %d not available", table, order, precision);
  CU(h, cudaSetDevice(h->device));
  const size_t es = precision == PENG_F32REF ? 4 : 8;
  CU(h, cudaMemcpyAsync(out, p, (h->tab_hi - h->tab_lo) * es, cudaMemcpyDeviceToHost, h->stream));  // sliced tables: the slice
  CU(h, cudaStreamSynchronize(h->stream));
  return PENG_OK;
}

// sparse read-back of a float / double table: out[i] = table[ids[i]] (the host driver prints a handful of rows)
template <typename R>
__global__ void k_gather_real(const R* __restrict__ table, const uint64_t* __restrict__ ids, uint64_t n, R* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = table[ids[i]];
}

extern "C" int peng_cuda_gather_table(peng_handle_t h, int table, int order, int precision, const uint64_t* ids, uint64_t n,
                                      void* out) {
  if (!h || !h->stats_ready || (n && (!ids || !out))) return fail(h, PENG_E_INVALID, "gather_table: stats not computed");
  const void* p = table_ptr(h, table, order, precision);
  if (!p) return fail(h, PENG_E_INVALID, "gather_table: table %d order %d precision %d not available", table, order, precision);
  if (n == 0) return PENG_OK;
  for (uint64_t i = 0; i < n; i++)
    if (ids[i] < h->tab_lo || ids[i] >= h->tab_hi) return fail(h, PENG_E_INVALID, "gather_table: id out of range (of this rank's slice)");
  CU(h, cudaSetDevice(h->device));
  const size_t es = precision == PENG_F32REF ? 4 : 8;
  TRY(need_scratch(h, n * 8 + n * es));
  uint64_t* d_ids = (uint64_t*)h->d_scratch;
  void* d_out = (void*)(d_ids + n);
  CU(h, cudaMemcpyAsync(d_ids, ids, n * 8, cudaMemcpyHostToDevice, h->stream));
  if (precision == PENG_F32REF) k_gather_real<float><<<grid_for(n, 256), 256, 0, h->stream>>>(vbase((const float*)p, h->tab_lo), d_ids, n, (float*)d_out);
  else k_gather_real<double><<<grid_for(n, 256), 256, 0, h->stream>>>(vbase((const double*)p, h->tab_lo), d_ids, n, (double*)d_out);
  LAUNCH_CHECK(h);
  CU(h, cudaMemcpyAsync(out, d_out, n * es, cudaMemcpyDeviceToHost, h->stream));
  CU(h, cudaStreamSynchronize(h->stream));
  return PENG_OK;
}

// ---- seed selection: the top partition levels of libstdc++'s introsort on the device (host_ingest.h explains the list
// formulation; tests/test_kernel_logic_sim.py checks it against std::__unguarded_partition_pivot) ----
struct SelState {
  float pivot_z;
  uint32_t nan, total_a, total_b, m, cut;
};
#define SEL_NT 256
#define SEL_PER 8
#define SEL_BLOCK (SEL_NT * SEL_PER)

__global__ void k_sel_init(uint32_t* __restrict__ order, const float* __restrict__ z, uint32_t n, SelState* __restrict__ st) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    order[i] = i;
    const float v = z[i];
    if (v != v) st->nan = 1u;
  }
}
// std::__move_median_to_first(first, first + 1, first + n / 2, last - 1) with comp(a, b) = z[a] > z[b]
__global__ void k_sel_pivot(uint32_t* __restrict__ order, const float* __restrict__ z, uint32_t n_cur, SelState* __restrict__ st) {
  const uint32_t a = 1, b = n_cur / 2, c = n_cur - 1;
  const float za = z[order[a]], zb = z[order[b]], zc = z[order[c]];
  uint32_t sel;
  if (za > zb) sel = zb > zc ? b : (za > zc ? c : a);
  else sel = za > zc ? a : (zb > zc ? c : b);
  const uint32_t t = order[0];
  order[0] = order[sel];
  order[sel] = t;
  st->pivot_z = z[order[0]];
}
// flags of the 8 consecutive positions of a thread: bit j = position stops the left scan (A), bit 8 + j = the right scan (B)
__device__ __forceinline__ uint32_t sel_flags(const uint32_t* __restrict__ order, const float* __restrict__ z, uint32_t n_cur, float zp,
                                              uint32_t i0) {
  uint32_t f = 0;
#pragma unroll
  for (int j = 0; j < SEL_PER; j++) {
    const uint32_t i = i0 + j;
    if (i >= 1 && i < n_cur) {
      const float v = z[order[i]];
      f |= (!(v > zp) ? 1u : 0u) << j;
      f |= (!(zp > v) ? 1u : 0u) << (8 + j);
    }
  }
  return f;
}
__device__ __forceinline__ uint2 sel_block_sum(uint32_t ca, uint32_t cb, uint2* warp_tot) {  // -> exclusive prefix of (ca, cb) within the CTA
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t ia = ca, ib = cb;
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t ya = __shfl_up_sync(0xffffffffu, ia, o), yb = __shfl_up_sync(0xffffffffu, ib, o);
    if (lane >= (uint32_t)o) { ia += ya; ib += yb; }
  }
  if (lane == 31) warp_tot[warp] = make_uint2(ia, ib);
  __syncthreads();
  uint32_t oa = ia - ca, ob = ib - cb;
  for (uint32_t w = 0; w < warp; w++) { oa += warp_tot[w].x; ob += warp_tot[w].y; }
  return make_uint2(oa, ob);
}
__global__ void __launch_bounds__(SEL_NT) k_sel_flags(const uint32_t* __restrict__ order, const float* __restrict__ z, uint32_t n_cur,
                                                      const SelState* __restrict__ st, uint32_t* __restrict__ block_a,
                                                      uint32_t* __restrict__ block_b) {
  __shared__ uint2 warp_tot[SEL_NT / 32];
  const uint32_t f = sel_flags(order, z, n_cur, st->pivot_z, blockIdx.x * SEL_BLOCK + threadIdx.x * SEL_PER);
  const uint2 off = sel_block_sum(__popc(f & 0xFFu), __popc(f >> 8), warp_tot);
  if (threadIdx.x == SEL_NT - 1) {
    block_a[blockIdx.x] = off.x + __popc(f & 0xFFu);
    block_b[blockIdx.x] = off.y + __popc(f >> 8);
  }
}
// exclusive scan of the per-block counts, in place (one CTA; a thread takes a contiguous run of blocks)
__global__ void __launch_bounds__(1024) k_sel_scan(uint32_t* __restrict__ block_a, uint32_t* __restrict__ block_b, uint32_t n_blocks,
                                                   SelState* __restrict__ st) {
  __shared__ uint2 warp_tot[32];
  const uint32_t per = (n_blocks + 1023) / 1024, b0 = threadIdx.x * per, b1 = min(n_blocks, b0 + per);
  uint32_t ca = 0, cb = 0;
  for (uint32_t b = b0; b < b1; b++) { ca += block_a[b]; cb += block_b[b]; }
  const uint2 off = sel_block_sum(ca, cb, warp_tot);
  uint32_t ra = off.x, rb = off.y;
  for (uint32_t b = b0; b < b1; b++) {
    const uint32_t va = block_a[b], vb = block_b[b];
    block_a[b] = ra;
    block_b[b] = rb;
    ra += va;
    rb += vb;
  }
  if (threadIdx.x == 1023) { st->total_a = ra; st->total_b = rb; }
}
__global__ void __launch_bounds__(SEL_NT) k_sel_compact(const uint32_t* __restrict__ order, const float* __restrict__ z, uint32_t n_cur,
                                                        const SelState* __restrict__ st, const uint32_t* __restrict__ block_a,
                                                        const uint32_t* __restrict__ block_b, uint32_t* __restrict__ list_a,
                                                        uint32_t* __restrict__ list_b) {
  __shared__ uint2 warp_tot[SEL_NT / 32];
  const uint32_t i0 = blockIdx.x * SEL_BLOCK + threadIdx.x * SEL_PER;
  const uint32_t f = sel_flags(order, z, n_cur, st->pivot_z, i0);
  const uint2 off = sel_block_sum(__popc(f & 0xFFu), __popc(f >> 8), warp_tot);
  uint32_t ra = block_a[blockIdx.x] + off.x, rb = block_b[blockIdx.x] + off.y;
  const uint32_t last_b = st->total_b - 1u;
#pragma unroll
  for (int j = 0; j < SEL_PER; j++) {
    if ((f >> j) & 1u) list_a[ra++] = i0 + j;                // ascending positions
    if ((f >> (8 + j)) & 1u) list_b[last_b - rb++] = i0 + j;  // descending positions
  }
}
// m = #{k : A[k] < B[k]} (monotone), cut = min(A[m], B[m - 1])
__global__ void k_sel_find(const uint32_t* __restrict__ list_a, const uint32_t* __restrict__ list_b, SelState* __restrict__ st) {
  uint32_t lo = 0, hi = min(st->total_a, st->total_b);
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (list_a[mid] < list_b[mid]) lo = mid + 1; else hi = mid;
  }
  uint32_t cut = 0xFFFFFFFFu;
  if (lo < st->total_a) cut = list_a[lo];
  if (lo >= 1) cut = min(cut, list_b[lo - 1]);
  st->m = lo;
  st->cut = cut;
}
__global__ void k_sel_swap(uint32_t* __restrict__ order, const uint32_t* __restrict__ list_a, const uint32_t* __restrict__ list_b,
                           const SelState* __restrict__ st) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < st->m) {
    const uint32_t i = list_a[k], j = list_b[k];
    const uint32_t t = order[i];
    order[i] = order[j];
    order[j] = t;
  }
}
__global__ void k_sel_pairs(const uint32_t* __restrict__ order, const float* __restrict__ z, uint32_t n, peng_host::ZId* __restrict__ out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const uint32_t id = order[i];
    out[i] = peng_host::ZId{z[id], id};
  }
}

struct DeviceLevelEngine {
  peng_handle_t h;
  uint32_t *order, *list_a, *list_b, *block_a, *block_b;
  SelState* st;
  int rc = PENG_OK;
  bool nan = false;
  int run_level(size_t n_cur, SelState* out) {
    const uint32_t n = (uint32_t)n_cur, nb = (n + SEL_BLOCK - 1) / SEL_BLOCK;
    k_sel_pivot<<<1, 1, 0, h->stream>>>(order, h->d_z32, n, st);
    k_sel_flags<<<nb, SEL_NT, 0, h->stream>>>(order, h->d_z32, n, st, block_a, block_b);
    k_sel_scan<<<1, 1024, 0, h->stream>>>(block_a, block_b, nb, st);
    k_sel_compact<<<nb, SEL_NT, 0, h->stream>>>(order, h->d_z32, n, st, block_a, block_b, list_a, list_b);
    k_sel_find<<<1, 1, 0, h->stream>>>(list_a, list_b, st);
    k_sel_swap<<<grid_for(n / 2 + 1, 256), 256, 0, h->stream>>>(order, list_a, list_b, st);
    h->launches += 5;
    LAUNCH_CHECK(h);
    CU(h, cudaMemcpyAsync(out, st, sizeof(SelState), cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return PENG_OK;
  }
  size_t level(size_t n_cur, float* pivot_z) {
    SelState s{};
    if (rc == PENG_OK) rc = run_level(n_cur, &s);
    if (rc != PENG_OK || s.nan) {  // stop the walk down the spine; the caller looks at rc / nan
      nan = s.nan != 0;
      *pivot_z = 0.f;
      return 0;
    }
    *pivot_z = s.pivot_z;
    return s.cut;
  }
  int run_fetch(size_t n, peng_host::ZId* out) {
    if (n == 0) return PENG_OK;
    TRY(need_scratch(h, n * sizeof(peng_host::ZId)));
    k_sel_pairs<<<grid_for(n, 256), 256, 0, h->stream>>>(order, h->d_z32, (uint32_t)n, (peng_host::ZId*)h->d_scratch);
    LAUNCH_CHECK(h);
    CU(h, cudaMemcpyAsync(out, h->d_scratch, n * sizeof(peng_host::ZId), cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return PENG_OK;
  }
  void fetch(size_t n, peng_host::ZId* out) {
    if (rc == PENG_OK && !nan) rc = run_fetch(n, out);
  }
};

// how far down the left spine the device goes: a level costs ~6 launches + one 24-byte read-back whatever its length, the
// host's pass over a range of n ids ~2.5 ns * n
#define SEL_STOP_BELOW 16384

static int select_host_whole_table(peng_handle_t h, float z_threshold, uint64_t count_threshold, int single_stranded,
                                   int filter_neighbors, std::vector<uint64_t>* sel) {
  std::vector<float> z(h->n_patterns);
  std::vector<uint64_t> counts(h->n_patterns);
  CU(h, cudaMemcpyAsync(z.data(), h->d_z32, h->n_patterns * 4, cudaMemcpyDeviceToHost, h->stream));
  CU(h, cudaMemcpyAsync(counts.data(), h->d_counts, h->n_patterns * 8, cudaMemcpyDeviceToHost, h->stream));
  CU(h, cudaStreamSynchronize(h->stream));
  *sel = peng_host::select_base_patterns(z.data(), counts.data(), h->W, z_threshold, count_threshold, single_stranded != 0,
                                         filter_neighbors != 0);
  return PENG_OK;
}

extern "C" int peng_cuda_select_base_patterns(peng_handle_t h, float z_threshold, uint64_t count_threshold,
                                              int single_stranded, int filter_neighbors, uint64_t* out, uint64_t cap,
                                              uint64_t* n_selected) {
  if (!h || !h->stats_ready || !h->d_z32) return fail(h, PENG_E_INVALID, "select: float32 stats not computed");
  if (!tables_whole(h)) return fail(h, PENG_E_INVALID, "select: needs the whole z table on one device (id-sliced tables are set)");
  CU(h, cudaSetDevice(h->device));
  const uint64_t n = h->n_patterns;
  std::vector<uint64_t> sel;
  bool on_device = false;
#if defined(__GLIBCXX__)
  // the device path indexes positions with 32 bits (W <= 14) and is the pruned replay of libstdc++'s introsort
  on_device = n > 4096 && h->W <= 14 && z_threshold == z_threshold && !h->select_on_host;
#endif
  if (on_device) {
#if defined(__GLIBCXX__)
    const uint32_t nb = (uint32_t)((n + SEL_BLOCK - 1) / SEL_BLOCK);
    const size_t words = 3 * (size_t)n + 2 * (size_t)nb + 16;
    if (h->sel_words < words) {
      TRY(dev_alloc(h, &h->d_sel, words));
      h->sel_words = words;
    }
    DeviceLevelEngine eng;
    eng.h = h;
    eng.order = h->d_sel;
    eng.list_a = eng.order + n;
    eng.list_b = eng.list_a + n;
    eng.block_a = eng.list_b + n;
    eng.block_b = eng.block_a + nb;
    eng.st = reinterpret_cast<SelState*>(eng.block_b + nb);
    CU(h, cudaMemsetAsync(eng.st, 0, sizeof(SelState), h->stream));
    k_sel_init<<<grid_for(n, 256), 256, 0, h->stream>>>(eng.order, h->d_z32, (uint32_t)n, eng.st);
    LAUNCH_CHECK(h);
    std::vector<peng_host::ZId> lead;
    const size_t n_lead = peng_host::sort_leading_by_levels(eng, (size_t)n, z_threshold, SEL_STOP_BELOW, &lead);
    if (h->sel_words > ((size_t)64 << 20)) {  // long pattern lengths: give the 12 bytes per id back
      dev_free(&h->d_sel);
      h->sel_words = 0;
    }
    TRY(eng.rc);
    if (eng.nan) {
      on_device = false;  // a NaN z-score: std::sort's behaviour on it is only reproduced by the plain sort of the whole table
    } else {
      size_t k = 0;
      while (k < n_lead && !(lead[k].z < z_threshold)) k++;
      std::vector<uint64_t> ids(k), cnt(k);
      for (size_t r = 0; r < k; r++) ids[r] = lead[r].id;
      if (k) TRY(peng_cuda_gather_counts(h, ids.data(), k, cnt.data()));
      sel = peng_host::select_walk(lead.data(), cnt.data(), k, h->W, count_threshold, single_stranded != 0, filter_neighbors != 0);
    }
#endif
  }
  if (!on_device) TRY(select_host_whole_table(h, z_threshold, count_threshold, single_stranded, filter_neighbors, &sel));
  if (n_selected) *n_selected = sel.size();
  if (out) std::copy(sel.begin(), sel.begin() + std::min<uint64_t>(cap, sel.size()), out);
  return PENG_OK;
}

// =====================================================================================================
// C ABI: IUPAC
// =====================================================================================================
static int iupac_masks_to_device(peng_handle_t h, const uint64_t* iupac_ids, int n, uint64_t** d_masks, size_t extra_bytes,
                                 unsigned char** extra) {
  std::vector<uint64_t> masks(n);
  for (int i = 0; i < n; i++) {
    if (!peng_host::iupac_letter_mask(iupac_ids[i], h->W, &masks[i]))
      return fail(h, PENG_E_INVALID, "iupac id %llu is not a valid length-%d pattern", (unsigned long long)iupac_ids[i], h->W);
  }
  const size_t mask_bytes = ((size_t)n * 8 + 255) & ~(size_t)255;
  TRY(need_scratch(h, mask_bytes + extra_bytes));
  *d_masks = (uint64_t*)h->d_scratch;
  if (extra) *extra = (unsigned char*)h->d_scratch + mask_bytes;
  CU(h, cudaMemcpyAsync(*d_masks, masks.data(), (size_t)n * 8, cudaMemcpyHostToDevice, h->stream));
  CU(h, cudaStreamSynchronize(h->stream));  // masks is a local
  return PENG_OK;
}

extern "C" int peng_cuda_iupac_count(peng_handle_t h, const uint64_t* iupac_ids, int n, uint64_t* counts_out) {
  TRY(ensure_final(h));
  if (!h || !h->counts_final || !iupac_ids || !counts_out || n < 0) return fail(h, PENG_E_INVALID, "iupac_count: bad call");
  if (n == 0) return PENG_OK;
  CU(h, cudaSetDevice(h->device));
  uint64_t* d_masks;
  unsigned char* extra;
  TRY(iupac_masks_to_device(h, iupac_ids, n, &d_masks, (size_t)n * 8, &extra));
  unsigned long long* d_out = (unsigned long long*)extra;
  if (h->strand == PENG_STRAND_BOTH)
    k_iupac_any<float, true, false><<<n, 256, 0, h->stream>>>(d_masks, h->W, h->d_counts, nullptr, nullptr, d_out, nullptr, nullptr, counts_lo(h), counts_hi(h));
  else
    k_iupac_any<float, false, false><<<n, 256, 0, h->stream>>>(d_masks, h->W, h->d_counts, nullptr, nullptr, d_out, nullptr, nullptr, counts_lo(h), counts_hi(h));
  LAUNCH_CHECK(h);
  CU(h, cudaMemcpyAsync(counts_out, d_out, (size_t)n * 8, cudaMemcpyDeviceToHost, h->stream));
  CU(h, cudaStreamSynchronize(h->stream));
  return PENG_OK;
}

extern "C" int peng_cuda_iupac_aggregate(peng_handle_t h, const uint64_t* iupac_ids, int n, int precision,
                                         uint64_t* n_sites, void* bg_p, void* expected) {
  if (!h || !h->stats_ready || !iupac_ids || n < 0 || !n_sites || !bg_p || !expected)
    return fail(h, PENG_E_INVALID, "iupac_aggregate: bad call (stats must be computed first)");
  if (n == 0) return PENG_OK;
  CU(h, cudaSetDevice(h->device));
  const bool both = h->strand == PENG_STRAND_BOTH;
  const int k = h->k_used;
  uint64_t* d_masks;
  unsigned char* extra;
  if (precision == PENG_F64) {
    if (!h->d_bg64[k]) return fail(h, PENG_E_INVALID, "iupac_aggregate: FP64 tables not computed");
    TRY(iupac_masks_to_device(h, iupac_ids, n, &d_masks, (size_t)n * 24, &extra));
    unsigned long long* d_n = (unsigned long long*)extra;
    double* d_b = (double*)(extra + (size_t)n * 8);
    double* d_e = (double*)(extra + (size_t)n * 16);
    // id-sliced tables: the sums over this rank's ids (the caller adds the ranks' partial sums)
    if (both) k_iupac_any<double, true, true><<<n, 256, 0, h->stream>>>(d_masks, h->W, h->d_counts, vbase(h->d_bg64[k], h->tab_lo), vbase(h->d_exp64, h->tab_lo), d_n, d_b, d_e, h->tab_lo, h->tab_hi);
    else k_iupac_any<double, false, true><<<n, 256, 0, h->stream>>>(d_masks, h->W, h->d_counts, vbase(h->d_bg64[k], h->tab_lo), vbase(h->d_exp64, h->tab_lo), d_n, d_b, d_e, h->tab_lo, h->tab_hi);
    LAUNCH_CHECK(h);
    CU(h, cudaMemcpyAsync(n_sites, d_n, (size_t)n * 8, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaMemcpyAsync(bg_p, d_b, (size_t)n * 8, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaMemcpyAsync(expected, d_e, (size_t)n * 8, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
    return PENG_OK;
  }
  if (!h->d_bg32[k]) return fail(h, PENG_E_INVALID, "iupac_aggregate: float32 tables not computed");
  if (!tables_whole(h)) return fail(h, PENG_E_INVALID, "iupac_aggregate: the float32 arm adds in the reference's order and needs whole tables (id-sliced tables: PENG_F64)");
  const size_t bm_words = h->W >= 3 ? ((size_t)1 << (2 * h->W - 5)) : 1;
  const size_t out_bytes = (size_t)n * 16;
  const size_t bm_bytes = both ? (size_t)n * bm_words * 4 : 0;
  TRY(iupac_masks_to_device(h, iupac_ids, n, &d_masks, out_bytes + 256 + bm_bytes, &extra));
  unsigned long long* d_n = (unsigned long long*)extra;
  float* d_b = (float*)(extra + (size_t)n * 8);
  float* d_e = (float*)(extra + (size_t)n * 12);
  uint32_t* d_bm = (uint32_t*)(extra + ((out_bytes + 255) & ~(size_t)255));
  if (both) {
    k_iupac_any<float, true, false><<<n, 256, 0, h->stream>>>(d_masks, h->W, h->d_counts, nullptr, nullptr, d_n, nullptr, nullptr, 0, h->n_patterns);
    LAUNCH_CHECK(h);
    k_iupac_f32_both<<<n, IUPAC_TILE_WORDS, 0, h->stream>>>(d_masks, h->W, h->d_bg32[k], h->d_exp32, d_bm, d_b, d_e);
    LAUNCH_CHECK(h);
  } else {
    k_iupac_any<float, false, false><<<n, 256, 0, h->stream>>>(d_masks, h->W, h->d_counts, nullptr, nullptr, d_n, nullptr, nullptr, 0, h->n_patterns);
    LAUNCH_CHECK(h);
    k_iupac_f32_plus<<<n, 256, 0, h->stream>>>(d_masks, h->W, h->d_bg32[k], h->d_exp32, d_b, d_e);
    LAUNCH_CHECK(h);
  }
  CU(h, cudaMemcpyAsync(n_sites, d_n, (size_t)n * 8, cudaMemcpyDeviceToHost, h->stream));
  CU(h, cudaMemcpyAsync(bg_p, d_b, (size_t)n * 4, cudaMemcpyDeviceToHost, h->stream));
  CU(h, cudaMemcpyAsync(expected, d_e, (size_t)n * 4, cudaMemcpyDeviceToHost, h->stream));
  CU(h, cudaStreamSynchronize(h->stream));
  return PENG_OK;
}

// =====================================================================================================
// C ABI: EM
// =====================================================================================================
// geometry of k_em_accumulate_all for a group of P PWMs
struct EmGeom {
  int R;
  unsigned grid;
  size_t smem;
};
static EmGeom em_geom(peng_handle_t h, int P, uint32_t n_units) {
  EmGeom g;
  g.R = std::max(1, std::min(EM_MAX_R, EM_NT / P));
  g.smem = (size_t)EM_STAGES * g.R * 16 * sizeof(double2) + (size_t)h->W * 4 * P * 8 + (size_t)EM_ROWS(h->W) * EM_NT * 8;
  const uint32_t per_sm = g.smem + 1024 <= (227u << 10) / 2 ? 2u : 1u;  // two CTAs per SM where their shared memory fits (W <= 14)
  g.grid = (unsigned)std::max<uint32_t>(1, std::min<uint32_t>((uint32_t)h->sm_count * per_sm, (n_units + g.R - 1) / g.R));
  return g;
}
template <int W>
static int launch_em_acc(peng_handle_t h, int first, int P, uint32_t ub, uint32_t ue) {
  auto& E = h->em;
  const EmGeom g = em_geom(h, P, ue - ub);
  auto kern = k_em_accumulate_all<W>;
  CU(h, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.smem));
  const int WA = W * 4;
  kern<<<g.grid, EM_NT, g.smem, h->stream>>>(E.d_pairs, E.d_old + (size_t)first * WA, E.d_active + first, P, g.R, 1.0 / E.sat, ub, ue,
                                           E.d_part + (size_t)first * E.nbx * WA, E.nbx);
  LAUNCH_CHECK(h);
  return PENG_OK;
}

extern "C" int peng_cuda_em_begin(peng_handle_t h, int n_pwms, const double* pwm_in, double saturation, double threshold,
                                  int max_iterations, int bg_order, int part, int n_parts, void** acc_ptr) {
  if (!h || !h->stats_ready || !pwm_in || n_pwms <= 0) return fail(h, PENG_E_INVALID, "em: bad call (stats must be computed first)");
  if (bg_order < 0 || bg_order > h->max_k || !h->d_bg64[bg_order]) return fail(h, PENG_E_INVALID, "em: FP64 bgprob of order %d not available", bg_order);
  if (n_parts < 1 || part < 0 || part >= n_parts) return fail(h, PENG_E_INVALID, "em: bad partition %d/%d", part, n_parts);
  if (!tables_whole(h)) {  // id-sliced tables: the EM slice is the table slice
    if (part != h->slice_part || n_parts != h->slice_parts) return fail(h, PENG_E_INVALID, "em: partition %d/%d is not this rank's table slice %d/%d", part, n_parts, h->slice_part, h->slice_parts);
    if (h->W < 6) return fail(h, PENG_E_INVALID, "em: id-sliced tables need W >= 6");
  }
  CU(h, cudaSetDevice(h->device));
  free_em(h);
  auto& E = h->em;
  const int WA = h->W * 4;
  E.n = n_pwms; E.max_it = max_iterations; E.order = bg_order; E.part = part; E.n_parts = n_parts;
  E.sat = saturation; E.thr = threshold;
  E.nbx = 1;
  if (h->W >= 6) {
    // this rank's slice of the id range, in tiles of 1024 ids (= 4 units)
    const uint32_t n_tiles = (uint32_t)(h->n_patterns >> 10);
    E.ub = 4u * (uint32_t)((uint64_t)n_tiles * part / n_parts);
    E.ue = 4u * (uint32_t)((uint64_t)n_tiles * (part + 1) / n_parts);
    E.nbx = (int)em_geom(h, std::min(n_pwms, EM_MAX_GROUP), E.ue - E.ub).grid;  // the largest grid of any group
    for (int first = 0; first < n_pwms; first += EM_MAX_GROUP)
      E.nbx = std::max(E.nbx, (int)em_geom(h, std::min(EM_MAX_GROUP, n_pwms - first), E.ue - E.ub).grid);
    const uint64_t n_ids = (uint64_t)(E.ue - E.ub) << EM_UNIT_BITS;
    if (E.cap_pairs < n_ids) {
      TRY(dev_alloc(h, &E.d_pairs, (size_t)std::max<uint64_t>(n_ids, 1)));
      E.cap_pairs = (size_t)std::max<uint64_t>(n_ids, 1);
      E.pairs_gen = 0;
    }
    if (n_ids && (E.pairs_gen != h->table_gen || E.pairs_order != bg_order || E.pairs_ub != E.ub || E.pairs_ue != E.ue)) {
      k_em_pairs<<<grid_for(n_ids, 256), 256, 0, h->stream>>>(h->d_counts, vbase(h->d_bg64[bg_order], h->tab_lo), (uint64_t)E.ub << EM_UNIT_BITS, n_ids, E.d_pairs);
      LAUNCH_CHECK(h);
      E.pairs_gen = h->table_gen; E.pairs_order = bg_order; E.pairs_ub = E.ub; E.pairs_ue = E.ue;
    }
  }
  if (E.cap_small < (size_t)n_pwms * WA) {
    const size_t c = (size_t)n_pwms * WA;  // (n_pwms <= c for the per-PWM arrays)
    TRY(dev_alloc(h, &E.d_old, c));
    TRY(dev_alloc(h, &E.d_acc, c));
    TRY(dev_alloc(h, &E.d_change, c));
    TRY(dev_alloc(h, &E.d_iter, c));
    TRY(dev_alloc(h, &E.d_active, c));
    TRY(dev_alloc(h, &E.d_nactive, (size_t)1));
    E.cap_small = c;
  }
  if (E.cap_part < (size_t)n_pwms * WA * E.nbx) {
    TRY(dev_alloc(h, &E.d_part, (size_t)n_pwms * WA * E.nbx));
    E.cap_part = (size_t)n_pwms * WA * E.nbx;
  }
  CU(h, cudaMemcpyAsync(E.d_old, pwm_in, (size_t)n_pwms * WA * 8, cudaMemcpyHostToDevice, h->stream));
  CU(h, cudaMemsetAsync(E.d_acc, 0, (size_t)n_pwms * WA * 8, h->stream));
  CU(h, cudaMemsetAsync(E.d_part, 0, (size_t)n_pwms * WA * E.nbx * 8, h->stream));  // groups with a smaller grid leave rows unwritten
  k_em_init<<<grid_for(n_pwms, 64), 64, 0, h->stream>>>(E.d_change, E.d_iter, E.d_active, n_pwms, h->W, threshold, max_iterations);
  LAUNCH_CHECK(h);
  CU(h, cudaStreamSynchronize(h->stream));
  E.open = true;
  if (acc_ptr) *acc_ptr = E.d_acc;
  return PENG_OK;
}

// the per-CTA partial rows of one iteration (everything of peng_cuda_em_accumulate but the final reduction)
static int em_accumulate_partials(peng_handle_t h) {
  auto& E = h->em;
  if (h->W >= 6) {
    for (int first = 0; first < E.n; first += EM_MAX_GROUP) {
      const int P = std::min(EM_MAX_GROUP, E.n - first);
      int rc = PENG_OK;
      switch (h->W) {
        case 6: rc = launch_em_acc<6>(h, first, P, E.ub, E.ue); break;
        case 8: rc = launch_em_acc<8>(h, first, P, E.ub, E.ue); break;
        case 10: rc = launch_em_acc<10>(h, first, P, E.ub, E.ue); break;
        case 12: rc = launch_em_acc<12>(h, first, P, E.ub, E.ue); break;
        case 14: rc = launch_em_acc<14>(h, first, P, E.ub, E.ue); break;
        case 16: rc = launch_em_acc<16>(h, first, P, E.ub, E.ue); break;
      }
      TRY(rc);
    }
  } else {
    const uint32_t n = (uint32_t)h->n_patterns;
    const uint32_t ib = (uint32_t)((uint64_t)n * E.part / E.n_parts), ie = (uint32_t)((uint64_t)n * (E.part + 1) / E.n_parts);
    k_em_accumulate_small<<<E.n, 64, 0, h->stream>>>(h->d_counts, h->d_bg64[E.order], E.d_old, E.d_active, h->W, E.sat, ib, ie, E.d_part);
    LAUNCH_CHECK(h);
  }
  return PENG_OK;
}

extern "C" int peng_cuda_em_accumulate(peng_handle_t h) {
  if (!h || !h->em.open) return fail(h, PENG_E_INVALID, "em_accumulate: no EM session");
  CU(h, cudaSetDevice(h->device));
  auto& E = h->em;
  TRY(em_accumulate_partials(h));
  k_em_reduce<<<E.n, PENG_MAX_W * 4 * 16, 0, h->stream>>>(E.d_part, E.d_active, E.nbx, h->W * 4, E.d_acc);
  LAUNCH_CHECK(h);
  return PENG_OK;
}

extern "C" int peng_cuda_em_update(peng_handle_t h, int* n_active) {
  if (!h || !h->em.open) return fail(h, PENG_E_INVALID, "em_update: no EM session");
  CU(h, cudaSetDevice(h->device));
  auto& E = h->em;
  CU(h, cudaMemsetAsync(E.d_nactive, 0, 4, h->stream));
  k_em_update_f64<<<E.n, 32, 0, h->stream>>>(E.d_acc, E.d_old, E.d_change, E.d_iter, E.d_active, h->W, E.thr, E.max_it, E.d_nactive);
  LAUNCH_CHECK(h);
  if (n_active) {
    CU(h, cudaMemcpyAsync(n_active, E.d_nactive, 4, cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
  }
  return PENG_OK;
}

extern "C" int peng_cuda_em_end(peng_handle_t h, double* pwm_out, int* iterations_out) {
  if (!h || !h->em.open) return fail(h, PENG_E_INVALID, "em_end: no EM session");
  CU(h, cudaSetDevice(h->device));
  auto& E = h->em;
  const int WA = h->W * 4;
  k_em_final_normalize<double><<<grid_for((uint64_t)E.n * h->W, 64), 64, 0, h->stream>>>(E.d_old, E.n, h->W);
  LAUNCH_CHECK(h);
  if (pwm_out) CU(h, cudaMemcpyAsync(pwm_out, E.d_old, (size_t)E.n * WA * 8, cudaMemcpyDeviceToHost, h->stream));
  if (iterations_out) CU(h, cudaMemcpyAsync(iterations_out, E.d_iter, (size_t)E.n * 4, cudaMemcpyDeviceToHost, h->stream));
  CU(h, cudaStreamSynchronize(h->stream));
  free_em(h);
  return PENG_OK;
}

static int em_f32ref(peng_handle_t h, int n_pwms, const float* pwm_in, float saturation, float threshold, int max_iterations,
                     int bg_order, float* pwm_out, int* iterations_out) {
  if (!h->d_bg32[bg_order]) return fail(h, PENG_E_INVALID, "em: float32 bgprob of order %d not available", bg_order);
  if (!tables_whole(h)) return fail(h, PENG_E_INVALID, "em: the float32 arm adds in the reference's order and needs whole tables (id-sliced tables: PENG_F64)");
  const int W = h->W, WA = W * 4;
  const uint64_t n = h->n_patterns;
  // terms buffer: n floats per PWM; batch the PWMs if that would exceed ~8 GB
  const int batch = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)n_pwms, (8ull << 30) / (n * 4)));
  float *d_old = nullptr, *d_acc = nullptr, *d_terms = nullptr;
  int *d_iter = nullptr, *d_active = nullptr, *d_nact = nullptr;
  int rc = PENG_OK;
  auto cleanup = [&]() { dev_free(&d_old); dev_free(&d_acc); dev_free(&d_terms); dev_free(&d_iter); dev_free(&d_active); dev_free(&d_nact); };
#define EM_TRY(x) do { rc = (x); if (rc != PENG_OK) { cleanup(); return rc; } } while (0)
#define EM_CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { cleanup(); return fail(h, PENG_E_CUDA, "%s failed: %s", #call, cudaGetErrorString(e_)); } } while (0)
  EM_TRY(dev_alloc(h, &d_old, (size_t)batch * WA));
  EM_TRY(dev_alloc(h, &d_acc, (size_t)batch * WA));
  EM_TRY(dev_alloc(h, &d_terms, (size_t)batch * n));
  EM_TRY(dev_alloc(h, &d_iter, (size_t)batch));
  EM_TRY(dev_alloc(h, &d_active, (size_t)batch));
  EM_TRY(dev_alloc(h, &d_nact, (size_t)1));
  for (int first = 0; first < n_pwms; first += batch) {
    const int nb = std::min(batch, n_pwms - first);
    EM_CU(cudaMemcpyAsync(d_old, pwm_in + (size_t)first * WA, (size_t)nb * WA * 4, cudaMemcpyHostToDevice, h->stream));
    k_em_init_f32<<<grid_for(nb, 64), 64, 0, h->stream>>>(d_iter, d_active, nb, W, threshold, max_iterations);
    h->launches++;
    for (int it = 0; it < max_iterations; it++) {
      EM_CU(cudaMemsetAsync(d_nact, 0, 4, h->stream));
      k_em_terms_f32<<<dim3(grid_for(n, 256), nb), 256, 0, h->stream>>>(h->d_counts, h->d_bg32[bg_order], d_old, d_active, W, saturation, n, d_terms);
      k_em_chains_f32<<<dim3(W, nb), 128, 0, h->stream>>>(d_terms, d_active, W, n, d_acc);
      k_em_update_f32<<<nb, 32, 0, h->stream>>>(d_acc, d_old, d_iter, d_active, W, threshold, max_iterations, d_nact);
      h->launches += 3;
      EM_CU(cudaGetLastError());
      if ((it & 3) == 3) {  // look at the convergence flags every few iterations only
        int na = 0;
        EM_CU(cudaMemcpyAsync(&na, d_nact, 4, cudaMemcpyDeviceToHost, h->stream));
        EM_CU(cudaStreamSynchronize(h->stream));
        if (na == 0) break;
      }
    }
    k_em_final_normalize<float><<<grid_for((uint64_t)nb * W, 64), 64, 0, h->stream>>>(d_old, nb, W);
    h->launches++;
    EM_CU(cudaMemcpyAsync(pwm_out + (size_t)first * WA, d_old, (size_t)nb * WA * 4, cudaMemcpyDeviceToHost, h->stream));
    if (iterations_out) EM_CU(cudaMemcpyAsync(iterations_out + first, d_iter, (size_t)nb * 4, cudaMemcpyDeviceToHost, h->stream));
    EM_CU(cudaStreamSynchronize(h->stream));
  }
  cleanup();
  return PENG_OK;
#undef EM_TRY
#undef EM_CU
}

extern "C" int peng_cuda_em(peng_handle_t h, int n_pwms, const void* pwm_in, int precision, double saturation, double threshold,
                            int max_iterations, int bg_order, void* pwm_out, int* iterations_out) {
  if (!h || !h->stats_ready || !pwm_in || !pwm_out || n_pwms < 0) return fail(h, PENG_E_INVALID, "em: bad call (stats must be computed first)");
  if (n_pwms == 0) return PENG_OK;
  if (bg_order < 0 || bg_order > h->max_k) return fail(h, PENG_E_INVALID, "em: bgprob order %d not available", bg_order);
  CU(h, cudaSetDevice(h->device));
  if (precision == PENG_F32REF)
    return em_f32ref(h, n_pwms, (const float*)pwm_in, (float)saturation, (float)threshold, max_iterations, bg_order,
                     (float*)pwm_out, iterations_out);
  if (!tables_whole(h))
    return fail(h, PENG_E_INVALID, "em: id-sliced tables — run peng_cuda_em_begin / _em_accumulate / (sum over the ranks) / _em_update on every rank");
  TRY(peng_cuda_em_begin(h, n_pwms, (const double*)pwm_in, saturation, threshold, max_iterations, bg_order, 0, 1, nullptr));
  auto& E = h->em;
  for (int it = 0; it < max_iterations; it++) {
    TRY(em_accumulate_partials(h));
    CU(h, cudaMemsetAsync(E.d_nactive, 0, 4, h->stream));
    k_em_reduce_update<<<E.n, PENG_MAX_W * 4 * EM_RED_PARTS, 0, h->stream>>>(E.d_part, E.nbx, E.d_acc, E.d_old, E.d_change, E.d_iter, E.d_active, h->W, E.thr,
                                                  E.max_it, E.d_nactive);
    LAUNCH_CHECK(h);
    if ((it & 3) == 3) {  // look at the convergence flags every 4th iteration only (converged PWMs cost nothing meanwhile)
      int na = -1;
      CU(h, cudaMemcpyAsync(&na, E.d_nactive, 4, cudaMemcpyDeviceToHost, h->stream));
      CU(h, cudaStreamSynchronize(h->stream));
      if (na == 0) break;
    }
  }
  return peng_cuda_em_end(h, (double*)pwm_out, iterations_out);
}
