// host_pack.cpp — see host_pack.h.  Compiled by g++ (-O3), linked into libpeng_cuda.so.
#include "host_pack.h"

#include <immintrin.h>

#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdlib>
#include <mutex>
#include <thread>

#include <unistd.h>

namespace peng_hostpack {

bool have_avx512() {
  static const bool ok = __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw");
  return ok;
}

// ---- up to 64 letters -> 128 code bits + 64 validity bits (bits of positions >= n are zero) --------------------------
// codes: 0 = other, 1..4 = ACGT (Alphabet::getCode, src/shared/Alphabet.cpp:33-41); ascii: FASTA letters, either case.
static inline void chunk_scalar(const uint8_t* p, uint32_t n, bool ascii, uint64_t& lo, uint64_t& hi, uint64_t& v) {
  uint64_t c[2] = {0, 0}, vv = 0;
  for (uint32_t i = 0; i < n; i++) {
    const uint32_t ch = p[i];
    uint32_t ok, code;
    if (ascii) {
      const uint32_t up = ch & 0xDFu;
      ok = (up == 'A') | (up == 'C') | (up == 'G') | (up == 'T');
      code = (up >> 1) & 3u;
      code ^= code >> 1;
    } else {
      ok = (ch >= 1u) & (ch <= 4u);
      code = (ch - 1u) & 3u;
    }
    code = ok ? code : 0u;
    c[i >> 5] |= (uint64_t)code << (2 * (i & 31));
    vv |= (uint64_t)ok << i;
  }
  lo = c[0];
  hi = c[1];
  v = vv;
}

// ---- sequential bit writer over 64-position blocks (scalar fallback) --------------------------------------------------------------------
struct BitWriter {
  uint64_t* codes;
  uint64_t* valid;
  uint64_t blk0, blk;
  uint32_t fill;  // positions of the current block already written
  uint64_t lo, hi, v;
  bool first_partial;  // the current block was entered in the middle (range boundary)
  std::vector<Partial>* out;

  void begin(uint64_t q) {
    blk = q >> 6;
    fill = (uint32_t)(q & 63);
    lo = hi = v = 0;
    first_partial = fill != 0;
  }
  uint64_t scratch[3];  // where the stores of a block entered in the middle go (it leaves as a Partial)
  // n in 1..64 positions; the bits of c_lo / c_hi / vv behind position n are zero.  Branch-free: the alignment changes
  // with every sequence, so neither tests on `fill` nor the "block complete" test would predict.  The current block is
  // stored after every append (same cache line again and again); completing it only moves the pointers on.
  inline void append(uint64_t c_lo, uint64_t c_hi, uint64_t vv, uint32_t n) {
    typedef unsigned __int128 u128;
    const u128 c = ((u128)c_hi << 64) | c_lo;
    const uint32_t sh = 2 * fill;  // 0 .. 126
    const u128 in = c << sh;
    const u128 over = (c >> 1) >> (127 - sh);  // what fell out at the top (0 for sh = 0)
    lo |= (uint64_t)in;
    hi |= (uint64_t)(in >> 64);
    v |= vv << fill;
    const uint64_t over_v = (vv >> 1) >> (63 - fill);
    const uint32_t tot = fill + n;
    const bool adv = tot >= 64;
    uint64_t* cp = first_partial ? scratch : codes + 2 * (blk - blk0);
    uint64_t* vp = first_partial ? scratch + 2 : valid + (blk - blk0);
    cp[0] = lo;
    cp[1] = hi;
    vp[0] = v;
    if (__builtin_expect(adv && first_partial, 0)) {
      out->push_back(Partial{blk, lo, hi, v});
      first_partial = false;
    }
    blk += adv;
    const uint64_t m = (uint64_t)0 - (uint64_t)adv;  // all ones: the block is complete, carry on with what fell out
    lo = ((uint64_t)over & m) | (lo & ~m);
    hi = ((uint64_t)(over >> 64) & m) | (hi & ~m);
    v = (over_v & m) | (v & ~m);
    fill = tot & 63u;
  }
  void zeros(uint64_t n) {
    while (n) {
      const uint32_t k = (uint32_t)std::min<uint64_t>(n, 64u - fill);
      append(0, 0, 0, k);
      n -= k;
    }
  }
  void end() {
    if (fill) out->push_back(Partial{blk, lo, hi, v});
  }
};

// ---- AVX-512 body: driven by the OUTPUT blocks ------------------------------------------------------------------------
// Block b holds the stream positions [64 b, 64 b + 64).  Its letters are gathered with one unaligned load where the block
// lies inside a sequence, else with one masked load per sequence that overlaps it (masked-off lanes are neither read nor
// can they fault; separator and padding positions stay 0 = "other"), converted once and stored where they belong: no bit
// writer, no shifts that depend on the alignment of a sequence.  Blocks the share covers only in part leave as Partials.
// Control flow follows the sequences (head in the open block, whole blocks inside, tail opens the next block), so the
// branches depend on a sequence's length class, not on its alignment.
template <bool ascii>
__attribute__((target("avx512f,avx512bw"))) static inline void block_convert(const __m512i z, __m128i& codes, __mmask64& ok) {
  __m512i code;
  if (ascii) {
    const __m512i up = _mm512_and_si512(z, _mm512_set1_epi8((char)0xDF));
    ok = _mm512_cmpeq_epi8_mask(up, _mm512_set1_epi8('A')) | _mm512_cmpeq_epi8_mask(up, _mm512_set1_epi8('C')) |
         _mm512_cmpeq_epi8_mask(up, _mm512_set1_epi8('G')) | _mm512_cmpeq_epi8_mask(up, _mm512_set1_epi8('T'));
    __m512i c = _mm512_and_si512(_mm512_srli_epi16(up, 1), _mm512_set1_epi8(3));                // A0 C1 G3 T2
    c = _mm512_xor_si512(c, _mm512_and_si512(_mm512_srli_epi16(c, 1), _mm512_set1_epi8(1)));  // A0 C1 G2 T3
    code = _mm512_maskz_mov_epi8(ok, c);
  } else {
    const __m512i t = _mm512_sub_epi8(z, _mm512_set1_epi8(1));
    ok = _mm512_cmplt_epu8_mask(t, _mm512_set1_epi8(4));
    code = _mm512_maskz_mov_epi8(ok, t);  // t is 0..3 wherever ok
  }
  const __m512i m1 = _mm512_maddubs_epi16(code, _mm512_set1_epi16(0x0401));
  const __m512i m2 = _mm512_madd_epi16(m1, _mm512_set1_epi32(0x00100001));
  codes = _mm512_cvtepi32_epi8(m2);
}

// a block that may be one of the two the share covers only in part
struct EmitCtx {
  uint64_t q_lo, q_hi;
  uint64_t *codes, *valid;  // indexed by the absolute block number
  std::vector<Partial>* partials;
  std::vector<ValidExc>* exc;
};
// zm: the lanes of z that hold a letter of a sequence (the implied validity of the block)
template <bool ascii>
__attribute__((target("avx512f,avx512bw,bmi2"), always_inline)) static inline void emit_block(const EmitCtx& cx, const __m512i z, const uint64_t zm,
                                                                                              const uint64_t Q) {
  __m128i c;
  __mmask64 ok;
  block_convert<ascii>(z, c, ok);
  if (__builtin_expect(Q >= cx.q_lo && Q + 64 <= cx.q_hi, 1)) {
    _mm_storeu_si128((__m128i*)(cx.codes + (Q >> 5)), c);
    cx.valid[Q >> 6] = (uint64_t)ok;
    if (__builtin_expect((uint64_t)ok != zm, 0) && cx.exc) cx.exc->push_back(ValidExc{Q >> 6, (uint64_t)ok});
  } else {
    cx.partials->push_back(Partial{Q >> 6, (uint64_t)_mm_cvtsi128_si64(c), (uint64_t)_mm_extract_epi64(c, 1), (uint64_t)ok});
  }
}

template <bool ascii>
__attribute__((target("avx512f,avx512bw,bmi2"))) static void run_blocks_avx512(const Share& sh, const uint64_t q_lo, const uint64_t q_hi) {
  if (q_hi <= q_lo) return;
  const uint8_t* const src = sh.src;
  const uint8_t* const* const seq_ptrs = sh.seq_ptrs;
  const uint64_t* const offsets = sh.offsets;
  const uint64_t s_end = sh.s_end, front_pad = sh.front_pad;
  uint64_t* const codes = sh.codes64 - 2 * sh.blk0;
  uint64_t* const valid = sh.valid64 - sh.blk0;
  const EmitCtx cx{q_lo, q_hi, codes, valid, sh.partials, sh.exc};
  // the OPEN block [Q, Q + 64): z holds the letters gathered so far, everything else is 0
  uint64_t Q = q_lo & ~(uint64_t)63;
  __m512i z = _mm512_setzero_si512();
  uint64_t zm = 0;
  uint64_t o = offsets[sh.s_begin];
  for (uint64_t s = sh.s_begin; s < s_end; s++) {
    const uint64_t o1 = offsets[s + 1];
    const uint64_t qs = front_pad + o + 2 * s, qe = qs + (o1 - o);  // stream positions of the sequence
    const uintptr_t base = (uintptr_t)(seq_ptrs ? seq_ptrs[s] : src + o) - (uintptr_t)qs;  // letter of position q at base + q
    o = o1;
    while (__builtin_expect(qs >= Q + 64, 0)) {  // front pad of the set
      emit_block<ascii>(cx, z, zm, Q);
      z = _mm512_setzero_si512();
      zm = 0;
      Q += 64;
    }
    // head: what of the sequence lies in the open block
    const uint64_t e = std::min(qe, Q + 64);
    if (e > qs) {
      const uint64_t m = _bzhi_u64(~(uint64_t)0 << (qs - Q), (uint32_t)(e - Q));
      z = _mm512_mask_loadu_epi8(z, (__mmask64)m, (const void*)(base + Q));
      zm |= m;
    }
    if (qe >= Q + 64) {  // the sequence reaches the end of the open block
      emit_block<ascii>(cx, z, zm, Q);
      Q += 64;
      for (const uint64_t Q_int = Q + ((qe - Q) & ~(uint64_t)63); Q < Q_int; Q += 64) {  // blocks inside the sequence: always whole blocks of the share
        __m128i c;
        __mmask64 ok;
        block_convert<ascii>(_mm512_loadu_si512((const void*)(base + Q)), c, ok);
        _mm_storeu_si128((__m128i*)(codes + (Q >> 5)), c);  // (streaming stores measured: 25 x slower — two 16 / 8-byte streams per thread never fill a write-combining line)
        valid[Q >> 6] = (uint64_t)ok;
        if (__builtin_expect((uint64_t)ok != ~(uint64_t)0, 0) && cx.exc) cx.exc->push_back(ValidExc{Q >> 6, (uint64_t)ok});
      }
      zm = _bzhi_u64(~(uint64_t)0, (uint32_t)(qe - Q));
      z = _mm512_maskz_loadu_epi8((__mmask64)zm, (const void*)(base + Q));  // tail: 0 .. 63 letters
    }
    if (__builtin_expect(qe + 2 >= Q + 64, 0)) {  // the two separator positions end the open block
      emit_block<ascii>(cx, z, zm, Q);
      z = _mm512_setzero_si512();
      zm = 0;
      Q += 64;
    }
  }
  for (; Q < q_hi; Q += 64) {  // the open block and the padding behind the last sequence
    emit_block<ascii>(cx, z, zm, Q);
    z = _mm512_setzero_si512();
    zm = 0;
  }
}
static void run_scalar(const Share& sh, BitWriter& w) {
  for (uint64_t s = sh.s_begin; s < sh.s_end; s++) {
    const uint8_t* p = sh.seq_ptrs ? sh.seq_ptrs[s] : sh.src + sh.offsets[s];
    const uint64_t L = sh.offsets[s + 1] - sh.offsets[s];
    uint32_t sep = 2;  // the two separator positions (core_seq.cuh) ride on the last chunk where they fit
    for (uint64_t i = 0; i < L; i += 64) {
      uint32_t n = (uint32_t)std::min<uint64_t>(64, L - i);
      uint64_t lo, hi, v;
      chunk_scalar(p + i, n, sh.ascii, lo, hi, v);
      if (i + 64 >= L && n <= 62) { n += 2; sep = 0; }
      w.append(lo, hi, v, n);
    }
    if (sep) w.append(0, 0, 0, 2);
  }
}

bool lists_exceptions() {
  static const bool no_avx = getenv("PENG_HOST_PACK_SCALAR") != nullptr;
  return have_avx512() && !no_avx;
}

void pack_share(const Share& sh) {
  if (lists_exceptions()) {
    const uint64_t q_lo = sh.lead_zero ? 0 : sh.front_pad + sh.offsets[sh.s_begin] + 2 * sh.s_begin;
    const uint64_t q_seq = sh.front_pad + sh.offsets[sh.s_end] + 2 * sh.s_end;  // behind the separators of the last sequence
    const uint64_t q_hi = std::max(q_seq, sh.q_end_pad);
    if (sh.ascii) run_blocks_avx512<true>(sh, q_lo, q_hi); else run_blocks_avx512<false>(sh, q_lo, q_hi);
    return;
  }
  BitWriter w;
  w.codes = sh.codes64;
  w.valid = sh.valid64;
  w.blk0 = sh.blk0;
  w.out = sh.partials;
  if (sh.lead_zero) {
    w.begin(0);
    w.zeros(sh.front_pad);
  } else {
    w.begin(sh.front_pad + sh.offsets[sh.s_begin] + 2 * sh.s_begin);
  }
  run_scalar(sh, w);
  if (sh.q_end_pad) {
    const uint64_t q = (w.blk << 6) + w.fill;
    if (sh.q_end_pad > q) w.zeros(sh.q_end_pad - q);
  }
  w.end();
}

// ---- worker pool ------------------------------------------------------------------------------------------------------
struct Pool::Impl {
  std::vector<std::thread> workers;
  std::mutex m;
  std::mutex job;  // held from start() to finish(): callers on several threads (one per device of a multi-GPU run) take turns
  std::condition_variable cv_work, cv_done;
  uint64_t generation = 0;
  bool stop = false;
  void (*fn)(int, void*) = nullptr;
  void* arg = nullptr;
  int n_tasks = 0;
  std::atomic<int> next{0};
  int busy = 0;  // workers still inside the current generation

  void drain() {
    for (;;) {
      const int i = next.fetch_add(1, std::memory_order_relaxed);
      if (i >= n_tasks) break;
      fn(i, arg);
    }
  }
  void worker() {
    uint64_t seen = 0;
    for (;;) {
      {
        std::unique_lock<std::mutex> lk(m);
        cv_work.wait(lk, [&] { return stop || generation != seen; });
        if (stop) return;
        seen = generation;
      }
      drain();
      {
        std::lock_guard<std::mutex> lk(m);
        if (--busy == 0) cv_done.notify_one();
      }
    }
  }
};

Pool::Pool(int n_threads) : p_(new Impl), n_(std::max(1, n_threads)) {
  for (int i = 1; i < n_; i++) p_->workers.emplace_back([this] { p_->worker(); });
}
Pool::~Pool() {
  {
    std::lock_guard<std::mutex> lk(p_->m);
    p_->stop = true;
  }
  p_->cv_work.notify_all();
  for (auto& t : p_->workers) t.join();
  delete p_;
}
void Pool::start(int n_tasks, void (*fn)(int, void*), void* arg) {
  p_->job.lock();  // released by finish(), on the same thread
  {
    std::lock_guard<std::mutex> lk(p_->m);
    p_->fn = fn;
    p_->arg = arg;
    p_->n_tasks = std::max(0, n_tasks);
    p_->next.store(0, std::memory_order_relaxed);
    p_->busy = (int)p_->workers.size();
    p_->generation++;
  }
  p_->cv_work.notify_all();
}
void Pool::finish() {
  p_->drain();  // the caller works too
  {
    std::unique_lock<std::mutex> lk(p_->m);
    p_->cv_done.wait(lk, [&] { return p_->busy == 0; });
  }
  p_->job.unlock();
}
void Pool::run(int n_tasks, void (*fn)(int, void*), void* arg) {
  if (n_tasks <= 0) return;
  start(n_tasks, fn, arg);
  finish();
}

Pool& global_pool() {
  static Pool* pool = nullptr;
  static pid_t owner = 0;
  static std::mutex m;
  std::lock_guard<std::mutex> lk(m);
  if (!pool || owner != getpid()) {  // a forked child has none of the parent's worker threads: it gets its own pool
    int n = 0;
    if (const char* e = getenv("PENG_HOST_THREADS")) n = atoi(e);
    if (n <= 0) {
      n = (int)std::thread::hardware_concurrency();
      if (const char* e = getenv("LOCAL_WORLD_SIZE")) {  // one process per GPU (torchrun): share the cores
        const int w = atoi(e);
        if (w > 1) n = std::max(1, n / w);
      }
    }
    pool = new Pool(std::min(std::max(n, 1), 32));  // never destroyed: worker threads must not be joined at exit
    owner = getpid();
  }
  return *pool;
}

// ---- one phase --------------------------------------------------------------------------------------------------------
void phase_blocks(const Phase& ph, uint64_t* blk0, uint64_t* blk1) {
  const uint64_t q0 = ph.first ? 0 : ph.front_pad + ph.offsets[ph.s_begin] + 2 * ph.s_begin;
  const uint64_t q1 = ph.last ? ph.q_total : ph.front_pad + ph.offsets[ph.s_end] + 2 * ph.s_end;
  *blk0 = q0 >> 6;
  *blk1 = q1 >> 6;
}

uint64_t plan_phase(const uint64_t* offsets, uint64_t n_seq, uint64_t s_begin, uint32_t front_pad, bool first, uint64_t max_blocks) {
  const uint64_t q0 = first ? 0 : front_pad + offsets[s_begin] + 2 * s_begin;
  const uint64_t blk0 = q0 >> 6;
  uint64_t lo = s_begin, hi = n_seq;  // invariant: lo fits
  while (lo < hi) {
    const uint64_t mid = (lo + hi + 1) >> 1;
    const uint64_t q1 = front_pad + offsets[mid] + 2 * mid;
    if ((q1 >> 6) + 1 - blk0 <= max_blocks) lo = mid; else hi = mid - 1;
  }
  return lo;
}

struct PhaseJob {
  const Phase* ph;
  uint64_t blk0, blk1;
  bool want_exc;
  std::vector<uint64_t> cut;  // share i = sequences [cut[i], cut[i + 1])
  std::vector<std::vector<Partial>> partials;
  std::vector<std::vector<ValidExc>> exc;
};
namespace {
void phase_task(int i, void* arg) {
  PhaseJob& r = *(PhaseJob*)arg;
  const Phase& ph = *r.ph;
  const int n = (int)r.cut.size() - 1;
  Share sh;
  sh.src = ph.src;
  sh.seq_ptrs = ph.seq_ptrs;
  sh.offsets = ph.offsets;
  sh.s_begin = r.cut[i];
  sh.s_end = r.cut[i + 1];
  sh.ascii = ph.ascii;
  sh.front_pad = ph.front_pad;
  sh.lead_zero = ph.first && i == 0;
  sh.q_end_pad = (ph.last && i == n - 1) ? ph.q_total : 0;
  sh.blk0 = r.blk0;
  sh.codes64 = ph.codes64;
  sh.valid64 = ph.valid64;
  sh.partials = &r.partials[i];
  sh.exc = r.exc.empty() ? nullptr : &r.exc[i];
  if (sh.s_begin == sh.s_end && !sh.lead_zero && !sh.q_end_pad) return;
  pack_share(sh);
}
}  // namespace

PhaseJob* pack_phase_begin(Pool& pool, const Phase& ph) {
  PhaseJob* job = new PhaseJob;
  PhaseJob& r = *job;
  phase_blocks(ph, &r.blk0, &r.blk1);
  r.ph = &ph;
  // shares of about equal numbers of bases, cut at sequence boundaries
  const uint64_t b0 = ph.offsets[ph.s_begin], b1 = ph.offsets[ph.s_end];
  // four shares per thread, handed out dynamically: a thread that loses its core for a moment does not hold up the phase
  int n = pool.size() * 4;
  if (b1 - b0 < (uint64_t)n * 65536) n = (int)std::max<uint64_t>(1, (b1 - b0) / 65536);
  r.cut.push_back(ph.s_begin);
  for (int i = 1; i < n; i++) {
    const uint64_t target = b0 + (b1 - b0) * (uint64_t)i / (uint64_t)n;
    const uint64_t* it = std::lower_bound(ph.offsets + r.cut.back(), ph.offsets + ph.s_end, target);
    r.cut.push_back((uint64_t)(it - ph.offsets));
  }
  r.cut.push_back(ph.s_end);
  r.partials.resize(n);
  r.want_exc = ph.exc && lists_exceptions();
  if (ph.exc) ph.exc->clear();
  if (ph.exc_complete) *ph.exc_complete = r.want_exc;
  if (r.want_exc) r.exc.resize(n);
  pool.start(n, phase_task, job);
  return job;
}

void pack_phase_end(Pool& pool, PhaseJob* job, bool* has_carry, Partial* carry) {
  pool.finish();
  PhaseJob& r = *job;
  const Phase& ph = *r.ph;
  const uint64_t blk0 = r.blk0, blk1 = r.blk1;
  const bool want_exc = r.want_exc;
  // boundary blocks: OR the contributions (at most two per share, plus the carry of the previous phase)
  std::vector<Partial> all;
  if (*has_carry) all.push_back(*carry);
  for (auto& v : r.partials) all.insert(all.end(), v.begin(), v.end());
  std::sort(all.begin(), all.end(), [](const Partial& a, const Partial& b) { return a.blk < b.blk; });
  *has_carry = false;
  for (size_t i = 0; i < all.size();) {
    Partial acc = all[i];
    size_t j = i + 1;
    for (; j < all.size() && all[j].blk == acc.blk; j++) {
      acc.lo |= all[j].lo;
      acc.hi |= all[j].hi;
      acc.v |= all[j].v;
    }
    if (acc.blk < blk1) {
      ph.codes64[2 * (acc.blk - blk0)] = acc.lo;
      ph.codes64[2 * (acc.blk - blk0) + 1] = acc.hi;
      ph.valid64[acc.blk - blk0] = acc.v;
      if (want_exc) ph.exc->push_back(ValidExc{acc.blk, acc.v});
    } else {
      *carry = acc;
      *has_carry = true;
    }
    i = j;
  }
  if (want_exc) {
    for (auto& v : r.exc) ph.exc->insert(ph.exc->end(), v.begin(), v.end());
    std::sort(ph.exc->begin(), ph.exc->end(), [](const ValidExc& a, const ValidExc& b) { return a.blk < b.blk; });
  }
  delete job;
}

void pack_phase(Pool& pool, const Phase& ph, bool* has_carry, Partial* carry) {
  pack_phase_end(pool, pack_phase_begin(pool, ph), has_carry, carry);
}

// ---- nibble packing ---------------------------------------------------------------------------------------------------
namespace {
__attribute__((target("avx512f,avx512bw"))) void nibble_avx512(const uint8_t* src, uint64_t n, uint8_t* dst) {
  uint64_t i = 0;
  const __m512i four = _mm512_set1_epi8(4), w = _mm512_set1_epi16(0x1001);  // even byte * 1 + odd byte * 16
  for (; i + 64 <= n; i += 64) {
    const __m512i z = _mm512_loadu_si512((const void*)(src + i));
    const __m512i c = _mm512_maskz_mov_epi8(_mm512_cmple_epu8_mask(z, four), z);
    _mm256_storeu_si256((__m256i*)(dst + (i >> 1)), _mm512_cvtepi16_epi8(_mm512_maddubs_epi16(c, w)));
  }
  for (; i < n; i += 2) {
    const uint32_t a = src[i] <= 4 ? src[i] : 0u, b = i + 1 < n && src[i + 1] <= 4 ? src[i + 1] : 0u;
    dst[i >> 1] = (uint8_t)(a | (b << 4));
  }
}
void nibble_scalar(const uint8_t* src, uint64_t n, uint8_t* dst) {
  for (uint64_t i = 0; i < n; i += 2) {
    const uint32_t a = src[i] <= 4 ? src[i] : 0u, b = i + 1 < n && src[i + 1] <= 4 ? src[i + 1] : 0u;
    dst[i >> 1] = (uint8_t)(a | (b << 4));
  }
}
struct NibbleRun {
  const uint8_t* src;
  uint8_t* dst;
  uint64_t n;
  int n_tasks;
};
void nibble_task(int i, void* arg) {
  const NibbleRun& r = *(const NibbleRun*)arg;
  // shares in multiples of 128 bases: every share starts on a whole destination byte (and a 64-byte source chunk)
  const uint64_t per = ((r.n / (uint64_t)r.n_tasks) + 127) & ~(uint64_t)127;
  const uint64_t b0 = std::min(r.n, per * (uint64_t)i), b1 = i == r.n_tasks - 1 ? r.n : std::min(r.n, per * (uint64_t)(i + 1));
  if (b1 <= b0) return;
  static const bool scalar = getenv("PENG_HOST_PACK_SCALAR") != nullptr;
  if (have_avx512() && !scalar) nibble_avx512(r.src + b0, b1 - b0, r.dst + (b0 >> 1));
  else nibble_scalar(r.src + b0, b1 - b0, r.dst + (b0 >> 1));
}
}  // namespace

void nibble_pack(Pool& pool, const uint8_t* src, uint64_t n, uint8_t* dst) {
  NibbleRun r{src, dst, n, std::max(1, std::min(pool.size() * 2, (int)(n / 65536 + 1)))};
  pool.run(r.n_tasks, nibble_task, &r);
}

}  // namespace peng_hostpack
