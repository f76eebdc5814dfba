// core_seq.cuh — per-thread bodies of the sequence-side kernels (ingest packing, k-mer start marking,
// exact non-overlapping k-mer counting, background (k+1)-mer counting).
//
// Every function here is PENG_HD (__host__ __device__) and free of warp intrinsics, so that the very same
// code that the __global__ wrappers in peng_cuda.cu run per thread can be driven thread-by-thread on a
// CPU by tests/sim (test-only shim, never shipped) on machines without a GPU.
//
// Device layout of an ingested sequence set ("stream"):
//   position q = PENG_FRONT_PAD + offsets[s] + 2*s + i   for base i of sequence s
//   i.e. all sequences concatenated, TWO invalid separator positions after each sequence, PENG_FRONT_PAD
//   invalid positions in front and >= PENG_TAIL_PAD behind.
//   codes : 2 bits per position, 16 per uint32 (position q at bits 2*(q&15) of word q>>4); A,C,G,T = 0..3
//   valid : 1 bit per position (ACGT letter of a sequence), 32 per uint32
//   sstart: 1 bit per position, set on the first base of every sequence
// Why two separators: the reference resumes scanning at n+2 after a run that produced k-mers ends at a
// non-ACGT index n (src/base_pattern.cpp:369,380,347) but starts every sequence at its first base; with
// two invalid positions between sequences the single stream rule "resume at n+2" reproduces both.
#pragma once
#include <stdint.h>

#ifndef PENG_HD
#ifdef __CUDACC__
#define PENG_HD __host__ __device__ __forceinline__
#else
#define PENG_HD inline
#endif
#endif

#define PENG_FRONT_PAD 64u
#define PENG_TAIL_PAD 128u

namespace peng {

// ---- small bit helpers (device intrinsics with host equivalents) -------------------------------------
PENG_HD uint32_t funnel_r(uint32_t lo, uint32_t hi, uint32_t s) {  // bits [s, s+32) of hi:lo, 0 <= s < 32
#if defined(__CUDA_ARCH__)
  return __funnelshift_r(lo, hi, s);
#else
  return s ? (lo >> s) | (hi << (32 - s)) : lo;
#endif
}
PENG_HD int popc32(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return __popc(x);
#else
  return __builtin_popcount(x);
#endif
}
PENG_HD uint32_t brev32(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return __brev(x);
#else
  x = ((x >> 1) & 0x55555555u) | ((x & 0x55555555u) << 1);
  x = ((x >> 2) & 0x33333333u) | ((x & 0x33333333u) << 2);
  x = ((x >> 4) & 0x0F0F0F0Fu) | ((x & 0x0F0F0F0Fu) << 4);
  x = ((x >> 8) & 0x00FF00FFu) | ((x & 0x00FF00FFu) << 8);
  return (x >> 16) | (x << 16);
#endif
}
template <typename T>
PENG_HD T ldg(const T* p) {
#if defined(__CUDA_ARCH__)
  return __ldg(p);
#else
  return *p;
#endif
}

// Reverse complement of a W-mer id (BasePattern::getRevCompId, src/base_pattern.cpp:119-134): digit p of
// the result is 3 - digit (W-1-p).  W <= 16, ids fit 32 bits.
PENG_HD uint32_t revcomp_id(uint32_t id, int W) {
  uint32_t r = brev32(id);                                          // reverses bits, also inside each pair
  r = ((r >> 1) & 0x55555555u) | ((r & 0x55555555u) << 1);          // swap the two bits of every pair back
  r >>= (32 - 2 * W);
  return (~r) & (W == 16 ? 0xFFFFFFFFu : ((1u << (2 * W)) - 1u));
}

PENG_HD bool bit_at(const uint32_t* mask, uint64_t q) { return (ldg(mask + (q >> 5)) >> (q & 31)) & 1u; }

// ---- (a) packing --------------------------------------------------------------------------------------
// One thread builds stream positions [32*t, 32*t+32): two code words, one valid word, one sstart word.
// src holds the sequences concatenated WITHOUT separators, either in the reference's letter codes
// (0 = other, 1..4 = ACGT; Alphabet::getCode, src/shared/Alphabet.cpp:33-41) or as raw FASTA letters.
// MODE 0: letter codes, one byte per base; 1: FASTA letters; 2: letter codes as NIBBLES (two bases per byte: base
// src_base + 2 k in the low, src_base + 2 k + 1 in the high nibble of src[k] — the host's cheap half-way packing of the part
// of a set it does not pack completely, host_pack.h: nibble_pack)
template <int MODE>
PENG_HD void pack_word(uint64_t t, const uint8_t* __restrict__ src, const uint64_t* __restrict__ offsets,
                       uint64_t n_seq, uint32_t* __restrict__ codes, uint32_t* __restrict__ valid,
                       uint32_t* __restrict__ sstart, uint32_t* code_words_out = nullptr, uint64_t src_base = 0) {
  constexpr bool ASCII = MODE == 1;
  const uint64_t q0 = t * 32;
  uint32_t c0 = 0, c1 = 0, v = 0, ss = 0;
  const uint64_t total = n_seq ? offsets[n_seq] : 0;
  const uint64_t stream_end = PENG_FRONT_PAD + total + 2 * n_seq;  // first position behind the last separator
  if (q0 + 32 > PENG_FRONT_PAD && q0 < stream_end) {
    // sequence s occupies [start(s), start(s) + len(s) + 2) with start(s) = FRONT + offsets[s] + 2 s.
    // Find the sequence holding the first position >= FRONT of this word: largest s with start(s) <= q.
    uint64_t q = q0 < PENG_FRONT_PAD ? PENG_FRONT_PAD : q0;
    uint64_t lo = 0, hi = n_seq - 1;
    while (lo < hi) {
      uint64_t mid = (lo + hi + 1) >> 1;
      if (PENG_FRONT_PAD + offsets[mid] + 2 * mid <= q) lo = mid; else hi = mid - 1;
    }
    uint64_t s = lo;
    uint64_t s_start = PENG_FRONT_PAD + offsets[s] + 2 * s;
    uint64_t s_len = offsets[s + 1] - offsets[s];
    for (; q < q0 + 32 && q < stream_end; q++) {
      while (q >= s_start + s_len + 2) {  // step over finished sequences (empty ones cannot occur)
        s++;
        s_start = PENG_FRONT_PAD + offsets[s] + 2 * s;
        s_len = offsets[s + 1] - offsets[s];
      }
      const uint64_t i = q - s_start;
      if (i < s_len) {
        uint32_t ch;
        if (MODE == 2) {
          const uint64_t rel = offsets[s] + i - src_base;
          ch = (src[rel >> 1] >> (4 * (rel & 1))) & 15u;
        } else {
          ch = src[offsets[s] + i];
        }
        uint32_t code, ok;
        if (ASCII) {
          const uint32_t up = ch & 0xDFu;  // fold case
          ok = (up == 'A') | (up == 'C') | (up == 'G') | (up == 'T');
          code = (up >> 1) & 3u;           // A0 C1 G3 T2
          code ^= code >> 1;               // A0 C1 G2 T3
        } else {
          ok = (ch >= 1u) & (ch <= 4u);
          code = (ch - 1u) & 3u;
        }
        code = ok ? code : 0u;
        const uint32_t b = (uint32_t)(q - q0);
        if (b < 16) c0 |= code << (2 * b); else c1 |= code << (2 * (b - 16));
        v |= ok << b;
        if (i == 0) ss |= 1u << b;
      }
    }
  }
  codes[2 * t] = c0;
  codes[2 * t + 1] = c1;
  valid[t] = v;
  sstart[t] = ss;
  if (code_words_out) { code_words_out[0] = c0; code_words_out[1] = c1; }  // for the caller's bit planes
}

// The validity word of stream positions [32 t, 32 t + 32) of a set WITHOUT letters other than ACGT: bit b is set iff position
// 32 t + b holds a letter of a sequence — a function of the offsets alone.  The packed ingest sends validity words only
// where they differ from this (host_pack.h: ValidExc).
PENG_HD uint32_t implied_valid_word(uint64_t t, const uint64_t* __restrict__ offsets, uint64_t n_seq) {
  const uint64_t q0 = t * 32, q1 = q0 + 32;
  const uint64_t total = n_seq ? offsets[n_seq] : 0;
  const uint64_t stream_end = PENG_FRONT_PAD + total + 2 * n_seq;
  if (q1 <= PENG_FRONT_PAD || q0 >= stream_end) return 0;
  const uint64_t q = q0 < PENG_FRONT_PAD ? PENG_FRONT_PAD : q0;
  uint64_t lo = 0, hi = n_seq - 1;  // largest s with start(s) <= q
  while (lo < hi) {
    const uint64_t mid = (lo + hi + 1) >> 1;
    if (PENG_FRONT_PAD + offsets[mid] + 2 * mid <= q) lo = mid; else hi = mid - 1;
  }
  uint32_t v = 0;
  uint64_t o = offsets[lo];
  for (uint64_t s = lo; s < n_seq; s++) {
    const uint64_t o1 = offsets[s + 1];
    const uint64_t a = PENG_FRONT_PAD + o + 2 * s, e = a + (o1 - o);  // letters at [a, e)
    o = o1;
    if (a >= q1) break;
    const uint64_t x = a > q0 ? a : q0, y = e < q1 ? e : q1;
    if (y > x) v |= (0xFFFFFFFFu >> (32 - (uint32_t)(y - x))) << (uint32_t)(x - q0);
  }
  return v;
}

// inverse of pack_word for one base (tests / CPU baseline export): reference letter code 0..4
PENG_HD uint8_t unpack_code(const uint32_t* codes, const uint32_t* valid, uint64_t q) {
  if (!bit_at(valid, q)) return 0;
  return (uint8_t)(((ldg(codes + (q >> 4)) >> (2 * (q & 15))) & 3u) + 1u);
}

// ---- synthetic sequence sets (BASELINE.json configs[2..4]) ---------------------------------------------
PENG_HD uint64_t mix64(uint64_t x) {  // splitmix64 finaliser: counter-based, any shard reproducible
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

struct SynthParams {
  uint64_t n_seq, first_seq, seed;
  uint32_t seq_len, motif_len;
  uint32_t plant_threshold;  // plant iff (hash & 0xFFFFFF) < threshold
  uint32_t n_threshold;      // base becomes N iff (hash & 0xFFFFFF) < threshold
  uint64_t motif_bits;       // 2 bits per motif base; base 6 (0-based) of a 13-mer may be degenerate, see below
  uint32_t degenerate_pos;   // motif position whose letter is C or G with p = 1/2 (S), 0xFFFFFFFF = none
};

// letter (0..3) / validity of base i of GLOBAL sequence g
PENG_HD uint32_t synth_base(const SynthParams& P, uint64_t g, uint32_t i, uint32_t* ok) {
  const uint64_t gi = g * (uint64_t)P.seq_len + i;
  const uint64_t h = mix64(P.seed ^ mix64(gi >> 4));  // 16 bases (+ 16 N draws) per hash
  uint32_t c = (uint32_t)(h >> (2 * (gi & 15))) & 3u;
  *ok = 1;
  if (P.n_threshold) {
    const uint64_t hn = mix64(h ^ 0xA5A5A5A5ull ^ (gi & 15));
    if ((uint32_t)(hn & 0xFFFFFFu) < P.n_threshold) *ok = 0;
  }
  if (P.motif_len && P.motif_len <= P.seq_len) {
    const uint64_t hs = mix64((P.seed + 0x51ED27ull) ^ mix64(g));
    if ((uint32_t)(hs & 0xFFFFFFu) < P.plant_threshold) {
      const uint32_t off = (uint32_t)((hs >> 24) % (uint64_t)(P.seq_len - P.motif_len + 1));
      const bool minus = (hs >> 60) & 1u;
      if (i >= off && i < off + P.motif_len) {
        uint32_t m = i - off;
        if (minus) m = P.motif_len - 1 - m;
        uint32_t mc = (uint32_t)(P.motif_bits >> (2 * m)) & 3u;
        if (m == P.degenerate_pos) mc = ((hs >> 61) & 1u) ? 1u : 2u;
        c = minus ? 3u - mc : mc;
        *ok = 1;
      }
    }
  }
  return c;
}

PENG_HD void synth_word(uint64_t t, const SynthParams& P, uint32_t* __restrict__ codes,
                        uint32_t* __restrict__ valid, uint32_t* __restrict__ sstart, uint32_t* code_words_out = nullptr) {
  const uint64_t q0 = t * 32;
  const uint64_t stride = (uint64_t)P.seq_len + 2;
  const uint64_t stream_end = PENG_FRONT_PAD + P.n_seq * stride;
  uint32_t c0 = 0, c1 = 0, v = 0, ss = 0;
  for (uint32_t b = 0; b < 32; b++) {
    const uint64_t q = q0 + b;
    if (q < PENG_FRONT_PAD || q >= stream_end) continue;
    const uint64_t s = (q - PENG_FRONT_PAD) / stride;
    const uint32_t i = (uint32_t)((q - PENG_FRONT_PAD) - s * stride);
    if (i >= P.seq_len) continue;
    uint32_t ok;
    uint32_t c = synth_base(P, P.first_seq + s, i, &ok);
    c = ok ? c : 0u;
    if (b < 16) c0 |= c << (2 * b); else c1 |= c << (2 * (b - 16));
    v |= ok << b;
    if (i == 0) ss |= 1u << b;
  }
  codes[2 * t] = c0;
  codes[2 * t + 1] = c1;
  valid[t] = v;
  sstart[t] = ss;
  if (code_words_out) { code_words_out[0] = c0; code_words_out[1] = c1; }  // for the caller's bit planes
}

// ---- (b1) which W-mers does the reference look at? -----------------------------------------------------
// The reference scans each sequence with a small state machine (src/base_pattern.cpp:347-381):
//   * a run opens where W consecutive letters are available; a failed attempt on index f resumes at f+1;
//   * a run that produced k-mers and stopped at a non-ACGT index n resumes at n+2: the k-mer STARTING at
//     n+1 is never looked at ("shifted" start).
// On the stream this means: the W-mer starting at position a is looked at iff valid[a..a+W-1] are all set
// and NOT (a is the first position of its valid segment and that segment is shifted), where
//   shifted(a) = valid[a-2] && ( len(prev segment) >= W+1 || (len == W && !shifted(start of prev segment)) )
// because the previous segment produced k-mers iff its effective length (len - shifted) is >= W.
PENG_HD bool segment_is_shifted(const uint32_t* valid, uint64_t a, int W) {
  bool flip = false;
  for (;;) {
    if (a < 2 || !bit_at(valid, a - 2)) return flip;  // base case false, xor flip
    // length of the valid segment ending at a-2, capped at W+1
    int len = 0;
    uint64_t p = a - 2;
    while (len <= W && bit_at(valid, p)) {  // the front pad is invalid, so p never underflows
      len++;
      p--;
    }
    if (len >= W + 1) return !flip;  // base case true
    if (len < W) return flip;
    a = a - 1 - (uint64_t)W;  // start of that exactly-W-long segment; shifted(a_old) = !shifted(a_new)
    flip = !flip;
  }
}

// kstart bits of positions [32t, 32t+32)
PENG_HD uint32_t kstart_word(uint64_t t, uint64_t n_words, const uint32_t* __restrict__ valid, int W) {
  const uint32_t v0 = ldg(valid + t);
  const uint32_t v1 = t + 1 < n_words ? ldg(valid + t + 1) : 0u;
  uint32_t win = v0;
  for (int d = 1; d < W; d++) win &= funnel_r(v0, v1, (uint32_t)d);  // bit b: valid[b .. b+W-1] all set
  if (win) {
    const uint32_t vm1 = t ? ldg(valid + t - 1) : 0u;
    const uint32_t prev1 = (v0 << 1) | (vm1 >> 31);  // bit b = valid[b-1]
    const uint32_t prev2 = (v0 << 2) | (vm1 >> 30);  // bit b = valid[b-2]
    uint32_t cand = win & ~prev1 & prev2;            // segment starts that might be shifted (rare)
    while (cand) {
      const uint32_t b = (uint32_t)popc32((cand & (0u - cand)) - 1u);
      cand &= cand - 1u;
      if (segment_is_shifted(valid, t * 32 + b, W)) win &= ~(1u << b);
    }
  }
  return win;
}

// One thread: kstart word t (positions [32t, 32t+32)); returns popcount (contribution to ltot).
PENG_HD int mark_kstart_word(uint64_t t, uint64_t n_words, const uint32_t* __restrict__ valid, int W,
                             uint32_t* __restrict__ kstart) {
  const uint32_t win = kstart_word(t, n_words, valid, W);
  kstart[t] = win;
  return popc32(win);
}

// ---- (b1') which looked-at W-mers MIGHT be suppressed? -------------------------------------------------
// The partition path (peng_cuda.cu: k_count_part) counts a looked-at W-mer unconditionally when no W-mer with
// the same key starts at any of the W-1 positions in front of it ("fast"), and hands everything else to the
// exact ring logic of count_chunk ("slow").  Two W-mers at distance d < W have the same key iff
//   (ff) they are equal:                   x[s+i] == x[s-d+i] for i = 0..W-1  (the stretch has period d), or
//   (rc, BOTH only) one is the reverse complement of the other: the stretch [s-d, s+W-1] is a reverse-
//       complement palindrome; its length W+d must be even (d = 2,4,..,W-2) and its centre lies between
//       m = s + (W-d)/2 - 1 and m+1.  Every such palindrome contains the one of half length W/2+1 around the
//       same centre, so  PAL(m) := x[m+1+i] == 3 - x[m-i] for i = 0..W/2  flags s = m-W/2+2 .. m (superset).
// Letters of invalid positions read as 0: such W-mers are not looked at anyway and a superset is all the
// fast path needs.  Everything is bit-parallel on the 2-bit code words (16 positions per word, bit 2k of a
// result word = position k).
PENG_HD uint32_t compress_even16(uint32_t w) {  // bits 0,2,4,..,30 -> bits 0..15
  w &= 0x55555555u;
  w = (w | (w >> 1)) & 0x33333333u;
  w = (w | (w >> 2)) & 0x0F0F0F0Fu;
  w = (w | (w >> 4)) & 0x00FF00FFu;
  w = (w | (w >> 8)) & 0x0000FFFFu;
  return w;
}

// cw[0..4] = code words 2u-1 .. 2u+3 (positions 32u-16 .. 32u+63).  The work is cut into independent steps so that
// a kernel can spread them between its shared-memory operations (k_count_part); conflict_candidates() runs them
// back to back.  c1 / c2 collect the spaced candidate flags of words 1 and 2 (positions 32u .. 32u+31).
struct CandState {
  uint32_t c1, c2;
  uint32_t p[3];  // PAL(m) at m = 16(j+1) + k, being narrowed
};
PENG_HD void cand_init(CandState& st) {
  st.c1 = 0; st.c2 = 0;
  st.p[0] = st.p[1] = st.p[2] = 0xFFFFFFFFu;
}
// (ff) A W-mer at s has an equal W-mer d < W positions in front of it iff x[p] == x[p-d] for p = s .. s+W-1 (W
// consecutive equalities at lag d).  Then x also repeats at every multiple D = m d < W of the lag, over the shorter
// stretch p = s+D-d .. s+W-1: at least R = W-D+d equalities at lag D ENDING at s+W-1.  So one "proxy" test
//     run of R equalities at lag D ending at s+W-1        (R <= W - D + d for every lag d | D it stands for)
// covers all those d at once; it is a superset test (false alarms at rate 4^-R per position, R >= 7), which is all the
// fast/slow split needs.  cand_ff_run(cw, D, R) ORs the flags of one such test into st; R = W is the exact test of lag D.
// The table FfLags<W> lists the tests per pattern length: 6 instead of 9 lags at W = 10, 7 instead of 13 at W = 14.
template <int W>
PENG_HD void cand_ff_run(const uint32_t* cw, int D, int R, CandState& st) {
  static_assert(W >= 4 && W <= 16 && (W % 2) == 0, "window arithmetic assumes W-1 <= 15 lags inside one word");
  uint32_t e[3];  // e[j] bit 2k: x[p] == x[p-D] at p = 16(j+1) + k
#pragma unroll
  for (int j = 0; j < 3; j++) {
    const uint32_t t = cw[j + 1] ^ funnel_r(cw[j], cw[j + 1], 32u - 2u * (uint32_t)D);
    e[j] = ~(t | (t >> 1)) & 0x55555555u;
  }
  // runs by doubling: r2, r4, r8 = runs of 2, 4, 8 starting at the bit's position; the third word only has to be right
  // for its low positions (zeros shifted in at its top): a run that starts at position 31 of the unit ends at 31 + 15
  // = 46 < 48.  sh(X, k) = X seen k positions further on.
#define PENG_SH(X, k, j) ((j) < 2 ? funnel_r(X[j], X[(j) + 1 < 3 ? (j) + 1 : 2], 2u * (uint32_t)(k)) : (X[2] >> (2u * (uint32_t)(k))))
  uint32_t r2[3], r4[3], r8[3], run[3];
#pragma unroll
  for (int j = 0; j < 3; j++) r2[j] = e[j] & PENG_SH(e, 1, j);
#pragma unroll
  for (int j = 0; j < 3; j++) r4[j] = r2[j] & PENG_SH(r2, 2, j);
#pragma unroll
  for (int j = 0; j < 3; j++) r8[j] = r4[j] & PENG_SH(r4, 4, j);
#pragma unroll
  for (int j = 0; j < 3; j++) {
    uint32_t v;
    if (R == 4) v = r4[j];
    else if (R == 6) v = r4[j] & PENG_SH(r2, 4, j);
    else if (R == 7) v = r4[j] & PENG_SH(r4, 3, j);
    else if (R == 8) v = r8[j];
    else if (R == 9) v = r8[j] & PENG_SH(e, 8, j);
    else if (R == 10) v = r8[j] & PENG_SH(r2, 8, j);
    else if (R == 11) v = r8[j] & PENG_SH(r4, 7, j);
    else if (R == 12) v = r8[j] & PENG_SH(r4, 8, j);
    else if (R == 14) v = r8[j] & PENG_SH(r4, 8, j) & PENG_SH(r2, 12, j);
    else v = r8[j] & PENG_SH(r8, 8, j);  // R == 16
    run[j] = v;
  }
  const int off = W - R;  // the run starts at s + off
  st.c1 |= off ? PENG_SH(run, off, 0) : run[0];
  st.c2 |= off ? PENG_SH(run, off, 1) : run[1];
#undef PENG_SH
}
// exact test of lag d (all W equalities)
template <int W>
PENG_HD void cand_ff_lag(const uint32_t* cw, int d, CandState& st) { cand_ff_run<W>(cw, d, W, st); }

// the lag tests of a pattern length: (D, R) pairs that together cover every lag d = 1 .. W-1 (checked at compile time)
template <int W>
struct FfLags {  // short patterns: every lag exactly
  static constexpr int N = W - 1;
  PENG_HD static constexpr int D(int i) { return i + 1; }
  PENG_HD static constexpr int R(int) { return W; }
};
template <>
struct FfLags<10> {  // 4 (R 7) covers 1, 2, 4; 6 (R 7) covers 3, 6
  static constexpr int N = 6;
  PENG_HD static constexpr int D(int i) { return i == 0 ? 4 : i == 1 ? 6 : i == 2 ? 5 : i + 4; }  // 4 6 5 7 8 9
  PENG_HD static constexpr int R(int i) { return i < 2 ? 7 : 10; }
};
template <>
struct FfLags<12> {  // 6 (R 7): 1, 2, 3, 6; 8 (R 8): 4, 8; 10 (R 7): 5, 10
  static constexpr int N = 6;
  PENG_HD static constexpr int D(int i) { return i == 0 ? 6 : i == 1 ? 8 : i == 2 ? 10 : 2 * i + 1; }  // 6 8 10 7 9 11
  PENG_HD static constexpr int R(int i) { return i == 0 ? 7 : i == 1 ? 8 : i == 2 ? 7 : 12; }
};
template <>
struct FfLags<14> {  // 8 (R 7): 1, 2, 4, 8; 9 (R 8): 3, 9; 12 (R 8): 6, 12; 10 (R 9): 5, 10
  static constexpr int N = 7;
  PENG_HD static constexpr int D(int i) { return i == 0 ? 8 : i == 1 ? 9 : i == 2 ? 12 : i == 3 ? 10 : i == 4 ? 7 : i == 5 ? 11 : 13; }
  PENG_HD static constexpr int R(int i) { return i == 0 ? 7 : i == 1 ? 8 : i == 2 ? 8 : i == 3 ? 9 : 14; }
};
template <>
struct FfLags<16> {  // 10 (R 7): 1, 2, 5, 10; 12 (R 7): 3, 4, 6, 12; 14 (R 9): 7, 14
  static constexpr int N = 8;
  PENG_HD static constexpr int D(int i) { return i == 0 ? 10 : i == 1 ? 12 : i == 2 ? 14 : i == 3 ? 8 : i == 4 ? 9 : 2 * i + 1; }  // 10 12 14 8 9 11 13 15
  PENG_HD static constexpr int R(int i) { return i < 2 ? 7 : i == 2 ? 9 : 16; }
};
template <int W>
constexpr bool ff_lags_cover_all() {
  for (int d = 1; d < W; d++) {
    bool ok = false;
    for (int i = 0; i < FfLags<W>::N; i++) {
      const int D = FfLags<W>::D(i), R = FfLags<W>::R(i);
      if (D < W && D % d == 0 && R <= W - D + d) ok = true;
    }
    if (!ok) return false;
  }
  return true;
}
static_assert(ff_lags_cover_all<4>() && ff_lags_cover_all<6>() && ff_lags_cover_all<8>() && ff_lags_cover_all<10>() &&
              ff_lags_cover_all<12>() && ff_lags_cover_all<14>() && ff_lags_cover_all<16>(), "a lag is not covered");
// (rc) arm i of the palindromes: x[m+1+i] complementary to x[m-i]; i = 0 .. W/2
PENG_HD void cand_rc_arm(const uint32_t* cw, int i, CandState& st) {
#pragma unroll
  for (int j = 0; j < 3; j++) {
    const uint32_t a = funnel_r(cw[j + 1], cw[j + 2], 2u * (uint32_t)(1 + i));                  // x[m+1+i]
    const uint32_t b = i ? funnel_r(cw[j], cw[j + 1], 32u - 2u * (uint32_t)i) : cw[j + 1];       // x[m-i]
    const uint32_t t = a ^ b;
    st.p[j] &= t & (t >> 1);  // both bits set <=> complementary letters
  }
}
// (rc) s is flagged by the centres m = s .. s + W/2 - 2
template <int W>
PENG_HD void cand_rc_finish(CandState& st) {
#pragma unroll
  for (int j = 0; j < 3; j++) st.p[j] &= 0x55555555u;
#pragma unroll
  for (int i = 0; i <= W / 2 - 2; i++) {
    st.c1 |= i ? funnel_r(st.p[0], st.p[1], 2u * (uint32_t)i) : st.p[0];
    st.c2 |= i ? funnel_r(st.p[1], st.p[2], 2u * (uint32_t)i) : st.p[1];
  }
}
PENG_HD uint32_t cand_result(const CandState& st) { return compress_even16(st.c1) | (compress_even16(st.c2) << 16); }

// ---- the same tests on BIT PLANES --------------------------------------------------------------------------
// The 2-bit layout spends two bit lanes per position, so every shift / AND of the scan above moves 16 positions.  With the
// two bits of a code in separate words (plane H = bit 1, plane L = bit 0, position 32u + b at bit b of word u; written
// once per ingest by k_make_planes) the same tests move 32 positions per instruction, the three 16-position words of a
// unit and its look-ahead become two words, and the flags come out one bit per position (no compress at the end).
// pl[0..5] = H[u-1], H[u], H[u+1], L[u-1], L[u], L[u+1].  The flags are identical to conflict_candidates() (tests/sim).
struct CandStatePl {
  uint32_t c;     // candidate flags of positions 32u .. 32u+31
  uint32_t p[2];  // PAL(m) at m = 32u + 32j + k, being narrowed
};
PENG_HD void cand_pl_init(CandStatePl& st) { st.c = 0; st.p[0] = st.p[1] = 0xFFFFFFFFu; }
template <int W>
PENG_HD void cand_pl_ff_run(const uint32_t* pl, int D, int R, CandStatePl& st) {
  static_assert(W >= 4 && W <= 16 && (W % 2) == 0, "window arithmetic assumes W-1 <= 15 lags");
  uint32_t e[2];  // e[j] bit k: x[p] == x[p-D] at p = 32u + 32j + k
#pragma unroll
  for (int j = 0; j < 2; j++) {
    const uint32_t th = pl[j + 1] ^ funnel_r(pl[j], pl[j + 1], 32u - (uint32_t)D);
    e[j] = ~(th | (pl[j + 4] ^ funnel_r(pl[j + 3], pl[j + 4], 32u - (uint32_t)D)));
  }
  // runs by doubling; the look-ahead word only has to be right for its low positions (a run that starts at position 31
  // of the unit ends at 31 + 15 = 46 < 64)
#define PENG_SHP(X, k, j) ((j) == 0 ? funnel_r(X[0], X[1], (uint32_t)(k)) : (X[1] >> (uint32_t)(k)))
  uint32_t r2[2], r4[2], r8[2], run[2];
#pragma unroll
  for (int j = 0; j < 2; j++) r2[j] = e[j] & PENG_SHP(e, 1, j);
#pragma unroll
  for (int j = 0; j < 2; j++) r4[j] = r2[j] & PENG_SHP(r2, 2, j);
#pragma unroll
  for (int j = 0; j < 2; j++) r8[j] = r4[j] & PENG_SHP(r4, 4, j);
#pragma unroll
  for (int j = 0; j < 2; j++) {
    uint32_t v;
    if (R == 4) v = r4[j];
    else if (R == 6) v = r4[j] & PENG_SHP(r2, 4, j);
    else if (R == 7) v = r4[j] & PENG_SHP(r4, 3, j);
    else if (R == 8) v = r8[j];
    else if (R == 9) v = r8[j] & PENG_SHP(e, 8, j);
    else if (R == 10) v = r8[j] & PENG_SHP(r2, 8, j);
    else if (R == 11) v = r8[j] & PENG_SHP(r4, 7, j);
    else if (R == 12) v = r8[j] & PENG_SHP(r4, 8, j);
    else if (R == 14) v = r8[j] & PENG_SHP(r4, 8, j) & PENG_SHP(r2, 12, j);
    else v = r8[j] & PENG_SHP(r8, 8, j);  // R == 16
    run[j] = v;
  }
  const int off = W - R;  // the run starts at s + off
  st.c |= off ? PENG_SHP(run, off, 0) : run[0];
#undef PENG_SHP
}
// (rc) arm i: x[m+1+i] complementary to x[m-i] <=> both planes differ
PENG_HD void cand_pl_rc_arm(const uint32_t* pl, int i, CandStatePl& st) {
#pragma unroll
  for (int j = 0; j < 2; j++) {
    const uint32_t ah = j == 0 ? funnel_r(pl[1], pl[2], (uint32_t)(1 + i)) : (pl[2] >> (uint32_t)(1 + i));  // x[m+1+i]
    const uint32_t al = j == 0 ? funnel_r(pl[4], pl[5], (uint32_t)(1 + i)) : (pl[5] >> (uint32_t)(1 + i));
    const uint32_t bh = i ? funnel_r(pl[j], pl[j + 1], 32u - (uint32_t)i) : pl[j + 1];                       // x[m-i]
    const uint32_t bl = i ? funnel_r(pl[j + 3], pl[j + 4], 32u - (uint32_t)i) : pl[j + 4];
    const uint32_t t = st.p[j] & (ah ^ bh);
    st.p[j] = t & (al ^ bl);
  }
}
template <int W>
PENG_HD void cand_pl_rc_finish(CandStatePl& st) {
#pragma unroll
  for (int i = 0; i <= W / 2 - 2; i++) st.c |= i ? funnel_r(st.p[0], st.p[1], (uint32_t)i) : st.p[0];
}
template <int W, bool BOTH>
struct CandStepsPl {
  static constexpr int N_FF = FfLags<W>::N, N_ARM = BOTH ? W / 2 + 1 : 0, N = N_FF + N_ARM + (BOTH ? 1 : 0);
  PENG_HD static void run(int I, const uint32_t* pl, CandStatePl& st) {
    if (I < N_FF) cand_pl_ff_run<W>(pl, FfLags<W>::D(I), FfLags<W>::R(I), st);
    else if (I < N_FF + N_ARM) cand_pl_rc_arm(pl, I - N_FF, st);
    else if (BOTH && I == N_FF + N_ARM) cand_pl_rc_finish<W>(st);
  }
};
template <int W, bool BOTH>
PENG_HD uint32_t conflict_candidates_planes(const uint32_t* pl) {
  CandStatePl st;
  cand_pl_init(st);
#pragma unroll
  for (int I = 0; I < CandStepsPl<W, BOTH>::N; I++) CandStepsPl<W, BOTH>::run(I, pl, st);
  return st.c;
}
// planes of one unit from its two code words (what k_make_planes stores): hi = bit 1 of every code, lo = bit 0
PENG_HD void planes_of_unit(uint32_t c0, uint32_t c1, uint32_t* hi, uint32_t* lo) {
  *lo = compress_even16(c0) | (compress_even16(c1) << 16);
  *hi = compress_even16(c0 >> 1) | (compress_even16(c1 >> 1) << 16);
}

// number of steps and step I of the whole computation (I is a compile-time constant after unrolling)
template <int W, bool BOTH>
struct CandSteps {
  static constexpr int N_FF = FfLags<W>::N, N_ARM = BOTH ? W / 2 + 1 : 0, N = N_FF + N_ARM + (BOTH ? 1 : 0);
  PENG_HD static void run(int I, const uint32_t* cw, CandState& st) {
    if (I < N_FF) cand_ff_run<W>(cw, FfLags<W>::D(I), FfLags<W>::R(I), st);
    else if (I < N_FF + N_ARM) cand_rc_arm(cw, I - N_FF, st);
    else if (BOTH && I == N_FF + N_ARM) cand_rc_finish<W>(st);
  }
};

// result bit b = candidate flag of the W-mer starting at 32u + b
template <int W, bool BOTH>
PENG_HD uint32_t conflict_candidates(const uint32_t* cw) {
  CandState st;
  cand_init(st);
#pragma unroll
  for (int I = 0; I < CandSteps<W, BOTH>::N; I++) CandSteps<W, BOTH>::run(I, cw, st);
  return cand_result(st);
}

// ---- (b2) exact greedy non-overlapping counting ---------------------------------------------------------
// BasePattern::count_patterns / count_patterns_single_strand (src/base_pattern.cpp:331-441): an occurrence
// is counted iff the last COUNTED occurrence of the same key (BOTH: min(id, revcomp id)) lies >= W
// positions back; suppression never crosses an invalid position.  Equivalent local rule:
//     counted(s) = looked_at(s) && no counted occurrence of key(s) among starts s-W+1 .. s-1.
// A thread owns the k-mers whose LAST base e lies in [e0, e0 + 16*n_words) and keeps the keys of the
// counted occurrences of the previous 15 positions in a 16-slot register ring (slot = e & 15).
// Warm-up: it starts 48 positions early with unknown history, records every looked-at key, and is
// "synced" once W-1 consecutive positions were clean (no equal key in their window): from then on all
// recorded keys are truly counted and the recursion is exact.  If it is not synced when it reaches e0
// (tandem repeats / homopolymers), count_resolve_slow() walks further back.
template <int W>
struct KeyT { typedef uint32_t type; };
template <>
struct KeyT<16> { typedef uint64_t type; };  // 32-bit ids: the "empty slot" sentinel needs a 33rd bit

template <int W, bool BOTH>
struct CountState {
  typedef typename KeyT<W>::type key_t;
  static constexpr key_t EMPTY = ~(key_t)0;
  static constexpr uint32_t MASK = W == 16 ? 0xFFFFFFFFu : ((1u << (2 * (W & 15))) - 1u);
  key_t slot[16];
  uint32_t fwd, rc;

  PENG_HD void reset() {
#pragma unroll
    for (int i = 0; i < 16; i++) slot[i] = EMPTY;
    fwd = 0;
    rc = 0;
  }
  // shift in base c (position e); afterwards fwd / rc are the ids of the W-mer ending at e
  PENG_HD key_t push(uint32_t c) {
    fwd = (fwd >> 2) | (c << (2 * W - 2));
    rc = ((rc << 2) | (3u - c)) & MASK;
    return (key_t)(BOTH ? (fwd < rc ? fwd : rc) : fwd);
  }
  template <int I>
  PENG_HD bool window_has(key_t key) const {  // any of the W-1 previous slots equal to key?
    bool m = false;
#pragma unroll
    for (int d = 1; d < W; d++) m |= (slot[(I - d) & 15] == key);
    return m;
  }
};

// kstart bits for the 16 k-mers ENDING at e = 16*we + i (i = 0..15): start positions e - W + 1
PENG_HD uint32_t kstart_bits_for_end_word(const uint32_t* __restrict__ kstart, uint64_t we, int W) {
  const int64_t s0 = (int64_t)(we * 16) - (W - 1);  // >= 0 because the stream has a front pad
  const uint64_t w = (uint64_t)s0 >> 5;
  const uint32_t sh = (uint32_t)s0 & 31u;
  const uint32_t lo = ldg(kstart + w);
  const uint32_t hi = sh > 16 ? ldg(kstart + w + 1) : 0u;
  return funnel_r(lo, hi, sh) & 0xFFFFu;
}

// Slow path: establish the ring state in front of end position e0 by scanning from further and further
// back until a sync point is found (the front pad guarantees one).  Plain loops, local-memory ring.
template <int W, bool BOTH>
PENG_HD void count_resolve_slow(const uint32_t* __restrict__ codes, const uint32_t* __restrict__ kstart,
                                uint64_t e0, CountState<W, BOTH>& st) {
  typedef typename CountState<W, BOTH>::key_t key_t;
  for (uint64_t back = 128;; back *= 2) {
    const uint64_t start = back >= e0 ? 0 : e0 - back;  // first base pushed
    st.reset();
    key_t ring[16];
    for (int i = 0; i < 16; i++) ring[i] = CountState<W, BOTH>::EMPTY;
    bool synced = (start == 0);  // stream start: empty history is exact
    int streak = 0;
    for (uint64_t e = start; e < e0; e++) {
      const uint32_t c = (ldg(codes + (e >> 4)) >> (2 * (e & 15))) & 3u;
      const key_t key = st.push(c);
      const bool complete = (e - start) >= (uint64_t)(W - 1);  // fwd/rc hold W real bases
      const bool ks = complete && e + 1 >= (uint64_t)W && bit_at(kstart, e - (W - 1));
      bool match = false;
      for (int d = 1; d < W; d++) match |= (ring[(e - d) & 15] == key);
      if (synced) {
        ring[e & 15] = (ks && !match) ? key : CountState<W, BOTH>::EMPTY;
      } else {
        ring[e & 15] = ks ? key : CountState<W, BOTH>::EMPTY;
        // the window of this position is fully recorded once 2(W-1) bases have been pushed
        const bool window_known = (e - start) >= (uint64_t)(2 * (W - 1));
        streak = (window_known && !(ks && match)) ? streak + 1 : 0;
        if (streak >= W - 1) synced = true;
      }
    }
    if (synced) {
      for (int i = 0; i < 16; i++) st.slot[i] = ring[i];
      return;
    }
  }
}

// The per-thread body.  Emits one histogram increment per counted occurrence through `emit(key, i)`,
// i = index of the occurrence's END position relative to e0.
// e0 must be a multiple of 16 and >= PENG_FRONT_PAD (so the 48-position warm-up stays inside the buffer).
template <int W, bool BOTH, typename Emit>
PENG_HD void count_chunk(const uint32_t* __restrict__ codes, const uint32_t* __restrict__ kstart, uint64_t e0,
                         uint32_t n_words, Emit emit) {
  typedef CountState<W, BOTH> State;
  typedef typename State::key_t key_t;
  State st;
  st.reset();
  const uint64_t we0 = e0 >> 4;

  // ---- warm-up over the 3 code words in front of e0 ----
  bool synced = false;
  int streak = 0;
  for (int wi = 0; wi < 3; wi++) {
    const uint64_t we = we0 - 3 + wi;
    const uint32_t cw = ldg(codes + we);
    const uint32_t ksw = kstart_bits_for_end_word(kstart, we, W);
#define PENG_WARM_STEP(I)                                                                     \
  {                                                                                           \
    const key_t key = st.push((cw >> (2 * I)) & 3u);                                          \
    const int step = wi * 16 + I;                                                             \
    const bool ks = ((ksw >> I) & 1u) && step >= W - 1;                                       \
    const bool match = st.template window_has<I>(key);                                        \
    if (synced) {                                                                             \
      st.slot[I] = (ks && !match) ? key : State::EMPTY;                                       \
    } else {                                                                                  \
      st.slot[I] = ks ? key : State::EMPTY;                                                   \
      streak = (step >= 2 * (W - 1) && !(ks && match)) ? streak + 1 : 0;                      \
      synced = streak >= W - 1;                                                               \
    }                                                                                         \
  }
    PENG_WARM_STEP(0) PENG_WARM_STEP(1) PENG_WARM_STEP(2) PENG_WARM_STEP(3)
    PENG_WARM_STEP(4) PENG_WARM_STEP(5) PENG_WARM_STEP(6) PENG_WARM_STEP(7)
    PENG_WARM_STEP(8) PENG_WARM_STEP(9) PENG_WARM_STEP(10) PENG_WARM_STEP(11)
    PENG_WARM_STEP(12) PENG_WARM_STEP(13) PENG_WARM_STEP(14) PENG_WARM_STEP(15)
#undef PENG_WARM_STEP
  }
  if (!synced) {
    const uint32_t f = st.fwd, r = st.rc;  // ids are already exact (48 >= W-1 bases pushed)
    count_resolve_slow<W, BOTH>(codes, kstart, e0, st);
    st.fwd = f;
    st.rc = r;
  }

  // ---- owned positions ----
  for (uint32_t w = 0; w < n_words; w++) {
    const uint64_t we = we0 + w;
    const uint32_t cw = ldg(codes + we);
    const uint32_t ksw = kstart_bits_for_end_word(kstart, we, W);
#define PENG_MAIN_STEP(I)                                        \
  {                                                              \
    const key_t key = st.push((cw >> (2 * I)) & 3u);             \
    const bool counted = ((ksw >> I) & 1u) && !st.template window_has<I>(key); \
    st.slot[I] = counted ? key : State::EMPTY;                   \
    if (counted) emit((uint32_t)key, w * 16u + I);               \
  }
    PENG_MAIN_STEP(0) PENG_MAIN_STEP(1) PENG_MAIN_STEP(2) PENG_MAIN_STEP(3)
    PENG_MAIN_STEP(4) PENG_MAIN_STEP(5) PENG_MAIN_STEP(6) PENG_MAIN_STEP(7)
    PENG_MAIN_STEP(8) PENG_MAIN_STEP(9) PENG_MAIN_STEP(10) PENG_MAIN_STEP(11)
    PENG_MAIN_STEP(12) PENG_MAIN_STEP(13) PENG_MAIN_STEP(14) PENG_MAIN_STEP(15)
#undef PENG_MAIN_STEP
  }
}

// ---- background (k+1)-mer counts -----------------------------------------------------------------------
// BackgroundModel ctor (src/shared/BackgroundModel.cpp:60-84) on Sequence::kmer_ (src/shared/Sequence.cpp:28-33):
// at base i of a sequence the window is the last min(i,8)+1 letters; the (k+1)-mer ending at i (i >= k) is
// counted under its BaMM id (first letter most significant) if every letter of the WINDOW is ACGT; if the
// window holds a non-ACGT letter it is counted as id 0 when each of its last k+1 letters is 'A' or
// non-ACGT, and dropped otherwise (C remainder of a negative number, :75-80).
// Returns for stream position q: bit k of the result = "order k is counted", ids[k] = id.
PENG_HD uint32_t background_ids(const uint32_t* __restrict__ codes, const uint32_t* __restrict__ valid,
                                const uint32_t* __restrict__ sstart, uint64_t q, int K, uint32_t* ids) {
  // gather the 9 positions q-8..q (front pad keeps q-8 >= 0 for real bases)
  uint32_t letters = 0, vmask = 0;  // index j = distance back from q
  int reach = 8;                    // how far back the window extends (stops at the sequence start)
  for (int j = 0; j <= 8; j++) {
    const uint64_t p = q - (uint64_t)j;
    const uint32_t ok = bit_at(valid, p);
    const uint32_t c = (ldg(codes + (p >> 4)) >> (2 * (p & 15))) & 3u;
    letters |= (ok ? c : 0u) << (2 * j);
    vmask |= ok << j;
    if (bit_at(sstart, p)) { reach = j; break; }
  }
  const uint32_t window = (2u << reach) - 1u;  // reach+1 positions
  const bool all_ok = (vmask & window) == window;
  uint32_t res = 0;
  for (int k = 0; k <= K; k++) {
    if (k > reach) break;  // i >= k (loop bound :71)
    const uint32_t m = (1u << (2 * (k + 1))) - 1u;
    // letters has the most recent base in its lowest digit = BaMM order (first letter most significant)
    const uint32_t y = letters & m;
    if (all_ok) {
      ids[k] = y;
      res |= 1u << k;
    } else if (y == 0) {  // invalid letters contribute 0 as well: "A or other" in all k+1 places
      ids[k] = 0;
      res |= 1u << k;
    }
  }
  return res;
}

}  // namespace peng
