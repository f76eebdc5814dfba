// part_bench.cu — prototype of the two-stage "partition, then privatised shared-memory histogram" counting path
// for W = 10: measures stage P (keys -> per-class buckets), stage H (bucket -> shared-memory
// bins -> global table) and checks the result against plain global atomics on the same random stream.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

constexpr int W = 10;
constexpr uint32_t KEYMASK = (1u << (2 * W)) - 1u;
constexpr int NCLASS = 37;
constexpr uint32_t NBINS = 1u << (2 * W);

__host__ __device__ inline uint32_t class_start(uint32_t c) { return (uint32_t)((((uint64_t)c << 20) + NCLASS - 1) / NCLASS); }

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
__global__ void k_gen(uint32_t* codes, uint32_t* kfast, uint64_t n_words32) {
  uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_words32) return;
  uint64_t h = mix64(t);
  codes[2 * t] = (uint32_t)h;
  codes[2 * t + 1] = (uint32_t)(h >> 32);
  // sequences of 200 + 2 separators: k-mer starts at offsets 0..190 of every 202
  uint32_t m = 0;
  for (int b = 0; b < 32; b++) {
    uint64_t q = t * 32 + b;
    if (q % 202 <= 190) m |= 1u << b;
  }
  kfast[t] = m;
}

// reference: one global atomic per k-mer start
__global__ void k_ref(const uint32_t* __restrict__ codes, const uint32_t* __restrict__ kfast, uint32_t* hist, uint64_t n_words16) {
  uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_words16) return;
  uint32_t a = codes[t], b = codes[t + 1];
  uint32_t m = (kfast[t >> 1] >> (16 * (t & 1))) & 0xFFFFu;
#pragma unroll
  for (int i = 0; i < 16; i++) {
    uint32_t key = __funnelshift_r(a, b, 2 * i) & KEYMASK;
    if ((m >> i) & 1u) atomicAdd(hist + key, 1u);
  }
}

// ---- stage P -------------------------------------------------------------------------------------------
// CTA of NT threads; a round = every thread handles WPT code words (16 positions each).  Keys are appended to
// per-class CTA queues in shared memory (slot = returning shared atomic on the class counter), after each
// round whole 128-entry blocks are copied to the CTA's private region of the class bucket in global memory.
// A queue that would overflow in a round (skewed input) makes the CTA redo that round with direct global atomics.
__device__ __forceinline__ void push_key(uint32_t key, uint32_t pred, uint32_t cnt_s, uint32_t q_s, uint32_t cap) {
  uint32_t c = (key * NCLASS) >> 20;
  c = pred ? c : NCLASS;  // positions that are not counted go to a scratch queue: no branch around the atomic
  uint32_t slot;
  asm volatile("atom.shared.add.u32 %0, [%1], 1;" : "=r"(slot) : "r"(cnt_s + c * 4u) : "memory");
  const uint32_t addr = q_s + (c * cap + slot) * 2u;
  asm volatile("{\n\t.reg .pred p;\n\tsetp.lt.u32 p, %0, %1;\n\t@p st.shared.u16 [%2], %3;\n\t}"
               :: "r"(slot), "r"(cap), "r"(addr), "h"((uint16_t)key) : "memory");
}

template <int NT, int WPT, int CAP, int MODE>
__global__ void __launch_bounds__(NT) k_part(const uint32_t* __restrict__ codes, const uint32_t* __restrict__ kfast,
                                             uint64_t n_words16, uint16_t* __restrict__ buckets, uint64_t bucket_cap,
                                             unsigned long long* __restrict__ cursor, uint32_t* __restrict__ hist) {
  extern __shared__ uint16_t q[];  // [NCLASS][CAP]
  __shared__ uint32_t cnt[NCLASS + 1], prev[NCLASS];
  __shared__ uint32_t fill[NCLASS];
  __shared__ uint32_t overflow;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x < NCLASS) { cnt[threadIdx.x] = 0; prev[threadIdx.x] = 0; fill[threadIdx.x] = 0; }
  if (threadIdx.x == 0) { overflow = 0; cnt[NCLASS] = 0; }
  __syncthreads();
  const uint32_t cnt_s = (uint32_t)__cvta_generic_to_shared(cnt), q_s = (uint32_t)__cvta_generic_to_shared(q);
  const uint64_t words_per_round = (uint64_t)NT * WPT;
  const uint64_t n_rounds = (n_words16 + words_per_round - 1) / words_per_round;
  for (uint64_t r = blockIdx.x; r < n_rounds; r += gridDim.x) {
    const uint64_t w0 = r * words_per_round;
#pragma unroll
    for (int j = 0; j < WPT; j++) {
      const uint64_t wi = w0 + (uint64_t)j * NT + threadIdx.x;  // coalesced across the CTA
      if (wi < n_words16) {
        const uint32_t a = __ldg(codes + wi), b = __ldg(codes + wi + 1);
        const uint32_t m = (__ldg(kfast + (wi >> 1)) >> (16 * (wi & 1))) & 0xFFFFu;
#pragma unroll
        for (int i = 0; i < 16; i++) push_key(__funnelshift_r(a, b, 2 * i) & KEYMASK, m & (1u << i), cnt_s, q_s, CAP);
      }
    }
    __syncthreads();
    for (int c = warp; c < NCLASS; c += NT / 32) {
      const uint32_t n = cnt[c];
      if (c == 0 && lane == 0) cnt[NCLASS] = 0;
      if (n > CAP) { if (lane == 0) overflow = 1; continue; }
      const uint32_t nfl = n & ~127u;
      if (nfl) {
        const uint32_t base = fill[c];
        uint16_t* dst = buckets + ((uint64_t)c * gridDim.x + blockIdx.x) * bucket_cap + base;
        const uint16_t* src = q + c * CAP;
        if (base + nfl <= bucket_cap) {
          for (uint32_t i = lane * 4; i < nfl; i += 128) *(uint2*)(dst + i) = *(const uint2*)(src + i);
        } else {  // private region full: straight to the global table
          const uint32_t start = class_start(c);
          for (uint32_t i = lane; i < nfl; i += 32) atomicAdd(hist + start + ((src[i] - start) & 0xFFFFu), 1u);
        }
        __syncwarp();
        if (lane == 0 && base + nfl <= bucket_cap) fill[c] = base + nfl;
        const uint32_t rem = n - nfl;
        uint2 v = make_uint2(0, 0);
        if (lane * 4 < rem) v = *(const uint2*)(src + nfl + lane * 4);
        __syncwarp();
        if (lane * 4 < rem) *(uint2*)(q + c * CAP + lane * 4) = v;
        __syncwarp();
        if (lane == 0) { cnt[c] = rem; prev[c] = rem; }
      }
    }
    __syncthreads();
    if (overflow) {  // rare: drop what this round queued (flushed classes keep their consistent state), redo it directly
      // a class that was flushed above already took this round's keys; only classes still holding them are rolled back
      // -> simplest exact rule: this prototype never reaches here on random input; the product handles it per class.
      if (threadIdx.x == 0) overflow = 0;
      __syncthreads();
    }
  }
  // leftovers (< 128 per class): straight to the global table
  for (int c = warp; c < NCLASS; c += NT / 32) {
    if (lane == 0) cursor[(uint64_t)c * gridDim.x + blockIdx.x] = fill[c];
    const uint32_t n = min(cnt[c], (uint32_t)CAP);
    const uint32_t start = class_start(c);
    for (uint32_t i = lane; i < n; i += 32) {
      const uint32_t lo = q[c * CAP + i];
      const uint32_t key = start + ((lo - start) & 0xFFFFu);
      atomicAdd(hist + key, 1u);
    }
  }
}

// ---- stage H -------------------------------------------------------------------------------------------
// CPC CTAs per class; private uint32 bins of the class in shared memory, non-returning atomics; every warp
// walks whole regions (one region = what one stage-P CTA wrote for this class).
template <int NT, int CPC>
__global__ void __launch_bounds__(NT) k_hist(const uint16_t* __restrict__ buckets, uint64_t bucket_cap, uint32_t n_regions,
                                             const unsigned long long* __restrict__ cursor, uint32_t* __restrict__ hist) {
  extern __shared__ uint32_t bins[];
  const uint32_t c = blockIdx.x / CPC, part = blockIdx.x % CPC;
  const uint32_t start = class_start(c), nb = class_start(c + 1) - start;
  for (uint32_t i = threadIdx.x; i < nb; i += NT) bins[i] = 0;
  __syncthreads();
  const uint32_t s16 = start & 0xFFFFu;
  const uint32_t bins_s = (uint32_t)__cvta_generic_to_shared(bins);
  const uint32_t lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (uint32_t reg = part + CPC * warp; reg < n_regions; reg += CPC * (NT / 32)) {
    const uint32_t n8 = (uint32_t)(cursor[(uint64_t)c * n_regions + reg] >> 3);  // 16-byte groups
    const uint4* src = reinterpret_cast<const uint4*>(buckets + ((uint64_t)c * n_regions + reg) * bucket_cap);
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
  for (uint32_t i = threadIdx.x; i < nb; i += NT) {
    const uint32_t v = bins[i];
    if (v) atomicAdd(hist + start + i, v);
  }
}

template <typename F>
static float time_it(F f, int reps = 5) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  f();
  CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int i = 0; i < reps; i++) {
    CK(cudaEventRecord(a));
    f();
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    if (ms < best) best = ms;
  }
  return best;
}

template <int NT, int WPT, int CAP, int MODE>
static void run_variant(const uint32_t* codes, const uint32_t* kfast, uint64_t n_words16, uint16_t* buckets, uint64_t bucket_bytes,
                        unsigned long long* cursor, uint32_t* hist, const uint32_t* hist_ref, int ctas_per_sm, double npos) {
  const size_t smem = (size_t)(NCLASS + 1) * CAP * 2;
  CK(cudaFuncSetAttribute(k_part<NT, WPT, CAP, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const size_t hsmem = (size_t)(class_start(1) + 1) * 4;
  CK(cudaFuncSetAttribute(k_hist<512, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hsmem));
  CK(cudaFuncSetAttribute(k_hist<1024, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hsmem));
  const uint32_t n_regions = 148 * ctas_per_sm;
  const uint64_t region_cap = (bucket_bytes / 2 / NCLASS / n_regions) & ~127ull;
  auto P = [&]() {
    CK(cudaMemsetAsync(hist, 0, NBINS * 4));
    k_part<NT, WPT, CAP, MODE><<<n_regions, NT, smem>>>(codes, kfast, n_words16, buckets, region_cap, cursor, hist);
  };
  auto H512 = [&]() { k_hist<512, 4><<<NCLASS * 4, 512, hsmem>>>(buckets, region_cap, n_regions, cursor, hist); };
  auto H1024 = [&]() { k_hist<1024, 4><<<NCLASS * 4, 1024, hsmem>>>(buckets, region_cap, n_regions, cursor, hist); };
  float tp = time_it(P);
  float th5 = time_it(H512);
  float th10 = time_it(H1024);
  P();
  H1024();
  CK(cudaDeviceSynchronize());
  std::vector<uint32_t> a(NBINS), b(NBINS);
  CK(cudaMemcpy(a.data(), hist, NBINS * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(b.data(), hist_ref, NBINS * 4, cudaMemcpyDeviceToHost));
  uint64_t bad = 0;
  for (uint32_t i = 0; i < NBINS; i++) bad += a[i] != b[i];
  printf("P<NT=%d,WPT=%d,CAP=%d,MODE=%d> x%d/SM smem=%zuKB: P %.3f ms (%.0f Gpos/s)  H512 %.3f ms  H1024 %.3f ms  total(best) %.3f ms = %.0f Gpos/s  mismatches=%llu\n",
         NT, WPT, CAP, MODE, ctas_per_sm, smem >> 10, tp, npos / tp / 1e6, th5, th10, tp + fminf(th5, th10), npos / (tp + fminf(th5, th10)) / 1e6,
         (unsigned long long)bad);
}

int main() {
  const uint64_t npos = 202000000ull;
  const uint64_t n_words32 = npos / 32 + 1;
  const uint64_t n_words16 = n_words32 * 2;
  uint32_t *codes, *kfast, *hist, *hist_ref;
  uint16_t* buckets;
  unsigned long long* cursor;
  CK(cudaMalloc(&codes, (n_words32 * 2 + 8) * 4));
  CK(cudaMemset(codes, 0, (n_words32 * 2 + 8) * 4));
  CK(cudaMalloc(&kfast, (n_words32 + 8) * 4));
  CK(cudaMalloc(&hist, NBINS * 4));
  CK(cudaMalloc(&hist_ref, NBINS * 4));
  CK(cudaMalloc(&cursor, (size_t)NCLASS * 148 * 8 * 8));
  const uint64_t bucket_bytes = npos * 2 * 3 / 2;
  CK(cudaMalloc(&buckets, bucket_bytes));
  k_gen<<<(unsigned)((n_words32 + 255) / 256), 256>>>(codes, kfast, n_words32);
  CK(cudaDeviceSynchronize());
  CK(cudaMemset(hist_ref, 0, NBINS * 4));
  float tr = time_it([&]() { k_ref<<<(unsigned)((n_words16 + 255) / 256), 256>>>(codes, kfast, hist_ref, n_words16); }, 1);
  CK(cudaMemset(hist_ref, 0, NBINS * 4));
  k_ref<<<(unsigned)((n_words16 + 255) / 256), 256>>>(codes, kfast, hist_ref, n_words16);
  CK(cudaDeviceSynchronize());
  printf("reference global atomics: %.3f ms (%.0f Gpos/s)\n", tr, npos / tr / 1e6);
#define RV(NT, WPT, CAP, MODE, PER) run_variant<NT, WPT, CAP, MODE>(codes, kfast, n_words16, buckets, bucket_bytes, cursor, hist, hist_ref, PER, (double)npos)
  RV(512, 1, 512, 0, 2);
  RV(512, 1, 512, 0, 4);
  RV(512, 2, 1024, 0, 2);
  RV(1024, 1, 1024, 0, 2);
  RV(1024, 1, 1024, 0, 1);
  RV(1024, 2, 2048, 0, 1);
  RV(256, 2, 512, 0, 4);
  return 0;
}
