// atomic_bench.cu — microbenchmarks that decide the histogram design:
// throughput of random-address increments into a 4^W table through
//   (1) global RED (L2 atomics), uint32 bins, table sizes 256 KB .. 1 GB
//   (2) shared-memory atomics, uint32 bins and packed 16-bit bins with carry detection
//   (3) distributed shared memory: red.shared::cluster into the tables of a 8/16-CTA cluster
// Random addresses from a per-thread xorshift; N increments per thread.  Prints G increments/s.
#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
namespace cg = cooperative_groups;

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ uint32_t xs(uint32_t& s) { s ^= s << 13; s ^= s >> 17; s ^= s << 5; return s; }

__global__ void k_global(uint32_t* hist, uint32_t mask, int n) {
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
#pragma unroll 4
  for (int i = 0; i < n; i++) atomicAdd(hist + (xs(s) & mask), 1u);
}
__global__ void k_global_u64(unsigned long long* hist, uint32_t mask, int n) {
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
#pragma unroll 4
  for (int i = 0; i < n; i++) atomicAdd(hist + (xs(s) & mask), 1ull);
}
// same-warp aggregation with match_any
__global__ void k_global_match(uint32_t* hist, uint32_t mask, int n) {
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
  for (int i = 0; i < n; i++) {
    uint32_t a = xs(s) & mask;
    uint32_t peers = __match_any_sync(0xffffffffu, a);
    if ((__ffs(peers) - 1) == (threadIdx.x & 31)) atomicAdd(hist + a, __popc(peers));
  }
}
__global__ void k_smem(uint32_t* out, uint32_t mask, int n) {
  extern __shared__ uint32_t sh[];
  for (uint32_t i = threadIdx.x; i <= mask; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
#pragma unroll 4
  for (int i = 0; i < n; i++) atomicAdd(sh + (xs(s) & mask), 1u);
  __syncthreads();
  if (threadIdx.x == 0) out[blockIdx.x] = sh[0];
}
__global__ void k_smem16(uint32_t* out, uint32_t* hist, uint32_t mask, int n) {
  extern __shared__ uint32_t sh[];
  for (uint32_t i = threadIdx.x; i <= (mask >> 1); i += blockDim.x) sh[i] = 0;
  __syncthreads();
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
#pragma unroll 4
  for (int i = 0; i < n; i++) {
    uint32_t key = xs(s) & mask;
    uint32_t sft = (key & 1u) * 16u;
    uint32_t old = atomicAdd(sh + (key >> 1), 1u << sft);
    if (((old >> sft) & 0xFFFFu) == 0x7FFFu) { atomicSub(sh + (key >> 1), 0x8000u << sft); atomicAdd(hist + key, 0x8000u); }
  }
  __syncthreads();
  if (threadIdx.x == 0) out[blockIdx.x] = sh[0];
}
// cluster-distributed table: bins_per_cta bins in each CTA's shared memory, cluster holds csize * bins_per_cta
__global__ void k_dsmem(uint32_t* out, uint32_t bins_per_cta_log2, int n) {
  extern __shared__ uint32_t sh[];
  cg::cluster_group cl = cg::this_cluster();
  const uint32_t bpc = 1u << bins_per_cta_log2;
  for (uint32_t i = threadIdx.x; i < bpc; i += blockDim.x) sh[i] = 0;
  cl.sync();
  const uint32_t csize = cl.num_blocks();
  const uint32_t total_mask = bpc * csize - 1;  // csize power of two
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
#pragma unroll 4
  for (int i = 0; i < n; i++) {
    uint32_t key = xs(s) & total_mask;
    uint32_t* remote = cl.map_shared_rank(sh, key >> bins_per_cta_log2);
    atomicAdd(remote + (key & (bpc - 1)), 1u);
  }
  cl.sync();
  if (threadIdx.x == 0) out[blockIdx.x] = sh[0];
}

template <typename F>
static float time_ms(F f, int reps = 5) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  f(); CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < reps; r++) {
    CK(cudaEventRecord(a)); f(); CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b));
    float ms; CK(cudaEventElapsedTime(&ms, a, b)); if (ms < best) best = ms;
  }
  CK(cudaGetLastError());
  return best;
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  const int sms = p.multiProcessorCount;
  printf("device %s, %d SMs\n", p.name, sms);
  uint32_t* hist; CK(cudaMalloc(&hist, 1u << 30)); CK(cudaMemset(hist, 0, 1u << 30));
  uint32_t* out; CK(cudaMalloc(&out, 1 << 20));
  const int n = 2048;
  for (int threads : {128, 256}) for (int bps : {4, 8, 16}) {
    const int grid = sms * bps;
    const double total = (double)grid * threads * n;
    for (int logb : {16, 20, 24, 28}) {
      uint32_t mask = (1u << logb) - 1;
      float ms = time_ms([&] { k_global<<<grid, threads>>>(hist, mask, n); });
      printf("global RED u32  bins=2^%d  grid=%d x %d : %.1f G/s\n", logb, grid, threads, total / ms / 1e6);
    }
  }
  {
    const int grid = sms * 8, threads = 256; const double total = (double)grid * threads * n;
    float ms = time_ms([&] { k_global_u64<<<grid, threads>>>((unsigned long long*)hist, (1u << 20) - 1, n); });
    printf("global RED u64  bins=2^20 : %.1f G/s\n", total / ms / 1e6);
    ms = time_ms([&] { k_global_match<<<grid, threads>>>(hist, (1u << 20) - 1, n); });
    printf("global RED u32 + match_any bins=2^20 : %.1f G/s\n", total / ms / 1e6);
  }
  for (int threads : {256, 512, 1024}) {
    for (int logb : {12, 14, 15}) {  // 16 KB, 64 KB, 128 KB of u32 bins
      size_t smem = (size_t)4 << logb;
      CK(cudaFuncSetAttribute(k_smem, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      int bps = smem <= 32768 ? 2 : 1;
      const int grid = sms * bps; const double total = (double)grid * threads * n;
      float ms = time_ms([&] { k_smem<<<grid, threads, smem>>>(out, (1u << logb) - 1, n); });
      printf("smem atomics u32 bins=2^%d threads=%d grid=%d : %.1f G/s\n", logb, threads, grid, total / ms / 1e6);
    }
    {
      size_t smem = (size_t)2 << 16;  // 65536 packed 16-bit bins = 128 KB
      CK(cudaFuncSetAttribute(k_smem16, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      const int grid = sms; const double total = (double)grid * threads * n;
      float ms = time_ms([&] { k_smem16<<<grid, threads, smem>>>(out, hist, (1u << 16) - 1, n); });
      printf("smem atomics 16-bit packed (returning) bins=2^16 threads=%d : %.1f G/s\n", threads, total / ms / 1e6);
    }
  }
  // DSMEM: cluster sizes 2..16, 2^15 u32 bins (128 KB) per CTA
  for (int csize : {2, 4, 8, 16}) {
    for (int threads : {256, 512, 1024}) {
      const int logb = 15; size_t smem = (size_t)4 << logb;
      CK(cudaFuncSetAttribute(k_dsmem, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      if (csize > 8) CK(cudaFuncSetAttribute(k_dsmem, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
      cudaLaunchConfig_t cfg = {};
      int grid = (sms / csize) * csize;
      cfg.gridDim = dim3(grid); cfg.blockDim = dim3(threads); cfg.dynamicSmemBytes = smem;
      cudaLaunchAttribute at[1]; at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = csize; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.attrs = at; cfg.numAttrs = 1;
      int maxc = 0;
      cudaError_t qe = cudaOccupancyMaxActiveClusters(&maxc, k_dsmem, &cfg);
      if (qe != cudaSuccess || maxc == 0) { printf("dsmem cluster=%d threads=%d : not launchable (%s, max clusters %d)\n", csize, threads, cudaGetErrorString(qe), maxc); cudaGetLastError(); continue; }
      grid = maxc * csize; cfg.gridDim = dim3(grid);
      const double total = (double)grid * threads * n;
      uint32_t lb = logb; int nn = n;
      cudaError_t le = cudaSuccess;
      float ms = time_ms([&] { le = cudaLaunchKernelEx(&cfg, k_dsmem, out, lb, nn); });
      if (le != cudaSuccess) { printf("dsmem cluster=%d launch failed: %s\n", csize, cudaGetErrorString(le)); continue; }
      printf("dsmem red cluster=%d threads=%d active clusters=%d (grid %d) : %.1f G/s\n", csize, threads, maxc, grid, total / ms / 1e6);
    }
  }
  return 0;
}
