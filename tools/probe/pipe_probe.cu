// Throughput probe for the integer instructions stage P is made of (sm_100a): warp-instructions per cycle and SMSP.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_probe pipe_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define N_IT 4096
template <int MODE>
__global__ void __launch_bounds__(1024) k(uint32_t* out, uint32_t m, uint32_t zero, long long* cyc) {
  uint32_t x[8];
#pragma unroll
  for (int i = 0; i < 8; i++) x[i] = threadIdx.x * 2654435761u + i + zero;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < N_IT; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (MODE == 0) asm volatile("shf.r.wrap.b32 %0, %0, %1, 7;" : "+r"(x[i]) : "r"(x[(i + 1) & 7]));
      if (MODE == 1) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(x[i]) : "r"(m));
      if (MODE == 2) asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(m), "r"(x[(i + 1) & 7]));
      if (MODE == 3) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(x[(i + 1) & 7]), "r"(m));
      if (MODE == 4) {  // SHF + IMAD.HI alternating (do the two pipes overlap?)
        if (i & 1) asm volatile("shf.r.wrap.b32 %0, %0, %1, 7;" : "+r"(x[i]) : "r"(x[(i + 1) & 7]));
        else asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(x[i]) : "r"(m));
      }
      if (MODE == 5) {  // LOP3 + mad.lo alternating
        if (i & 1) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(x[(i + 1) & 7]), "r"(m));
        else asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(m), "r"(x[(i + 1) & 7]));
      }
      if (MODE == 6) asm volatile("mad.hi.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(m), "r"(x[(i + 1) & 7]));
      if (MODE == 7) {  // funnel as mul.hi + mad.lo
        uint32_t t;
        asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(t) : "r"(x[i]), "r"(m));
        asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(x[i]) : "r"(x[(i + 1) & 7]), "r"(m), "r"(t));
      }
      if (MODE == 8) {  // 2 LOP3 : 1 mul.hi : 1 mad.lo
        if ((i & 3) < 2) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[i]) : "r"(x[(i + 1) & 7]), "r"(m));
        else if ((i & 3) == 2) asm volatile("mul.hi.u32 %0, %0, %1;" : "+r"(x[i]) : "r"(m));
        else asm volatile("mad.lo.u32 %0, %0, %1, %2;" : "+r"(x[i]) : "r"(m), "r"(x[(i + 1) & 7]));
      }
    }
  }
  const long long t1 = clock64();
  uint32_t s = 0;
#pragma unroll
  for (int i = 0; i < 8; i++) s ^= x[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
template <int MODE>
void run(const char* name, int per_it) {
  uint32_t* out; long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
  k<MODE><<<148, 1024>>>(out, 0x02000001u, 0, cyc);
  k<MODE><<<148, 1024>>>(out, 0x02000001u, 0, cyc);
  cudaDeviceSynchronize();
  long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  // 32 warps per SM = 8 warps per SMSP, each N_IT * per_it instructions
  const double inst = 8.0 * N_IT * per_it;
  printf("%-28s %8.3f warp-instr / cycle / SMSP   (%lld cycles)  %s\n", name, inst / c, c, cudaGetErrorString(cudaGetLastError()));
}
int main() {
  run<0>("SHF", 8); run<1>("IMAD.HI (mul.hi)", 8); run<2>("IMAD (mad.lo)", 8); run<3>("LOP3", 8);
  run<4>("SHF + mul.hi 1:1", 8); run<5>("LOP3 + mad.lo 1:1", 8); run<6>("mad.hi", 8);
  run<7>("funnel = mul.hi + mad.lo", 16); run<8>("2 LOP3 : mul.hi : mad.lo", 8);
  return 0;
}
