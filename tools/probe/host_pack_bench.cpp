// Host packing rate (host_pack.cpp) without a GPU: g++ -O2 -std=c++17 -I peng-motif_b200/csrc tools/probe/host_pack_bench.cpp
//   peng-motif_b200/csrc/host_pack.o -lpthread -o /tmp/host_pack_bench ; PENG_HOST_THREADS=8 /tmp/host_pack_bench [n_seq] [len]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "host_pack.h"
using namespace peng_hostpack;
int main(int argc, char** argv) {
  const uint64_t n = argc > 1 ? strtoull(argv[1], 0, 10) : 1000000, L = argc > 2 ? strtoull(argv[2], 0, 10) : 200;
  std::vector<uint8_t> src(n * L);
  uint32_t x = 12345;
  for (auto& c : src) { x = x * 1664525u + 1013904223u; c = 1 + (x >> 30); }
  std::vector<uint64_t> offs(n + 1);
  for (uint64_t s = 0; s <= n; s++) offs[s] = s * L;
  const uint64_t max_blocks = (16u << 20) / 24;
  std::vector<uint64_t> buf(3 * max_blocks + 64);
  Pool& pool = global_pool();
  auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
  for (int rep = 0; rep < 5; rep++) {
    const double t0 = now();
    bool has_carry = false, first = true;
    Partial carry{};
    int phases = 0;
    for (uint64_t s = 0; s < n;) {
      const uint64_t e = plan_phase(offs.data(), n, s, 64, first, max_blocks - 4);
      Phase ph{src.data(), offs.data(), s, e, false, 64, first, false, 0, buf.data(), buf.data() + 2 * max_blocks};
      pack_phase(pool, ph, &has_carry, &carry);
      s = e; first = false; phases++;
    }
    const double t = now() - t0;
    printf("threads %d: %.2f ms, %.2f GB/s (%d phases)\n", pool.size(), 1e3 * t, n * L / t / 1e9, phases);
  }
  return 0;
}
