// Host memory bandwidth as the ingest sees it: T threads each copy their share of a buffer (memcpy: 1 B read + 1 B written
// per byte) — the ceiling of everything the host threads do to sequence bytes (packing reads 1 B and writes 0.375 B per base).
//   g++ -O2 -std=c++17 -pthread tools/probe/host_bw.cpp -o /tmp/host_bw && /tmp/host_bw
#include <chrono>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>
int main() {
  const size_t N = (size_t)1 << 30;
  std::vector<char> a(N, 1), b(N, 0);
  const unsigned hw = std::thread::hardware_concurrency();
  printf("hardware threads: %u\n", hw);
  for (unsigned T : {1u, 2u, 4u, 8u, 16u, 32u, 64u}) {
    if (T > hw) break;
    double best = 0;
    for (int rep = 0; rep < 3; rep++) {
      const auto t0 = std::chrono::steady_clock::now();
      std::vector<std::thread> th;
      for (unsigned i = 0; i < T; i++) th.emplace_back([&, i] { memcpy(b.data() + N / T * i, a.data() + N / T * i, N / T); });
      for (auto& t : th) t.join();
      const double s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
      best = std::max(best, 2.0 * N / s / 1e9);
    }
    printf("threads %2u: %.1f GB/s (read + write)\n", T, best);
  }
  return 0;
}
