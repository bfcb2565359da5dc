// Host memory-copy microbenchmark (no GPU): what the staging threads of hostcopy.cu and the permuted gather
// of crest_cuda_cregen_cre can reach on this box.  g++ -O2 -mavx2 -pthread hostcopy_bench.cpp
#include <immintrin.h>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <random>
#include <thread>
#include <vector>

static void nt_copy(void* dst, const void* src, size_t n) {   // dst 32-byte aligned, n multiple of 32
  const __m256i* s = (const __m256i*)src;
  __m256i* d = (__m256i*)dst;
  for (size_t i = 0; i < n / 32; ++i) _mm256_stream_si256(d + i, _mm256_loadu_si256(s + i));
}

template <class F>
static double run(int T, size_t nitems, F f) {
  std::atomic<size_t> next{0};
  auto t0 = std::chrono::steady_clock::now();
  std::vector<std::thread> th;
  auto work = [&] { for (;;) { size_t i = next.fetch_add(1); if (i >= nitems) break; f(i); } _mm_sfence(); };
  for (int t = 1; t < T; ++t) th.emplace_back(work);
  work();
  for (auto& x : th) x.join();
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

int main() {
  const size_t rec = 1440, n = 100000, bytes = rec * n;
  char* a = (char*)aligned_alloc(4096, bytes);
  char* b = (char*)aligned_alloc(4096, bytes);
  memset(a, 1, bytes); memset(b, 2, bytes);
  std::vector<int> perm(n);
  std::iota(perm.begin(), perm.end(), 0);
  std::mt19937 g(1);
  // kept-first partition-like permutation: 15 % kept in order, the rest reversed at the back
  std::vector<int> src(n);
  { int j = 0, l = (int)n; for (size_t i = 0; i < n; ++i) { bool dup = (g() % 100) < 85; if (!dup) src[j++] = (int)i; else src[--l] = (int)i; } }
  printf("hardware_concurrency %u\n", std::thread::hardware_concurrency());
  for (int T : {1, 4, 8, 12, 16, 24, 31}) {
    for (size_t chunk : {(size_t)256 << 10, (size_t)1 << 20, (size_t)4 << 20}) {
      const size_t nch = (bytes + chunk - 1) / chunk;
      double best_m = 1e9, best_n = 1e9;
      for (int r = 0; r < 4; ++r) {
        best_m = std::min(best_m, run(T, nch, [&](size_t i) { size_t o = i * chunk, len = std::min(chunk, bytes - o); memcpy(b + o, a + o, len); }));
        best_n = std::min(best_n, run(T, nch, [&](size_t i) { size_t o = i * chunk, len = std::min(chunk, bytes - o); nt_copy(b + o, a + o, len); }));
      }
      printf("T=%2d chunk=%4zu KB  memcpy %6.2f ms (%5.1f GB/s)  nt %6.2f ms (%5.1f GB/s)\n", T, chunk >> 10, best_m, bytes / best_m / 1e6, best_n, bytes / best_n / 1e6);
    }
    // permuted gather of 1440-byte records in blocks of 256 records
    const size_t blk = 256, nb = (n + blk - 1) / blk;
    double gm = 1e9, gn = 1e9;
    for (int r = 0; r < 4; ++r) {
      gm = std::min(gm, run(T, nb, [&](size_t i) { for (size_t p = i * blk; p < std::min(n, (i + 1) * blk); ++p) memcpy(b + rec * p, a + rec * (size_t)src[p], rec); }));
      gn = std::min(gn, run(T, nb, [&](size_t i) { for (size_t p = i * blk; p < std::min(n, (i + 1) * blk); ++p) nt_copy(b + rec * p, a + rec * (size_t)src[p], rec); }));
    }
    printf("T=%2d gather 1440-B records: memcpy %6.2f ms (%5.1f GB/s)  nt %6.2f ms (%5.1f GB/s)\n", T, gm, bytes / gm / 1e6, gn, bytes / gn / 1e6);
  }
  return 0;
}
