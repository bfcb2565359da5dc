// Large transfers between CALLER memory and the devices.
//
// A Fortran caller's arrays are pageable.  cudaMemcpy stages pageable memory through one driver thread
// (measured 11 GB/s H2D, 17 GB/s D2H on the B200 box for the 144 MB of C3).  Here T host threads share an
// ordered list of 4 MB chunks; each thread owns two pinned slots and one stream per device: the memcpy
// into / out of one slot overlaps the DMA of the other.  A transfer is a list of SEGMENTS (device, device
// pointer, host pointer, bytes, block): one call can feed several GPUs (each over its own PCIe link) and
// reports progress block by block through a BlockGate, which is what lets the one-call driver start the
// pair kernels of the first row blocks while later ones are still on the bus (driver.cu).
// Pinned or registered caller memory is not staged: one cudaMemcpyAsync per segment.
// CREST_CUDA_COPY_THREADS=n sets T (default: the host's hardware threads minus two, at most 16); 0 = plain
// cudaMemcpy.  The threads only copy memory and live for the duration of the call (no pool that would
// compete with the caller's OpenMP teams).
// The host-side copies use non-temporal stores (fast_copy): neither a pinned slot that the DMA engine reads
// next nor the caller's 145 MB array is wanted in the cache, and without the read-for-ownership of regular
// stores one thread moves 13 instead of 9 GB/s, sixteen 86 instead of 66 GB/s, and the permuted gather of
// 1440-byte records runs 1.6-2x faster (tools/microbench/hostcopy_bench.cpp on the B200 box's 16-core Xeon,
// profiles/r02_hostcopy_microbench.txt).
#include "ctx.cuh"
#include "hostcopy.cuh"
#include <immintrin.h>
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <thread>

namespace crest {

// memcpy with non-temporal stores for anything that is not small.  The stores are weakly ordered: callers
// that hand the destination to a DMA engine or another thread call fast_copy_fence() first.
__attribute__((target("avx2"))) static void nt_copy_avx2(char* d, const char* s, size_t n) {
  const size_t head = (32 - ((uintptr_t)d & 31)) & 31;
  if (head) { std::memcpy(d, s, head); d += head; s += head; n -= head; }
  const size_t nv = n / 32;
  const __m256i* sv = reinterpret_cast<const __m256i*>(s);
  __m256i* dv = reinterpret_cast<__m256i*>(d);
  size_t i = 0;
  for (; i + 4 <= nv; i += 4) {
    const __m256i a = _mm256_loadu_si256(sv + i), b = _mm256_loadu_si256(sv + i + 1);
    const __m256i c = _mm256_loadu_si256(sv + i + 2), e = _mm256_loadu_si256(sv + i + 3);
    _mm256_stream_si256(dv + i, a); _mm256_stream_si256(dv + i + 1, b);
    _mm256_stream_si256(dv + i + 2, c); _mm256_stream_si256(dv + i + 3, e);
  }
  for (; i < nv; ++i) _mm256_stream_si256(dv + i, _mm256_loadu_si256(sv + i));
  if (n & 31) std::memcpy(d + nv * 32, s + nv * 32, n & 31);
}
void fast_copy(void* dst, const void* src, size_t n) {
  static const bool avx2 = __builtin_cpu_supports("avx2");
  if (avx2 && n >= 512) nt_copy_avx2(static_cast<char*>(dst), static_cast<const char*>(src), n);
  else std::memcpy(dst, src, n);
}
void fast_copy_fence() { _mm_sfence(); }

namespace {
constexpr size_t kChunk = 1u << 20;   // 1 MB: a block of the upload pipeline is shared evenly by <= 16 threads
constexpr int kMaxThreads = 16;
constexpr int kMaxDev = 16;
void* g_slot[kMaxThreads][2] = {};
// streams and events of the staging threads live as long as the slots: creating them costs ~0.2 ms each
// when several threads do it at once (driver lock), which used to delay the first block of every call
cudaStream_t g_stream[kMaxThreads][kMaxDev] = {};
cudaEvent_t g_event[kMaxThreads][2][kMaxDev] = {};

bool ensure_slots(int T) {
  for (int t = 0; t < T; ++t)
    for (int k = 0; k < 2; ++k)
      if (!g_slot[t][k] &&
          cudaHostAlloc(&g_slot[t][k], kChunk, cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        return false;
      }
  return true;
}
}  // namespace

void free_copy_slots() {
  for (auto& s : g_slot)
    for (void*& q : s)
      if (q) { cudaFreeHost(q); q = nullptr; }
  for (int t = 0; t < kMaxThreads; ++t)
    for (int d = 0; d < kMaxDev; ++d) {
      if (!g_stream[t][d] && !g_event[t][0][d] && !g_event[t][1][d]) continue;
      cudaSetDevice(d);
      if (g_stream[t][d]) { cudaStreamDestroy(g_stream[t][d]); g_stream[t][d] = nullptr; }
      for (int k = 0; k < 2; ++k)
        if (g_event[t][k][d]) { cudaEventDestroy(g_event[t][k][d]); g_event[t][k][d] = nullptr; }
    }
}

int copy_threads() {
  static int env_t = -2;
  if (env_t == -2) {
    const char* e = std::getenv("CREST_CUDA_COPY_THREADS");
    env_t = e ? std::atoi(e) : -1;
  }
  int t = tune().copy_threads >= 0 ? tune().copy_threads : env_t;   // CREST_CUDA_TUNE copythr=<n>
  return t;
}

bool host_is_pageable(const void* p) {
  cudaPointerAttributes at{};
  const bool pageable = cudaPointerGetAttributes(&at, p) != cudaSuccess || at.type == cudaMemoryTypeUnregistered;
  cudaGetLastError();
  return pageable;
}

void BlockGate::reset(int nblocks, int nthreads) {
  std::lock_guard<std::mutex> lk(mu);
  done.assign((size_t)nblocks, 0);
  need = nthreads;
  failed = false;
}
void BlockGate::pass(int block) {
  {
    std::lock_guard<std::mutex> lk(mu);
    if (block >= 0 && block < (int)done.size()) ++done[block];
  }
  cv.notify_all();
}
void BlockGate::fail() {
  { std::lock_guard<std::mutex> lk(mu); failed = true; }
  cv.notify_all();
}
bool BlockGate::wait(int block) {
  std::unique_lock<std::mutex> lk(mu);
  cv.wait(lk, [&] { return failed || done[block] >= need; });
  return !failed;
}

int staged_copy(const std::vector<CopySeg>& segs, int dir, int nthreads, int nblocks, BlockGate* gate) {
  const cudaMemcpyKind kind = dir ? cudaMemcpyDeviceToHost : cudaMemcpyHostToDevice;
  size_t total = 0;
  bool pageable = false;
  int ndev_used = 0;
  {
    bool seen[64] = {};
    for (const CopySeg& s : segs) {
      total += s.bytes;
      if (s.bytes && host_is_pageable(s.host)) pageable = true;
      if (s.device >= 0 && s.device < 64 && !seen[s.device]) { seen[s.device] = true; ++ndev_used; }
    }
  }
  int T = nthreads;
  const int hw = (int)std::thread::hardware_concurrency();
  // default: every hardware thread but two (the caller's thread and one for the rest of the system); measured
  // on the 16-core host of the B200 box the copy rate still grows from 8 to 12 threads and is flat beyond
  if (T < 0) T = hw > 3 ? hw - 2 : 1;
  if (hw > 1 && T > hw - 1) T = hw - 1;
  if (T > kMaxThreads) T = kMaxThreads;
  if (total < (gate ? (8u << 20) : (32u << 20))) T = 0;   // small transfers: the staging threads cost more than they save
  if (!pageable || T <= 0 || !ensure_slots(T)) {
    // direct path: one async copy per segment on the device's null stream, blocks signalled in order
    if (gate) gate->reset(nblocks, 1);
    int blk = segs.empty() ? nblocks : segs[0].block;
    for (int b = 0; b < blk && gate; ++b) gate->pass(b);
    for (size_t i = 0; i < segs.size(); ++i) {
      const CopySeg& s = segs[i];
      if (s.block != blk) {
        // leaving a block: its segments are complete once the devices they went to have drained
        for (size_t j = 0; j < i; ++j)
          if (segs[j].block == blk) { cudaSetDevice(segs[j].device); cudaStreamSynchronize(cudaStreamPerThread); }
        for (int b = blk; b < s.block && gate; ++b) gate->pass(b);
        blk = s.block;
      }
      if (!s.bytes) continue;
      if (cudaSetDevice(s.device) != cudaSuccess ||
          cudaMemcpyAsync(dir ? s.host : s.dev, dir ? s.dev : s.host, s.bytes, kind, cudaStreamPerThread) != cudaSuccess) {
        set_error("%s copy of %zu bytes failed: %s", dir ? "D2H" : "H2D", s.bytes, cudaGetErrorString(cudaGetLastError()));
        if (gate) gate->fail();
        return 1;
      }
    }
    for (const CopySeg& s : segs) {
      cudaSetDevice(s.device);
      if (cudaStreamSynchronize(cudaStreamPerThread) != cudaSuccess) {
        set_error("%s copy failed: %s", dir ? "D2H" : "H2D", cudaGetErrorString(cudaGetLastError()));
        if (gate) gate->fail();
        return 1;
      }
    }
    for (int b = blk < 0 ? 0 : blk; b < nblocks && gate; ++b) gate->pass(b);
    return 0;
  }

  struct Chunk { int seg; size_t off, len; };
  std::vector<Chunk> chunks;
  for (size_t i = 0; i < segs.size(); ++i)
    for (size_t off = 0; off < segs[i].bytes; off += kChunk)
      chunks.push_back({(int)i, off, std::min(kChunk, segs[i].bytes - off)});
  const size_t nchunk = chunks.size();
  std::atomic<size_t> next{0};
  std::atomic<int> err{0};
  if (gate) gate->reset(nblocks, T);
  int maxdev = 0;
  for (const CopySeg& s : segs) maxdev = std::max(maxdev, s.device);

  if (maxdev >= kMaxDev) { set_error("device ordinal %d beyond the staging tables", maxdev); if (gate) gate->fail(); return 1; }
  auto work = [&](int t) {
    bool ok = true;
    bool used[kMaxDev] = {};
    auto stream_of = [&](int dev) -> cudaStream_t {
      if (!g_stream[t][dev]) {
        if (cudaSetDevice(dev) != cudaSuccess ||
            cudaStreamCreateWithFlags(&g_stream[t][dev], cudaStreamNonBlocking) != cudaSuccess) ok = false;
      }
      used[dev] = true;
      return g_stream[t][dev];
    };
    auto drain = [&]() {
      for (int d = 0; d <= maxdev; ++d)
        if (used[d]) { cudaSetDevice(d); if (cudaStreamSynchronize(g_stream[t][d]) != cudaSuccess) ok = false; }
    };
    // the event that guards slot k while its DMA targets device dev
    cudaEvent_t cur_ev[2] = {nullptr, nullptr};
    auto slot_event = [&](int k, int dev) -> cudaEvent_t {
      if (!g_event[t][k][dev]) {
        cudaSetDevice(dev);
        if (cudaEventCreateWithFlags(&g_event[t][k][dev], cudaEventDisableTiming) != cudaSuccess) ok = false;
      }
      return g_event[t][k][dev];
    };
    int cur_block = -1, k = 0;
    // D2H keeps one DMA in flight: `pend` is the chunk whose DMA was issued into slot pk
    long pend = -1; int pk = 0;
    auto finish_pending = [&]() {
      if (pend < 0) return;
      const Chunk& c = chunks[(size_t)pend];
      if (cudaEventSynchronize(cur_ev[pk]) != cudaSuccess) ok = false;
      fast_copy((char*)segs[c.seg].host + c.off, g_slot[t][pk], c.len);
      pend = -1;
    };
    for (;;) {
      const size_t idx = next.fetch_add(1);
      const int blk = idx < nchunk ? segs[chunks[idx].seg].block : nblocks;
      if (blk != cur_block) {
        if (dir) { finish_pending(); fast_copy_fence(); }
        drain();
        if (gate && ok) for (int b = std::max(cur_block, 0); b < blk; ++b) gate->pass(b);
        cur_block = blk;
      }
      if (idx >= nchunk || !ok) break;
      const Chunk& c = chunks[idx];
      const CopySeg& s = segs[c.seg];
      cudaStream_t stream = stream_of(s.device);
      if (!ok) break;
      if (dir == 0) {
        if (cur_ev[k] && cudaEventSynchronize(cur_ev[k]) != cudaSuccess) ok = false;   // slot k free again
        fast_copy(g_slot[t][k], (const char*)s.host + c.off, c.len);
        fast_copy_fence();                         // the DMA engine reads the slot next
        cur_ev[k] = slot_event(k, s.device);
        cudaSetDevice(s.device);
        ok = ok && cudaMemcpyAsync((char*)s.dev + c.off, g_slot[t][k], c.len, kind, stream) == cudaSuccess;
        ok = ok && cudaEventRecord(cur_ev[k], stream) == cudaSuccess;
        k ^= 1;
      } else {
        // issue this chunk's DMA into the free slot, then copy the previous chunk out
        const int kk = pend >= 0 ? (pk ^ 1) : 0;
        cur_ev[kk] = slot_event(kk, s.device);
        if (!ok) break;
        cudaSetDevice(s.device);
        ok = ok && cudaMemcpyAsync(g_slot[t][kk], (const char*)s.dev + c.off, c.len, kind, stream) == cudaSuccess;
        ok = ok && cudaEventRecord(cur_ev[kk], stream) == cudaSuccess;
        finish_pending();
        pend = (long)idx; pk = kk;
      }
    }
    if (dir) { finish_pending(); fast_copy_fence(); }
    drain();
    if (!ok) { err.store(1); if (gate) gate->fail(); }
  };
  std::vector<std::thread> th;
  for (int t = 1; t < T; ++t) th.emplace_back(work, t);
  work(0);
  for (auto& x : th) x.join();
  if (err.load()) {
    set_error("staged %s copy of %zu bytes failed: %s", dir ? "D2H" : "H2D", total,
              cudaGetErrorString(cudaGetLastError()));
    return 1;
  }
  return 0;
}

int big_copy(void* dst, const void* src, size_t bytes, int dir, int device) {
  std::vector<CopySeg> segs(1);
  segs[0].device = device;
  segs[0].dev = dir ? const_cast<void*>(src) : dst;
  segs[0].host = dir ? dst : const_cast<void*>(src);
  segs[0].bytes = bytes;
  segs[0].block = 0;
  int cur = 0;
  cudaGetDevice(&cur);
  const int rc = staged_copy(segs, dir, copy_threads(), 1, nullptr);
  cudaSetDevice(cur);
  return rc;
}

}  // namespace crest
