// One-call entry points of the C ABI (what the Fortran shim binds) on a process-global ENGINE:
// 1-8 GPUs driven from ONE host thread (CREST is one process with OpenMP and no MPI: the caller is
// newcregen, reference src/cregen.f90:221-235, from its serial region).
//
//   crest_cuda_init(ngpu, &used) / crest_cuda_finalize()       SURVEY.md 8b
//   crest_cuda_cregen_dup    K0-K4 (cregen.f90:1222-1376)
//   crest_cuda_cregen_cre    whole cregen_CRE (cregen.f90:1195-1474)
//   crest_cuda_cregen_group  K5-K6 (cregen.f90:1421-1469)
//
// Stage A is a PIPELINE over row blocks: the ensemble is energy sorted and pair (i, j) needs rows i and
// j < i only, so after the first b rows are on the device a fraction (b/N)^2 of all pairs can run.  A
// staging thread streams xyz block by block (hostcopy.cu: pageable caller memory through pinned slots,
// every GPU pulling its share of a block over its own PCIe link); as a block lands each GPU pulls the
// other GPUs' shares over NVLink (peer copies), packs the block (pack_axis + TF32 operands, redundantly -
// cheaper than an all-gather of the operands) and runs ITS shard of the block's work list
// (pairs_stage).  Links are combined without a collective: every GPU's kernels atomicMax_system straight
// into GPU 0's dup[] over NVLink peer memory (reads of dup[] for the largest-j-wins pruning go the same
// way), so there is nothing to reduce afterwards; GPU 0 waits for the other streams and chases roots.
// Outputs come back sliced over the GPUs' PCIe links as well (every GPU holds the whole xyz_axis).
#include "ctx.cuh"
#include "hostcopy.cuh"
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <string>
#include <thread>

using namespace crest;

namespace {

constexpr int kMaxGpu = 8;

struct Engine {
  bool inited = false;
  int n = 0;
  crest_cuda_ctx* c[kMaxGpu] = {};
  cudaEvent_t ev_zero = nullptr;             // GPU 0: dup[] zeroed
  cudaEvent_t ev_done[kMaxGpu] = {};         // per GPU: its pair kernels of the call are done
  cudaEvent_t ev_packed[kMaxGpu] = {};       // per GPU: xyz_axis complete
  void* h_stage = nullptr;                   // pinned staging of the whole-CREGEN call's geometry
  size_t h_stage_bytes = 0;
  cudaStream_t copy_st[kMaxGpu] = {};        // D2H of xyz_axis into h_stage under the pair stage
  bool hint_dense = false;                   // the previous call met candidate-dense work items (plan_blocks)
};
Engine g_eng;
std::mutex g_mu;                             // the one-call entry points are serialised

void engine_destroy() {
  Engine& E = g_eng;
  for (int d = 0; d < kMaxGpu; ++d) {
    if (!E.c[d]) continue;
    cudaSetDevice(E.c[d]->device);
    cudaDeviceSynchronize();
    if (E.ev_done[d]) cudaEventDestroy(E.ev_done[d]);
    if (E.ev_packed[d]) cudaEventDestroy(E.ev_packed[d]);
    if (E.copy_st[d]) cudaStreamDestroy(E.copy_st[d]);
    if (d == 0 && E.ev_zero) cudaEventDestroy(E.ev_zero);
    crest_cuda_ctx_destroy(E.c[d]);
    E.c[d] = nullptr; E.ev_done[d] = nullptr; E.ev_packed[d] = nullptr; E.copy_st[d] = nullptr;
  }
  E.ev_zero = nullptr;
  if (E.h_stage) { cudaFreeHost(E.h_stage); E.h_stage = nullptr; E.h_stage_bytes = 0; }
  free_copy_slots();
  E.n = 0;
  E.inited = false;
  E.hint_dense = false;
}

// devices: n == 1 -> the CURRENT device (a caller that has selected its GPU keeps it); n > 1 -> 0..n-1
int engine_init(int ngpu_requested) {
  Engine& E = g_eng;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
    cudaGetLastError();
    set_error("no CUDA device available (there is no CPU fallback)");
    return 1;
  }
  int n = ngpu_requested <= 0 ? ndev : std::min(ngpu_requested, ndev);
  n = std::min(n, kMaxGpu);
  if (n > 1) {
    // all-to-all peer access (NVLink / NVSwitch); without it the engine stays on one GPU
    for (int a = 0; a < n && n > 1; ++a)
      for (int b = 0; b < n; ++b) {
        if (a == b) continue;
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, a, b) != cudaSuccess || !can) { n = 1; break; }
      }
  }
  int cur = 0;
  CREST_CUDA_TRY(cudaGetDevice(&cur));
  if (E.inited) {
    const bool same = E.n == n && (n > 1 || E.c[0]->device == cur);
    if (same) return 0;
    engine_destroy();
  }
  for (int d = 0; d < n; ++d) {
    const int dev = n == 1 ? cur : d;
    if (crest_cuda_ctx_create(&E.c[d], dev)) { engine_destroy(); return 1; }
    CREST_CUDA_TRY(cudaSetDevice(dev));
    CREST_CUDA_TRY(cudaStreamCreateWithFlags(&E.c[d]->stream, cudaStreamNonBlocking));
    CREST_CUDA_TRY(cudaStreamCreateWithFlags(&E.copy_st[d], cudaStreamNonBlocking));
    CREST_CUDA_TRY(cudaEventCreateWithFlags(&E.ev_done[d], cudaEventDisableTiming));
    CREST_CUDA_TRY(cudaEventCreateWithFlags(&E.ev_packed[d], cudaEventDisableTiming));
    if (d == 0) CREST_CUDA_TRY(cudaEventCreateWithFlags(&E.ev_zero, cudaEventDisableTiming));
  }
  if (n > 1)
    for (int a = 0; a < n; ++a) {
      CREST_CUDA_TRY(cudaSetDevice(a));
      for (int b = 0; b < n; ++b) {
        if (a == b) continue;
        const cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
          set_error("cudaDeviceEnablePeerAccess(%d -> %d) failed: %s", a, b, cudaGetErrorString(e));
          cudaGetLastError();
          engine_destroy();
          return 1;
        }
        cudaGetLastError();
      }
    }
  CREST_CUDA_TRY(cudaSetDevice(n == 1 ? cur : 0));
  E.n = n;
  E.inited = true;
  return 0;
}

int engine(Engine** out) {
  if (!g_eng.inited) {
    // lazily: CREST_CUDA_NGPU gpus (default 1 = the current device); crest_cuda_init overrides
    const char* e = std::getenv("CREST_CUDA_NGPU");
    if (engine_init(e ? std::atoi(e) : 1)) return 1;
  }
  *out = &g_eng;
  return 0;
}

struct DeviceGuard {   // every entry point leaves the caller's current device as it found it
  int dev = -1;
  DeviceGuard() { if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); dev = -1; } }
  ~DeviceGuard() { if (dev >= 0) cudaSetDevice(dev); }
};

struct Phase {   // CREST_CUDA_TIMING=1: phase wall times on stderr (developer aid; adds syncs)
  bool on;
  std::chrono::steady_clock::time_point t;
  Phase() : on(std::getenv("CREST_CUDA_TIMING") && std::atoi(std::getenv("CREST_CUDA_TIMING")) > 0),
            t(std::chrono::steady_clock::now()) {}
  void lap(const Engine& E, const char* what) {
    if (!on) return;
    for (int d = 0; d < E.n; ++d) { cudaSetDevice(E.c[d]->device); cudaDeviceSynchronize(); }
    const auto now = std::chrono::steady_clock::now();
    std::fprintf(stderr, "[cregen] %-34s %8.3f ms\n", what,
                 std::chrono::duration<double, std::milli>(now - t).count());
    t = now;
  }
};

// Row blocks of the upload pipeline (whole 128-row tiles); one block when the FP64-only kernels run (their
// work lists are not staged) or the ensemble is small.  The ensemble is energy sorted and pair (i, j) needs
// rows i and j < i only: after a fraction f of the rows is on the device, f^2 of the pair work can run.
// With rho = (pair time per GPU) / (upload time) the GPU that started on a first block f1 stays busy if the
// next boundary is at most f + rho f^2 (the work of the rows so far lasts until the next block has landed).
//   compute bound (rho >= 1.5; C3 on 1-2 GPUs): a small first block (0.20) and boundaries growing by
//     f <- f + 2 rho f^2 - C3 on one GPU: 0.20, 0.46, 1.  The rule f + rho f^2 itself (0.10, 0.14, 0.20, 0.31,
//     0.61, 1) starts the GPU earlier and is 0.1 ms faster on C3, but every stage ends in the tail of its
//     slowest rows, and on a collapsed ensemble (C3dense: rows that burn their evaluation budget) six stages
//     cost 4 ms more than three (profiles/r02_d_block_schedule.txt; knobs blockf1=<%>, blockgain=<%>);
//   upload bound (rho < 1.5; many GPUs, large molecules): equal blocks of ~36 MB x (GPUs / 2), 2-8 of them;
//     every stage costs ~0.1 ms of tails and, on 8 GPUs, 56 peer copies (profiles/r02_b_timeline_n8.txt).
// The time estimates only shape the schedule; results do not depend on it.
std::vector<int> plan_blocks(const crest_cuda_ctx* c, bool staged, int ngpu, bool hint_dense) {
  std::vector<int> b{0, c->ntile};
  const size_t bytes = (size_t)24 * c->nat * c->nall;
  if (tune().blocks > 0) {
    const int nb = std::min(tune().blocks, c->ntile);
    if (nb <= 1) return b;
    b.clear();
    for (int k = 0; k <= nb; ++k) b.push_back((int)((long long)c->ntile * k / nb));
    return b;
  }
  if (!staged || bytes < ((size_t)24 << 20)) return b;
  // An ensemble with candidate-dense items (collapsed ensembles, floppy macrocycles) pays for every stage: each
  // one ends in the tail of its slowest rows and re-classifies its items - C4ext takes 23.7 ms as one block,
  // 27.6 as two, 34.5 as three (profiles/r02_i_c4ext_blocks.txt).  Density is only known afterwards, so the
  // engine remembers it: CREST calls CREGEN 5-20 times per search on ensembles of the same kind, and a call
  // that met dense items makes the next one run as a single block (first call: the schedule below).
  if (hint_dense) return b;
  // rough rates: staged upload ~40 GB/s (PCIe 5 x16 with the staging threads keeping up), screen kernel
  // ~4.2e11 pairs/s per GPU up to 88 atoms, ~300 TFLOP/s of 18 A flop per pair on the slab kernel above
  const double t_up = (double)bytes / 40e9;
  const double pairs = 0.5 * (double)c->nall * c->nall;
  const double t_pair = (c->use_col ? pairs / 4.2e11 : pairs * 18.0 * c->nsel / 3.0e14) / std::max(1, ngpu);
  const double rho = t_pair / t_up;
  if (rho >= 1.5 && !tune().even_blocks) {
    std::vector<double> f;
    const double f1_0 = tune().block_f1 > 0 ? 0.01 * tune().block_f1 : 0.20;
    const double gain = tune().block_gain > 0 ? 0.01 * tune().block_gain : 2.0;
    for (double f1 = f1_0; f1 < 0.5; f1 += 0.05) {          // a larger first block if 8 stages do not reach the end
      f.assign(1, f1);
      while (f.back() < 1.0 && f.size() < 9)
        f.push_back(f.back() + std::max(gain * rho * f.back() * f.back(), 0.4 * f1_0));
      if (f.back() >= 1.0) break;
    }
    if (f.back() >= 1.0 && f.size() >= 2) {
      b.assign(1, 0);
      for (size_t k = 0; k + 1 < f.size(); ++k) {
        const int t = (int)(f[k] * c->ntile + 0.5);
        if (t > b.back() && t < c->ntile) b.push_back(t);
      }
      b.push_back(c->ntile);
      return b;
    }
  }
  int nb = (int)std::min<size_t>(8, std::max<size_t>(2, bytes / (((size_t)36 << 20) * (size_t)std::max(1, ngpu / 2))));
  if (nb > c->ntile) nb = c->ntile;
  if (nb <= 1) return b;
  b.clear();
  for (int k = 0; k <= nb; ++k) b.push_back((int)((long long)c->ntile * k / nb));
  return b;
}

// Stage A (K0-K4) on the engine's GPUs.  On return every stream is idle, GPU 0 holds rot / aniso / dup
// (roots) and every GPU the whole xyz_axis.  If stage_xyz != NULL the principal-axis geometry is also
// copied into that PINNED buffer (pure DMA, sliced over the GPUs, under the pair kernels).
// rot_out / xyz_axis_out (caller memory, may be NULL): copied out UNDER the pair kernels as well - both
// are final once the last block is packed, and the host thread has nothing else to do by then.
int stage_a(Engine& E, int nat, int nall, const int* at, const double* xyz, const double* er,
            const int* rmsd_mask, int rot_from_mask, int enantio, const crest_cuda_thresholds* thr,
            long long* npairs, double* stage_xyz, double* rot_out, double* xyz_axis_out, Phase& ph) {
  const int n = E.n;
  const bool tl = std::getenv("CREST_CUDA_TIMING") && std::atoi(std::getenv("CREST_CUDA_TIMING")) == 2;
  const auto t_host0 = std::chrono::steady_clock::now();
  auto hms = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_host0).count(); };
  // One worker thread per GPU beyond the first issues that GPU's copies and launches: a single host
  // thread spends ~5 us per launch, and 8 GPUs x (7 peer copies + pack + 5 pair-stage launches) per block
  // made the HOST the bottleneck of an 8-GPU call (profiles/r02_a_timeline_n8.txt: 6 ms of set-up and
  // 0.6-0.9 ms of enqueueing per block for 0.1-0.5 ms of GPU work).  Errors travel back as text.
  std::vector<std::string> werr((size_t)n);
  auto run_parallel = [&](const std::function<int(int)>& f) -> int {
    std::vector<std::thread> th;
    std::vector<int> rcs((size_t)n, 0);
    for (int d = 1; d < n; ++d)
      th.emplace_back([&, d] { rcs[d] = f(d); if (rcs[d]) werr[d] = crest_cuda_last_error(); });
    rcs[0] = f(0);
    for (auto& t : th) t.join();
    for (int d = 0; d < n; ++d)
      if (rcs[d]) { if (d > 0) set_error("GPU %d: %s", E.c[d]->device, werr[d].c_str()); return 1; }
    return 0;
  };
  crest_cuda_ctx* c0 = E.c[0];
  if (npairs) *npairs = 0;
  // phase 1 (per GPU): reserve buffers, small uploads, zero the accumulators, this GPU's work lists
  std::vector<int> tb;
  bool staged = false;
  (void)tune();                                              // parse the knobs before any worker starts
  if (run_parallel([&](int d) -> int {
        if (crest_cuda_ctx_upload(E.c[d], nat, nall, at, nullptr, er, rmsd_mask, rot_from_mask)) return 1;
        return 0;
      })) return 1;
  const double t_reserve = hms();
  if (nall == 0) return 0;
  staged = tf32_path(c0, 0, nullptr) && nall >= 2;
  tb = plan_blocks(c0, staged, n, E.hint_dense);
  const int nb = (int)tb.size() - 1;
  const size_t rec = (size_t)24 * nat;                       // bytes per structure

  // copy plan: block k = rows [128 tb[k], 128 tb[k+1]) cut into n shares, one per GPU
  std::vector<CopySeg> segs;
  std::vector<size_t> share_off((size_t)nb * (n + 1));
  for (int k = 0; k < nb; ++k) {
    const size_t r0 = (size_t)tb[k] * 128, r1 = std::min<size_t>((size_t)tb[k + 1] * 128, nall);
    for (int d = 0; d <= n; ++d) share_off[(size_t)k * (n + 1) + d] = (r0 + (r1 - r0) * d / n) * rec;
    for (int d = 0; d < n; ++d) {
      const size_t o0 = share_off[(size_t)k * (n + 1) + d], o1 = share_off[(size_t)k * (n + 1) + d + 1];
      if (o1 > o0) segs.push_back({E.c[d]->device, (char*)E.c[d]->d_xyz + o0, (char*)const_cast<double*>(xyz) + o0, o1 - o0, k});
    }
  }
  // The upload starts NOW, beside the per-GPU set-up below: d_xyz is reserved, every stream of the engine was
  // idle when the previous call returned, and nothing of the set-up touches d_xyz.
  BlockGate gate;
  gate.reset(nb, 1 << 30);                                   // armed before the copier starts
  int copy_rc = 0;
  const int T = copy_threads();
  std::thread copier([&] { copy_rc = staged_copy(segs, 0, T, nb, &gate); });
  if (run_parallel([&](int d) -> int {
        crest_cuda_ctx* c = E.c[d];
        CREST_CUDA_TRY(cudaSetDevice(c->device));
        // ctx_upload's small copies and memsets ran on the legacy default stream, which the engine's
        // non-blocking streams do not order against (a pageable cudaMemcpy returns before its DMA has landed)
        CREST_CUDA_TRY(cudaStreamSynchronize(cudaStreamLegacy));
        c->dup_target = c0->d_dup;
        if (begin_call(c, c->stream)) return 1;
        if (staged && plan_stages(c, thr->ethr, n, d, tb, c->stream)) return 1;
        if (d == 0) CREST_CUDA_TRY(cudaEventRecord(E.ev_zero, c->stream));
        return 0;
      })) {
    copier.join();
    for (int d = 0; d < n; ++d) { cudaSetDevice(E.c[d]->device); cudaDeviceSynchronize(); }
    return 1;
  }
  const double t_plan = hms();
  std::vector<cudaEvent_t> tev;
  std::vector<double> t_gate((size_t)nb, 0.0), t_enq((size_t)nb, 0.0);
  if (tl) {
    cudaSetDevice(c0->device);
    tev.resize(2 * (size_t)nb + 1);
    for (auto& e : tev) cudaEventCreate(&e);
    cudaEventRecord(tev[0], c0->stream);
  }
  // phase 2 (per GPU): as block k lands, pull the other GPUs' shares, pack, run this GPU's rows of it
  const int rc = run_parallel([&](int d) -> int {
    crest_cuda_ctx* c = E.c[d];
    for (int k = 0; k < nb; ++k) {
      if (!gate.wait(k)) { set_error("upload failed"); return 1; }
      if (tl && d == 0) t_gate[k] = hms();
      const int r0 = tb[k] * 128, r1 = std::min(tb[k + 1] * 128, nall);
      CREST_CUDA_TRY(cudaSetDevice(c->device));
      for (int s = 0; s < n; ++s) {                          // the other GPUs' shares of the block over NVLink
        if (s == d) continue;
        const size_t o0 = share_off[(size_t)k * (n + 1) + s], o1 = share_off[(size_t)k * (n + 1) + s + 1];
        if (o1 > o0)
          CREST_CUDA_TRY(cudaMemcpyPeerAsync((char*)c->d_xyz + o0, c->device, (char*)E.c[s]->d_xyz + o0,
                                             E.c[s]->device, o1 - o0, c->stream));
      }
      if (pack_rows(c, r0, r1, c->stream)) return 1;
      if (k == nb - 1) {
        CREST_CUDA_TRY(cudaEventRecord(E.ev_packed[d], c->stream));
        if (stage_xyz) {
          // this GPU's slice of the finished xyz_axis -> pinned staging, under the pair kernels
          const size_t o0 = (size_t)nall * d / n * rec, o1 = (size_t)nall * (d + 1) / n * rec;
          CREST_CUDA_TRY(cudaStreamWaitEvent(E.copy_st[d], E.ev_packed[d], 0));
          if (o1 > o0)
            CREST_CUDA_TRY(cudaMemcpyAsync((char*)stage_xyz + o0, (char*)c->d_xyz_axis + o0, o1 - o0,
                                           cudaMemcpyDeviceToHost, E.copy_st[d]));
        }
      }
      if (d > 0 && k == 0) CREST_CUDA_TRY(cudaStreamWaitEvent(c->stream, E.ev_zero, 0));
      if (tl && d == 0) cudaEventRecord(tev[2 * (size_t)k + 1], c->stream);
      if (pairs_stage(c, enantio, thr, d, n, staged ? k : 0, 0, nullptr, c->stream)) return 1;
      if (tl && d == 0) { cudaEventRecord(tev[2 * (size_t)k + 2], c->stream); t_enq[k] = hms(); }
    }
    if (d > 0) CREST_CUDA_TRY(cudaEventRecord(E.ev_done[d], c->stream));
    return 0;
  });
  if (rc) gate.fail();
  copier.join();
  if (rc || copy_rc) {
    for (int d = 0; d < n; ++d) { cudaSetDevice(E.c[d]->device); cudaDeviceSynchronize(); }
    return 1;
  }
  // GPU 0 chases the roots once every GPU's links have arrived
  CREST_CUDA_TRY(cudaSetDevice(c0->device));
  for (int d = 1; d < n; ++d) CREST_CUDA_TRY(cudaStreamWaitEvent(c0->stream, E.ev_done[d], 0));
  if (crest_cuda_ctx_sweep(c0, c0->stream)) return 1;
  // everything is enqueued: outputs that are final after the last pack leave while the pair kernels run
  if (rot_out) {
    CREST_CUDA_TRY(cudaStreamWaitEvent(E.copy_st[0], E.ev_packed[0], 0));
    CREST_CUDA_TRY(cudaMemcpyAsync(rot_out, c0->d_rot, sizeof(double) * 3 * (size_t)nall, cudaMemcpyDeviceToHost, E.copy_st[0]));
    CREST_CUDA_TRY(cudaStreamSynchronize(E.copy_st[0]));
  }
  if (xyz_axis_out) {
    // every GPU holds the whole principal-axis geometry: each returns a slice over its own PCIe link
    std::vector<CopySeg> out;
    for (int d = 0; d < n; ++d) {
      CREST_CUDA_TRY(cudaSetDevice(E.c[d]->device));
      CREST_CUDA_TRY(cudaEventSynchronize(E.ev_packed[d]));
      const size_t o0 = (size_t)nall * d / n * rec, o1 = (size_t)nall * (d + 1) / n * rec;
      if (o1 > o0) out.push_back({E.c[d]->device, (char*)E.c[d]->d_xyz_axis + o0, (char*)xyz_axis_out + o0, o1 - o0, 0});
    }
    const int crc = staged_copy(out, 1, copy_threads(), 1, nullptr);
    cudaSetDevice(c0->device);
    if (crc) return 1;
  }
  long long np = 0;
  for (int d = 0; d < n; ++d) {
    CREST_CUDA_TRY(cudaSetDevice(E.c[d]->device));
    long long v = 0;
    CREST_CUDA_TRY(cudaMemcpyAsync(&v, E.c[d]->d_npairs, sizeof(v), cudaMemcpyDeviceToHost, E.c[d]->stream));
    CREST_CUDA_TRY(cudaStreamSynchronize(E.c[d]->stream));
    if (stage_xyz) CREST_CUDA_TRY(cudaStreamSynchronize(E.copy_st[d]));
    np += v;
  }
  if (npairs) *npairs = np;
  CREST_CUDA_TRY(cudaSetDevice(c0->device));
  if (staged) {                                              // density of this call = the next call's block schedule
    unsigned h[16] = {};
    CREST_CUDA_TRY(cudaMemcpy(h, c0->tf.cand_count, sizeof(h), cudaMemcpyDeviceToHost));
    E.hint_dense = h[5] != 0u;
  }
  ph.lap(E, "upload | pack | pairs | early D2H");
  if (tl) {
    std::fprintf(stderr, "[timeline] host: reserve done %.3f, begin+plan done %.3f, all synced at %.3f ms (%d blocks)\n",
                 t_reserve, t_plan, hms(), nb);
    for (int k = 0; k < nb; ++k) {
      float a = 0, b = 0;
      cudaEventElapsedTime(&a, tev[0], tev[2 * (size_t)k + 1]);
      cudaEventElapsedTime(&b, tev[0], tev[2 * (size_t)k + 2]);
      std::fprintf(stderr, "[timeline] block %d: host gate %.3f enqueued %.3f | GPU0 pairs %.3f .. %.3f ms after first launch\n",
                   k, t_gate[k], t_enq[k], a, b);
    }
    for (auto& e : tev) cudaEventDestroy(e);
  }
  return 0;
}

int ensure_stageb(crest_cuda_ctx* c, size_t n) {
  if (n <= c->cap_stageb) return 0;
  cudaFree(c->d_src); cudaFree(c->d_group); cudaFree(c->d_er2); cudaFree(c->d_rot2); cudaFree(c->d_enuc);
  c->d_src = c->d_group = nullptr; c->d_er2 = c->d_rot2 = c->d_enuc = nullptr; c->cap_stageb = 0;
  CREST_CUDA_TRY(cudaMalloc(&c->d_src, sizeof(int) * n));
  CREST_CUDA_TRY(cudaMalloc(&c->d_group, sizeof(int) * (n + 1)));
  CREST_CUDA_TRY(cudaMalloc(&c->d_er2, sizeof(double) * n));
  CREST_CUDA_TRY(cudaMalloc(&c->d_rot2, sizeof(double) * 3 * n));
  CREST_CUDA_TRY(cudaMalloc(&c->d_enuc, sizeof(double) * n));
  c->cap_stageb = n;
  return 0;
}

// generic scratch: grows, never shrinks (the 5-20 CREGEN calls of one search reuse it)
struct DevScratch {
  void* p = nullptr;
  ~DevScratch() { cudaFree(p); }
  int get(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1) == cudaSuccess ? 0 : 1; }
};

}  // namespace

extern "C" {

int crest_cuda_init(int ngpu_requested, int* ngpu_used) {
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceGuard guard;
  if (engine_init(ngpu_requested)) { if (ngpu_used) *ngpu_used = 0; return 1; }
  if (ngpu_used) *ngpu_used = g_eng.n;
  return 0;
}

void crest_cuda_finalize(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceGuard guard;
  engine_destroy();
}

int crest_cuda_ngpu(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  return g_eng.inited ? g_eng.n : 0;
}

int crest_cuda_cregen_dup(int nat, int nall, const int* at, const double* xyz, const double* er,
                          const int* rmsd_mask, int rot_from_mask, int enantio,
                          const crest_cuda_thresholds* thr, double* rot, double* xyz_axis,
                          int* dup, long long* npairs) {
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceGuard guard;
  if (check_thr(thr)) return 1;
  if (nat <= 0 || nall < 0 || (nall > 0 && !xyz)) { set_error("bad argument"); return 1; }
  Engine* E;
  if (engine(&E)) return 1;
  Phase ph;
  if (stage_a(*E, nat, nall, at, xyz, er, rmsd_mask, rot_from_mask, enantio, thr, npairs, nullptr, rot, xyz_axis, ph)) return 1;
  if (nall == 0) return 0;
  crest_cuda_ctx* c0 = E->c[0];
  if (dup) CREST_CUDA_TRY(cudaMemcpy(dup, c0->d_dup, sizeof(int) * (size_t)nall, cudaMemcpyDeviceToHost));
  ph.lap(*E, "D2H dup");
  return 0;
}

int crest_cuda_cregen_group(int nat, int nkept, const int* at, const double* xyz,
                            const double* er, const double* rot,
                            const crest_cuda_thresholds* thr, int* group, double* enuc) {
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceGuard guard;
  if (check_thr(thr)) return 1;
  if (nat <= 0 || nkept < 0 || !at || !group) { set_error("bad argument"); return 1; }
  Engine* E;
  if (engine(&E)) return 1;
  crest_cuda_ctx* c = E->c[0];
  CREST_CUDA_TRY(cudaSetDevice(c->device));
  if (nkept == 0) { group[0] = 0; return 0; }
  if (!xyz || !er || !rot) { set_error("NULL input array"); return 1; }
  // K6 finds its energy window by binary search: the energies of the kept structures must ascend
  for (int i = 1; i < nkept; ++i)
    if (er[i] < er[i - 1]) {
      set_error("energies must ascend (cregen_esort post-condition): er[%d] < er[%d]", i, i - 1);
      return 1;
    }
  const size_t n = nkept;
  if (ensure_stageb(c, n)) return 1;
  DevScratch d_at, d_xyz;
  if (d_at.get(sizeof(int) * nat) || d_xyz.get(sizeof(double) * 3 * nat * n)) {
    set_error("cudaMalloc failed: %s", cudaGetErrorString(cudaGetLastError()));
    return 1;
  }
  CREST_CUDA_TRY(cudaMemcpy(d_at.p, at, sizeof(int) * nat, cudaMemcpyHostToDevice));
  if (big_copy(d_xyz.p, xyz, sizeof(double) * 3 * nat * n, 0, c->device)) return 1;
  CREST_CUDA_TRY(cudaMemcpy(c->d_er2, er, sizeof(double) * n, cudaMemcpyHostToDevice));
  CREST_CUDA_TRY(cudaMemcpy(c->d_rot2, rot, sizeof(double) * 3 * n, cudaMemcpyHostToDevice));
  if (launch_enuc(nat, nkept, (const int*)d_at.p, (const double*)d_xyz.p, c->d_enuc, 0)) return 2;
  if (launch_group(nkept, c->d_er2, c->d_rot2, c->d_enuc, thr->ethr, thr->bthr, thr->bthrmax,
                   thr->bthrshift, c->d_group, 0)) return 2;
  CREST_CUDA_TRY(cudaMemcpy(group, c->d_group, sizeof(int) * (n + 1), cudaMemcpyDeviceToHost));
  if (enuc) CREST_CUDA_TRY(cudaMemcpy(enuc, c->d_enuc, sizeof(double) * n, cudaMemcpyDeviceToHost));
  return 0;
}

int crest_cuda_cregen_cre(int nat, int nall, const int* at, double* xyz, double* er,
                          const int* rmsd_mask, int rot_from_mask, int enantio, int nosort,
                          const crest_cuda_thresholds* thr, double* rot, int* order, int* group,
                          int* nallout, long long* npairs) {
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceGuard guard;
  if (!rot || !order || !group || !nallout) { set_error("NULL output array"); return 1; }
  if (check_thr(thr)) return 1;
  if (nat <= 0 || nall < 0 || (nall > 0 && !xyz)) { set_error("bad argument"); return 1; }
  Engine* E;
  if (engine(&E)) return 1;
  Phase ph;
  const size_t n = (size_t)nall, rec = (size_t)3 * nat;
  // pinned staging of the principal-axis geometry: filled by DMA under the pair kernels; the kept-first
  // permutation is then applied on the way into the caller's array by host threads
  double* stage = nullptr;
  const size_t need = sizeof(double) * rec * n;
  if (need >= ((size_t)8 << 20) && !tune().no_stage) {
    if (need > E->h_stage_bytes) {
      if (E->h_stage) cudaFreeHost(E->h_stage);
      E->h_stage = nullptr; E->h_stage_bytes = 0;
      if (cudaHostAlloc(&E->h_stage, need, cudaHostAllocPortable) == cudaSuccess) E->h_stage_bytes = need;
      else cudaGetLastError();
    }
    stage = (double*)E->h_stage;
  }
  if (stage_a(*E, nat, nall, at, xyz, er, rmsd_mask, rot_from_mask, enantio, thr, npairs, stage, nullptr, nullptr, ph)) return 1;
  for (int i = 0; i <= nall; ++i) group[i] = 0;
  if (nall == 0) { *nallout = 0; return 0; }
  crest_cuda_ctx* c = E->c[0];
  CREST_CUDA_TRY(cudaSetDevice(c->device));
  std::vector<int> dup(n);
  CREST_CUDA_TRY(cudaMemcpy(dup.data(), c->d_dup, sizeof(int) * n, cudaMemcpyDeviceToHost));
  // kept-first partition, duplicates appended in reverse encounter order (cregen.f90:1397-1419)
  int ndup = 0;
  for (int i = 0; i < nall; ++i) ndup += dup[i] != 0;
  int nout = nall - ndup;
  std::vector<int> src(n);
  if (nosort) {
    nout = nall;
    for (int i = 0; i < nall; ++i) { order[i] = i + 1; src[i] = i; }
  } else {
    int j = 0, l = nall + 1;
    for (int i = 0; i < nall; ++i) {
      order[i] = (dup[i] == 0) ? ++j : --l;
      src[order[i] - 1] = i;
    }
  }
  ph.lap(*E, "backtrack, D2H dup, partition");
  // permutation, K5 and K6 on GPU 0; with the staging buffer the geometry reaches the caller's array by a
  // host-side permuted gather that runs beside them, else by one D2H of the permuted device copy
  if (ensure_stageb(c, n)) return 1;
  cudaStream_t st = c->stream;
  CREST_CUDA_TRY(cudaMemcpyAsync(c->d_src, src.data(), sizeof(int) * n, cudaMemcpyHostToDevice, st));
  // d_xyz (the input copy) is free now: it receives the permuted principal-axis geometries
  if (launch_permute(nat, nall, c->d_src, c->d_xyz_axis, c->d_er, c->d_rot, c->d_xyz, c->d_er2, c->d_rot2, st)) return 2;
  CREST_CUDA_TRY(cudaMemsetAsync(c->d_group, 0, sizeof(int) * (n + 1), st));
  if (launch_enuc(nat, nout, c->d_at, c->d_xyz, c->d_enuc, st)) return 2;
  int gather_threads = 0;
  std::vector<std::thread> gth;
  if (stage) {
    // non-temporal stores, 12-16 threads: 1.9 ms for the 144 MB of C3 against 4.2 ms with 8 memcpy threads
    // (profiles/r02_hostcopy_microbench.txt)
    gather_threads = std::max(1, std::min(16, (int)std::thread::hardware_concurrency() - 1));
    for (int t = 0; t < gather_threads; ++t)
      gth.emplace_back([&, t] {
        const size_t p0 = n * t / gather_threads, p1 = n * (t + 1) / gather_threads;
        for (size_t p = p0; p < p1; ++p)
          fast_copy(xyz + rec * p, stage + rec * (size_t)src[p], rec * sizeof(double));
        fast_copy_fence();
      });
  }
  int rc = launch_group(nout, c->d_er2, c->d_rot2, c->d_enuc, thr->ethr, thr->bthr, thr->bthrmax, thr->bthrshift,
                        c->d_group, st);
  if (!rc) {
    rc = cudaMemcpyAsync(er, c->d_er2, sizeof(double) * n, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
         cudaMemcpyAsync(rot, c->d_rot2, sizeof(double) * 3 * n, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
         cudaMemcpyAsync(group, c->d_group, sizeof(int) * ((size_t)nout + 1), cudaMemcpyDeviceToHost, st) != cudaSuccess ||
         cudaStreamSynchronize(st) != cudaSuccess;
    if (rc) set_error("CUDA error in cregen_cre: %s", cudaGetErrorString(cudaGetLastError()));
  }
  for (auto& t : gth) t.join();
  ph.lap(*E, "permute, enuc, group | host gather");
  if (rc) return rc;
  if (!stage) {
    if (big_copy(xyz, c->d_xyz, sizeof(double) * rec * n, 1, c->device)) return 2;
    ph.lap(*E, "D2H xyz");
  }
  c->launches += 3;
  if (nosort) nout = group[0];
  *nallout = nout;
  return 0;
}

int crest_cuda_rotconst(int nat, int n, const int* at, const double* xyz, double* rot) {
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceGuard guard;
  if (nat <= 0 || n < 0 || !at || !rot || (!xyz && n)) { set_error("bad argument"); return 1; }
  Engine* E;
  if (engine(&E)) return 1;
  CREST_CUDA_TRY(cudaSetDevice(E->c[0]->device));
  if (n == 0) return 0;
  DevScratch d_at, d_xyz, d_rot;
  if (d_at.get(sizeof(int) * nat) || d_xyz.get(sizeof(double) * 3 * nat * (size_t)n) ||
      d_rot.get(sizeof(double) * 3 * (size_t)n)) {
    set_error("cudaMalloc failed: %s", cudaGetErrorString(cudaGetLastError()));
    return 1;
  }
  CREST_CUDA_TRY(cudaMemcpy(d_at.p, at, sizeof(int) * nat, cudaMemcpyHostToDevice));
  CREST_CUDA_TRY(cudaMemcpy(d_xyz.p, xyz, sizeof(double) * 3 * nat * (size_t)n, cudaMemcpyHostToDevice));
  if (launch_rotconst(nat, n, (const int*)d_at.p, (const double*)d_xyz.p, (double*)d_rot.p, 0)) return 1;
  CREST_CUDA_TRY(cudaMemcpy(rot, d_rot.p, sizeof(double) * 3 * (size_t)n, cudaMemcpyDeviceToHost));
  return 0;
}

int crest_cuda_rmsd_pairs(int A, int n, const double* packed, long long npairs, const int* i_idx,
                          const int* j_idx, double* rmsd, double* rmsd_mirror) {
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceGuard guard;
  if (A <= 0 || n <= 0 || npairs < 0 || !packed || !rmsd || !rmsd_mirror) {
    set_error("bad argument"); return 1;
  }
  Engine* E;
  if (engine(&E)) return 1;
  CREST_CUDA_TRY(cudaSetDevice(E->c[0]->device));
  if (npairs == 0) return 0;
  for (long long p = 0; p < npairs; ++p)
    if (i_idx[p] < 0 || i_idx[p] >= n || j_idx[p] < 0 || j_idx[p] >= n) {
      set_error("pair %lld index out of range", p); return 1;
    }
  DevScratch d_p, d_r, d_m, d_i, d_j;
  if (d_p.get(sizeof(double) * 3 * A * (size_t)n) || d_r.get(sizeof(double) * npairs) ||
      d_m.get(sizeof(double) * npairs) || d_i.get(sizeof(int) * npairs) || d_j.get(sizeof(int) * npairs)) {
    set_error("cudaMalloc failed: %s", cudaGetErrorString(cudaGetLastError()));
    return 1;
  }
  CREST_CUDA_TRY(cudaMemcpy(d_p.p, packed, sizeof(double) * 3 * A * (size_t)n, cudaMemcpyHostToDevice));
  CREST_CUDA_TRY(cudaMemcpy(d_i.p, i_idx, sizeof(int) * npairs, cudaMemcpyHostToDevice));
  CREST_CUDA_TRY(cudaMemcpy(d_j.p, j_idx, sizeof(int) * npairs, cudaMemcpyHostToDevice));
  if (launch_rmsd_pairs(A, n, (const double*)d_p.p, npairs, (const int*)d_i.p, (const int*)d_j.p,
                        (double*)d_r.p, (double*)d_m.p, 0)) return 1;
  CREST_CUDA_TRY(cudaMemcpy(rmsd, d_r.p, sizeof(double) * npairs, cudaMemcpyDeviceToHost));
  CREST_CUDA_TRY(cudaMemcpy(rmsd_mirror, d_m.p, sizeof(double) * npairs, cudaMemcpyDeviceToHost));
  return 0;
}

int crest_cuda_rmsdalign(int nat, int nall, const int* at, double* xyz) {
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceGuard guard;
  if (nat <= 0 || nall < 0 || !at || (!xyz && nall)) { set_error("bad argument"); return 1; }
  Engine* E;
  if (engine(&E)) return 1;
  crest_cuda_ctx* c = E->c[0];
  CREST_CUDA_TRY(cudaSetDevice(c->device));
  if (nall <= 1) return 0;
  const size_t bytes = sizeof(double) * 3 * nat * (size_t)nall;
  DevScratch d_at, d_xyz;
  if (d_at.get(sizeof(int) * nat) || d_xyz.get(bytes)) {
    set_error("cudaMalloc failed: %s", cudaGetErrorString(cudaGetLastError()));
    return 1;
  }
  CREST_CUDA_TRY(cudaMemcpy(d_at.p, at, sizeof(int) * nat, cudaMemcpyHostToDevice));
  if (big_copy(d_xyz.p, xyz, bytes, 0, c->device)) return 1;
  if (launch_rmsdalign(nat, nall, (const int*)d_at.p, (double*)d_xyz.p, 0)) return 2;
  CREST_CUDA_TRY(cudaDeviceSynchronize());
  if (big_copy(xyz, d_xyz.p, bytes, 1, c->device)) return 1;
  return 0;
}

int crest_cuda_rmat_allpairs(int nat, int nall, const int* at, const double* xyz,
                             const int* rmsd_mask, int enantio, float* rmat) {
  std::lock_guard<std::mutex> lk(g_mu);
  DeviceGuard guard;
  if (!rmat) { set_error("rmat is NULL"); return 1; }
  Engine* E;
  if (engine(&E)) return 1;
  crest_cuda_ctx* c = E->c[0];
  std::vector<double> er((size_t)(nall > 0 ? nall : 1), 0.0);
  if (crest_cuda_ctx_upload(c, nat, nall, at, xyz, er.data(), rmsd_mask, 0)) return 1;
  if (crest_cuda_ctx_pack(c, nullptr)) return 1;
  const size_t np = (size_t)nall * (nall > 0 ? nall - 1 : 0) / 2;
  if (np == 0) return 0;
  DevScratch d_rmat;
  if (d_rmat.get(sizeof(float) * np)) { set_error("cudaMalloc failed: %s", cudaGetErrorString(cudaGetLastError())); return 1; }
  crest_cuda_thresholds t{1e300, 0.0, 0.0, 0.0, 0.0};
  if (pairs_stage(c, enantio, &t, 0, 1, 0, 0, (float*)d_rmat.p, nullptr)) return 1;
  CREST_CUDA_TRY(cudaMemcpy(rmat, d_rmat.p, sizeof(float) * np, cudaMemcpyDeviceToHost));
  return 0;
}

}  // extern "C"
