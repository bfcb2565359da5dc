// Staged transfers between caller memory and one or more devices (hostcopy.cu).
#pragma once
#include <condition_variable>
#include <cstddef>
#include <mutex>
#include <vector>

namespace crest {

struct CopySeg {
  int device;       // CUDA device the segment lives on
  void* dev;        // device pointer
  void* host;       // caller memory (pageable, pinned or registered)
  size_t bytes;
  int block;        // progress unit: segments must be listed in ascending block order
};

// Progress of a staged copy, block by block: wait(k) returns once every chunk of the blocks <= k is on
// the device (H2D) / in caller memory (D2H); false if the copy failed.
struct BlockGate {
  std::mutex mu;
  std::condition_variable cv;
  std::vector<int> done;
  int need = 1;
  bool failed = false;
  void reset(int nblocks, int nthreads);
  void pass(int block);
  void fail();
  bool wait(int block);
};

// memcpy with non-temporal stores (AVX2; plain memcpy below 512 bytes or without AVX2).  Weakly ordered:
// fast_copy_fence() before another agent (DMA engine, another thread) reads the destination.
void fast_copy(void* dst, const void* src, size_t n);
void fast_copy_fence();

// dir 0: host -> device, 1: device -> host.  nthreads < 0: default (hardware threads - 2, <= 16).  Blocking.
// Returns 0 ok, 1 error (text set).  Leaves an arbitrary current device.
int staged_copy(const std::vector<CopySeg>& segs, int dir, int nthreads, int nblocks, BlockGate* gate);
// one segment on `device`; restores the current device
int big_copy(void* dst, const void* src, size_t bytes, int dir, int device);
bool host_is_pageable(const void* p);
void free_copy_slots();

}  // namespace crest
