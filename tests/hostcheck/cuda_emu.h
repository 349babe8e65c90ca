// cuda_emu.h -- TEST-ONLY: runs CUDA kernels written against a small subset of the device API on
// CPU threads, so that kernels can be checked in the build container (no GPU there).  One CTA at
// a time, one std::thread per CUDA thread; __syncthreads is a std::barrier over the CTA, warp
// shuffles exchange through a per-warp slot array guarded by a per-warp barrier (so a shuffle
// that not all 32 lanes reach hangs instead of passing silently).  Include BEFORE the kernel
// headers, compile with g++ -std=c++20 -pthread (not nvcc).  Not part of the product.
#pragma once
#include <cuda_runtime.h>   // vector types, host_defines (the __global__ ... attributes are ignored by g++)
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <barrier>
#include <memory>
#include <thread>
#include <vector>

#undef __shared__
#define __shared__ static
#ifndef __launch_bounds__
#define __launch_bounds__(...)
#endif
#define PYN_CUDA_EMU 1

namespace cuda_emu {

struct Cta {
  explicit Cta(int nthreads)
      : bar(nthreads), slots((size_t)((nthreads + 31) / 32) * 32, 0ull) {
    const int nw = (nthreads + 31) / 32;
    for (int w = 0; w < nw; ++w) {
      const int lanes = nthreads - w * 32 < 32 ? nthreads - w * 32 : 32;
      wbar.emplace_back(new std::barrier<>(lanes));
    }
  }
  std::barrier<> bar;
  std::vector<std::unique_ptr<std::barrier<>>> wbar;
  std::vector<unsigned long long> slots;
};

struct ThreadState {
  uint3 tid{0, 0, 0}, bid{0, 0, 0};
  dim3 bdim{1, 1, 1}, gdim{1, 1, 1};
  Cta* cta = nullptr;
  void* dyn_smem = nullptr;
};
inline thread_local ThreadState ts;

template <typename T>
inline T shfl(T v, int src_lane) {
  static_assert(sizeof(T) <= 8, "shuffle of at most 8 bytes");
  const int t = (int)ts.tid.x, w = t >> 5;
  unsigned long long bits = 0;
  memcpy(&bits, &v, sizeof(T));
  ts.cta->slots[(size_t)t] = bits;
  ts.cta->wbar[(size_t)w]->arrive_and_wait();
  const unsigned long long got = ts.cta->slots[(size_t)(w * 32 + (src_lane & 31))];
  ts.cta->wbar[(size_t)w]->arrive_and_wait();
  T out;
  memcpy(&out, &got, sizeof(T));
  return out;
}

// launch<<<grid, block, smem>>>: CTAs one after the other (a grid-wide ticket still works: the last
// CTA to run is the last to arrive)
template <typename F>
inline void launch(int grid, int block, size_t smem_bytes, F&& kernel_call) {
  std::vector<unsigned long long> smem((smem_bytes + 7) / 8 + 2, 0ull);
  for (int b = 0; b < grid; ++b) {
    Cta cta(block);
    std::vector<std::thread> th;
    th.reserve((size_t)block);
    for (int t = 0; t < block; ++t)
      th.emplace_back([&, t, b]() {
        ts.tid = uint3{(unsigned)t, 0, 0};
        ts.bid = uint3{(unsigned)b, 0, 0};
        ts.bdim = dim3((unsigned)block, 1, 1);
        ts.gdim = dim3((unsigned)grid, 1, 1);
        ts.cta = &cta;
        ts.dyn_smem = smem.data();
        kernel_call();
      });
    for (auto& x : th) x.join();
  }
}

}  // namespace cuda_emu

#define threadIdx (cuda_emu::ts.tid)
#define blockIdx (cuda_emu::ts.bid)
#define blockDim (cuda_emu::ts.bdim)
#define gridDim (cuda_emu::ts.gdim)

inline void __syncthreads() { cuda_emu::ts.cta->bar.arrive_and_wait(); }
inline void __syncwarp(unsigned = 0xffffffffu) {
  cuda_emu::ts.cta->wbar[(size_t)(cuda_emu::ts.tid.x >> 5)]->arrive_and_wait();
}
inline void __threadfence() { __atomic_thread_fence(__ATOMIC_SEQ_CST); }
template <typename T>
inline T __shfl_xor_sync(unsigned, T v, int lane_mask) {
  return cuda_emu::shfl(v, (int)(cuda_emu::ts.tid.x & 31) ^ lane_mask);
}
template <typename T>
inline T __shfl_sync(unsigned, T v, int src) { return cuda_emu::shfl(v, src); }
inline unsigned atomicAdd(unsigned* p, unsigned v) { return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST); }
inline unsigned long long atomicAdd(unsigned long long* p, unsigned long long v) {
  return __atomic_fetch_add(p, v, __ATOMIC_SEQ_CST);
}
inline unsigned long long atomicMax(unsigned long long* p, unsigned long long v) {
  unsigned long long old = __atomic_load_n(p, __ATOMIC_SEQ_CST);
  while (old < v && !__atomic_compare_exchange_n(p, &old, v, false, __ATOMIC_SEQ_CST, __ATOMIC_SEQ_CST)) {
  }
  return old;
}
inline long long __double_as_longlong(double v) {
  long long b;
  memcpy(&b, &v, 8);
  return b;
}
