// common.cuh -- context, error plumbing, device buffers shared by all kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <cmath>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/pynama_b200.h"

#ifndef PYN_HD
#ifdef __CUDACC__
#define PYN_HD __host__ __device__ __forceinline__
#else
#define PYN_HD inline
#endif
#endif

namespace pyn {

void set_error(const char* fmt, ...);

#define PYN_CUDA(call)                                                                 \
  do {                                                                                 \
    cudaError_t _e = (call);                                                           \
    if (_e != cudaSuccess) {                                                           \
      pyn::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(_e)); \
      return PYN_E_CUDA;                                                               \
    }                                                                                  \
  } while (0)

#define PYN_CHECK(call)         \
  do {                          \
    int _r = (call);            \
    if (_r != PYN_OK) return _r; \
  } while (0)

#define PYN_REQUIRE(cond, ...)      \
  do {                              \
    if (!(cond)) {                  \
      pyn::set_error(__VA_ARGS__);  \
      return PYN_E_INVALID;         \
    }                               \
  } while (0)

// after a kernel launch
#define PYN_LAUNCHED(ctx)             \
  do {                                \
    (ctx)->launches++;                \
    PYN_CUDA(cudaGetLastError());     \
  } while (0)

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs
extern int g_ksp_check_every;
extern bool g_force_idx64;
extern bool g_asm_coarse_unfused;
extern bool g_halo_wait_at_start;
extern bool g_ksp_adaptive_chunks;
extern bool g_ksp_graphs;

}  // namespace pyn

struct pyn_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  int64_t launches = 0;
  int64_t bytes_in_use = 0;
  int num_sms = pyn::kNumSMs;
  size_t smem_optin = 0;
  // reduction scratch: per-CTA partials + ticket + pinned result
  double* red_partials = nullptr;   // [kRedMaxBlocks * kRedMaxVals]
  unsigned int* red_ticket = nullptr;
  double* red_result_dev = nullptr; // [kRedMaxVals]
  double* red_result_host = nullptr;  // pinned
  void* flush_buf = nullptr;
  size_t flush_bytes = 0;
  // multi-GPU (comm.cu): one process per GPU
  int rank = 0, nranks = 1;
  void* nccl_comm = nullptr;
  int64_t collectives = 0;
  // peer-memory reduction windows (comm.cuh): own + peers' (CUDA IPC), sequence counter
  void* red_win = nullptr;
  void* red_peer[8] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
  bool p2p_on = false;
  unsigned long long p2p_seq = 0;
  int* p2p_timeout = nullptr;         // device flag: a bounded spin on a peer flag ran out
  // per-kernel-class CUDA-event profiler (bench.py's live roofline timing)
  bool prof_on = false;
  std::vector<cudaEvent_t> prof_ev;     // pairs (begin, end)
  std::vector<int> prof_cls;            // class of pending pair i
  int prof_pending = 0;
  double prof_ms[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  int64_t prof_count[8] = {0, 0, 0, 0, 0, 0, 0, 0};
};

namespace pyn {

constexpr int kRedMaxBlocks = 1184;  // 8 CTAs per SM
constexpr int kRedMaxVals = 8;

// Device buffer with byte accounting on the owning context.
template <typename T>
struct DevBuf {
  static constexpr size_t kPoolMaxBytes = (size_t)64 << 20;
  T* p = nullptr;
  int64_t n = 0;
  bool pooled = false;
  pyn_ctx* ctx = nullptr;
  DevBuf() {}
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  ~DevBuf() { release(); }
  int alloc(pyn_ctx* c, int64_t count) {
    release();
    ctx = c;
    n = count;
    if (count <= 0) return PYN_OK;
    // Small buffers (vectors, Krylov work, scatter plans: anything the time loop creates and
    // drops) come from the device's stream-ordered pool, whose release threshold pyn_init raised:
    // a freed block is reused by the next request and neither call synchronises the device.
    // Large one-time buffers (matrices, element block stores) use cudaMalloc: growing the pool
    // costs ~16-90 ms per GB on B200 against ~4 ms per GB for cudaMalloc (measured).
    const size_t bytes = sizeof(T) * (size_t)count;
    pooled = bytes < kPoolMaxBytes;
    cudaError_t e = pooled ? cudaMallocAsync((void**)&p, bytes, ctx->stream) : cudaMalloc((void**)&p, bytes);
    if (e != cudaSuccess) {
      p = nullptr;
      n = 0;
      set_error("device allocation of %lld bytes failed: %s", (long long)bytes, cudaGetErrorString(e));
      return PYN_E_NOMEM;
    }
    ctx->bytes_in_use += (int64_t)sizeof(T) * count;
    return PYN_OK;
  }
  int zero() {
    if (n > 0) PYN_CUDA(cudaMemsetAsync(p, 0, sizeof(T) * (size_t)n, ctx->stream));
    return PYN_OK;
  }
  int upload(const T* host, int64_t count) {
    if (count > 0)
      PYN_CUDA(cudaMemcpyAsync(p, host, sizeof(T) * (size_t)count, cudaMemcpyHostToDevice, ctx->stream));
    return PYN_OK;
  }
  int download(T* host, int64_t count) const {
    if (count > 0) {
      PYN_CUDA(cudaMemcpyAsync(host, p, sizeof(T) * (size_t)count, cudaMemcpyDeviceToHost, ctx->stream));
      PYN_CUDA(cudaStreamSynchronize(ctx->stream));
    }
    return PYN_OK;
  }
  void release() {
    if (p) {
      if (ctx && pooled)
        cudaFreeAsync(p, ctx->stream);
      else
        cudaFree(p);
      if (ctx) ctx->bytes_in_use -= (int64_t)sizeof(T) * n;
    }
    p = nullptr;
    n = 0;
  }
  void swap(DevBuf& o) {
    std::swap(pooled, o.pooled);
    std::swap(p, o.p);
    std::swap(n, o.n);
    std::swap(ctx, o.ctx);
  }
};

inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

// grid size for a grid-stride kernel: a multiple of the SM count
inline int grid_for(const pyn_ctx* ctx, int64_t work_items, int block, int ctas_per_sm = 8) {
  int64_t need = ceil_div(work_items, block);
  int64_t cap = (int64_t)ctx->num_sms * ctas_per_sm;
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

// Kernel-class profiler: brackets one launch with a CUDA-event pair on the
// context's stream when enabled (pyn_prof_enable); no-ops otherwise.
enum ProfClass { PROF_SPMV_KSP = 0, PROF_SPMV = 1, PROF_CG_UPDATE = 2, PROF_CG_DIR = 3,
                 PROF_ELEM_KLE = 4, PROF_NUMERIC = 5, PROF_COUNT = 8 };
void prof_begin(pyn_ctx* ctx, int cls);
void prof_end(pyn_ctx* ctx);
void prof_drain(pyn_ctx* ctx);

// exclusive scan of int64 counts on the device (single CTA chained; symbolic phase only)
int exclusive_scan_i64(pyn_ctx* ctx, const int64_t* in, int64_t* out, int64_t n, int64_t* total_host);
int exclusive_scan_i32_to_i64(pyn_ctx* ctx, const int32_t* in, int64_t* out, int64_t n, int64_t* total_host);

}  // namespace pyn

struct pyn_vec {
  pyn_ctx* ctx = nullptr;
  int64_t n = 0;
  // row-partitioned Vec: entries [0, owned) belong to this rank (reductions run over them and
  // are summed over ranks), [owned, n) are ghost copies.  owned < 0: not distributed.
  int64_t owned = -1;
  pyn::DevBuf<double> d;
};
