// vec.cuh -- device vector primitives and deterministic reductions.
#pragma once
#include "common.cuh"

namespace pyn {

// ---- block / grid reductions (fixed summation order -> run-to-run reproducible) ----
template <int NV>
__device__ __forceinline__ void block_reduce(double (&v)[NV], double* smem /* [NV*32] */) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int k = 0; k < NV; ++k)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
  __syncthreads();
  if (lane == 0)
#pragma unroll
    for (int k = 0; k < NV; ++k) smem[k * 32 + wid] = v[k];
  __syncthreads();
  if (wid == 0) {
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      double t = lane < nw ? smem[k * 32 + lane] : 0.0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
      v[k] = t;  // valid in warp 0
    }
  }
}

// Every CTA deposits NV partials; the last CTA to arrive (ticket) folds them in
// block order and writes out[0..NV).  Call with all threads of the CTA;
// v[] must hold the CTA totals in warp 0 (after block_reduce).
template <int NV>
__device__ __forceinline__ void grid_reduce_finish(double (&v)[NV], double* partials,
                                                   unsigned int* ticket, double* out) {
  __shared__ bool is_last;
  if (threadIdx.x == 0) {
#pragma unroll
    for (int k = 0; k < NV; ++k) partials[(size_t)blockIdx.x * NV + k] = v[k];
    __threadfence();
    const unsigned int t = atomicAdd(ticket, 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (wid < NV) {
      double t = 0.0;
      for (unsigned int b = lane; b < gridDim.x; b += 32) t += partials[(size_t)b * NV + wid];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
      if (lane == 0) out[wid] = t;
    }
    if (threadIdx.x == 0) *ticket = 0u;
  }
}

int vec_dot_dev(pyn_ctx* ctx, const double* x, const double* y, int64_t n, double* out_dev);
// dist: n counts the owned entries; the result is summed over ranks
int vec_dot(pyn_ctx* ctx, const double* x, const double* y, int64_t n, double* out_host, bool dist = false);
int vec_sum(pyn_ctx* ctx, const double* x, int64_t n, double* out_host, bool dist = false);
int vec_set(pyn_ctx* ctx, double* x, int64_t n, double a);
int vec_scale(pyn_ctx* ctx, double* x, int64_t n, double a);
// y = a*x + b*y
int vec_axpby(pyn_ctx* ctx, double* y, double a, const double* x, double b, int64_t n);
// w = a*x + y
int vec_waxpy(pyn_ctx* ctx, double* w, double a, const double* x, const double* y, int64_t n);
int vec_pointwise_mult(pyn_ctx* ctx, double* w, const double* x, const double* y, int64_t n);
int vec_reciprocal(pyn_ctx* ctx, double* x, int64_t n);
int vec_maxpy(pyn_ctx* ctx, double* y, int k, const double* a_host, const double* const* x_host_ptrs, int64_t n);
int vec_scatter(pyn_ctx* ctx, double* v, int64_t k, const int64_t* idx_dev, const double* vals_dev, int add);
int vec_gather(pyn_ctx* ctx, const double* v, int64_t k, const int64_t* idx_dev, double* vals_dev);

}  // namespace pyn
