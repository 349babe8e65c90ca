// scan.cu -- device exclusive scan (int32/int64 counts -> int64 offsets).
// Used only by the symbolic phase (row pointers, contribution pointers).
#include "common.cuh"

namespace pyn {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 16;
constexpr int kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ int64_t block_exclusive_scan(int64_t v, int64_t* total) {
  __shared__ int64_t warp_sums[kScanThreads / 32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int64_t incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int64_t t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) warp_sums[wid] = incl;
  __syncthreads();
  if (wid == 0) {
    int64_t w = lane < kScanThreads / 32 ? warp_sums[lane] : 0;
    int64_t wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int64_t t = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += t;
    }
    if (lane < kScanThreads / 32) warp_sums[lane] = wi - w;  // exclusive warp offsets
    if (lane == kScanThreads / 32 - 1) *total = wi;
  }
  __syncthreads();
  int64_t r = incl - v + warp_sums[wid];
  __syncthreads();
  return r;
}

template <typename Tin>
__global__ void __launch_bounds__(kScanThreads)
scan_tile_reduce(const Tin* __restrict__ in, int64_t n, int64_t* __restrict__ tile_sums) {
  __shared__ int64_t tot;
  const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
  int64_t s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i)
    if (base + i < n) s += (int64_t)in[base + i];
  block_exclusive_scan(s, &tot);
  if (threadIdx.x == 0) tile_sums[blockIdx.x] = tot;
}

template <typename Tin>
__global__ void __launch_bounds__(kScanThreads)
scan_tile_apply(const Tin* __restrict__ in, int64_t n, const int64_t* __restrict__ tile_offsets,
                int64_t* __restrict__ out) {
  __shared__ int64_t tot;
  const int64_t base = (int64_t)blockIdx.x * kScanTile + (int64_t)threadIdx.x * kScanItems;
  int64_t v[kScanItems];
  int64_t s = 0;
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    v[i] = (base + i < n) ? (int64_t)in[base + i] : 0;
    s += v[i];
  }
  int64_t off = block_exclusive_scan(s, &tot) + (tile_offsets ? tile_offsets[blockIdx.x] : 0);
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    if (base + i < n) out[base + i] = off;
    off += v[i];
  }
  // the entry one past the end carries the grand total
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == kScanThreads - 1) {
    out[n] = (tile_offsets ? tile_offsets[blockIdx.x] : 0) + tot;
  }
}

template <typename Tin>
static int scan_impl(pyn_ctx* ctx, const Tin* in, int64_t* out, int64_t n) {
  if (n <= 0) {
    PYN_CUDA(cudaMemsetAsync(out, 0, sizeof(int64_t), ctx->stream));
    return PYN_OK;
  }
  const int64_t ntiles = ceil_div(n, kScanTile);
  if (ntiles == 1) {
    scan_tile_apply<Tin><<<1, kScanThreads, 0, ctx->stream>>>(in, n, nullptr, out);
    PYN_LAUNCHED(ctx);
    return PYN_OK;
  }
  DevBuf<int64_t> sums, offs;
  PYN_CHECK(sums.alloc(ctx, ntiles));
  PYN_CHECK(offs.alloc(ctx, ntiles + 1));
  scan_tile_reduce<Tin><<<(unsigned)ntiles, kScanThreads, 0, ctx->stream>>>(in, n, sums.p);
  PYN_LAUNCHED(ctx);
  PYN_CHECK(scan_impl<int64_t>(ctx, sums.p, offs.p, ntiles));
  scan_tile_apply<Tin><<<(unsigned)ntiles, kScanThreads, 0, ctx->stream>>>(in, n, offs.p, out);
  PYN_LAUNCHED(ctx);
  // temporaries are freed by cudaFree, which synchronises with the device
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  return PYN_OK;
}

static int fetch_total(pyn_ctx* ctx, const int64_t* out, int64_t n, int64_t* total_host) {
  if (!total_host) return PYN_OK;
  PYN_CUDA(cudaMemcpyAsync(total_host, out + n, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  PYN_CUDA(cudaStreamSynchronize(ctx->stream));
  return PYN_OK;
}

int exclusive_scan_i64(pyn_ctx* ctx, const int64_t* in, int64_t* out, int64_t n, int64_t* total_host) {
  PYN_CHECK(scan_impl<int64_t>(ctx, in, out, n));
  return fetch_total(ctx, out, n < 0 ? 0 : n, total_host);
}

int exclusive_scan_i32_to_i64(pyn_ctx* ctx, const int32_t* in, int64_t* out, int64_t n,
                              int64_t* total_host) {
  PYN_CHECK(scan_impl<int32_t>(ctx, in, out, n));
  return fetch_total(ctx, out, n < 0 ? 0 : n, total_host);
}

}  // namespace pyn
