// noalloc.cu -- pyn_vec_dot_noalloc: the inner product of the Krylov loop for callers that may not
// allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// No counterpart in the reference.  It is a variant of
// dot_kernel (vec.cu), whose per-CTA partial sums and arrival ticket live in the context's
// cudaMalloc'ed reduction scratch; here the scratch is a file-scope device array of fixed size, so
// the call path contains no cudaMalloc / cudaFree / allocating library routine at all.
//
// Scratch: kChunks partial sums (one per chunk of kChunk entries) + one 256-byte slot for the ticket
//   = 16384 * 8 B + 256 B = 131 328 B per GPU that loads the library, held until the process exits.
// Limit: n <= kChunks * kChunk = 268 435 456 entries (PYN_NOALLOC_MAX_N); larger sizes are refused.
// Every call on a device shares the one array and it keeps the last call's contents: the ticket is
// cleared on the caller's stream before each launch, the partials are overwritten before they are
// read.  Two calls that overlap on the GPU (different streams, no ordering between them) RACE on it.
//
// One partial per fixed chunk (not per resident CTA) also makes the result independent of the grid
// size and of the device's SM count: sums are formed in chunk order.
#include <mutex>

#include "vec.cuh"

namespace pyn {

constexpr int kChunk = 16384;    // entries per CTA: 256 threads x 64
constexpr int kChunks = 16384;   // partial sums the scratch holds
constexpr int kTicketSlot = 32;  // doubles (256 B) after the partials, first 4 bytes = ticket
static_assert((int64_t)kChunk * kChunks == PYN_NOALLOC_MAX_N, "header and scratch size disagree");

// In host code this name is a host-side stand-in, NOT a device address: the address the kernel
// gets comes from cudaGetSymbolAddress (noalloc_scratch below).
static __device__ __align__(256) double g_noalloc_scratch[kChunks + kTicketSlot];

__global__ void __launch_bounds__(256)
dot_chunk_kernel(const double* __restrict__ x, const double* __restrict__ y, int64_t n,
                 double* scratch, double* __restrict__ out) {
  __shared__ double sm[32];
  __shared__ bool is_last;
  double* partials = scratch;
  unsigned int* ticket = reinterpret_cast<unsigned int*>(scratch + kChunks);
  const int64_t base = (int64_t)blockIdx.x * kChunk;
  const int64_t end = base + kChunk < n ? base + kChunk : n;
  double v[1] = {0.0};
  for (int64_t i = base + threadIdx.x; i < end; i += 256) v[0] += x[i] * y[i];
  block_reduce<1>(v, sm);
  if (threadIdx.x == 0) {
    partials[blockIdx.x] = v[0];
    __threadfence();
    is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  // last CTA to arrive: partials in chunk order, strided over the threads (fixed order)
  double t[1] = {0.0};
  for (unsigned int b = threadIdx.x; b < gridDim.x; b += 256) t[0] += __ldcg(partials + b);
  block_reduce<1>(t, sm);
  if (threadIdx.x == 0) {
    out[0] = t[0];
    *ticket = 0u;
  }
}

// device address of the scratch, looked up once per device with that device current
static int noalloc_scratch(int device, double** out) {
  constexpr int kMaxDevices = 64;
  static double* ptr[kMaxDevices] = {};
  static std::mutex mu;
  PYN_REQUIRE(device >= 0 && device < kMaxDevices, "pyn_vec_dot_noalloc: device %d out of range", device);
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[device]) {
    double* p = nullptr;
    PYN_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
    PYN_REQUIRE(p, "pyn_vec_dot_noalloc: cudaGetSymbolAddress returned a null scratch address");
    ptr[device] = p;
  }
  *out = ptr[device];
  return PYN_OK;
}

}  // namespace pyn

using namespace pyn;

extern "C" int pyn_vec_dot_noalloc(pyn_ctx* ctx, const double* x, const double* y, int64_t n, double* out,
                                   void* stream) {
  PYN_REQUIRE(ctx && out, "pyn_vec_dot_noalloc: null context or output");
  PYN_REQUIRE(n >= 0, "pyn_vec_dot_noalloc: negative size %lld", (long long)n);
  PYN_REQUIRE(n <= (int64_t)PYN_NOALLOC_MAX_N,
              "pyn_vec_dot_noalloc: %lld entries exceed the fixed scratch (PYN_NOALLOC_MAX_N = %lld); "
              "use pyn_vec_dot", (long long)n, (long long)PYN_NOALLOC_MAX_N);
  PYN_REQUIRE(n == 0 || (x && y), "pyn_vec_dot_noalloc: null input");
  int device = -1;
  PYN_CUDA(cudaGetDevice(&device));
  PYN_REQUIRE(device == ctx->device, "pyn_vec_dot_noalloc: current device %d is not the context's (%d)",
              device, ctx->device);
  double* scratch = nullptr;
  PYN_CHECK(noalloc_scratch(device, &scratch));
  cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
  // the array keeps whatever the last call (or an aborted one) left: the ticket must start at zero
  PYN_CUDA(cudaMemsetAsync(scratch + kChunks, 0, sizeof(unsigned int), s));
  const int grid = n == 0 ? 1 : (int)((n + kChunk - 1) / kChunk);
  dot_chunk_kernel<<<grid, 256, 0, s>>>(x, y, n, scratch, out);
  PYN_LAUNCHED(ctx);
  return PYN_OK;
}
