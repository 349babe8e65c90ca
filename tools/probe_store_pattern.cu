// store-pattern floor: warp per row, nb blocks per row, P planes of nb doubles each (row-plane layout)
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void __launch_bounds__(256, 2) k(long nrows, int nb, int P, double* __restrict__ val) {
  const int lane = threadIdx.x & 31;
  const long warp = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((long)gridDim.x * blockDim.x) >> 5;
  for (long r = warp; r < nrows; r += nw) {
    const long p0 = r * nb;
    for (int u0 = 0; u0 < nb; u0 += 32) {
      const int u = u0 + lane;
      if (u < nb) {
        double* dst = val + p0 * P + u;
        for (int q = 0; q < P; ++q, dst += nb) {
          if (MODE == 0) __stcs(dst, 1.0 + q); else if (MODE == 1) *dst = 1.0 + q; else __stwt(dst, 1.0 + q);
        }
      }
    }
  }
}
int main() {
  const long nrows = 117649; const int P = 18;
  for (int nb : {119, 120, 128, 64, 343}) {
    double* v; cudaMalloc(&v, nrows * nb * P * 8);
    for (int mode = 0; mode < 3; ++mode) for (int grid : {296, 1184, 4736}) {
      cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
      for (int it = 0; it < 3; ++it) {
        cudaEventRecord(a);
        if (mode == 0) k<0><<<grid, 256>>>(nrows, nb, P, v); else if (mode == 1) k<1><<<grid, 256>>>(nrows, nb, P, v); else k<2><<<grid, 256>>>(nrows, nb, P, v);
        cudaEventRecord(b); cudaEventSynchronize(b);
      }
      float ms; cudaEventElapsedTime(&ms, a, b);
      printf("nb %3d mode %d (0 cs,1 plain,2 wt) grid %4d: %.3f ms  %.0f GB/s\n", nb, mode, grid, ms, nrows * nb * P * 8.0 / ms / 1e6);
    }
    cudaFree(v);
  }
  return 0;
}
