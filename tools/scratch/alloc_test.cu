#include <cuda_runtime.h>
#include <chrono>
#include <cstdio>
static double now() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
int main() {
  cudaFree(0);
  cudaStream_t s; cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking);
  size_t sz[4] = {1200u << 20, 1200u << 20, 400u << 20, 900u << 20};
  void* p[4];
  for (int rep = 0; rep < 2; ++rep) {
    double t0 = now();
    for (int i = 0; i < 4; ++i) cudaMalloc(&p[i], sz[i]);
    double t1 = now();
    for (int i = 0; i < 4; ++i) cudaMemsetAsync(p[i], 0, sz[i], s);
    cudaStreamSynchronize(s);
    double t2 = now();
    for (int i = 0; i < 4; ++i) cudaFree(p[i]);
    double t3 = now();
    printf("cudaMalloc x4 (3.7GB): %.2f ms, memset %.2f ms, cudaFree %.2f ms\n", t1 - t0, t2 - t1, t3 - t2);
  }
  cudaMemPool_t pool; cudaDeviceGetDefaultMemPool(&pool, 0);
  unsigned long long keep = ~0ull; cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
  for (int rep = 0; rep < 3; ++rep) {
    double t0 = now();
    for (int i = 0; i < 4; ++i) cudaMallocAsync(&p[i], sz[i], s);
    cudaStreamSynchronize(s);
    double t1 = now();
    for (int i = 0; i < 4; ++i) cudaMemsetAsync(p[i], 0, sz[i], s);
    cudaStreamSynchronize(s);
    double t2 = now();
    for (int i = 0; i < 4; ++i) cudaFreeAsync(p[i], s);
    cudaStreamSynchronize(s);
    double t3 = now();
    printf("cudaMallocAsync x4: %.2f ms, memset %.2f ms, free %.2f ms\n", t1 - t0, t2 - t1, t3 - t2);
  }
  // one big then small ones
  double t0 = now();
  void* big; cudaMallocAsync(&big, (size_t)8 << 30, s); cudaStreamSynchronize(s);
  double t1 = now();
  cudaFreeAsync(big, s); cudaStreamSynchronize(s);
  double t2 = now();
  for (int i = 0; i < 4; ++i) cudaMallocAsync(&p[i], sz[i], s);
  cudaStreamSynchronize(s);
  double t3 = now();
  printf("pool grow to 8GB: %.2f ms, free %.2f, 4 allocs after: %.2f ms\n", t1 - t0, t2 - t1, t3 - t2);
  return 0;
}
