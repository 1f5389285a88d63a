// Probe: throughput of 1-D bulk copies (global -> shared) by copy size,
// source alignment and number of issuing lanes.
// usage: bulk_rate
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, 100000;\n@!p bra W_%=;\n}\n" ::"r"(bar), "r"(parity) : "memory");
}
// each CTA: `depth` copies in flight (ring of depth buffers of `size` bytes), total `n` copies, one thread issues
__global__ void k(const char *src, size_t span, int size, int depth, int n, int misalign, int stride) {
  extern __shared__ __align__(128) unsigned char sm[];
  __shared__ __align__(8) unsigned long long bars[64];
  uint32_t bar0 = (uint32_t)__cvta_generic_to_shared(bars);
  uint32_t dst0 = (uint32_t)__cvta_generic_to_shared(sm);
  if (threadIdx.x == 0) {
    for (int i = 0; i < depth; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8 * i));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    size_t off = ((size_t)blockIdx.x * 7919 * 4096) % (span - (size_t)n * stride - 65536);
    off = off / 128 * 128 + misalign;
    for (int i = 0; i < n; ++i) {
      int s = i % depth;
      if (i >= depth) mbar_wait(bar0 + 8 * s, ((i / depth) - 1) & 1);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar0 + 8 * s), "r"(size) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(dst0 + s * ((size + 127) / 128 * 128)), "l"(src + off + (size_t)i * stride), "r"(size), "r"(bar0 + 8 * s) : "memory");
    }
    for (int i = n; i < n + depth; ++i) { int s = i % depth; if (i >= depth) mbar_wait(bar0 + 8 * s, ((i / depth) - 1) & 1); }
  }
}
int main() {
  size_t span = 8ull << 30;
  char *d; cudaMalloc(&d, span); cudaMemset(d, 1, span);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int sizes[] = {512, 1024, 1072, 2048, 4096, 8192, 16384};
  for (int mis = 0; mis <= 16; mis += 16)
    for (int size : sizes) {
      int depth = 96 * 1024 / ((size + 127) / 128 * 128); if (depth > 64) depth = 64;
      int n = (int)((64ull << 20) / size);  // 64 MiB per CTA
      if (n > 60000) n = 60000;
      int stride = size >= 1072 && size < 2048 ? 1056 : size;
      k<<<296, 32, 100 * 1024>>>(d, span, size, depth, 64, mis, stride);
      cudaEventRecord(e0);
      k<<<296, 32, 100 * 1024>>>(d, span, size, depth, n, mis, stride);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      double gb = 296.0 * n * size / 1e9;
      printf("size %6d misalign %2d depth %2d: %8.1f GB/s  (%.1f ns per copy per CTA) %s\n", size, mis, depth, gb / (ms * 1e-3), ms * 1e6 / n, cudaGetErrorString(cudaGetLastError()));
    }
  return 0;
}
