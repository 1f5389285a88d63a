// Probe: which tensor-map shapes / coordinates the copy engine accepts.
// usage: tma_probe <mode> <box> <c0> [rank]
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
struct Maps { CUtensorMap m[2]; };
__global__ void k(const __grid_constant__ Maps mp, int which, int c0, int c1, int bytes, uint16_t *out, int rank) {
  extern __shared__ __align__(128) unsigned char sm[];
  __shared__ __align__(8) unsigned long long bar;
  uint32_t bar_sa = (uint32_t)__cvta_generic_to_shared(&bar);
  uint32_t dst = (uint32_t)__cvta_generic_to_shared(sm);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_sa));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_sa), "r"(bytes) : "memory");
    const CUtensorMap *map = &mp.m[which];
    if (rank == 1)
      asm volatile("cp.async.bulk.tensor.1d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2}], [%3];"
        ::"r"(dst), "l"((uint64_t)map), "r"(c0), "r"(bar_sa) : "memory");
    else if (rank == 2)
      asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(dst), "l"((uint64_t)map), "r"(c0), "r"(c1), "r"(bar_sa) : "memory");
    else
      asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(dst), "l"((uint64_t)map), "r"(c0), "r"(0), "r"(c1), "r"(bar_sa) : "memory");
  }
  asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(bar_sa) : "memory");
  for (int i = threadIdx.x; i < bytes / 2; i += blockDim.x) out[i] = ((uint16_t *)sm)[i];
}
typedef CUresult (*Enc)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main(int argc, char **argv) {
  int rank = argc > 1 ? atoi(argv[1]) : 2;
  int box = argc > 2 ? atoi(argv[2]) : 192;
  int c0 = argc > 3 ? atoi(argv[3]) : 0;
  int rows = argc > 4 ? atoi(argv[4]) : 1;     // rank 3: box1 rows of `box` elements, row stride box*2 bytes (overlapping dim0)
  long long N = 1470000;
  std::vector<uint16_t> h(2 * N);
  for (long long i = 0; i < 2 * N; ++i) h[i] = (uint16_t)(i * 7 + 1);
  uint16_t *d, *out;
  cudaMalloc(&d, 2 * N * 2); cudaMalloc(&out, 65536);
  cudaMemcpy(d, h.data(), 2 * N * 2, cudaMemcpyHostToDevice);
  void *p = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  Enc enc = (Enc)p;
  Maps mp;
  CUresult r;
  if (rank == 1) {
    cuuint64_t dims[1] = {(cuuint64_t)N}; cuuint64_t str[1] = {0}; cuuint32_t bx[1] = {(cuuint32_t)box}; cuuint32_t es[1] = {1};
    r = enc(&mp.m[0], CU_TENSOR_MAP_DATA_TYPE_UINT16, 1, d, dims, str, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  } else if (rank == 2) {
    cuuint64_t dims[2] = {(cuuint64_t)N, 2}; cuuint64_t str[1] = {(cuuint64_t)N * 2}; cuuint32_t bx[2] = {(cuuint32_t)box, 1}; cuuint32_t es[2] = {1, 1};
    r = enc(&mp.m[0], CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, d, dims, str, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  } else {
    // dim0: flat elements of a group; dim1: chunks of `box` elements, stride box * 2 bytes (overlaps dim0); dim2: groups
    cuuint64_t dims[3] = {(cuuint64_t)N, (cuuint64_t)rows, 2}; cuuint64_t str[2] = {(cuuint64_t)box * 2, (cuuint64_t)N * 2};
    cuuint32_t bx[3] = {(cuuint32_t)box, (cuuint32_t)rows, 1}; cuuint32_t es[3] = {1, 1, 1};
    r = enc(&mp.m[0], CU_TENSOR_MAP_DATA_TYPE_UINT16, 3, d, dims, str, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  printf("rank %d box %d c0 %d rows %d: encode rc=%d\n", rank, box, c0, rows, (int)r);
  if (r != CUDA_SUCCESS) return 1;
  mp.m[1] = mp.m[0];
  int bytes = box * 2 * (rank == 3 ? rows : 1);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  k<<<1, 128, 65536>>>(mp, 1, c0, 1, bytes, out, rank);
  cudaError_t e = cudaDeviceSynchronize();
  printf("  kernel: %s\n", cudaGetErrorString(e));
  if (e != cudaSuccess) return 2;
  std::vector<uint16_t> o(bytes / 2);
  cudaMemcpy(o.data(), out, bytes, cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int i = 0; i < bytes / 2; ++i) {
    long long src = N + c0 + i;   // group 1 (c1 = 1)
    uint16_t want = (c0 + i < 0 || c0 + i >= N) ? 0 : h[src];
    if (o[i] != want) { if (bad < 4) printf("  mismatch at %d: %u want %u\n", i, o[i], want); ++bad; }
  }
  printf("  %s (%d mismatches of %d)\n", bad ? "WRONG" : "ok", bad, bytes / 2);
  return 0;
}
