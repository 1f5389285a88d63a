// K2: arg-max over classes + privatised confusion matrix.
// Contract: include/mlc.h (mlc_confusion).
#include "mlc_common.cuh"

namespace mlc {

// numpy arg-max order: a NaN beats everything, otherwise larger wins, ties
// (and NaN vs NaN) go to the lower index.
__device__ __forceinline__ bool beats(float v, int i, float bv, int bi) {
  const bool vn = v != v, bn = bv != bv;
  if (vn != bn) return vn;
  if (!vn && v != bv) return v > bv;
  return i < bi;
}

struct ConfArgs {
  const void *preds;
  const int64_t *labels;
  unsigned long long *cm;
  unsigned long long *stats;
  int32_t *argmax;
  long long n;
  int C;
};

constexpr int kSmallThreads = 256;
constexpr int kSmallMaxC = 64;

// Small C (<= 64): a CTA stages 256 rows of the [n][C] matrix through shared
// memory with coalesced loads, one thread then scans one row; counts go to a
// CTA-private C x C table in shared memory and are flushed once per CTA.
template <typename T>
__global__ void __launch_bounds__(kSmallThreads)
confusion_small(const __grid_constant__ ConfArgs a) {
  extern __shared__ uint32_t smem[];
  const int C = a.C;
  const int pitch = C | 1;  // odd pitch: row-per-thread reads hit 32 banks
  uint32_t *cm = smem;                                      // [C*C]
  float *tile = reinterpret_cast<float *>(smem + C * C);    // [256][pitch]
  for (int i = threadIdx.x; i < C * C; i += blockDim.x) cm[i] = 0;
  __syncthreads();
  const T *preds = static_cast<const T *>(a.preds);
  unsigned correct = 0, dropped = 0;
  const long long n_tiles = (a.n + kSmallThreads - 1) / kSmallThreads;
  for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    const long long r0 = t * kSmallThreads;
    const int rows = (int)min((long long)kSmallThreads, a.n - r0);
    const T *src = preds + r0 * C;
    const int elems = rows * C;
    for (int e = threadIdx.x; e < elems; e += blockDim.x) {
      const int r = e / C;
      tile[r * pitch + (e - r * C)] = to_f32<T>(src[e]);
    }
    __syncthreads();
    if ((int)threadIdx.x < rows) {
      const float *row = tile + threadIdx.x * pitch;
      float bv = row[0];
      int bi = 0;
      for (int c = 1; c < C; ++c) {
        const float v = row[c];
        if (beats(v, c, bv, bi)) { bv = v; bi = c; }
      }
      const long long y = a.labels[r0 + threadIdx.x];
      if (a.argmax) a.argmax[r0 + threadIdx.x] = bi;
      if (y >= 0 && y < C) {
        atomicAdd(&cm[(int)y * C + bi], 1u);
        correct += (y == bi);
      } else {
        ++dropped;
      }
    }
    __syncthreads();
  }
  for (int i = threadIdx.x; i < C * C; i += blockDim.x)
    if (cm[i]) atomicAdd(a.cm + i, (unsigned long long)cm[i]);
  // stats: warp reduce then one atomic per warp
  for (int o = 16; o; o >>= 1) {
    correct += __shfl_xor_sync(0xffffffffu, correct, o);
    dropped += __shfl_xor_sync(0xffffffffu, dropped, o);
  }
  if ((threadIdx.x & 31) == 0) {
    if (correct) atomicAdd(a.stats, (unsigned long long)correct);
    if (dropped) atomicAdd(a.stats + 1, (unsigned long long)dropped);
  }
}

// Large C: one warp per row, lanes stride over the classes.
template <typename T>
__global__ void __launch_bounds__(256)
confusion_wide(const __grid_constant__ ConfArgs a) {
  const int lane = threadIdx.x & 31;
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long n_warps = ((long long)gridDim.x * blockDim.x) >> 5;
  const T *preds = static_cast<const T *>(a.preds);
  unsigned correct = 0, dropped = 0;
  for (long long r = warp; r < a.n; r += n_warps) {
    const T *row = preds + r * a.C;
    float bv = 0.f;
    int bi = 0x7fffffff;  // "no candidate yet"
    for (int c = lane; c < a.C; c += 32) {
      const float v = to_f32<T>(row[c]);
      if (bi == 0x7fffffff || beats(v, c, bv, bi)) { bv = v; bi = c; }
    }
    for (int o = 16; o; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (oi != 0x7fffffff && (bi == 0x7fffffff || beats(ov, oi, bv, bi))) {
        bv = ov;
        bi = oi;
      }
    }
    if (lane == 0) {
      const long long y = a.labels[r];
      if (a.argmax) a.argmax[r] = bi;
      if (y >= 0 && y < a.C) {
        atomicAdd(a.cm + y * a.C + bi, 1ull);
        correct += (y == bi);
      } else {
        ++dropped;
      }
    }
  }
  if (lane == 0) {
    if (correct) atomicAdd(a.stats, (unsigned long long)correct);
    if (dropped) atomicAdd(a.stats + 1, (unsigned long long)dropped);
  }
}

}  // namespace mlc

using namespace mlc;

extern "C" int mlc_confusion(const void *preds_dev, int dtype, int64_t n,
                             int n_cls, const int64_t *labels_dev,
                             int64_t *cm_dev, int64_t *stats_dev,
                             int32_t *argmax_dev, void *stream) {
  MLC_REQUIRE(dtype == MLC_F16 || dtype == MLC_F32,
              "dtype %d is not MLC_F16/MLC_F32", dtype);
  MLC_REQUIRE(n >= 0 && n_cls >= 1, "bad shape n=%lld n_cls=%d", (long long)n,
              n_cls);
  MLC_REQUIRE((long long)n_cls * n_cls < (1ll << 31), "n_cls=%d too large",
              n_cls);
  MLC_REQUIRE(preds_dev && labels_dev && cm_dev && stats_dev,
              "preds/labels/cm/stats pointer is null");
  if (n == 0) return MLC_OK;
  ConfArgs a;
  a.preds = preds_dev;
  a.labels = labels_dev;
  a.cm = reinterpret_cast<unsigned long long *>(cm_dev);
  a.stats = reinterpret_cast<unsigned long long *>(stats_dev);
  a.argmax = argmax_dev;
  a.n = n;
  a.C = n_cls;
  cudaStream_t st = (cudaStream_t)stream;
  const int sms = sm_count();
  if (n_cls <= kSmallMaxC) {
    const size_t smem = (size_t)n_cls * n_cls * 4 +
                        (size_t)kSmallThreads * (n_cls | 1) * 4;
    const long long tiles = (n + kSmallThreads - 1) / kSmallThreads;
    const int grid = (int)min(tiles, 4ll * sms);
    if (smem > 48 * 1024) {
      if (dtype == MLC_F16)
        MLC_CUDA(cudaFuncSetAttribute(
            confusion_small<__half>,
            cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      else
        MLC_CUDA(cudaFuncSetAttribute(
            confusion_small<float>,
            cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    if (dtype == MLC_F16)
      confusion_small<__half><<<grid, kSmallThreads, smem, st>>>(a);
    else
      confusion_small<float><<<grid, kSmallThreads, smem, st>>>(a);
  } else {
    const long long blocks = (n + 7) / 8;  // 8 warps (rows) per CTA sweep
    const int grid = (int)min(blocks, 8ll * sms);
    if (dtype == MLC_F16)
      confusion_wide<__half><<<grid, 256, 0, st>>>(a);
    else
      confusion_wide<float><<<grid, 256, 0, st>>>(a);
  }
  MLC_LAUNCH_CHECK("confusion");
  return MLC_OK;
}
