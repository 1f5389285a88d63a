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
constexpr int kSmallMaxC = 64;    // CTA-private C x C table in shared memory
constexpr int kRowsMaxC = 256;    // thread-per-row staging (table in global)

// Small C (<= 64): a WARP stages 32 consecutive rows - one contiguous,
// 16-byte aligned block of the [n][C] matrix - through its own slice of
// shared memory with 128-bit loads (no CTA-wide barrier), then lane r scans
// row r.  The scan starts at class (r mod C) and wraps, which spreads the
// lanes over the banks for any C (beats() compares indices, so the order of
// the scan does not matter for ties).  Counts go to a CTA-private C x C table
// in shared memory, flushed once per CTA.  VEC: preds is 16-byte aligned.
// CM: the C x C table is CTA-private in shared memory (C <= 64); otherwise
// (C <= 256) the counts go straight to the global table - one 64-bit atomic
// per row, spread over C*C addresses - and the CTA runs as many warps as
// 32 x C fp32 tiles fit in shared memory.
template <typename T, bool VEC, bool CM>
__global__ void __launch_bounds__(kSmallThreads)
confusion_small(const __grid_constant__ ConfArgs a) {
  extern __shared__ __align__(16) uint32_t smem[];
  const int C = a.C;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int tile_words = (32 * C + 3) & ~3;  // 16-byte multiple per warp
  uint32_t *cm = smem;                       // [C*C] (padded to 16 bytes)
  float *tile = reinterpret_cast<float *>(smem + (CM ? (C * C + 3) & ~3 : 0)) +
                warp * tile_words;           // [32][C] fp32, this warp's
  if (CM) {
    for (int i = threadIdx.x; i < C * C; i += blockDim.x) cm[i] = 0;
    __syncthreads();
  }
  const int wpc = blockDim.x >> 5;           // warps per CTA
  const T *preds = static_cast<const T *>(a.preds);
  unsigned correct = 0, dropped = 0;
  const long long n_blocks = (a.n + 31) >> 5;
  const long long warps = (long long)gridDim.x * wpc;
  for (long long rb = (long long)blockIdx.x * wpc + warp;
       rb < n_blocks; rb += warps) {
    const long long r0 = rb << 5;
    const int rows = (int)min(32ll, a.n - r0);
    const int elems = rows * C;
    const T *src = preds + r0 * C;
    // the label load is issued before the tile is touched
    long long y = -1;
    if (lane < rows) y = a.labels[r0 + lane];
    constexpr int PER = 16 / sizeof(T);  // elements per 128-bit load
    const int nvec = VEC ? elems / PER : 0;
    for (int v = lane; v < nvec; v += 32) {
      const uint4 q = ldg_stream16(src + v * PER);
      if (sizeof(T) == 4) {
        *reinterpret_cast<uint4 *>(tile + v * 4) = q;
      } else {
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
        float f[8];
        unpack8(w, f);
        *reinterpret_cast<float4 *>(tile + v * 8) =
            make_float4(f[0], f[1], f[2], f[3]);
        *reinterpret_cast<float4 *>(tile + v * 8 + 4) =
            make_float4(f[4], f[5], f[6], f[7]);
      }
    }
    for (int e = nvec * PER + lane; e < elems; e += 32)
      tile[e] = to_f32<T>(src[e]);
    __syncwarp();
    if (lane < rows) {
      const float *row = tile + lane * C;
      int c = lane % C;
      float bv = row[c];
      int bi = c;
      for (int j = 1; j < C; ++j) {
        c = c + 1 == C ? 0 : c + 1;
        const float v = row[c];
        if (beats(v, c, bv, bi)) { bv = v; bi = c; }
      }
      if (a.argmax) a.argmax[r0 + lane] = bi;
      if (y >= 0 && y < C) {
        if (CM) atomicAdd(&cm[(int)y * C + bi], 1u);
        else atomicAdd(a.cm + y * C + bi, 1ull);
        correct += (y == bi);
      } else {
        ++dropped;
      }
    }
    __syncwarp();
  }
  if (CM) {
    __syncthreads();
    for (int i = threadIdx.x; i < C * C; i += blockDim.x)
      if (cm[i]) atomicAdd(a.cm + i, (unsigned long long)cm[i]);
  }
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
    if (sizeof(T) == 2 && (a.C & 1) == 0 &&
        (reinterpret_cast<uintptr_t>(preds) & 3) == 0) {
      // fp16 rows of even length start on a word: two classes per load, eight
      // loads per lane issued before the first compare
      const uint32_t *row2 = reinterpret_cast<const uint32_t *>(row);
      const int C2 = a.C >> 1;
      for (int c0 = lane; c0 < C2; c0 += 256) {
        uint32_t w[8];
#pragma unroll
        for (int u = 0; u < 8; ++u)
          w[u] = c0 + 32 * u < C2 ? ldg_nc32(row2 + c0 + 32 * u) : 0u;
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int c = 2 * (c0 + 32 * u);
          if (c < a.C) {
            const float2 f =
                __half22float2(*reinterpret_cast<const __half2 *>(&w[u]));
            if (bi == 0x7fffffff || beats(f.x, c, bv, bi)) { bv = f.x; bi = c; }
            if (beats(f.y, c + 1, bv, bi)) { bv = f.y; bi = c + 1; }
          }
        }
      }
    } else {  // (batching eight loads here measured slower: 1.22 vs 1.03 ms)
#pragma unroll 4
      for (int c = lane; c < a.C; c += 32) {
        const float v = to_f32<T>(row[c]);
        if (bi == 0x7fffffff || beats(v, c, bv, bi)) { bv = v; bi = c; }
      }
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
#ifdef MLC_BOUNDS_CHECK
  {
    BoundsList bl;
    bl.add(preds_dev, (size_t)n * n_cls * (dtype == MLC_F16 ? 2 : 4));
    MLC_CUDA(bounds_commit(bl, st));
  }
#endif
  const int sms = sm_count();
  if (n_cls <= kRowsMaxC) {
    const bool cmem = n_cls <= kSmallMaxC;
    const size_t tile_b = (size_t)((32 * n_cls + 3) / 4 * 4) * 4;
    const size_t cm_b = cmem ? (size_t)((n_cls * n_cls + 3) / 4 * 4) * 4 : 0;
    // warps per CTA: 8, fewer when the tiles are large
    const int wpc = (int)max((size_t)1, min((size_t)(kSmallThreads / 32),
                                            (96 * 1024 - cm_b) / tile_b));
    const size_t smem = cm_b + wpc * tile_b;
    const long long ctas = ((n + 31) / 32 + wpc - 1) / wpc;
    const int per_sm = (int)max((size_t)1, min((size_t)8, (200 * 1024) / smem));
    const int grid = (int)min(ctas, (long long)per_sm * sms);
    const bool vec = ((uintptr_t)preds_dev % 16) == 0;
    void (*k)(ConfArgs);
    if (dtype == MLC_F16)
      k = cmem ? (vec ? confusion_small<__half, true, true>
                      : confusion_small<__half, false, true>)
               : (vec ? confusion_small<__half, true, false>
                      : confusion_small<__half, false, false>);
    else
      k = cmem ? (vec ? confusion_small<float, true, true>
                      : confusion_small<float, false, true>)
               : (vec ? confusion_small<float, true, false>
                      : confusion_small<float, false, false>);
    if (smem > 48 * 1024)
      MLC_CUDA(cudaFuncSetAttribute(
          k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k<<<grid, wpc * 32, smem, st>>>(a);
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
