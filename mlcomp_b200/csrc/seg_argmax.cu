// Segmentation arg-max: per pixel, the first maximal class of the fused
// (un-flipped, activated, averaged) class scores - the multiclass deployment
// of the same executors.  Contract: include/mlc.h (mlc_seg_argmax).
//
// Reference arithmetic: apply_activation(x, 'softmax') = torch.softmax(x, 1)
// per source (utils/torch.py:16-17), TtaWrap.inverse (tta.py:20-31), the
// builder-defined fp32 mean over sources (as everywhere in this library), then
// numpy's arg-max over the class axis as in worker/reports/classification.py
// :111,136 - a NaN beats everything, ties go to the lowest index.  The class
// scores are computed exactly as mlc_seg_prob computes them (same max / sum /
// divide order, same mean rule), so label == argmax(mlc_seg_prob) bit for bit.
//
// One thread per pixel.  Per source the three steps of the soft-max (running
// maximum over the classes, sum of expf(x - max) in class order, the divide)
// re-read the pixel's class values; they stay in L1 (the working set of a CTA
// is 256 px x S x C values), so DRAM sees every logit once.
#include "mlc_common.cuh"

namespace mlc {

constexpr int kAmThreads = 256;
constexpr int kAmMaxSmemClasses = 64;  // C x C table in shared memory
constexpr int kAmRegClasses = 8;       // class values of a pixel in registers

struct ArgmaxArgs {
  SrcSet src;
  const uint8_t *gt_onehot;  // [n][C][H][W] or null
  const uint8_t *gt_label;   // [n][H][W] or null
  uint8_t *label;            // [n][H][W] or null
  uint8_t *onehot;           // [n][C][H][W] or null
  unsigned long long *counts;  // [n][C][3] = {T, P, I} or null
  unsigned long long *cm;      // [C][C] rows = truth, cols = prediction, or null
  long long n_img;
  int C, H, W;
  int act;
  int chunks_per_img;
};

// element (class plane `pl`, output pixel (y, x)) of source k, un-transformed
template <typename T>
__device__ __forceinline__ float am_fetch(const ArgmaxArgs &a, int k,
                                          long long pl, int y, int x) {
  const int xf = a.src.xf[k];
  const T *p = static_cast<const T *>(a.src.ptr[k]) +
               (size_t)pl * a.H * a.W;
  if (!(xf & MLC_XF_TRANSPOSE)) {
    const int ys = (xf & MLC_XF_VFLIP) ? a.H - 1 - y : y;
    const int xs = (xf & MLC_XF_HFLIP) ? a.W - 1 - x : x;
    return to_f32<T>(p[(size_t)ys * a.W + xs]);
  }
  // source plane is [W][H]: out[y][x] = src[ys][xs], (u, v) = (x, y)
  const int ys = (xf & MLC_XF_VFLIP) ? a.W - 1 - x : x;
  const int xs = (xf & MLC_XF_HFLIP) ? a.H - 1 - y : y;
  return to_f32<T>(p[(size_t)ys * a.H + xs]);
}

// offset of output pixel (y, x) inside a class plane of source k
__device__ __forceinline__ size_t am_offset(const ArgmaxArgs &a, int k, int y,
                                            int x) {
  const int xf = a.src.xf[k];
  if (!(xf & MLC_XF_TRANSPOSE)) {
    const int ys = (xf & MLC_XF_VFLIP) ? a.H - 1 - y : y;
    const int xs = (xf & MLC_XF_HFLIP) ? a.W - 1 - x : x;
    return (size_t)ys * a.W + xs;
  }
  const int ys = (xf & MLC_XF_VFLIP) ? a.W - 1 - x : x;
  const int xs = (xf & MLC_XF_HFLIP) ? a.H - 1 - y : y;
  return (size_t)ys * a.H + xs;
}

// Class scores of one pixel for exactly CT classes (CT <= kAmRegClasses):
// the pixel's class values of one source live in registers - every logit is
// loaded once, every exponential taken once (soft-max steps in torch's order:
// maximum, sum of expf(x - max) in class order, divide); sum over the sources
// left to right.
// numpy arg-max order over mean scores: a NaN beats everything, otherwise
// larger wins, ties (and NaN vs NaN) go to the lower index
struct AmBest {
  float best = 0.f;
  int best_c = 0;
  bool best_nan = false;
  __device__ __forceinline__ void consider(int c, float v, int S) {
    float w[8] = {v, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    mean8(w, S);
    v = w[0];
    const bool is_nan = !(v == v);
    if (c == 0 || (!best_nan && (is_nan || v > best))) {
      best = v;
      best_c = c;
      best_nan = is_nan;
    }
  }
};

template <typename T, int CT>
__device__ __forceinline__ int am_scores(const ArgmaxArgs &a, long long pl0,
                                         size_t px, int y, int x) {
  float acc[CT];
  const int S = a.src.n;
  for (int k = 0; k < S; ++k) {
    const T *p = static_cast<const T *>(a.src.ptr[k]) + (size_t)pl0 * px +
                 am_offset(a, k, y, x);
    float v[CT];
#pragma unroll
    for (int c = 0; c < CT; ++c) v[c] = to_f32<T>(p[(size_t)c * px]);
    if (a.act == MLC_ACT_SOFTMAX) {
      float m = -3.402823466e+38f, sum = 0.f;
#pragma unroll
      for (int c = 0; c < CT; ++c) m = m < v[c] ? v[c] : m;
#pragma unroll
      for (int c = 0; c < CT; ++c) {
        v[c] = expf(v[c] - m);
        sum += v[c];
      }
#pragma unroll
      for (int c = 0; c < CT; ++c) v[c] = v[c] / sum;
    } else {
#pragma unroll
      for (int c = 0; c < CT; ++c) v[c] = activate(v[c], a.act);
    }
#pragma unroll
    for (int c = 0; c < CT; ++c) acc[c] = k == 0 ? v[c] : acc[c] + v[c];
  }
  AmBest b;
#pragma unroll
  for (int c = 0; c < CT; ++c) b.consider(c, acc[c], S);
  return b.best_c;
}

// CT: the class count when it is 2..8 (scores in registers, see am_scores),
// 0 = any.  A kernel per count: one kernel holding all eight variants needed
// 128 registers and ran latency-bound at 16 warps per SM.
template <typename T, int CT>
__global__ void __launch_bounds__(kAmThreads, 4)
seg_argmax_kernel(const __grid_constant__ ArgmaxArgs a) {
  extern __shared__ unsigned int am_smem[];  // [C][3] (+ [C][C])
  const int C = a.C;
  unsigned int *s_cnt = am_smem;
  unsigned int *s_cm = am_smem + 3 * C;
  const bool cm_smem = a.cm && C <= kAmMaxSmemClasses;
  const long long px = (long long)a.H * a.W;
  const long long n_chunks = a.n_img * a.chunks_per_img;
  const long long per = (px + a.chunks_per_img - 1) / a.chunks_per_img;
  const int S = a.src.n;

  for (long long chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
    const long long img = chunk / a.chunks_per_img;
    const long long p0 = (chunk - img * a.chunks_per_img) * per;
    const long long p1 = min(px, p0 + per);
    for (int i = threadIdx.x; i < 3 * C + (cm_smem ? C * C : 0);
         i += blockDim.x)
      am_smem[i] = 0;
    __syncthreads();
    for (long long p = p0 + threadIdx.x; p < p1; p += blockDim.x) {
      const int y = (int)(p / a.W), x = (int)(p - (long long)y * a.W);
      const long long pl0 = img * C;  // class-0 plane of this image
      int best_c = 0;
      if constexpr (CT > 0) {
        best_c = am_scores<T, CT>(a, pl0, (size_t)px, y, x);
      } else {
        // many classes: per source the running maximum and the sum first,
        // then class by class (the class values are re-read from L1)
        float mk[MLC_MAX_SOURCES], zk[MLC_MAX_SOURCES];
        if (a.act == MLC_ACT_SOFTMAX) {
          for (int k = 0; k < S; ++k) {
            float m = -3.402823466e+38f;
            for (int j = 0; j < C; ++j) {
              const float t = am_fetch<T>(a, k, pl0 + j, y, x);
              m = m < t ? t : m;
            }
            float sum = 0.f;
            for (int j = 0; j < C; ++j)
              sum += expf(am_fetch<T>(a, k, pl0 + j, y, x) - m);
            mk[k] = m;
            zk[k] = sum;
          }
        }
        AmBest b;
        for (int c = 0; c < C; ++c) {
          float v = 0.f;
          for (int k = 0; k < S; ++k) {
            const float t = am_fetch<T>(a, k, pl0 + c, y, x);
            const float q = a.act == MLC_ACT_SOFTMAX
                                ? expf(t - mk[k]) / zk[k]
                                : activate(t, a.act);
            v = k == 0 ? q : v + q;
          }
          b.consider(c, v, S);
        }
        best_c = b.best_c;
      }
      if (a.label) a.label[img * px + p] = (uint8_t)best_c;
      if (a.onehot)
        for (int c = 0; c < C; ++c)
          a.onehot[(img * C + c) * px + p] = c == best_c ? 1 : 0;
      // Counting: the lanes of a warp that agree on a key elect one lane,
      // which adds their number - a plain atomic per pixel would put 256
      // threads on a dozen shared-memory words (measured: 30x slower).
      const unsigned active = __activemask();
      const int lane = threadIdx.x & 31;
      auto add_by_key = [&](bool valid, int key, unsigned int *table) {
        const unsigned voters = __ballot_sync(active, valid);
        if (valid) {
          const unsigned peers = __match_any_sync(voters, key);
          if (lane == __ffs(peers) - 1) atomicAdd(&table[key], __popc(peers));
        }
      };
      if (a.counts) {
        add_by_key(true, best_c * 3 + 1, s_cnt);  // P
        if (a.gt_onehot) {
          // all class bytes first (independent loads), then the votes: a load
          // followed by its vote per class serialised C memory latencies
          auto vote = [&](int c, bool g) {
            const unsigned bt = __ballot_sync(active, g);
            const unsigned bi = __ballot_sync(active, g && c == best_c);
            if (lane == __ffs(active) - 1) {
              if (bt) atomicAdd(&s_cnt[c * 3 + 0], __popc(bt));  // T
              if (bi) atomicAdd(&s_cnt[c * 3 + 2], __popc(bi));  // I
            }
          };
          if constexpr (CT > 0) {
            uint8_t gv[CT];
#pragma unroll
            for (int c = 0; c < CT; ++c)
              gv[c] = a.gt_onehot[(img * C + c) * px + p];
#pragma unroll
            for (int c = 0; c < CT; ++c) vote(c, gv[c] != 0);
          } else {
            for (int c = 0; c < C; ++c)
              vote(c, a.gt_onehot[(img * C + c) * px + p] != 0);
          }
        } else if (a.gt_label) {
          const int g = a.gt_label[img * px + p];
          add_by_key(g < C, g * 3 + 0, s_cnt);
          add_by_key(g < C && g == best_c, g * 3 + 2, s_cnt);
        }
      }
      if (a.cm && a.gt_label) {
        const int g = a.gt_label[img * px + p];
        // labels outside [0, C) are dropped (sklearn's rule)
        if (cm_smem) {
          add_by_key(g < C, g * C + best_c, s_cm);
        } else if (g < C) {
          atomicAdd(&a.cm[(size_t)g * C + best_c], 1ull);
        }
      }
    }
    __syncthreads();
    if (a.counts)
      for (int i = threadIdx.x; i < 3 * C; i += blockDim.x)
        if (s_cnt[i]) atomicAdd(&a.counts[img * 3 * C + i], (unsigned long long)s_cnt[i]);
    if (cm_smem)
      for (int i = threadIdx.x; i < C * C; i += blockDim.x)
        if (s_cm[i]) atomicAdd(&a.cm[i], (unsigned long long)s_cm[i]);
    __syncthreads();
  }
}

}  // namespace mlc

using namespace mlc;

extern "C" int mlc_seg_argmax(const void *const *src_dev_ptrs,
                              const int32_t *xform, int n_src, int dtype,
                              int64_t n_img, int n_cls, int H, int W, int act,
                              const uint8_t *gt_onehot_dev,
                              const uint8_t *gt_label_dev, uint8_t *label_dev,
                              uint8_t *onehot_dev, int64_t *counts_dev,
                              int64_t *cm_dev, void *stream) {
  MLC_REQUIRE(dtype == MLC_F16 || dtype == MLC_F32,
              "dtype %d is not MLC_F16/MLC_F32", dtype);
  MLC_REQUIRE(act == MLC_ACT_NONE || act == MLC_ACT_SIGMOID ||
                  act == MLC_ACT_SOFTMAX,
              "unknown activation code %d", act);
  MLC_REQUIRE(n_img >= 0 && n_cls >= 1 && n_cls <= 255 && H >= 1 && W >= 1,
              "bad shape n_img=%lld n_cls=%d (max 255) H=%d W=%d",
              (long long)n_img, n_cls, H, W);
  MLC_REQUIRE((long long)H * W < (1ll << 31), "plane of %d x %d too large", H,
              W);
  MLC_REQUIRE(label_dev || onehot_dev || counts_dev || cm_dev,
              "no output requested");
  MLC_REQUIRE(!cm_dev || gt_label_dev,
              "the pixel confusion matrix needs gt_label_dev");
  MLC_REQUIRE(!(gt_onehot_dev && gt_label_dev),
              "pass the truth as one-hot planes or as a label map, not both");
  ArgmaxArgs a;
  if (int rc = fill_srcset(a.src, src_dev_ptrs, xform, n_src,
                           dtype == MLC_F16 ? 2 : 4, H, W))
    return rc;
  cudaStream_t st = (cudaStream_t)stream;
  if (counts_dev)
    MLC_CUDA(cudaMemsetAsync(counts_dev, 0, (size_t)n_img * n_cls * 3 * 8, st));
  if (cm_dev)
    MLC_CUDA(cudaMemsetAsync(cm_dev, 0, (size_t)n_cls * n_cls * 8, st));
  if (n_img == 0) return MLC_OK;
  a.gt_onehot = gt_onehot_dev;
  a.gt_label = gt_label_dev;
  a.label = label_dev;
  a.onehot = onehot_dev;
  a.counts = reinterpret_cast<unsigned long long *>(counts_dev);
  a.cm = reinterpret_cast<unsigned long long *>(cm_dev);
  a.n_img = n_img;
  a.C = n_cls;
  a.H = H;
  a.W = W;
  a.act = act;
  // chunks of >= 16 Ki pixels, enough of them to fill the machine
  const long long px = (long long)H * W;
  long long chunks = 1;
  const long long want = 8ll * sm_count();
  if (n_img < want) chunks = min((want + n_img - 1) / n_img, (px + 16383) / 16384);
  // a chunk's 32-bit shared counters must not overflow
  chunks = max(chunks, (px + (1ll << 30) - 1) >> 30);
  a.chunks_per_img = (int)max(1ll, chunks);
  const size_t smem =
      (size_t)(3 * n_cls + (cm_dev && n_cls <= kAmMaxSmemClasses
                                ? n_cls * n_cls : 0)) * 4;
  const long long n_chunks = n_img * a.chunks_per_img;
  const int grid = (int)min(n_chunks, 16ll * sm_count());
  void (*k)(ArgmaxArgs);
#define MLC_AM(T)                                                      \
  (n_cls == 2   ? seg_argmax_kernel<T, 2>                               \
   : n_cls == 3 ? seg_argmax_kernel<T, 3>                               \
   : n_cls == 4 ? seg_argmax_kernel<T, 4>                               \
   : n_cls == 5 ? seg_argmax_kernel<T, 5>                               \
   : n_cls == 6 ? seg_argmax_kernel<T, 6>                               \
   : n_cls == 7 ? seg_argmax_kernel<T, 7>                               \
   : n_cls == 8 ? seg_argmax_kernel<T, 8>                               \
                : seg_argmax_kernel<T, 0>)
  if (dtype == MLC_F16) k = MLC_AM(__half);
  else k = MLC_AM(float);
#undef MLC_AM
  k<<<grid, kAmThreads, smem, st>>>(a);
  MLC_LAUNCH_CHECK("seg_argmax_kernel");
  return MLC_OK;
}
