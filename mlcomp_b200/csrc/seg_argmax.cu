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
// Two kernels.  seg_argmax_kernel: one thread per pixel, any geometry, dtype,
// activation and class count - the exact arithmetic throughout (one accurate
// expf and one IEEE divide per logit), thread-per-pixel 2-byte loads; 0.15 of
// the HBM peak.  seg_argmax_items (fp16, 2..8 classes, soft-max or no
// activation, rows of whole 16-byte items, no transpose): a thread takes an
// 8-pixel item with one 128-bit load per class and source and decides the
// arg-max from a GUARDED ESTIMATE of the class scores, as seg_hist_guarded does
// for the threshold bins:
//   per source   d_c = x_c - max_c x    (the reference's own fp32 subtraction)
//                t_c = ex2.approx(d_c * log2 e),  p_c = t_c * rcp.approx(sum t)
//   score_c = sum over the sources of p_c
// Error of p_c against the true soft-max, C <= 8: t_c
// has the relative error |d_c| 2^-23.7 (rounded product with the rounded
// constant) + 2^-22 (ex2.approx), and since p_c |d_c| <= e^d |d| <= 1 / e the
// subtraction of the maximum makes that small exactly for the classes that
// matter; with the sum's roundings (7 x 2^-24), one rcp ulp and the product:
// |p_c - true| < 2^-19.6, the reference's own expf / sum / divide < 2^-20.3,
// so |score_c - reference sum| < S * 2^-19 and the difference of two scores
// is off by less than S * 2^-18.  (An estimate WITHOUT the maximum was tried:
// 18 instructions per pixel fewer, but its error grows with |x| itself, the
// margin had to be S * 2^-13 and four times as many pixels took the exact
// path: 2.73 ms against 2.54.)  The mean is a monotone map of the sum.  So
// when the largest estimated score exceeds the second largest by more than
// S * 2^-16 (a 4-fold margin), the reference's arg-max is that class.  Every
// other pixel - near-ties, exact ties (which go to the lowest index), NaN /
// inf anywhere (every estimated score of the pixel is then NaN) - is evaluated by
// the exact code of the first kernel (am_scores).  With no activation the
// scores are plain fp32 sums: computed exactly, no estimate.
// Counts: per class, the item's eight labels are compared as packed bytes and
// the T / P / I counts of a thread stay in registers until its chunk ends.
#include <type_traits>

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
  int items_per_row;  // seg_argmax_items: W / 8
  float margin;       // seg_argmax_items: S * 2^-16
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

// ---- the item kernel ---------------------------------------------------------
// exact arg-max of pixel i of the item whose raw vectors lie in the calling
// thread's shared-memory slots (source k, class c at slot_sa + (k * CT + c) *
// stride, in SOURCE order): the arithmetic of am_scores / seg_argmax_kernel,
// without going back to global memory
template <int CT>
__device__ __noinline__ int am_exact_pixel(const ArgmaxArgs &a,
                                           uint32_t slot_sa, uint32_t stride,
                                           int i) {
  float acc[CT];
  const int S = a.src.n;
  for (int k = 0; k < S; ++k) {
    const int e = (a.src.xf[k] & MLC_XF_HFLIP) ? 7 - i : i;
    float v[CT];
#pragma unroll
    for (int c = 0; c < CT; ++c) {
      unsigned short hbits;
      asm volatile("ld.shared.u16 %0, [%1];"
                   : "=h"(hbits)
                   : "r"(slot_sa + (uint32_t)(k * CT + c) * stride + e * 2u));
      v[c] = __half2float(__ushort_as_half(hbits));
    }
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
    }
#pragma unroll
    for (int c = 0; c < CT; ++c) acc[c] = k == 0 ? v[c] : acc[c] + v[c];
  }
  AmBest b;
#pragma unroll
  for (int c = 0; c < CT; ++c) b.consider(c, acc[c], S);
  return b.best_c;
}

// 0x80 in every byte of v that is not zero
__device__ __forceinline__ uint32_t nonzero_bytes(uint32_t v) {
  return (((v & 0x7f7f7f7fu) + 0x7f7f7f7fu) | v) & 0x80808080u;
}

constexpr float kAmLog2e = 1.4426950408889634f;

__device__ __forceinline__ void am_cp16(uint32_t dst, const void *src) {
  MLC_CHECK_LOAD(src, 16);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src)
               : "memory");
}
__device__ __forceinline__ void am_cp8(uint32_t dst, const void *src) {
  MLC_CHECK_LOAD(src, 8);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src)
               : "memory");
}
__device__ __forceinline__ uint4 am_lds16(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "r"(addr));
  return v;
}
__device__ __forceinline__ uint2 am_lds8(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y)
               : "r"(addr));
  return v;
}

// Every thread copies its NEXT item - one 16-byte vector per class and source,
// the truth bytes - with cp.async into its own shared-memory slots while it
// works on the current one (two stages; no registers are held by loads in
// flight: with direct loads the kernel ran at 4 warps per scheduler, 128
// registers each, and 60 % of the issue slots had no eligible warp).
template <int CT, int ACT>
__global__ void __launch_bounds__(kAmThreads, 2)
seg_argmax_items(const __grid_constant__ ArgmaxArgs a) {
  extern __shared__ __align__(16) unsigned int am_smem[];
  // [C][3], then [C * C][32] (with a confusion matrix), then the stages
  constexpr int C = CT;
  constexpr int NT = kAmThreads;
  unsigned int *s_cnt = am_smem;
  unsigned int *s_cm = am_smem + 3 * C;  // per-lane columns
  const long long px = (long long)a.H * a.W;
  const int ipr = a.items_per_row;
  const long long items = px / 8;
  const long long n_chunks = a.n_img * a.chunks_per_img;
  const long long per = (items + a.chunks_per_img - 1) / a.chunks_per_img;
  const int S = a.src.n;
  const int lane = threadIdx.x & 31;
  const bool want_cm = a.cm && a.gt_label;
  // a thread's slots of one stage: S x C vectors at a stride of NT x 16 bytes,
  // then the truth bytes (C x 8 one-hot, or 8 of the label map) at NT x 8
  const uint32_t n_head = 3 * C + (want_cm ? C * C * 32 : 0);
  const uint32_t stage_sa =
      ((uint32_t)__cvta_generic_to_shared(am_smem) + n_head * 4u + 15u) & ~15u;
  const uint32_t gt_words = a.gt_onehot ? C : (a.gt_label ? 1 : 0);
  const uint32_t gt_base = (uint32_t)(S * C) * NT * 16u;
  const uint32_t stage_bytes = gt_base + gt_words * NT * 8u;
  const uint32_t my16 = stage_sa + threadIdx.x * 16u;
  const uint32_t my8 = stage_sa + gt_base + threadIdx.x * 8u;

  for (long long chunk = blockIdx.x; chunk < n_chunks; chunk += gridDim.x) {
    const long long img = chunk / a.chunks_per_img;
    const long long i0 = (chunk - img * a.chunks_per_img) * per;
    const long long i1 = min(items, i0 + per);
    const long long pl0 = img * C;  // class-0 plane of this image
    for (int i = threadIdx.x; i < (int)n_head; i += blockDim.x) am_smem[i] = 0;
    __syncthreads();
    uint32_t nT[C], nP[C], nI[C];
#pragma unroll
    for (int c = 0; c < C; ++c) nT[c] = nP[c] = nI[c] = 0u;
    auto issue = [&](long long it_, uint32_t st) {
      // (32-bit: a plane has fewer than 2^31 pixels)
      const int y_ = (int)((uint32_t)it_ / (uint32_t)ipr),
                xi_ = (int)((uint32_t)it_ - (uint32_t)y_ * (uint32_t)ipr);
      for (int k = 0; k < S; ++k) {
        const int xf = a.src.xf[k];
        const int ys = (xf & MLC_XF_VFLIP) ? a.H - 1 - y_ : y_;
        const int xs = (xf & MLC_XF_HFLIP) ? ipr - 1 - xi_ : xi_;
        const uint4 *p = reinterpret_cast<const uint4 *>(
                             static_cast<const __half *>(a.src.ptr[k]) +
                             (size_t)pl0 * px) +
                         ((size_t)ys * ipr + xs);
#pragma unroll
        for (int c = 0; c < C; ++c)
          am_cp16(my16 + st + (uint32_t)(k * C + c) * NT * 16u,
                  p + (size_t)c * (px / 8));
      }
      if (a.gt_onehot) {
#pragma unroll
        for (int c = 0; c < C; ++c)
          am_cp8(my8 + st + (uint32_t)c * NT * 8u,
                 a.gt_onehot + (size_t)(pl0 + c) * px + (size_t)it_ * 8);
      } else if (a.gt_label) {
        am_cp8(my8 + st, a.gt_label + (size_t)img * px + (size_t)it_ * 8);
      }
    };
    uint32_t cur = 0u;  // byte offset of the stage that holds the current item
    if (i0 + threadIdx.x < i1) issue(i0 + threadIdx.x, cur);
    asm volatile("cp.async.commit_group;" ::: "memory");
    for (long long it = i0 + threadIdx.x; it < i1; it += blockDim.x) {
      if (it + blockDim.x < i1) issue(it + blockDim.x, cur ^ stage_bytes);
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group 1;" ::: "memory");
      float acc[8][C];
      // (the first source assigns, the others add: a compile-time flag - as a
      // run-time test of k it cost a select per class and pixel)
      auto do_source = [&](int k, auto first_tag) {
        constexpr bool kFirst = decltype(first_tag)::value;
        const bool hf = a.src.xf[k] & MLC_XF_HFLIP;
        uint32_t w[C][4];
#pragma unroll
        for (int c = 0; c < C; ++c) {
          const uint4 r = am_lds16(my16 + cur + (uint32_t)(k * C + c) * NT * 16u);
          w[c][0] = r.x; w[c][1] = r.y; w[c][2] = r.z; w[c][3] = r.w;
        }
        if (hf) {  // the item's pixels come in reverse order
#pragma unroll
          for (int c = 0; c < C; ++c) {
            const uint32_t t0 = prmt32(w[c][3], 0u, 0x1032u),
                           t1 = prmt32(w[c][2], 0u, 0x1032u),
                           t2 = prmt32(w[c][1], 0u, 0x1032u),
                           t3 = prmt32(w[c][0], 0u, 0x1032u);
            w[c][0] = t0; w[c][1] = t1; w[c][2] = t2; w[c][3] = t3;
          }
        }
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          float v[2][C];
#pragma unroll
          for (int c = 0; c < C; ++c) {
            const float2 f = __half22float2(
                *reinterpret_cast<const __half2 *>(&w[c][h]));
            v[0][c] = f.x;
            v[1][c] = f.y;
          }
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            if (ACT == MLC_ACT_SOFTMAX) {
              float m = v[e][0];
#pragma unroll
              for (int c = 1; c < C; ++c) m = fmaxf(m, v[e][c]);
              float z = 0.f;
#pragma unroll
              for (int c = 0; c < C; ++c) {
                float t;
                asm("ex2.approx.ftz.f32 %0, %1;"
                    : "=f"(t)
                    : "f"((v[e][c] - m) * kAmLog2e));
                v[e][c] = t;
                z += t;
              }
              float r;
              asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(z));
#pragma unroll
              for (int c = 0; c < C; ++c) v[e][c] *= r;
            }
#pragma unroll
            for (int c = 0; c < C; ++c)
              acc[2 * h + e][c] = kFirst ? v[e][c] : acc[2 * h + e][c] + v[e][c];
          }
        }
      };
      do_source(0, std::true_type{});
      for (int k = 1; k < S; ++k) do_source(k, std::false_type{});
      // labels of the eight pixels, one per byte
      uint32_t lab[2] = {0u, 0u};
      if (ACT == MLC_ACT_SOFTMAX) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float best = acc[i][0], second = -1.f;
          int bc = 0;
#pragma unroll
          for (int c = 1; c < C; ++c) {
            const float v = acc[i][c];
            const bool up = v > best;
            second = up ? best : fmaxf(second, v);
            bc = up ? c : bc;
            best = up ? v : best;
          }
          // A NaN or +inf logit, or a pixel whose logits are all -inf, makes
          // the sum of a source's exponentials NaN and with it the score of
          // EVERY class of the pixel: best = NaN, the test is false
          if (!(best - second > a.margin))
            bc = am_exact_pixel<CT>(a, my16 + cur, NT * 16u, i);
          lab[i >> 2] |= (uint32_t)bc << (8 * (i & 3));
        }
      } else {
        // plain sums: the reference's own arithmetic (mean8, then numpy's
        // arg-max order)
#pragma unroll
        for (int c = 0; c < C; ++c) {
          float wv[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) wv[i] = acc[i][c];
          mean8(wv, S);
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[i][c] = wv[i];
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float best = acc[i][0];
          int bc = 0;
          bool best_nan = !(best == best);
#pragma unroll
          for (int c = 1; c < C; ++c) {
            const float v = acc[i][c];
            const bool is_nan = !(v == v);
            const bool up = !best_nan && (is_nan || v > best);
            bc = up ? c : bc;
            best = up ? v : best;
            best_nan = best_nan || (up && is_nan);
          }
          lab[i >> 2] |= (uint32_t)bc << (8 * (i & 3));
        }
      }
      const size_t at = (size_t)img * px + (size_t)it * 8;
      if (a.label)
        *reinterpret_cast<uint2 *>(a.label + at) = make_uint2(lab[0], lab[1]);
      uint2 gl = make_uint2(0u, 0u);
      if (a.gt_label) gl = am_lds8(my8 + cur);
#pragma unroll
      for (int c = 0; c < C; ++c) {
        // bytes equal to c: x ^ c has a zero byte there (labels are < 128)
        const uint32_t pat = 0x01010101u * (uint32_t)c;
        const uint32_t e0 = ~nonzero_bytes(lab[0] ^ pat) & 0x80808080u,
                       e1 = ~nonzero_bytes(lab[1] ^ pat) & 0x80808080u;
        if (a.onehot)
          *reinterpret_cast<uint2 *>(a.onehot + (size_t)(pl0 + c) * px +
                                     (size_t)it * 8) =
              make_uint2(e0 >> 7, e1 >> 7);
        if (a.counts) {
          nP[c] += __popc(e0 | (e1 >> 1));
          uint32_t t0 = 0u, t1 = 0u;
          if (a.gt_onehot) {
            const uint2 g = am_lds8(my8 + cur + (uint32_t)c * NT * 8u);
            t0 = nonzero_bytes(g.x);
            t1 = nonzero_bytes(g.y);
          } else if (a.gt_label) {
            t0 = ~nonzero_bytes(gl.x ^ pat) & 0x80808080u;
            t1 = ~nonzero_bytes(gl.y ^ pat) & 0x80808080u;
          }
          nT[c] += __popc(t0 | (t1 >> 1));
          nI[c] += __popc((t0 & e0) | ((t1 & e1) >> 1));
        }
      }
      if (want_cm) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t g = ((i < 4 ? gl.x : gl.y) >> (8 * (i & 3))) & 0xFFu;
          const uint32_t b = (lab[i >> 2] >> (8 * (i & 3))) & 0xFFu;
          // labels outside [0, C) are dropped (sklearn's rule)
          if (g < (uint32_t)C) atomicAdd(&s_cm[(g * C + b) * 32 + lane], 1u);
        }
      }
      cur ^= stage_bytes;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    if (a.counts) {
#pragma unroll
      for (int c = 0; c < C; ++c) {
        uint32_t t = nT[c], p = nP[c], q = nI[c];
#pragma unroll
        for (int d = 16; d >= 1; d >>= 1) {
          t += __shfl_xor_sync(0xffffffffu, t, d);
          p += __shfl_xor_sync(0xffffffffu, p, d);
          q += __shfl_xor_sync(0xffffffffu, q, d);
        }
        if (lane == 0) {
          if (t) atomicAdd(&s_cnt[c * 3 + 0], t);
          if (p) atomicAdd(&s_cnt[c * 3 + 1], p);
          if (q) atomicAdd(&s_cnt[c * 3 + 2], q);
        }
      }
    }
    __syncthreads();
    if (a.counts)
      for (int i = threadIdx.x; i < 3 * C; i += blockDim.x)
        if (s_cnt[i])
          atomicAdd(&a.counts[img * 3 * C + i], (unsigned long long)s_cnt[i]);
    if (want_cm)
      for (int i = threadIdx.x; i < C * C; i += blockDim.x) {
        unsigned int v = 0;
        for (int l = 0; l < 32; ++l) v += s_cm[i * 32 + ((l + i) & 31)];
        if (v) atomicAdd(&a.cm[i], (unsigned long long)v);
      }
    __syncthreads();
  }
}

template <int CT>
static cudaError_t launch_items(const ArgmaxArgs &a, int grid, size_t smem,
                                cudaStream_t st) {
  void (*k)(ArgmaxArgs) = a.act == MLC_ACT_SOFTMAX
                              ? seg_argmax_items<CT, MLC_ACT_SOFTMAX>
                              : seg_argmax_items<CT, MLC_ACT_NONE>;
  const cudaError_t e = cudaFuncSetAttribute(
      k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  k<<<grid, kAmThreads, smem, st>>>(a);
  return cudaGetLastError();
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
  // the item kernel: fp16, 2..8 classes, rows of whole aligned items
  bool items = dtype == MLC_F16 && n_cls >= 2 && n_cls <= 8 &&
               (act == MLC_ACT_SOFTMAX || act == MLC_ACT_NONE) && (W % 8) == 0 &&
               !(a.src.any_xf & MLC_XF_TRANSPOSE) &&
               ((uintptr_t)gt_onehot_dev % 8) == 0 &&
               ((uintptr_t)gt_label_dev % 8) == 0 &&
               ((uintptr_t)label_dev % 8) == 0 && ((uintptr_t)onehot_dev % 8) == 0;
  for (int i = 0; i < n_src; ++i)
    items = items && ((uintptr_t)src_dev_ptrs[i] % 16) == 0;
  if (items) {
#ifdef MLC_BOUNDS_CHECK
    {
      BoundsList bl;
      const size_t px_all = (size_t)n_img * H * W;
      for (int i = 0; i < n_src; ++i)
        bl.add(src_dev_ptrs[i], px_all * n_cls * 2);
      bl.add(gt_onehot_dev, px_all * n_cls);
      bl.add(gt_label_dev, px_all);
      MLC_CUDA(bounds_commit(bl, st));
    }
#endif
    a.items_per_row = W / 8;
    a.margin = (float)n_src * 1.52587890625e-5f;  // S * 2^-16
    // counters, then two stages of (S x C x 16 + truth bytes) per thread
    const size_t gt_words = gt_onehot_dev ? n_cls : (gt_label_dev ? 1 : 0);
    const size_t ismem =
        (size_t)(3 * n_cls + (cm_dev ? n_cls * n_cls * 32 : 0)) * 4 + 16 +
        2 * ((size_t)n_src * n_cls * 16 + gt_words * 8) * kAmThreads;
    items = ismem <= 227 * 1024;
    if (items) {
      const int igrid = (int)min(n_chunks, 2ll * sm_count());
      cudaError_t e;
      switch (n_cls) {
        case 2: e = launch_items<2>(a, igrid, ismem, st); break;
        case 3: e = launch_items<3>(a, igrid, ismem, st); break;
        case 4: e = launch_items<4>(a, igrid, ismem, st); break;
        case 5: e = launch_items<5>(a, igrid, ismem, st); break;
        case 6: e = launch_items<6>(a, igrid, ismem, st); break;
        case 7: e = launch_items<7>(a, igrid, ismem, st); break;
        default: e = launch_items<8>(a, igrid, ismem, st); break;
      }
      MLC_CUDA(e);
      return MLC_OK;
    }
  }
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
