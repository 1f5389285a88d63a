// Device pieces shared by the guarded-estimate histogram kernels
// (seg_hist_guarded.cu: aligned layouts and enclosing-vector rows;
// seg_hist_guarded_rows.cu: rows of any width staged by TMA).
#pragma once

#include "seg_hist.cuh"

namespace mlc {


// CTA size / resident CTAs per SM (compile-time so that A/B builds can sweep
// them: python -m mlcomp_b200.build --variant NAME --define MLC_G_THREADS=384)
#ifndef MLC_G_THREADS
#define MLC_G_THREADS 512
#endif
#ifndef MLC_G_MINB
#define MLC_G_MINB 2
#endif
constexpr int kGThreads = MLC_G_THREADS;
constexpr float kNegLog2e = -1.4426950408889634f;
// denominators (products of up to four 1 + e^-x) from this bit pattern on -
// 1e37f, +inf, every NaN - may have overflowed
constexpr uint32_t kDLimitBits = 0x7CF0BDC2u;

#ifdef MLC_DEV_KNOBS
static __device__ unsigned long long g_guard_slow_px;  // pixels sent to the exact
                                                         // path (per .cu file)
#endif

__device__ __forceinline__ float ex2_approx(float y) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(y));
  return r;
}
__device__ __forceinline__ float rcp_approx(float y) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(y));
  return r;
}

// Exact bin of one output pixel: what seg_hist_generic computes.
// stage_sa != 0: the item's raw source words still lie in this thread's
// shared-memory slots (source k at stage_sa + k * stage_stride, 16 bytes in
// SOURCE order) - read them there instead of going back to global memory.
static __device__ __noinline__ uint32_t guarded_exact_bin(
    const HistArgs &a,
                                                   long long plane, int it,
                                                   int i, bool rows,
                                                   uint32_t thr_sa,
                                                   uint32_t stage_sa,
                                                   uint32_t stage_stride) {
  const float *thr = static_cast<const float *>(
      __cvta_shared_to_generic((size_t)thr_sa));
  float p = 0.f;
  if (stage_sa) {
    for (int k = 0; k < a.src.n; ++k) {
      const int e = (rows && ((a.g_rev >> k) & 1)) ? 7 - i : i;
      const uint32_t h = lds_u16(stage_sa + k * stage_stride + e * 2);
      const float v = activate(
          __half2float(__ushort_as_half((unsigned short)h)), MLC_ACT_SIGMOID);
      p = k == 0 ? v : p + v;
    }
  } else {
    const size_t plane_off = (size_t)plane * a.plane_px;
    // it < 0: (-it - 1, i) is the pixel (y, x) itself (row tails)
    const int y = it < 0 ? -it - 1 : (rows ? it / a.items_per_row : 0);
    const int x =
        it < 0 ? i : (it - y * (rows ? a.items_per_row : 0)) * 8 + i;
    for (int k = 0; k < a.src.n; ++k) {
      const int xf = a.src.xf[k];
      const int ys = (xf & MLC_XF_VFLIP) ? a.H - 1 - y : y;
      const int xs = (xf & MLC_XF_HFLIP) ? a.W - 1 - x : x;
      const __half *pl = static_cast<const __half *>(a.src.ptr[k]) + plane_off;
      const float v =
          activate(__half2float(pl[(size_t)ys * a.W + xs]), MLC_ACT_SIGMOID);
      p = k == 0 ? v : p + v;
    }
  }
  float q[8] = {p, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  mean8(q, a.src.n);
#ifdef MLC_DEV_KNOBS
  atomicAdd(&g_guard_slow_px, 1ull);
#endif
  return (uint32_t)count_ge(q[0], thr, a.B, a.scale, a.bias);
}

// q[i] (+)= sum over the NG sources of the group of 1 / (1 + 2^(x * -log2 e))
// for the eight pixels of one item; dmax keeps the largest denominator as a
// bit pattern (unsigned order: finite positive < +inf < every NaN).
// REV: bit j set = source j holds the item's pixels in reverse order (a
// horizontal flip): pixel p is then element 7 - p of its words - resolved here
// by naming the right register, not by moving data.
template <int NG, int REV>
__device__ __forceinline__ void group_sum(const uint32_t (&w)[4][4], bool first,
                                          float (&q)[8], uint32_t &dmax) {
#pragma unroll
  for (int h = 0; h < 4; ++h) {
    float a[4][2];
#pragma unroll
    for (int j = 0; j < NG; ++j) {
      const bool rev = (REV >> j) & 1;
      const float2 x = __half22float2(
          *reinterpret_cast<const __half2 *>(&w[j][rev ? 3 - h : h]));
      a[j][0] = 1.0f + ex2_approx((rev ? x.y : x.x) * kNegLog2e);
      a[j][1] = 1.0f + ex2_approx((rev ? x.x : x.y) * kNegLog2e);
    }
#pragma unroll
    for (int e = 0; e < 2; ++e) {
      float n, d;
      if (NG == 4) {
        const float ab = a[0][e] * a[1][e], cd = a[2][e] * a[3][e];
        n = fmaf(a[2][e] + a[3][e], ab, (a[0][e] + a[1][e]) * cd);
        d = ab * cd;
      } else if (NG == 3) {
        const float ab = a[0][e] * a[1][e];
        n = fmaf(a[0][e] + a[1][e], a[2][e], ab);
        d = ab * a[2][e];
      } else if (NG == 2) {
        n = a[0][e] + a[1][e];
        d = a[0][e] * a[1][e];
      } else {
        n = 1.0f;
        d = a[0][e];
      }
      dmax = max(dmax, __float_as_uint(d));
      const float g = n * rcp_approx(d);
      q[2 * h + e] = first ? g : q[2 * h + e] + g;
    }
  }
}

__device__ __forceinline__ void reverse8(uint32_t (&w)[4]) {
  const uint32_t t0 = prmt32(w[3], 0u, 0x1032u), t1 = prmt32(w[2], 0u, 0x1032u),
                 t2 = prmt32(w[1], 0u, 0x1032u), t3 = prmt32(w[0], 0u, 0x1032u);
  w[0] = t0; w[1] = t1; w[2] = t2; w[3] = t3;
}

__device__ __forceinline__ uint4 lds_u4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "r"(addr));
  return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void st_shared_u32(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint2 lds_u2(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y)
               : "r"(addr));
  return v;
}

// From the sums of sigmoids of one item to its histogram updates (and bin
// bytes): grid position moved down / up by the error bound, exact evaluation
// of the pixels the two disagree on, one shared-memory atomic per pixel.
// slot: shared address of this lane's column of counters (16 KiB aligned base
// | lane * 4); the threshold table lies behind the counters.
template <bool ROWS, int BINS>  // BINS: 0 none, 1 aligned item, 2 any offset
__device__ __forceinline__ void finish_item(const HistArgs &a,
                                            const float (&q)[8], uint32_t dmax,
                                            uint2 g, uint32_t slot,
                                            long long plane, int it,
                                            size_t bins_item,
                                            uint32_t stage_sa = 0u,
                                            uint32_t stage_stride = 0u,
                                            int ex_it = 0, int ex_i0 = 0,
                                            int nvalid = 8) {
  // nvalid < 8 (a row's last item in the kernel for rows of any width): only
  // the first `nvalid` pixels exist, the others lie beyond the row.
  // ex_it < 0: the exact path is addressed as pixel (y, x) = (-ex_it - 1,
  // ex_i0 + i) instead of by item index
  const bool tail = ex_it < 0;
  const float kq = a.g_kq, fb_lo = a.g_fb_lo, fb_hi = a.g_fb_hi;
  uint32_t fx[8], differ = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    fx[i] = __float_as_uint(fmaf(q[i], kq, fb_lo));
    differ |= fx[i] ^ __float_as_uint(fmaf(q[i], kq, fb_hi));
  }
  // a product that may have overflowed, +inf or a NaN anywhere in the item (a
  // NaN input makes every later product NaN): nothing of the item is trusted
  const bool all = dmax >= kDLimitBits;
  if ((differ >> 16) != 0u || all) {
    // rare: the estimate cannot decide (or cannot be trusted for) some pixel
    // of this item - those pixels are evaluated exactly
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const uint32_t hi = __float_as_uint(fmaf(q[i], kq, fb_hi));
      if (i < nvalid && (all || ((fx[i] ^ hi) >> 16) != 0u))
        fx[i] = 0x43008000u |
                (guarded_exact_bin(a, plane, tail ? ex_it : it,
                                   tail ? ex_i0 + i : i, ROWS,
                                   (slot & ~16383u) + (a.B + 1) * 128u,
                                   stage_sa, stage_stride)
                 << 16);
    }
  }
  // bin = bits 16..22 of fx (the exponent field reads 0x4300 + bin)
  if (!BINS && nvalid < 8) {
    // pixels beyond the row go to bin 127: a row of counters that the caller
    // provides and never flushes (a few instructions for the lanes that hold a
    // row's last item, instead of a second counting path: that one was taken,
    // divergently, by every second warp)
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (i >= nvalid) fx[i] = 0x437F0000u;
    const uint64_t keep = ~0ull >> (64 - 8 * nvalid);  // nvalid >= 1
    g.x &= (uint32_t)keep;
    g.y &= (uint32_t)(keep >> 32);
    nvalid = 8;
  }
  if (nvalid < 8) {  // (with bin planes only)
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const uint32_t gb = ((i < 4 ? g.x : g.y) >> (8 * (i & 3))) & 0xFFu;
      if (i < nvalid)
        red_shared_add(((fx[i] >> 9) & 0x3F80u) | slot, gb ? 0x10000u : 1u);
    }
  } else if ((g.x | g.y) == 0u) {  // all background: the increment is 1
#pragma unroll
    for (int i = 0; i < 8; ++i)
      red_shared_add(((fx[i] >> 9) & 0x3F80u) | slot, 1u);
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const uint32_t gb = ((i < 4 ? g.x : g.y) >> (8 * (i & 3))) & 0xFFu;
      red_shared_add(((fx[i] >> 9) & 0x3F80u) | slot, gb ? 0x10000u : 1u);
    }
  }
  if (BINS) {
    // byte 2 of each fx is its bin (< 128)
    uint2 v;
    v.x = prmt32(prmt32(fx[0], fx[1], 0x0062u), prmt32(fx[2], fx[3], 0x0062u),
                 0x5410u);
    v.y = prmt32(prmt32(fx[4], fx[5], 0x0062u), prmt32(fx[6], fx[7], 0x0062u),
                 0x5410u);
    if (BINS == 1) {
      reinterpret_cast<uint2 *>(a.bins)[bins_item] = v;
    } else {  // bins_item is a byte offset of any alignment
      uint8_t *dst = a.bins + bins_item;
      if (nvalid < 8) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
          if (i < nvalid) dst[i] = (uint8_t)((i < 4 ? v.x : v.y) >> (8 * (i & 3)));
      } else if ((reinterpret_cast<uintptr_t>(dst) & 3) == 0) {
        reinterpret_cast<uint32_t *>(dst)[0] = v.x;
        reinterpret_cast<uint32_t *>(dst)[1] = v.y;
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          dst[i] = (uint8_t)(v.x >> (8 * i));
          dst[4 + i] = (uint8_t)(v.y >> (8 * i));
        }
      }
    }
  }
}

// The threshold grid must be uniform over the unit interval; fills the
// constants of the estimate.  False: this grid stays on the exact kernels.
static bool guarded_grid(HistArgs &a) {
  const int B = a.B, S = a.src.n;
  if (B < 2 || B > 120 || S < 1) return false;
  // f(p) = p*scale + bias puts thr[k] at k + 1, p = 0 at >= 0, p = 1 at <= B+1
  const double sc = (double)a.scale, bi = (double)a.bias;
  if (!(sc > 0.0) || bi < 0.0 || sc + bi > (double)B + 1.0 + 1e-6) return false;
  double g = 0.0;
  for (int k = 0; k < B; ++k) {
    if (k > 0 && !(a.thr.v[k] > a.thr.v[k - 1])) return false;
    const double d = (double)a.thr.v[k] * sc + bi - (double)(k + 1);
    g = d < 0 ? (-d > g ? -d : g) : (d > g ? d : g);
  }
  if (g > 1e-3) return false;
  const double u = 5.9604644775390625e-8;  // 2^-24
  // error of the estimated mean against the reference mean, in grid steps
  // (DESIGN.md section 3), + grid irregularity + rounding of scale / S +
  // rounding of the two fma (constant and result: 2^-17 each at [128, 256))
  const double eps = sc * (24.0 + 0.25 * S) * u + g + u * (B + 1) +
                     2.0 * 7.62939453125e-6;
  const double steps = eps * 65536.0 + 2.0;   // in units of 2^-16
  if (steps > 4096.0) return false;
  const double e = ((double)(long long)steps + 1.0) / 65536.0;
  a.g_kq = (float)(sc / (double)S);
  a.g_fb_lo = (float)(bi + 128.0 - e);
  a.g_fb_hi = (float)(bi + 128.0 + e);
  return true;
}

}  // namespace mlc
