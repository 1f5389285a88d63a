// K1 for fp16 sources under a sigmoid (TTA variants, ensemble members) in the
// aligned layouts: the bin of the mean probability from a GUARDED ESTIMATE
// instead of one exact sigmoid per source value.
//
// Why: the exact kernels are bound by work per pixel, not by HBM - an exact
// sigmoid is expf + an IEEE divide (~14 instructions, 2 MUFU) per source
// value, or a random look-up in a 152 KiB shared-memory table (3.5 LSU
// wavefronts each; seg_hist_tab runs at 82 % LSU utilisation and 53 % of the
// HBM peak on a 4-flip TTA set).  What the histogram needs is not the
// probability but cnt(p) = #{k : p >= thr[k]}, and that is decided by any
// estimate of p whose error is smaller than its distance to the nearest
// threshold:
//
//   t_i = 2^(-x_i * log2 e)       ex2.approx (MUFU), x_i exact in fp32
//   a_i = 1 + t_i                 sigmoid(x_i) = 1 / a_i
//   q   = sum_i 1/a_i = N / D     per group of four sources ONE reciprocal:
//         N = (a+b) cd + (c+d) ab,  D = ab cd      (5 MUFU per pixel, not 8)
//   f   = q * (scale / S) + bias  position on the threshold grid; computed as
//         f + 128 in [128, 256), where a float carries exactly 16 fractional
//         bits: the integer part of the bit pattern is the bin, the fraction
//         its distance to the grid lines
//
// |estimate - reference mean| <= 24 * 2^-24 (derivation in DESIGN.md section
// 3: ex2.approx 2 ulp, rcp.approx 1 ulp, 10 roundings in N/D, and on the
// reference side expf 2 ulp + its adds), so a pixel whose fraction lies in
// [g_lo, 65535 - g_lo] has the reference's bin for certain.  Every other pixel
// (about 4 in 10 000 on the default grid), every pixel of an item in which a
// product overflowed (|x| > ~20) or a NaN appeared, takes the exact path:
// activate() per source, fp32 sum left to right, one divide by S, table walk
// - the very arithmetic of seg_hist_generic / seg_hist_tab.  Results are
// therefore identical to those kernels for every input (tests: equality of
// the histograms and bin planes on random, on-threshold, NaN / inf inputs).
//
// The host enables the kernel only for a uniform threshold grid that spans
// the unit interval (every thr[k] within 1e-3 grid steps of its place, <= 120
// thresholds); anything else stays on the exact kernels.
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
constexpr float kDLimit = 1.0e37f;  // products above this may have overflowed

#ifdef MLC_DEV_KNOBS
__device__ unsigned long long g_guard_slow_px;  // pixels sent to the exact path
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
__device__ __noinline__ uint32_t guarded_exact_bin(const HistArgs &a,
                                                   size_t plane_off, int y,
                                                   int x, const float *thr) {
  float p = 0.f;
  for (int k = 0; k < a.src.n; ++k) {
    const int xf = a.src.xf[k];
    const int ys = (xf & MLC_XF_VFLIP) ? a.H - 1 - y : y;
    const int xs = (xf & MLC_XF_HFLIP) ? a.W - 1 - x : x;
    const __half *pl = static_cast<const __half *>(a.src.ptr[k]) + plane_off;
    const float v =
        activate(__half2float(pl[(size_t)ys * a.W + xs]), MLC_ACT_SIGMOID);
    p = k == 0 ? v : p + v;
  }
  float q[8] = {p, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  mean8(q, a.src.n);
#ifdef MLC_DEV_KNOBS
  atomicAdd(&g_guard_slow_px, 1ull);
#endif
  return (uint32_t)count_ge(q[0], thr, a.B, a.scale, a.bias);
}

// q[i] (+)= sum over the NG sources of the group of 1 / (1 + 2^(x * -log2 e))
// for the eight pixels of one item; dmax keeps the largest denominator.
template <int NG>
__device__ __forceinline__ void group_sum(const uint32_t (&w)[4][4], bool first,
                                          float (&q)[8], float &dmax) {
#pragma unroll
  for (int h = 0; h < 4; ++h) {
    float a[4][2];
#pragma unroll
    for (int j = 0; j < NG; ++j) {
      const float2 x =
          __half22float2(*reinterpret_cast<const __half2 *>(&w[j][h]));
      a[j][0] = 1.0f + ex2_approx(x.x * kNegLog2e);
      a[j][1] = 1.0f + ex2_approx(x.y * kNegLog2e);
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
      dmax = fmaxf(dmax, d);
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

// ROWS: flipped sources, rows a whole number of aligned 16-byte items (flips
// are item-index arithmetic); else no transforms, planes are flat arrays.
// BINS: also write every pixel's bin index (mlc_seg_hist_bins).
template <bool ROWS, bool BINS>
__global__ void __launch_bounds__(kGThreads, MLC_G_MINB)
seg_hist_guarded(const __grid_constant__ HistArgs a) {
  extern __shared__ __align__(16) uint32_t gsm[];
  const int B = a.B, nbins = B + 1;
  uint32_t *cnt = gsm;                                   // [nbins][32]
  float *thr = reinterpret_cast<float *>(gsm + nbins * 32);  // [B]
  for (int i = threadIdx.x; i < nbins * 32; i += blockDim.x) cnt[i] = 0;
  for (int i = threadIdx.x; i < B; i += blockDim.x) thr[i] = a.thr.v[i];
  __syncthreads();
  const uint32_t lane4 = (threadIdx.x & 31) * 4;
  const uint32_t cnt_sa = (uint32_t)__cvta_generic_to_shared(cnt);
  const long long n_units = a.n_planes * a.units_per_plane;
  const int S = a.src.n;
  const float kq = a.g_kq, fb = a.g_fb;
  const uint32_t lo = a.g_lo, span = a.g_span;
  const int ipr = a.items_per_row;  // ROWS: W / 8

  for (long long unit = blockIdx.x; unit < n_units; unit += gridDim.x) {
    const long long plane = unit / a.units_per_plane;
    const int sub = (int)(unit - plane * a.units_per_plane);
    const size_t plane_off = (size_t)plane * a.plane_px;
    const uint8_t *gtp = a.gt + plane_off;
    const int i0 = (int)((long long)sub * a.items_per_unit);
    const int i1 = (int)min(a.items_per_plane, (long long)i0 + a.items_per_unit);
    int it = i0 + (int)threadIdx.x;
    int y = 0, xi = it;
    int dy = 0, dx = (int)blockDim.x;
    if (ROWS) {
      y = it / ipr;
      xi = it - y * ipr;
      dy = (int)blockDim.x / ipr;
      dx = (int)blockDim.x - dy * ipr;
    }
    for (; it < i1; it += blockDim.x) {
      float q[8];
      float dmax = 0.f;
      for (int k0 = 0; k0 < S; k0 += 4) {
        const int ng = min(4, S - k0);
        uint32_t w[4][4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (j < ng) {
            size_t e0;
            if (ROWS) {
              const int xf = a.src.xf[k0 + j];
              const int ys = (xf & MLC_XF_VFLIP) ? a.H - 1 - y : y;
              const int xs = (xf & MLC_XF_HFLIP) ? ipr - 1 - xi : xi;
              e0 = plane_off + (size_t)(ys * ipr + xs) * 8;
            } else {
              e0 = plane_off + (size_t)it * 8;
            }
            const uint4 r = ldg_stream16(
                static_cast<const __half *>(a.src.ptr[k0 + j]) + e0);
            w[j][0] = r.x; w[j][1] = r.y; w[j][2] = r.z; w[j][3] = r.w;
          }
        }
        if (ROWS) {
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (j < ng && (a.src.xf[k0 + j] & MLC_XF_HFLIP)) reverse8(w[j]);
        }
        if (ng == 4) group_sum<4>(w, k0 == 0, q, dmax);
        else if (ng == 3) group_sum<3>(w, k0 == 0, q, dmax);
        else if (ng == 2) group_sum<2>(w, k0 == 0, q, dmax);
        else group_sum<1>(w, k0 == 0, q, dmax);
      }
      const uint2 g = ldg_stream8(gtp + (size_t)it * 8);
      // bin index and distance to the grid lines of every pixel
      uint32_t fx[8], worst = 0;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        fx[i] = __float_as_uint(fmaf(q[i], kq, fb));
        worst = max(worst, (fx[i] - lo) & 0xFFFFu);
      }
      if (worst > span || !(dmax < kDLimit)) {
        // rare: the estimate cannot decide (or cannot be trusted for) some
        // pixel of this item - those pixels are evaluated exactly
        const bool all = !(dmax < kDLimit);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          if (all || ((fx[i] - lo) & 0xFFFFu) > span) {
            const int px = ROWS ? xi * 8 + i : it * 8 + i;
            fx[i] = 0x43008000u |
                    (guarded_exact_bin(a, plane_off, ROWS ? y : 0, px, thr)
                     << 16);
          }
        }
      }
      // bin = bits 16..22 of fx (the exponent field reads 0x4300 + bin)
      if ((g.x | g.y) == 0u) {  // all background: the increment is 1
#pragma unroll
        for (int i = 0; i < 8; ++i)
          red_shared_add(cnt_sa + (((fx[i] >> 9) & 0x3F80u) | lane4), 1u);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t gb = ((i < 4 ? g.x : g.y) >> (8 * (i & 3))) & 0xFFu;
          red_shared_add(cnt_sa + (((fx[i] >> 9) & 0x3F80u) | lane4),
                         gb ? 0x10000u : 1u);
        }
      }
      if (BINS) {
        // byte 2 of each fx is its bin (< 128)
        uint2 v;
        v.x = prmt32(prmt32(fx[0], fx[1], 0x0062u), prmt32(fx[2], fx[3], 0x0062u),
                     0x5410u);
        v.y = prmt32(prmt32(fx[4], fx[5], 0x0062u), prmt32(fx[6], fx[7], 0x0062u),
                     0x5410u);
        *reinterpret_cast<uint2 *>(a.bins + plane_off + (size_t)it * 8) = v;
      }
      if (ROWS) {
        xi += dx;
        y += dy;
        if (xi >= ipr) {
          xi -= ipr;
          ++y;
        }
      }
    }
    uint32_t *out = a.hist + (size_t)plane * 2 * nbins;
    flush_bins(cnt, nbins, out, a.units_per_plane > 1);
  }
}

bool seg_hist_guarded_launch(HistArgs &a, bool rows, cudaStream_t st, int *rc) {
  *rc = MLC_OK;
  const int B = a.B, S = a.src.n;
  if (B < 2 || B > 120 || S < 1) return false;
  if (a.plane_px % 8 != 0 || a.items_per_plane != a.plane_px / 8) return false;
  if (a.plane_px / 8 >= (1ll << 31) - 2 * kGThreads) return false;
  // the grid must be uniform over the unit interval: f(p) = p*scale + bias
  // puts thr[k] at k + 1, p = 0 at >= 0 and p = 1 at <= B + 1
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
  // error of the estimated mean against the reference mean, in grid steps,
  // + grid irregularity + rounding of scale / S + rounding of the final fma
  const double eps = sc * (24.0 + 0.25 * S) * u + g + u * (B + 1) +
                     7.62939453125e-6;
  const double lo = eps * 65536.0 + 2.0;
  if (lo > 4096.0) return false;
  a.g_kq = (float)(sc / (double)S);
  a.g_fb = (float)(bi + 128.0);
  a.g_lo = (uint32_t)lo + 1u;
  a.g_span = 65535u - 2u * a.g_lo;
  const size_t smem = (size_t)(B + 1) * 32 * 4 + (size_t)B * 4;
  const long long n_units = a.n_planes * a.units_per_plane;
  const int grid = (int)(n_units < (1ll << 22) ? n_units : (1ll << 22));
  void (*k)(HistArgs) =
      a.bins ? (rows ? seg_hist_guarded<true, true> : seg_hist_guarded<false, true>)
             : (rows ? seg_hist_guarded<true, false>
                     : seg_hist_guarded<false, false>);
  k<<<grid, kGThreads, smem, st>>>(a);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) *rc = cuda_fail(e, "seg_hist_guarded");
  return true;
}

}  // namespace mlc

#ifdef MLC_DEV_KNOBS
// Developer builds only: pixels the guarded kernel sent to its exact path
// since the last call (reads and clears the counter; synchronises).
extern "C" long long mlc_dev_guard_slow_pixels(void) {
  unsigned long long v = 0, z = 0;
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(&v, mlc::g_guard_slow_px, sizeof(v));
  cudaMemcpyToSymbol(mlc::g_guard_slow_px, &z, sizeof(z));
  return (long long)v;
}
#endif
