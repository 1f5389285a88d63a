// K1: fused TTA un-flip + mean + sigmoid + label-conditioned threshold
// histogram, one HBM pass.  See include/mlc.h (mlc_seg_hist) for the contract.
//
// Shared-memory histogram layout (both kernels): word [bin][lane], 32 lanes
// per bin, so a lane only ever touches bank == its lane id: the atomics of
// one warp instruction never collide, whatever the probability distribution
// looks like (masks are mostly background, i.e. nearly every pixel of a warp
// lands in the same few bins).  Each word packs two 16-bit counters, low =
// pixels with gt 0, high = pixels with gt 1; a CTA flushes to global memory
// before a lane column can reach 65 535.
#include "seg_hist.cuh"

#include <cstdlib>

namespace mlc {

// ---- generic kernel: any number of sources, flips, f16/f32 ---------------
// Binning: cnt(p) = #{k : p >= thr[k]}.  A linear first guess c is checked
// against ONE 8-byte shared-memory entry pair[c] = (thr[c-1], thr[c]) (with
// -inf / +inf sentinels): cnt == c  <=>  thr[c-1] <= p < thr[c].  Only when
// the check fails (guess off by one, NaN, +inf, irregular table) does the
// pixel walk the table (count_ge), so the result is exact for any ascending
// table and costs one LDS.64 on a near-uniform one.
// VEC: flat mode with 16-byte-aligned tensors - every full item after the
// first of a plane is loaded with fused_flat8.
// MODE 0: any geometry; 1 (VEC): flat aligned planes; 2: flipped rows of
// whole aligned items (fused_rows8) - the fp32 / no-activation twin of
// seg_hist_tab<2>.
// BINS: also write the per-pixel bin index (a separate instantiation: the
// plain one pays nothing for it - measured 6-10 % on the table kernel when
// it was a run-time test).
template <typename T, int MODE, bool SM, bool BINS>
__global__ void __launch_bounds__(kHistThreads, 2)
seg_hist_generic(const __grid_constant__ HistArgs a) {
  constexpr bool VEC = MODE == 1;
  extern __shared__ uint32_t smem[];
  const int B = a.B, nbins = B + 1;
  const float scale = a.scale, bias = a.bias;
  uint32_t *cnt = smem;                                        // [nbins][32]
  float2 *pair = reinterpret_cast<float2 *>(smem + nbins * 32);  // [nbins]
  float *thr = reinterpret_cast<float *>(pair + nbins);          // [B]
  for (int i = threadIdx.x; i < nbins * 32; i += blockDim.x) cnt[i] = 0;
  for (int i = threadIdx.x; i < B; i += blockDim.x) thr[i] = a.thr.v[i];
  for (int i = threadIdx.x; i < nbins; i += blockDim.x)
    pair[i] = make_float2(i > 0 ? a.thr.v[i - 1] : -INFINITY,
                          i < B ? a.thr.v[i] : INFINITY);
  __syncthreads();
  const uint32_t lane4 = (threadIdx.x & 31) * 4;
  const uint32_t cnt_sa = (uint32_t)__cvta_generic_to_shared(cnt);
  const uint32_t pair_sa = (uint32_t)__cvta_generic_to_shared(pair);
  const long long n_units = a.n_planes * a.units_per_plane;

  for (long long unit = blockIdx.x; unit < n_units; unit += gridDim.x) {
    const long long plane = unit / a.units_per_plane;
    const int sub = (int)(unit - plane * a.units_per_plane);
    const size_t plane_off = (size_t)plane * a.plane_px;
    const bool inner = plane > 0 && plane + 1 < a.n_planes;
    const uint8_t *gtp = a.gt + plane_off;
    const int shift = a.flat ? (int)(plane_off & 7) : 0;
    const long long n_items =
        a.flat ? (a.plane_px + shift + 7) / 8 : a.items_per_plane;
    const long long i0 = (long long)sub * a.items_per_unit;
    const long long i1 = min(n_items, i0 + a.items_per_unit);

    // bins of one item's eight probabilities -> the warp-lane counters
    auto bin8 = [&](const float (&p)[8], const uint2 g, const int nv,
                    const size_t px_off) {
      int c[8];
      bool ok = nv == 8;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        c[i] = max(0, min(__float2int_rd(fmaf(p[i], scale, bias)), B));
        const float2 e = lds_f2(pair_sa + c[i] * 8);
        ok &= p[i] >= e.x && p[i] < e.y;
      }
      if (ok) {  // all 8 guesses verified: the common case
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t gb = ((i < 4 ? g.x : g.y) >> (8 * (i & 3))) & 0xFFu;
          red_shared_add(cnt_sa + ((c[i] << 7) | lane4),
                         gb ? 0x10000u : 1u);
        }
        if (BINS) store_bins8(a.bins + plane_off + px_off, c, 8);
      } else {  // partial item, guess off by one, NaN / +inf, odd table
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          if (i < nv) {
            const int cc = count_ge(p[i], thr, B, scale, bias);
            c[i] = cc;
            const uint32_t gb =
                ((i < 4 ? g.x : g.y) >> (8 * (i & 3))) & 0xFFu;
            red_shared_add(cnt_sa + ((cc << 7) | lane4), gb ? 0x10000u : 1u);
          }
        }
        if (BINS) store_bins8(a.bins + plane_off + px_off, c, nv);
      }
    };
    // one item through the bounds-checked loads (any mode, any alignment)
    auto item_checked = [&](const long long it) {
      int y, x0, nv;
      if (a.flat) {
        y = 0;
        x0 = it == 0 ? 0 : (int)(it * 8 - shift);
        nv = it == 0 ? min(8 - shift, a.W) : min(8, a.W - x0);
      } else if (it < a.full_items) {  // the W/8 full items of each row first
        y = (int)((unsigned)it / (unsigned)a.items_per_row);
        x0 = ((int)it - y * a.items_per_row) * 8;
        nv = 8;
      } else {  // then the H row tails together: warps do not mix the two
        y = (int)(it - a.full_items);
        x0 = a.items_per_row * 8;
        nv = a.W & 7;
      }
      float p[8];
      const PlaneCtx cx = {plane, a.n_planes, a.plane_px, a.n_cls};
      fused_prob8<T, SM>(a.src, plane_off, cx, a.H, a.W, y, x0, nv, inner,
                         a.act, p);
      bin8(p, fetch8_bytes(gtp, a.H, a.W, y, x0, nv, inner), nv,
           (size_t)y * a.W + x0);
    };

    if (VEC && !SM) {
      // Aligned flat planes: every item but the plane's head (when the plane
      // does not start on an 8-element boundary) and its partial tail is one
      // 16-byte load per source and 8 label bytes, indexed in 32 bits.
      const int f0 = shift ? 1 : 0;
      const int f1 = (int)((a.plane_px + shift) >> 3);  // first non-full item
      const int lo = max((int)i0, f0), hi = min((int)i1, f1);
      for (int it = lo + (int)threadIdx.x; it < hi; it += blockDim.x) {
        const int x0 = it * 8 - shift;
        float p[8];
        fused_flat8<T>(a.src, plane_off + x0, a.act, p);
        mean8(p, a.src.n);
        bin8(p, ldg_stream8(gtp + x0), 8, (size_t)x0);
      }
      if (threadIdx.x == 0 && f0 && i0 == 0) item_checked(0);
      if (threadIdx.x == 32 && f1 >= f0 && f1 < n_items && f1 >= i0 &&
          f1 < i1)
        item_checked(f1);
    } else if (MODE == 2 && !SM) {
      const int ipr = a.items_per_row;  // W / 8, every item full
      const int dy = (int)blockDim.x / ipr, dx = (int)blockDim.x - dy * ipr;
      int it = (int)i0 + (int)threadIdx.x;
      int y = it / ipr, xi = it - y * ipr;
      for (; it < (int)i1; it += blockDim.x) {
        float p[8];
        fused_rows8<T>(a.src, plane_off, a.H, ipr, y, xi, a.act, p);
        mean8(p, a.src.n);
        bin8(p, ldg_stream8(gtp + (size_t)it * 8), 8, (size_t)it * 8);
        xi += dx;
        y += dy;
        if (xi >= ipr) {
          xi -= ipr;
          ++y;
        }
      }
    } else {
      for (long long it = i0 + threadIdx.x; it < i1; it += blockDim.x)
        item_checked(it);
    }
    uint32_t *out = a.hist + (size_t)plane * 2 * nbins;
    flush_bins(cnt, nbins, out, a.units_per_plane > 1);
  }
}

// ---- activation-table kernel: any number of fp16 sources -----------------
// seg_hist_generic with act() of every source value read from the shared-
// memory table of mlc_common.cuh (act8_table) instead of evaluated: the
// multi-source configurations (4-flip TTA, 5-model ensemble) are otherwise
// bound by the sigmoid arithmetic (2 MUFU + ~14 instructions per value), not
// by HBM.  One 1024-thread CTA per SM; the counters, the (thr[c-1], thr[c])
// pairs and the threshold table live in the hole of the table's layout.
constexpr int kTabThreads = 1024;
// Row mode (MODE 0): sources whose enclosing vectors are loaded before the
// first is used, and the CTA size.  Measured on C3 (1024 x 4 x 350x525, 4
// flips): 1 source / 1024 threads 3.01 ms; 2 sources / 1024 threads 4.63 ms
// (spills at the 64-register cap); 4 sources / 512 threads (128 registers)
// 3.57 ms; 2 / 512: 3.89 ms.  The word-load version before it: 3.15 ms.
#ifndef MLC_ROW_BATCH
#define MLC_ROW_BATCH 1
#endif
constexpr int kRowBatch = MLC_ROW_BATCH;
#ifndef MLC_ROW_THREADS
#define MLC_ROW_THREADS 1024
#endif
constexpr int kTabRowThreads = MLC_ROW_THREADS;

// MODE 0: any geometry (window loads); 1: flat planes, aligned tensors;
// 2: flipped sources (no transpose) whose rows are a whole number of aligned
// 16-byte items - the flips are index arithmetic on items, every load is one
// 128-bit load and up to four sources are in flight together.
template <int MODE, bool BINS>
__global__ void __launch_bounds__(MODE == 0 ? kTabRowThreads : kTabThreads, 1)
seg_hist_tab(const __grid_constant__ HistArgs a) {
  constexpr bool VEC = MODE == 1;
  extern __shared__ __align__(16) unsigned char tab_smem[];
  const int B = a.B, nbins = B + 1;
  const float scale = a.scale, bias = a.bias;
  uint32_t *cnt = reinterpret_cast<uint32_t *>(tab_smem + kActTabHoleOff);
  float2 *pair = reinterpret_cast<float2 *>(cnt + nbins * 32);
  float *thr = reinterpret_cast<float *>(pair + nbins);
  act_table_build(tab_smem, a.act);
  for (int i = threadIdx.x; i < nbins * 32; i += blockDim.x) cnt[i] = 0;
  for (int i = threadIdx.x; i < B; i += blockDim.x) thr[i] = a.thr.v[i];
  for (int i = threadIdx.x; i < nbins; i += blockDim.x)
    pair[i] = make_float2(i > 0 ? a.thr.v[i - 1] : -INFINITY,
                          i < B ? a.thr.v[i] : INFINITY);
  __syncthreads();
  const uint32_t lane4 = (threadIdx.x & 31) * 4;
  const uint32_t tab_sa = (uint32_t)__cvta_generic_to_shared(tab_smem);
  const uint32_t cnt_sa = (uint32_t)__cvta_generic_to_shared(cnt);
  const uint32_t pair_sa = (uint32_t)__cvta_generic_to_shared(pair);
  const long long n_units = a.n_planes * a.units_per_plane;
  const int S = a.src.n;

  for (long long unit = blockIdx.x; unit < n_units; unit += gridDim.x) {
    const long long plane = unit / a.units_per_plane;
    const int sub = (int)(unit - plane * a.units_per_plane);
    const size_t plane_off = (size_t)plane * a.plane_px;
    const bool inner = plane > 0 && plane + 1 < a.n_planes;
    const uint8_t *gtp = a.gt + plane_off;
    const int shift = a.flat ? (int)(plane_off & 7) : 0;
    const long long n_items =
        a.flat ? (a.plane_px + shift + 7) / 8 : a.items_per_plane;
    const long long i0 = (long long)sub * a.items_per_unit;
    const long long i1 = min(n_items, i0 + a.items_per_unit);

    auto bin8 = [&](float (&p)[8], const uint2 g, const int nv,
                    const size_t px_off) {
      mean8(p, S);
      int c[8];
      bool ok = nv == 8;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        c[i] = max(0, min(__float2int_rd(fmaf(p[i], scale, bias)), B));
        const float2 e = lds_f2(pair_sa + c[i] * 8);
        ok &= p[i] >= e.x && p[i] < e.y;
      }
      if (ok) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t gb = ((i < 4 ? g.x : g.y) >> (8 * (i & 3))) & 0xFFu;
          red_shared_add(cnt_sa + ((c[i] << 7) | lane4),
                         gb ? 0x10000u : 1u);
        }
        if (BINS) store_bins8(a.bins + plane_off + px_off, c, 8);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          if (i < nv) {
            const int cc = count_ge(p[i], thr, B, scale, bias);
            c[i] = cc;
            const uint32_t gb =
                ((i < 4 ? g.x : g.y) >> (8 * (i & 3))) & 0xFFu;
            red_shared_add(cnt_sa + ((cc << 7) | lane4), gb ? 0x10000u : 1u);
          }
        }
        if (BINS) store_bins8(a.bins + plane_off + px_off, c, nv);
      }
    };
    // one item through the window / bounds-checked loads (rows, flips, edges)
    auto item_checked = [&](const long long it) {
      int y, x0, nv;
      if (a.flat) {
        y = 0;
        x0 = it == 0 ? 0 : (int)(it * 8 - shift);
        nv = it == 0 ? min(8 - shift, a.W) : min(8, a.W - x0);
      } else if (it < a.full_items) {  // the W/8 full items of each row first
        y = (int)((unsigned)it / (unsigned)a.items_per_row);
        x0 = ((int)it - y * a.items_per_row) * 8;
        nv = 8;
      } else {  // then the H row tails together: warps do not mix the two
        y = (int)(it - a.full_items);
        x0 = a.items_per_row * 8;
        nv = a.W & 7;
      }
      float p[8];
      for (int k = 0; k < S; ++k) {
        const __half *pl = static_cast<const __half *>(a.src.ptr[k]) +
                           plane_off;
        const int xf = a.src.xf[k];
        float v[8];
        bool done = false;
        if (!(xf & MLC_XF_TRANSPOSE) && nv == 8) {
          const int ys = (xf & MLC_XF_VFLIP) ? a.H - 1 - y : y;
          const bool rev = (xf & MLC_XF_HFLIP) != 0;
          const long long s0 =
              (long long)ys * a.W + (rev ? a.W - 8 - x0 : x0);
          if (inner || (s0 >= 8 && s0 + 16 <= a.plane_px)) {
            uint32_t w[4];
            window8_packed(pl + s0, rev, w);
            act8_table(w, tab_sa, a.act, v);
            done = true;
          }
        }
        if (!done) {
          fetch8<__half>(pl, xf, a.H, a.W, y, x0, nv, inner, v);
          activate8(v, a.act);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) p[i] = k == 0 ? v[i] : p[i] + v[i];
      }
      bin8(p, fetch8_bytes(gtp, a.H, a.W, y, x0, nv, inner), nv,
           (size_t)y * a.W + x0);
    };

    if (VEC) {
      // aligned flat planes: full items in a 32-bit-indexed loop, one 128-bit
      // load per source with 4 sources in flight; head and tail items apart
      const int f0 = shift ? 1 : 0;
      const int f1 = (int)((a.plane_px + shift) >> 3);  // first non-full item
      const int lo = max((int)i0, f0), hi = min((int)i1, f1);
      for (int it = lo + (int)threadIdx.x; it < hi; it += blockDim.x) {
        const int x0 = it * 8 - shift;
        const size_t e0 = plane_off + x0;
        float p[8];
        for (int k0 = 0; k0 < S; k0 += 4) {
          uint4 r[4];
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (k0 + j < S)
              r[j] = ldg_stream16(
                  static_cast<const __half *>(a.src.ptr[k0 + j]) + e0);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (k0 + j < S) {
              const uint32_t w[4] = {r[j].x, r[j].y, r[j].z, r[j].w};
              float v[8];
              act8_table(w, tab_sa, a.act, v);
#pragma unroll
              for (int i = 0; i < 8; ++i)
                p[i] = (k0 + j == 0) ? v[i] : p[i] + v[i];
            }
          }
        }
        bin8(p, ldg_stream8(gtp + x0), 8, (size_t)x0);
      }
      if (threadIdx.x == 0 && f0 && i0 == 0) item_checked(0);
      if (threadIdx.x == 32 && f1 >= f0 && f1 < n_items && f1 >= i0 &&
          f1 < i1)
        item_checked(f1);
    } else if (MODE == 2) {
      const int ipr = a.items_per_row;  // W / 8, every item full
      const int dy = (int)blockDim.x / ipr, dx = (int)blockDim.x - dy * ipr;
      int it = (int)i0 + (int)threadIdx.x;
      int y = it / ipr, xi = it - y * ipr;
      for (; it < (int)i1; it += blockDim.x) {
        float p[8];
        for (int k0 = 0; k0 < S; k0 += 4) {
          uint4 r[4];
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (k0 + j < S) {
              const int xf = a.src.xf[k0 + j];
              const int ys = (xf & MLC_XF_VFLIP) ? a.H - 1 - y : y;
              const int xs = (xf & MLC_XF_HFLIP) ? ipr - 1 - xi : xi;
              r[j] = ldg_stream16(
                  static_cast<const __half *>(a.src.ptr[k0 + j]) + plane_off +
                  (size_t)(ys * ipr + xs) * 8);
            }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (k0 + j < S) {
              uint32_t w[4] = {r[j].x, r[j].y, r[j].z, r[j].w};
              if (a.src.xf[k0 + j] & MLC_XF_HFLIP) {  // reverse the 8 halves
                const uint32_t t0 = prmt32(w[3], 0u, 0x1032u),
                               t1 = prmt32(w[2], 0u, 0x1032u),
                               t2 = prmt32(w[1], 0u, 0x1032u),
                               t3 = prmt32(w[0], 0u, 0x1032u);
                w[0] = t0; w[1] = t1; w[2] = t2; w[3] = t3;
              }
              float v[8];
              act8_table(w, tab_sa, a.act, v);
#pragma unroll
              for (int i = 0; i < 8; ++i)
                p[i] = (k0 + j == 0) ? v[i] : p[i] + v[i];
            }
          }
        }
        bin8(p, ldg_stream8(gtp + (size_t)it * 8), 8, (size_t)it * 8);
        xi += dx;
        y += dy;
        if (xi >= ipr) {
          xi -= ipr;
          ++y;
        }
      }
    } else if (!a.flat && !(a.src.any_xf & MLC_XF_TRANSPOSE) &&
               a.items_per_row > 0 && a.W >= 16) {
      // Rows of any width / alignment (350 x 525 ...).  Full items of the
      // rows whose neighbourhood stays inside the tensor: every source is
      // read as the two aligned 16-byte vectors enclosing its window, two
      // sources' loads in flight before the first is used (the per-source
      // word loads of item_checked keep 20 bytes per thread in flight: ncu
      // showed 40 % of the samples waiting on them).  Outermost rows and row
      // tails go through the checked path.
      const int ipr = a.items_per_row;
      const int dy = (int)blockDim.x / ipr, dx = (int)blockDim.x - dy * ipr;
      const int full_end = (int)min(i1, a.full_items);
      int it = (int)i0 + (int)threadIdx.x;
      int y = it / ipr, xi = it - y * ipr;
      for (; it < full_end; it += blockDim.x) {
        if (inner || (y >= 1 && y <= a.H - 2)) {
          const int x0 = xi * 8;
          float p[8];
          for (int k0 = 0; k0 < S; k0 += kRowBatch) {
            uint32_t t[kRowBatch][8];
            uint32_t off[kRowBatch];
#pragma unroll
            for (int j = 0; j < kRowBatch; ++j)
              if (k0 + j < S) {
                const int xf = a.src.xf[k0 + j];
                const int ys = (xf & MLC_XF_VFLIP) ? a.H - 1 - y : y;
                const int xs = (xf & MLC_XF_HFLIP) ? a.W - 8 - x0 : x0;
                const uintptr_t q = reinterpret_cast<uintptr_t>(
                    static_cast<const __half *>(a.src.ptr[k0 + j]) +
                    plane_off + (size_t)ys * a.W + xs);
                const uintptr_t b = q & ~(uintptr_t)15;
                off[j] = (uint32_t)(q & 15) >> 1;
                const uint4 lo = ldg_stream16(reinterpret_cast<const void *>(b));
                const uint4 hi =
                    ldg_stream16(reinterpret_cast<const void *>(b + 16));
                t[j][0] = lo.x; t[j][1] = lo.y; t[j][2] = lo.z; t[j][3] = lo.w;
                t[j][4] = hi.x; t[j][5] = hi.y; t[j][6] = hi.z; t[j][7] = hi.w;
              }
#pragma unroll
            for (int j = 0; j < kRowBatch; ++j) {
              if (k0 + j < S) {
                uint32_t w[4];
                window8_from_enclosing(
                    t[j], off[j], (a.src.xf[k0 + j] & MLC_XF_HFLIP) != 0, w);
                float v[8];
                act8_table(w, tab_sa, a.act, v);
#pragma unroll
                for (int i = 0; i < 8; ++i)
                  p[i] = (k0 + j == 0) ? v[i] : p[i] + v[i];
              }
            }
          }
          bin8(p, fetch8_bytes(gtp, a.H, a.W, y, x0, 8, true), 8,
               (size_t)y * a.W + x0);
        } else {
          item_checked(it);
        }
        xi += dx;
        y += dy;
        if (xi >= ipr) {
          xi -= ipr;
          ++y;
        }
      }
      for (long long jt = max(i0, a.full_items) + threadIdx.x; jt < i1;
           jt += blockDim.x)
        item_checked(jt);
    } else {
      for (long long it = i0 + threadIdx.x; it < i1; it += blockDim.x)
        item_checked(it);
    }
    uint32_t *out = a.hist + (size_t)plane * 2 * nbins;
    flush_bins(cnt, nbins, out, a.units_per_plane > 1);
  }
}

// ---- table kernel: ONE untransformed fp16 source --------------------------
// sigmoid is a function of the 16-bit code alone, so cnt(sigmoid(x)) is a
// 64 Ki-entry byte table (built once per threshold grid by lut_build with the
// very same activate()/compare code, hence identical results by
// construction).  No MUFU work, no divide: the kernel is a pure stream.
__global__ void lut_build(ThrTable t, int B, int act, uint8_t *lut) {
  const int code = blockIdx.x * blockDim.x + threadIdx.x;
  if (code >= 65536) return;
  const float x = __half2float(__ushort_as_half((unsigned short)code));
  const float p = activate(x, act);
  int c = 0;
  for (int k = 0; k < B; ++k) c += (p >= t.v[k]) ? 1 : 0;  // NaN -> 0
  lut[code] = (uint8_t)c;
}

constexpr int kLutUnroll = 4;

#ifdef MLC_DEV_KNOBS
__global__ void __launch_bounds__(kHistThreads, 2)
seg_hist_lut(const __grid_constant__ HistArgs a) {
  extern __shared__ uint32_t smem[];
  const int nbins = a.B + 1;
  uint8_t *lut = reinterpret_cast<uint8_t *>(smem);   // 64 KiB
  uint32_t *cnt = smem + 65536 / 4;                   // [nbins][32]
  {
    const uint4 *g = reinterpret_cast<const uint4 *>(a.lut);
    uint4 *s = reinterpret_cast<uint4 *>(lut);
    for (int i = threadIdx.x; i < 65536 / 16; i += blockDim.x) s[i] = g[i];
    for (int i = threadIdx.x; i < nbins * 32; i += blockDim.x) cnt[i] = 0;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31;
  uint32_t *col = cnt + lane;
  const long long n_units = a.n_planes * a.units_per_plane;
  const __half *src = static_cast<const __half *>(a.src.ptr[0]);

  for (long long unit = blockIdx.x; unit < n_units; unit += gridDim.x) {
    const long long plane = unit / a.units_per_plane;
    const int sub = (int)(unit - plane * a.units_per_plane);
    const size_t plane_off = (size_t)plane * a.plane_px;
    const uint4 *lg = reinterpret_cast<const uint4 *>(src + plane_off);
    const uint2 *gt = reinterpret_cast<const uint2 *>(a.gt + plane_off);
    const long long i0 = (long long)sub * a.items_per_unit;
    const long long i1 = min(a.items_per_plane, i0 + a.items_per_unit);

    for (long long base = i0; base < i1;
         base += (long long)kLutUnroll * kHistThreads) {
      uint4 v[kLutUnroll];
      uint2 g[kLutUnroll];
#pragma unroll
      for (int u = 0; u < kLutUnroll; ++u) {
        const long long it = base + u * kHistThreads + threadIdx.x;
        if (it < i1) {
          v[u] = ldg_stream16(lg + it);
          g[u] = ldg_stream8(gt + it);
        }
      }
#pragma unroll
      for (int u = 0; u < kLutUnroll; ++u) {
        const long long it = base + u * kHistThreads + threadIdx.x;
        if (it < i1) {
          const uint32_t w[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint32_t code = (i & 1) ? (w[i >> 1] >> 16)
                                          : (w[i >> 1] & 0xffffu);
            const uint32_t gb = ((i < 4 ? g[u].x : g[u].y) >> (8 * (i & 3))) &
                                0xffu;
            const uint32_t bin = lut[code];
            atomicAdd(col + bin * 32, gb ? 0x10000u : 1u);
          }
        }
      }
    }
    uint32_t *out = a.hist + (size_t)plane * 2 * nbins;
    flush_bins(cnt, nbins, out, a.units_per_plane > 1);
  }
}

#endif  // MLC_DEV_KNOBS

// ---- cell-table kernel: the same job with conflict-free look-ups ----------
// ncu on seg_hist_lut (profiles/README.md): the LSU data pipe is 95 % busy
// because 32 random byte look-ups hit 32 random banks (3.3 wavefronts per
// LDS).  Neighbouring fp16 codes almost always share a bin, so a table with
// one 16-bit entry per CELL of 32 consecutive codes (2 048 cells, 4 KiB) is
// enough: entry = base_bin * 32 + c, bin = (entry + pos) >> 5, where pos is
// the code's position inside the cell counted in the direction of increasing
// x (low five bits, inverted for negative codes) and c = 32 - (position of
// the single bin edge inside the cell), 0 if there is none.  One 32-bit word
// now covers 64 codes, the codes of one warp fall into a few dozen adjacent
// words, and adjacent words live in different banks: ~1 wavefront per LDS.
// Cells with more than one edge or a non-monotone sequence (the +-inf / NaN
// cells, or a very dense custom threshold table) hold 0xFFFF and send the
// pixel to the byte table in global memory, so results stay identical to
// seg_hist_lut / seg_hist_generic for ANY table.
constexpr int kCells = 2048;

__global__ void cell_build(uint8_t *lut) {
  const int cell = blockIdx.x * blockDim.x + threadIdx.x;
  if (cell >= kCells) return;
  const bool neg = cell >= kCells / 2;  // code bit 15
  const uint8_t *src = lut + cell * 32;
  const int base = src[neg ? 31 : 0];
  int edge = 32;
  bool ok = true;
  for (int j = 1; j < 32; ++j) {
    const int v = src[neg ? 31 - j : j];
    if (edge == 32) {
      if (v == base + 1) edge = j;
      else if (v != base) ok = false;
    } else if (v != base + 1) {
      ok = false;
    }
  }
  uint16_t *cells = reinterpret_cast<uint16_t *>(lut + 65536);
  cells[cell] = ok ? (uint16_t)(base * 32 + (edge < 32 ? 32 - edge : 0))
                   : (uint16_t)0xFFFF;
}

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t d;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
  return d;
}


// Static shared memory on purpose: its addresses are link-time constants, so
// every LDS / ATOMS below is [register + immediate] with no address add.
struct CellSmem {
  uint16_t cells[kCells];                 // entries x4, 0xFFFF = escape
  uint32_t cnt[(MLC_MAX_THR + 1) * 32];   // [bin][lane]
};

// Slow path of one 8-pixel item: at least one of its cells escaped to the
// exact byte table in global memory (inf/NaN cells, very dense tables).
// Taken per ITEM, not per pixel: a per-pixel test gets if-converted and its
// predicated instructions then sit in the issue stream of every pixel.
__device__ __forceinline__ void cell_item_slow(const uint8_t *lut, uint4 v,
                                            uint2 g, uint32_t cnt_sa,
                                            uint32_t lane4) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const uint32_t code = (i & 1) ? w[i >> 1] >> 16 : w[i >> 1] & 0xFFFFu;
    const uint32_t gb = ((i < 4 ? g.x : g.y) >> (8 * (i & 3))) & 0xFFu;
    const uint32_t bin = __ldg(lut + code);
    red_shared_add(cnt_sa + ((bin << 7) | lane4), gb ? 0x10000u : 1u);
  }
}

// Fast path of one item; ALL_BG = its 8 truth bytes are zero (the common
// case): the increment is the constant 1.
template <bool ALL_BG>
__device__ __forceinline__ void cell_item_fast(const uint32_t t[4],
                                               const uint32_t e[8], uint2 g,
                                               uint32_t cnt_sa,
                                               uint32_t lane4) {
  uint32_t nz[2] = {0u, 0u};
  if (!ALL_BG) {
    nz[0] = __vcmpne4(g.x, 0u) & 0x01010101u;
    nz[1] = __vcmpne4(g.y, 0u) & 0x01010101u;
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const uint32_t tt = t[i >> 1];
    const uint32_t pos4 = (i & 1) ? (tt >> 14) & 0x7Cu : (tt << 2) & 0x7Cu;
    const uint32_t x4 = e[i] + pos4;
    uint32_t inc = 1u;
    if (!ALL_BG)
      inc = prmt(nz[i >> 2], 0u, 0x4440u | (i & 3)) * 0xFFFFu + 1u;
    red_shared_add(cnt_sa + ((x4 & ~127u) | lane4), inc);
  }
}

template <int NT, int UNROLL, int MINB>
__global__ void __launch_bounds__(NT, MINB)
seg_hist_cell(const __grid_constant__ HistArgs a) {
  constexpr int kCellUnroll = UNROLL;
  constexpr int kHistThreads = NT;
  __shared__ __align__(16) CellSmem sm;
  const int nbins = a.B + 1;
  {
    // entries x4: entry4 + pos*4 = 4*(base*32 + c + pos), and
    // (that & ~127) = bin * 128 = byte offset of the bin's row of counters
    const uint16_t *g = reinterpret_cast<const uint16_t *>(a.lut + 65536);
    for (int i = threadIdx.x; i < kCells; i += blockDim.x) {
      const uint16_t e = g[i];
      sm.cells[i] = e == 0xFFFFu ? e : (uint16_t)(e * 4u);
    }
    for (int i = threadIdx.x; i < nbins * 32; i += blockDim.x) sm.cnt[i] = 0;
  }
  __syncthreads();
  const uint32_t lane4 = (threadIdx.x & 31) * 4;
  const uint32_t cells_sa = (uint32_t)__cvta_generic_to_shared(sm.cells);
  const uint32_t cnt_sa = (uint32_t)__cvta_generic_to_shared(sm.cnt);
  const long long n_units = a.n_planes * a.units_per_plane;
  const __half *src = static_cast<const __half *>(a.src.ptr[0]);

  for (long long unit = blockIdx.x; unit < n_units; unit += gridDim.x) {
    const long long plane = unit / a.units_per_plane;
    const int sub = (int)(unit - plane * a.units_per_plane);
    const size_t plane_off = (size_t)plane * a.plane_px;
    const uint4 *lg = reinterpret_cast<const uint4 *>(src + plane_off);
    const uint2 *gt = reinterpret_cast<const uint2 *>(a.gt + plane_off);
    const long long i0 = (long long)sub * a.items_per_unit;
    const long long i1 = min(a.items_per_plane, i0 + a.items_per_unit);

    for (long long base = i0; base < i1;
         base += (long long)kCellUnroll * kHistThreads) {
      uint4 v[kCellUnroll];
      uint2 g[kCellUnroll];
#pragma unroll
      for (int u = 0; u < kCellUnroll; ++u) {
        const long long it = base + u * kHistThreads + threadIdx.x;
        if (it < i1) {
          v[u] = ldg_stream16(lg + it);
          g[u] = ldg_stream8(gt + it);
        }
      }
#pragma unroll
      for (int u = 0; u < kCellUnroll; ++u) {
        const long long it = base + u * kHistThreads + threadIdx.x;
        if (it >= i1) continue;
        const uint32_t w[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
        uint32_t t[4], e[8];
#pragma unroll
        for (int h = 0; h < 4; ++h) {
          // sign mask per half, then invert the low 5 bits of negative
          // codes so that `pos` runs in the direction of increasing x
          const uint32_t m = prmt(w[h], 0u, 0xBB99u);
          t[h] = w[h] ^ (m & 0x001F001Fu);
          e[2 * h] = lds_u16(cells_sa + ((t[h] >> 4) & 0xFFEu));
          e[2 * h + 1] = lds_u16(cells_sa + ((t[h] >> 20) & 0xFFEu));
        }
        const uint32_t any = (e[0] | e[1] | e[2]) | (e[3] | e[4] | e[5]) |
                             (e[6] | e[7]);
        if (any & 0x8000u)
          cell_item_slow(a.lut, v[u], g[u], cnt_sa, lane4);
        else if ((g[u].x | g[u].y) == 0u)
          cell_item_fast<true>(t, e, g[u], cnt_sa, lane4);
        else
          cell_item_fast<false>(t, e, g[u], cnt_sa, lane4);
      }
    }
    uint32_t *out = a.hist + (size_t)plane * 2 * nbins;
    flush_bins(sm.cnt, nbins, out, a.units_per_plane > 1);
  }
}

#ifdef MLC_DEV_KNOBS
// ---- cell-table kernel fed by bulk async copies (TMA, 1-D) ---------------
// Same arithmetic as seg_hist_cell.  A producer warp streams the plane into a
// ring of shared-memory stages with cp.async.bulk (UBLKCP) signalled through
// mbarriers; the 512 consumer threads never wait on a global load, hold no
// staging registers, and the copy engine keeps kTmaStages x 12 KiB per CTA in
// flight whatever the consumers are doing.  Needs plane_px % 16 == 0.
// Kept as an opt-in variant (MLC_HIST_VARIANT=3): on the C2 workload it runs
// at 79-80 % of the measured HBM peak against 84-86 % for seg_hist_cell.
constexpr int kTmaStages = 4;
constexpr int kTmaConsumers = 512;
constexpr int kTmaThreads = kTmaConsumers + 32;

struct TmaStages {
  uint4 lg[kTmaStages][kTmaConsumers];  // 8 KiB of logits per stage
  uint2 gt[kTmaStages][kTmaConsumers];  // 4 KiB of truth per stage
  unsigned long long full[kTmaStages];
  unsigned long long empty[kTmaStages];
};

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "LAB_WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra LAB_DONE_%=;\n"
      "bra LAB_WAIT_%=;\n"
      "LAB_DONE_%=:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src,
                                         uint32_t bytes, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1], %2, [%3];" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}

__global__ void __launch_bounds__(kTmaThreads, 2)
seg_hist_cell_tma(const __grid_constant__ HistArgs a) {
  __shared__ __align__(16) CellSmem sm;
  extern __shared__ __align__(128) unsigned char dyn_smem[];
  TmaStages &st = *reinterpret_cast<TmaStages *>(dyn_smem);
  const int nbins = a.B + 1;
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int i = 0; i < kTmaStages; ++i) {
      mbar_init((uint32_t)__cvta_generic_to_shared(&st.full[i]), 1);
      mbar_init((uint32_t)__cvta_generic_to_shared(&st.empty[i]),
                kTmaConsumers / 32);  // one arrival per consumer warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  {
    const uint16_t *g = reinterpret_cast<const uint16_t *>(a.lut + 65536);
    for (int i = tid; i < kCells; i += blockDim.x) {
      const uint16_t e = g[i];
      sm.cells[i] = e == 0xFFFFu ? e : (uint16_t)(e * 4u);
    }
    for (int i = tid; i < nbins * 32; i += blockDim.x) sm.cnt[i] = 0;
  }
  __syncthreads();
  const long long n_units = a.n_planes * a.units_per_plane;
  const __half *src = static_cast<const __half *>(a.src.ptr[0]);
  int stage = 0;
  uint32_t phase = 0;

  if (tid >= kTmaConsumers) {
    // ---------------- producer warp: one elected lane issues the copies ----
    if (tid == kTmaConsumers) {
      for (long long unit = blockIdx.x; unit < n_units; unit += gridDim.x) {
        const long long plane = unit / a.units_per_plane;
        const int sub = (int)(unit - plane * a.units_per_plane);
        const size_t plane_off = (size_t)plane * a.plane_px;
        const long long i0 = (long long)sub * a.items_per_unit;
        const long long i1 = min(a.items_per_plane, i0 + a.items_per_unit);
        for (long long base = i0; base < i1; base += kTmaConsumers) {
          const uint32_t n = (uint32_t)min((long long)kTmaConsumers, i1 - base);
          const uint32_t full =
              (uint32_t)__cvta_generic_to_shared(&st.full[stage]);
          mbar_wait((uint32_t)__cvta_generic_to_shared(&st.empty[stage]),
                    phase ^ 1u);
          mbar_expect_tx(full, n * 24u);
          bulk_g2s((uint32_t)__cvta_generic_to_shared(&st.lg[stage][0]),
                   src + plane_off + base * 8, n * 16u, full);
          bulk_g2s((uint32_t)__cvta_generic_to_shared(&st.gt[stage][0]),
                   a.gt + plane_off + base * 8, n * 8u, full);
          if (++stage == kTmaStages) { stage = 0; phase ^= 1u; }
        }
      }
    }
    return;
  }

  // ---------------- 512 consumers -------------------------------------------
  const uint32_t lane4 = (tid & 31) * 4;
  const uint32_t cells_sa = (uint32_t)__cvta_generic_to_shared(sm.cells);
  const uint32_t cnt_sa = (uint32_t)__cvta_generic_to_shared(sm.cnt);
  for (long long unit = blockIdx.x; unit < n_units; unit += gridDim.x) {
    const long long plane = unit / a.units_per_plane;
    const int sub = (int)(unit - plane * a.units_per_plane);
    const long long i0 = (long long)sub * a.items_per_unit;
    const long long i1 = min(a.items_per_plane, i0 + a.items_per_unit);
    for (long long base = i0; base < i1; base += kTmaConsumers) {
      const bool mine = base + tid < i1;
      mbar_wait((uint32_t)__cvta_generic_to_shared(&st.full[stage]), phase);
      uint4 v = make_uint4(0, 0, 0, 0);
      uint2 g = make_uint2(0, 0);
      if (mine) {
        v = st.lg[stage][tid];
        g = st.gt[stage][tid];
      }
      // the stage is free once every lane holds its 24 bytes in registers;
      // one elected arrival per warp (32 same-address arrivals would
      // serialise in the LSU)
      __syncwarp();
      if ((tid & 31) == 0)
        mbar_arrive((uint32_t)__cvta_generic_to_shared(&st.empty[stage]));
      if (++stage == kTmaStages) { stage = 0; phase ^= 1u; }
      if (!mine) continue;
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
      uint32_t t[4], e[8];
#pragma unroll
      for (int h = 0; h < 4; ++h) {
        const uint32_t m = prmt32(w[h], 0u, 0xBB99u);
        t[h] = w[h] ^ (m & 0x001F001Fu);
        e[2 * h] = lds_u16(cells_sa + ((t[h] >> 4) & 0xFFEu));
        e[2 * h + 1] = lds_u16(cells_sa + ((t[h] >> 20) & 0xFFEu));
      }
      const uint32_t any = (e[0] | e[1] | e[2]) | (e[3] | e[4] | e[5]) |
                           (e[6] | e[7]);
      if (any & 0x8000u)
        cell_item_slow(a.lut, v, g, cnt_sa, lane4);
      else if ((g.x | g.y) == 0u)
        cell_item_fast<true>(t, e, g, cnt_sa, lane4);
      else
        cell_item_fast<false>(t, e, g, cnt_sa, lane4);
    }
    // flush this plane's counters: barrier over the consumers only
    uint32_t *out = a.hist + (size_t)plane * 2 * nbins;
    asm volatile("bar.sync 1, %0;" ::"n"(kTmaConsumers) : "memory");
    for (int j = tid; j < nbins; j += kTmaConsumers) {
      uint32_t lo = 0, hi = 0;
#pragma unroll 8
      for (int i = 0; i < 32; ++i) {
        const int col = (j + i) & 31;
        const uint32_t wv = sm.cnt[j * 32 + col];
        sm.cnt[j * 32 + col] = 0;
        lo += wv & 0xffffu;
        hi += wv >> 16;
      }
      if (a.units_per_plane > 1) {
        if (lo) atomicAdd(out + j, lo);
        if (hi) atomicAdd(out + nbins + j, hi);
      } else {
        out[j] = lo;
        out[nbins + j] = hi;
      }
    }
    asm volatile("bar.sync 1, %0;" ::"n"(kTmaConsumers) : "memory");
  }
}

#endif  // MLC_DEV_KNOBS

// ---- fused probabilities only --------------------------------------------
struct ProbArgs {
  SrcSet src;
  float *prob;
  long long n_planes;
  long long plane_px;
  int n_cls;
  int H, W;
  int act;
  int items_per_row;
  long long items_per_plane;
};

// Aligned layouts (planes a whole number of 8-element items, 16-byte aligned
// tensors): one 128-bit load per source and item, two 128-bit stores.
// ROWS: flipped sources, rows a whole number of items (fused_rows8);
// otherwise no transforms at all - the tensor is one flat array.
// ---- logit-domain kernel: ONE untransformed source, any element type -------
// With one source no mean is taken, and the activation is non-decreasing over
// its whole input type (identity trivially; the device sigmoid over ALL fp32
// values: exhaustive sweep in tests/test_gpu_fullsize.py), so
//   cnt(act(x)) = #{k : act(x) >= thr[k]} = #{k : x >= x'_k},
// x'_k = the smallest input with act(x'_k) >= thr[k], found per CTA by a
// 32-step bisection over the order-preserving keys with the kernels' own
// activate() (NaN if no input passes).  No activation is evaluated per pixel.
// Binning x against the x'_k: a table of kXCells cells over the span of the
// finite x'_k (affine map, clamped; cell 0 lies below them all), each cell
// holding (number of x'_k in lower cells, the one x'_k inside it or NaN); a
// cell that holds two or more is an escape.  One LDS.64 and one compare per
// pixel give the exact count; NaN pixels compare false everywhere -> 0.
constexpr int kXCells = 2048;
struct XSmem {
  float2 cell[kXCells];                  // (count below as int bits, x' inside)
  uint32_t cnt[(MLC_MAX_THR + 1) * 32];  // [bin][lane]
  float xthr[MLC_MAX_THR + 1];
  float map[2];                          // scale, bias of the cell index
};
constexpr int kXEscape = 0x7FFF;

__device__ __forceinline__ int x_cell(float x, float scale, float bias) {
  // monotone in x; NaN -> 0, +-inf -> the clamped ends
  return max(0, min(__float2int_rd(fmaf(x, scale, bias)), kXCells - 1));
}

template <typename T>
__global__ void __launch_bounds__(256, 3)
seg_hist_logit(const __grid_constant__ HistArgs a) {
  extern __shared__ __align__(16) unsigned char xraw[];
  XSmem &sm = *reinterpret_cast<XSmem *>(xraw);
  const int B = a.B, nbins = B + 1;
  // ---- per-CTA set-up ------------------------------------------------------
  for (int i = threadIdx.x; i < nbins * 32; i += blockDim.x) sm.cnt[i] = 0;
  if ((int)threadIdx.x < B) {
    const float thr = a.thr.v[threadIdx.x];
    const uint32_t k_ninf = 0x007FFFFFu, k_pinf = 0xFF800000u;
    uint32_t lo = k_ninf, hi = k_pinf + 1;  // keys of -inf, one past +inf
    while (lo < hi) {
      const uint32_t mid = lo + ((hi - lo) >> 1);
      const uint32_t b = mid ^ ((mid & 0x80000000u) ? 0x80000000u : 0xFFFFFFFFu);
      if (activate(__uint_as_float(b), a.act) >= thr) hi = mid;
      else lo = mid + 1;
    }
    sm.xthr[threadIdx.x] =
        lo > k_pinf ? __uint_as_float(0x7FC00000u)
                    : __uint_as_float(lo ^ ((lo & 0x80000000u) ? 0x80000000u
                                                               : 0xFFFFFFFFu));
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float x_lo = 0.f, x_hi = 0.f;
    bool any = false;
    for (int k = 0; k < B; ++k) {
      const float t = sm.xthr[k];
      if (t == t && fabsf(t) <= 3.0e38f) {  // finite (ascending with k)
        if (!any) x_lo = t;
        x_hi = t;
        any = true;
      }
    }
    // cells 1 .. kXCells-2 span [x_lo, x_hi]; cell 0 is everything below
    const float span = x_hi - x_lo;
    // any finite positive scale keeps x -> cell monotone (0 or inf would
    // turn +-inf / 0 into NaN products); exactness never depends on it
    const float scale =
        (any && span > 0.f && span < 3.0e38f)
            ? fminf(fmaxf((float)(kXCells - 3) / span, 1e-30f), 1e30f)
            : 1e-30f;
    sm.map[0] = scale;
    sm.map[1] = 1.5f - x_lo * scale;  // x_lo lands mid-cell 1, x_hi in the last but one
  }
  __syncthreads();
  const float scale = sm.map[0], bias = sm.map[1];
  // cell table: every threshold drops itself into its cell, then a block scan
  // over the cells turns the per-cell counts into "thresholds in lower cells"
  for (int i = threadIdx.x; i < kXCells; i += blockDim.x)
    sm.cell[i] = make_float2(__int_as_float(0), __uint_as_float(0x7FC00000u));
  __syncthreads();
  if ((int)threadIdx.x < B) {
    const float t = sm.xthr[threadIdx.x];
    if (t == t) {
      const int ck = x_cell(t, scale, bias);
      atomicAdd(reinterpret_cast<int *>(&sm.cell[ck].x), 1);
      sm.cell[ck].y = t;  // only read when it is the cell's single threshold
    }
  }
  __syncthreads();
  {
    constexpr int PER = kXCells / 256;  // cells per thread, blockDim.x == 256
    const int c0 = threadIdx.x * PER;
    int n[PER], sum = 0;
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      n[j] = __float_as_int(sm.cell[c0 + j].x);
      sum += n[j];
    }
    int inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, inc, o);
      if ((threadIdx.x & 31) >= o) inc += v;
    }
    __shared__ int wtot[8];
    if ((threadIdx.x & 31) == 31) wtot[threadIdx.x >> 5] = inc;
    __syncthreads();
    int before = inc - sum;
    for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) before += wtot[w];
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      sm.cell[c0 + j].x = __int_as_float(n[j] > 1 ? kXEscape : before);
      before += n[j];
    }
  }
  __syncthreads();
  const uint32_t lane4 = (threadIdx.x & 31) * 4;
  const uint32_t cnt_sa = (uint32_t)__cvta_generic_to_shared(sm.cnt);
  const uint32_t cell_sa = (uint32_t)__cvta_generic_to_shared(sm.cell);
  const long long n_units = a.n_planes * a.units_per_plane;
  const T *src = static_cast<const T *>(a.src.ptr[0]);
  auto exact = [&](float x) {  // escape cells, head / tail pixels
    int c = 0;
    while (c < B && x >= sm.xthr[c]) ++c;
    return c;
  };
  auto add = [&](int c, uint32_t gb) {
    red_shared_add(cnt_sa + ((c << 7) | lane4), gb ? 0x10000u : 1u);
  };

  for (long long unit = blockIdx.x; unit < n_units; unit += gridDim.x) {
    const long long plane = unit / a.units_per_plane;
    const int sub = (int)(unit - plane * a.units_per_plane);
    const size_t plane_off = (size_t)plane * a.plane_px;
    const T *lg = src + plane_off;
    const uint8_t *gtp = a.gt + plane_off;
    // 8-pixel items aligned to 8 elements of the TENSOR (16 / 32-byte loads);
    // up to 7 head and 7 tail pixels of the plane go one by one
    const int shift = (int)(plane_off & 7);
    const int head = shift ? 8 - shift : 0;          // pixels before item 0
    const long long n_it = (a.plane_px - min((long long)head, a.plane_px)) >> 3;
    const long long per = (n_it + a.units_per_plane - 1) / a.units_per_plane;
    const long long i0 = sub * per, i1 = min(n_it, i0 + per);
    constexpr int UN = sizeof(T) == 4 ? 2 : 4;  // items in flight per thread
    for (long long base = i0; base < i1; base += (long long)UN * blockDim.x) {
      uint4 q[UN][sizeof(T) == 4 ? 2 : 1];
      uint2 g[UN];
#pragma unroll
      for (int u = 0; u < UN; ++u) {
        const long long it = base + (long long)u * blockDim.x + threadIdx.x;
        if (it < i1) {
          const long long e0 = head + it * 8;
          q[u][0] = ldg_stream16(lg + e0);
          if (sizeof(T) == 4) q[u][sizeof(T) == 4 ? 1 : 0] = ldg_stream16(lg + e0 + 4);
          g[u] = ldg_stream8(gtp + e0);
        }
      }
#pragma unroll
      for (int u = 0; u < UN; ++u) {
        const long long it = base + (long long)u * blockDim.x + threadIdx.x;
        if (it >= i1) break;
        float x[8];
        if (sizeof(T) == 4) {
          const uint4 q0 = q[u][0], q1 = q[u][sizeof(T) == 4 ? 1 : 0];
          x[0] = __uint_as_float(q0.x); x[1] = __uint_as_float(q0.y);
          x[2] = __uint_as_float(q0.z); x[3] = __uint_as_float(q0.w);
          x[4] = __uint_as_float(q1.x); x[5] = __uint_as_float(q1.y);
          x[6] = __uint_as_float(q1.z); x[7] = __uint_as_float(q1.w);
        } else {
          const uint32_t w[4] = {q[u][0].x, q[u][0].y, q[u][0].z, q[u][0].w};
          unpack8(w, x);
        }
        int c[8];
        bool esc = false;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float2 e = lds_f2(cell_sa + x_cell(x[i], scale, bias) * 8);
          const int below = __float_as_int(e.x);
          esc |= below == kXEscape;
          c[i] = below + (x[i] >= e.y ? 1 : 0);
        }
        if (esc) {
#pragma unroll
          for (int i = 0; i < 8; ++i) c[i] = exact(x[i]);
        }
#pragma unroll
        for (int i = 0; i < 8; ++i)
          add(c[i], ((i < 4 ? g[u].x : g[u].y) >> (8 * (i & 3))) & 0xFFu);
      }
    }
    if (sub == 0) {  // head and tail pixels of the plane
      const long long tail0 = head + n_it * 8;
      if (threadIdx.x < 16) {  // at most 7 + 7 pixels
        const long long e = (long long)threadIdx.x < head
                                ? threadIdx.x
                                : tail0 + (threadIdx.x - head);
        if (e < a.plane_px) add(exact(to_f32<T>(lg[e])), gtp[e]);
      }
    }
    uint32_t *out = a.hist + (size_t)plane * 2 * nbins;
    flush_bins(sm.cnt, nbins, out, a.units_per_plane > 1);
  }
}

template <typename T, bool ROWS>
__global__ void __launch_bounds__(256)
seg_prob_vec_kernel(const __grid_constant__ ProbArgs a) {
  const long long ipp = a.plane_px >> 3;  // items per plane
  const long long total = a.n_planes * ipp;
  const int ipr = a.W >> 3;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
       g < total; g += (long long)gridDim.x * blockDim.x) {
    float p[8];
    if (ROWS) {
      const long long plane = g / ipp;
      const int it = (int)(g - plane * ipp);
      const int y = it / ipr;
      fused_rows8<T>(a.src, (size_t)plane * a.plane_px, a.H, ipr, y,
                     it - y * ipr, a.act, p);
    } else {
      fused_flat8<T>(a.src, (size_t)g * 8, a.act, p);
    }
    mean8(p, a.src.n);
    float4 *o = reinterpret_cast<float4 *>(a.prob + (size_t)g * 8);
    o[0] = make_float4(p[0], p[1], p[2], p[3]);
    o[1] = make_float4(p[4], p[5], p[6], p[7]);
  }
}

template <typename T, bool SM>
__global__ void __launch_bounds__(256)
seg_prob_kernel(const __grid_constant__ ProbArgs a) {
  const long long total = a.n_planes * a.items_per_plane;
  for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
       g < total; g += (long long)gridDim.x * blockDim.x) {
    const long long plane = g / a.items_per_plane;
    const long long it = g - plane * a.items_per_plane;
    const int y = (int)(it / a.items_per_row);
    const int x0 = (int)(it - (long long)y * a.items_per_row) * 8;
    const int nv = min(8, a.W - x0);
    float p[8];
    const size_t plane_off = (size_t)plane * a.plane_px;
    const bool inner = plane > 0 && plane + 1 < a.n_planes;
    const PlaneCtx cx = {plane, a.n_planes, a.plane_px, a.n_cls};
    fused_prob8<T, SM>(a.src, plane_off, cx, a.H, a.W, y, x0, nv, inner, a.act,
                       p);
    float *o = a.prob + plane_off + (size_t)y * a.W + x0;
    for (int i = 0; i < nv; ++i) o[i] = p[i];
  }
}

// Developer builds (-DMLC_DEV_KNOBS, `python -m mlcomp_b200.build
// --dev-knobs`) keep the retired schedules of the single-source kernel and an
// environment switch for A/B runs; the shipped library reads no environment.
//   MLC_HIST_VARIANT=1 byte-table kernel (seg_hist_lut), =3 cell table fed by
//   bulk async copies (seg_hist_cell_tma), =4 multi-source fp16 jobs evaluate
//   the sigmoid (seg_hist_generic), =5 no logit-domain kernel, =6 no guarded
//   estimate (seg_hist_tab serves the multi-source fp16 jobs)
static int hist_variant() {
#ifdef MLC_DEV_KNOBS
  const char *e = getenv("MLC_HIST_VARIANT");
  return e ? atoi(e) : 0;
#else
  return 0;
#endif
}

static int check_thr(const float *thr, int n_thr) {
  MLC_REQUIRE(thr != nullptr, "thr_host is null");
  MLC_REQUIRE(n_thr >= 1 && n_thr <= MLC_MAX_THR,
              "n_thr=%d outside [1, %d]", n_thr, MLC_MAX_THR);
  for (int k = 0; k < n_thr; ++k) {
    MLC_REQUIRE(thr[k] == thr[k], "thr[%d] is NaN", k);
    MLC_REQUIRE(k == 0 || thr[k] >= thr[k - 1],
                "thresholds must be ascending (thr[%d] < thr[%d])", k, k - 1);
  }
  return MLC_OK;
}

}  // namespace mlc

using namespace mlc;

extern "C" int mlc_seg_lut_build(const float *thr_host, int n_thr, int act,
                                 uint8_t *lut_dev, void *stream) {
  if (int rc = check_thr(thr_host, n_thr)) return rc;
  MLC_REQUIRE(lut_dev != nullptr, "lut_dev is null");
  MLC_REQUIRE(act == MLC_ACT_NONE || act == MLC_ACT_SIGMOID,
              "unknown activation code %d", act);
  ThrTable t;
  for (int k = 0; k < n_thr; ++k) t.v[k] = thr_host[k];
  lut_build<<<65536 / 256, 256, 0, (cudaStream_t)stream>>>(t, n_thr, act,
                                                           lut_dev);
  MLC_LAUNCH_CHECK("lut_build");
  cell_build<<<kCells / 256, 256, 0, (cudaStream_t)stream>>>(lut_dev);
  MLC_LAUNCH_CHECK("cell_build");
  return MLC_OK;
}

static int seg_hist_impl(const void *const *src_dev_ptrs, const int32_t *xform,
                         int n_src, int dtype, const uint8_t *gt_dev,
                         int64_t n_img, int n_cls, int H, int W,
                         const float *thr_host, int n_thr, int act,
                         const uint8_t *lut_dev, uint32_t *hist_dev,
                         uint8_t *bins_dev, void *stream) {
  if (int rc = check_thr(thr_host, n_thr)) return rc;
  MLC_REQUIRE(dtype == MLC_F16 || dtype == MLC_F32,
              "dtype %d is not MLC_F16/MLC_F32", dtype);
  MLC_REQUIRE(act == MLC_ACT_NONE || act == MLC_ACT_SIGMOID ||
                  act == MLC_ACT_SOFTMAX,
              "unknown activation code %d", act);
  MLC_REQUIRE(gt_dev && hist_dev, "gt_dev / hist_dev is null");
  MLC_REQUIRE(n_img >= 0 && n_cls >= 1 && H >= 1 && W >= 1,
              "bad shape n_img=%lld n_cls=%d H=%d W=%d", (long long)n_img,
              n_cls, H, W);
  MLC_REQUIRE((long long)H * W < (1ll << 31), "plane of %d x %d too large", H,
              W);
  HistArgs a;
  if (int rc = fill_srcset(a.src, src_dev_ptrs, xform, n_src,
                           dtype == MLC_F16 ? 2 : 4, H, W))
    return rc;
  if (n_img == 0) return MLC_OK;
  cudaStream_t st = (cudaStream_t)stream;
  for (int k = 0; k < n_thr; ++k) a.thr.v[k] = thr_host[k];
  a.gt = gt_dev;
  a.hist = hist_dev;
  a.lut = lut_dev;
  a.bins = bins_dev;
  a.n_planes = (long long)n_img * n_cls;
  a.plane_px = (long long)H * W;
  a.n_cls = n_cls;
  a.B = n_thr;
  a.act = act;
  a.flat = a.src.any_xf == 0;
  a.H = a.flat ? 1 : H;
  a.W = a.flat ? (int)a.plane_px : W;
  a.items_per_row = a.W / 8;
  a.full_items = (long long)a.items_per_row * H;
  // linear first guess for count_ge: k ~ (p - thr[0]) * (B-1)/(thr[B-1]-thr[0])
  const float span = thr_host[n_thr - 1] - thr_host[0];
  a.scale = (n_thr > 1 && span > 0) ? (float)(n_thr - 1) / span : 0.f;
  a.bias = 1.0f - thr_host[0] * a.scale;

#ifdef MLC_BOUNDS_CHECK
  {
    BoundsList bl;
    const size_t eb = dtype == MLC_F16 ? 2 : 4;
    for (int k = 0; k < n_src; ++k)
      bl.add(src_dev_ptrs[k], (size_t)a.n_planes * a.plane_px * eb);
    bl.add(gt_dev, (size_t)a.n_planes * a.plane_px);
    bl.add(lut_dev, MLC_LUT_BYTES);
    MLC_CUDA(bounds_commit(bl, st));
  }
#endif
  const int sms = sm_count();
  // (the per-code table kernels do not write bins: one source is cheap to
  // read again)
  const bool table = lut_dev && !bins_dev && act != MLC_ACT_SOFTMAX &&
                     n_src == 1 &&
                     dtype == MLC_F16 && a.flat &&
                     (a.plane_px % 8) == 0 &&
                     ((uintptr_t)src_dev_ptrs[0] % 16) == 0 &&
                     ((uintptr_t)gt_dev % 8) == 0;
  // flat mode may prepend one partial item to reach 16-byte alignment
  a.items_per_plane = a.flat ? (a.plane_px + 7 + 7) / 8
                             : a.full_items + ((a.W & 7) ? H : 0);
  if (table) a.items_per_plane = a.plane_px / 8;
  // split planes into units so that (a) a lane column stays below 65 536 and
  // (b) small jobs still fill the machine
  long long units = (a.plane_px + kMaxUnitPx - 1) / kMaxUnitPx;
  const long long want = 4ll * sms;
  if (a.n_planes * units < want) {
    const long long more = (want + a.n_planes - 1) / a.n_planes;
    const long long cap = (a.items_per_plane + 4095) / 4096;  // >= 32 Ki px
    units = max(units, min(more, max(1ll, cap)));
  }
  a.units_per_plane = (int)units;
  a.items_per_unit = (a.items_per_plane + units - 1) / units;
  if (table) {  // keep unit boundaries on whole unrolled CTA sweeps
    const long long q = (long long)kLutUnroll * kHistThreads;
    a.items_per_unit = (a.items_per_unit + q - 1) / q * q;
  }
  const size_t hist_bytes = (size_t)a.n_planes * 2 * (n_thr + 1) * 4;
  if (units > 1) MLC_CUDA(cudaMemsetAsync(hist_dev, 0, hist_bytes, st));
  const long long n_units = a.n_planes * units;

  // one untransformed source without the per-code table (fp32 logits, fp16
  // without activation or without a lut): binned in the logit domain
  const bool logit = !table && !bins_dev && n_src == 1 && a.flat &&
                     act != MLC_ACT_SOFTMAX && hist_variant() != 5 &&
                     ((uintptr_t)src_dev_ptrs[0] % 16) == 0 &&
                     ((uintptr_t)gt_dev % 8) == 0;
  if (logit) {
    void (*lk)(HistArgs) = dtype == MLC_F16 ? seg_hist_logit<__half>
                                            : seg_hist_logit<float>;
    MLC_CUDA(cudaFuncSetAttribute(
        lk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(XSmem)));
    // persistent, 6 CTAs per SM (two waves of the 3 resident ones): the
    // per-CTA bisection makes one CTA per unit 4 % slower here
    // (fp32 C2 1.79 vs 1.73 ms); MLC_LOGIT_GRID overrides, 0 = per unit
    int lgs = 6;
#ifdef MLC_DEV_KNOBS
    if (const char *e_lg = getenv("MLC_LOGIT_GRID")) lgs = atoi(e_lg);
#endif
    const long long lgrid = lgs > 0 ? min(n_units, (long long)lgs * sms)
                                    : min(n_units, (long long)1 << 22);
    lk<<<(int)lgrid, 256, sizeof(XSmem), st>>>(a);
    MLC_LAUNCH_CHECK("seg_hist_logit");
#ifdef MLC_DEV_KNOBS
  } else if (table && hist_variant() == 3 && (a.plane_px % 16) == 0) {
    // unit boundaries on whole 512-item chunks (already guaranteed: the
    // table path rounds items_per_unit to a multiple of 2048)
    MLC_CUDA(cudaFuncSetAttribute(seg_hist_cell_tma,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)sizeof(TmaStages) + 128));
    const int grid = (int)min(n_units, 2ll * sms);
    seg_hist_cell_tma<<<grid, kTmaThreads, sizeof(TmaStages) + 128, st>>>(a);
    MLC_LAUNCH_CHECK("seg_hist_cell_tma");
#endif
  } else if (table && hist_variant() != 1) {  // always, unless a dev knob
    // One CTA per unit, not a persistent grid: measured on C2 (26 664
    // planes), 4 resident CTAs per SM looping over a static share of the
    // planes reach 86 % of the measured HBM peak, the same CTAs handed out by
    // the hardware scheduler as SMs free up reach 96 % (no tail, no lock-step
    // between co-resident CTAs); the per-CTA set-up is 4 KiB of table and
    // 13 KiB of zeroed counters against 1.2 MB streamed.
    seg_hist_cell<256, 4, 4>
        <<<(int)min(n_units, (long long)1 << 22), 256, 0, st>>>(a);
    MLC_LAUNCH_CHECK("seg_hist_cell");
#ifdef MLC_DEV_KNOBS
  } else if (table) {
    const size_t smem = 65536 + (size_t)(n_thr + 1) * 32 * 4;
    // per-device attribute, cheap: set on every call rather than cache
    MLC_CUDA(cudaFuncSetAttribute(seg_hist_lut,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  65536 + (MLC_MAX_THR + 1) * 128));
    const int grid = (int)min(n_units, 2ll * sms);
    seg_hist_lut<<<grid, kHistThreads, smem, st>>>(a);
    MLC_LAUNCH_CHECK("seg_hist_lut");
#endif
  } else {
    const size_t smem = (size_t)(n_thr + 1) * (32 * 4 + 8) + (size_t)n_thr * 4;
    // one CTA per unit (see seg_hist_cell above)
    const int grid = (int)min(n_units, (long long)1 << 22);
    bool vec = a.flat && ((uintptr_t)gt_dev % 16) == 0;
    for (int i = 0; i < n_src; ++i)
      vec = vec && ((uintptr_t)src_dev_ptrs[i] % 16) == 0;
    // flipped rows of whole aligned items (e.g. 256 x 1600 TTA)
    bool vrows = !a.flat && !(a.src.any_xf & MLC_XF_TRANSPOSE) &&
                 (W % 8) == 0 && ((uintptr_t)gt_dev % 8) == 0 &&
                 a.items_per_plane < (1ll << 31) - kTabThreads;
    for (int i = 0; i < n_src; ++i)
      vrows = vrows && ((uintptr_t)src_dev_ptrs[i] % 16) == 0;
    if (dtype == MLC_F16 && act == MLC_ACT_SIGMOID && (vec || vrows) &&
        (a.plane_px % 8) == 0 && hist_variant() != 4 && hist_variant() != 6) {
      // guarded estimate (seg_hist_guarded.cu); declines irregular grids
      const long long ipp = a.items_per_plane, ipu = a.items_per_unit;
      a.items_per_plane = a.plane_px / 8;
      a.items_per_unit = (a.items_per_plane + units - 1) / units;
      int rc = MLC_OK;
      if (seg_hist_guarded_launch(a, !vec, st, &rc)) return rc;
      a.items_per_plane = ipp;
      a.items_per_unit = ipu;
    }
    if (dtype == MLC_F16 && act == MLC_ACT_SIGMOID &&
        ((!vec && !vrows) || (a.plane_px % 8) != 0) &&
        hist_variant() != 4 && hist_variant() != 6) {
      // rows of any width / alignment: the same estimate, rows staged by
      // bulk copies (declines transposes, unaligned tensors, irregular grids)
      int rc = MLC_OK;
      HistArgs b = a;
      if (seg_hist_guarded_rows_launch(b, st, &rc)) return rc;
    }
    if (dtype == MLC_F16 && act == MLC_ACT_SIGMOID && hist_variant() != 4) {
      // activation table in shared memory: one 1024-thread CTA per SM
      const size_t tab_smem = kActTabBytes;
      static_assert((MLC_MAX_THR + 1) * (32 * 4 + 8) + MLC_MAX_THR * 4 <=
                        kActTabHoleBytes,
                    "counters do not fit the hole of the activation table");
      // persistent here: every CTA builds the 152 KiB table first, and the
      // kernel is LSU-bound, not balance-bound (grids of 2x / 4x / one CTA
      // per unit measured within 1 %)
      const int tgrid = (int)min(n_units, (long long)sms);
      void (*tk)(HistArgs) =
          bins_dev ? (vec ? seg_hist_tab<1, true>
                      : vrows ? seg_hist_tab<2, true> : seg_hist_tab<0, true>)
                   : (vec ? seg_hist_tab<1, false>
                      : vrows ? seg_hist_tab<2, false>
                              : seg_hist_tab<0, false>);
      MLC_CUDA(cudaFuncSetAttribute(
          tk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tab_smem));
      tk<<<tgrid, (vec || vrows) ? kTabThreads : kTabRowThreads, tab_smem, st>>>(a);
    } else {
      const bool sm = act == MLC_ACT_SOFTMAX;
      void (*gk)(HistArgs);
#define MLC_GK(T, B)                                              \
  (sm ? seg_hist_generic<T, 0, true, B>                           \
   : vec ? seg_hist_generic<T, 1, false, B>                       \
   : vrows ? seg_hist_generic<T, 2, false, B>                     \
           : seg_hist_generic<T, 0, false, B>)
      if (dtype == MLC_F16)
        gk = bins_dev ? MLC_GK(__half, true) : MLC_GK(__half, false);
      else
        gk = bins_dev ? MLC_GK(float, true) : MLC_GK(float, false);
#undef MLC_GK
      gk<<<grid, kHistThreads, smem, st>>>(a);
    }
    MLC_LAUNCH_CHECK("seg_hist_generic");
  }
  return MLC_OK;
}

extern "C" int mlc_seg_hist(const void *const *src_dev_ptrs,
                            const int32_t *xform, int n_src, int dtype,
                            const uint8_t *gt_dev, int64_t n_img, int n_cls,
                            int H, int W, const float *thr_host, int n_thr,
                            int act, const uint8_t *lut_dev,
                            uint32_t *hist_dev, void *stream) {
  return seg_hist_impl(src_dev_ptrs, xform, n_src, dtype, gt_dev, n_img, n_cls,
                       H, W, thr_host, n_thr, act, lut_dev, hist_dev, nullptr,
                       stream);
}

extern "C" int mlc_seg_hist_bins(const void *const *src_dev_ptrs,
                                 const int32_t *xform, int n_src, int dtype,
                                 const uint8_t *gt_dev, int64_t n_img,
                                 int n_cls, int H, int W,
                                 const float *thr_host, int n_thr, int act,
                                 uint32_t *hist_dev, uint8_t *bins_dev,
                                 void *stream) {
  MLC_REQUIRE(bins_dev, "bins_dev is null");
  return seg_hist_impl(src_dev_ptrs, xform, n_src, dtype, gt_dev, n_img, n_cls,
                       H, W, thr_host, n_thr, act, nullptr, hist_dev, bins_dev,
                       stream);
}

extern "C" int mlc_seg_prob(const void *const *src_dev_ptrs,
                            const int32_t *xform, int n_src, int dtype,
                            int64_t n_img, int n_cls, int H, int W, int act,
                            float *prob_dev, void *stream) {
  MLC_REQUIRE(dtype == MLC_F16 || dtype == MLC_F32,
              "dtype %d is not MLC_F16/MLC_F32", dtype);
  MLC_REQUIRE(act == MLC_ACT_NONE || act == MLC_ACT_SIGMOID ||
                  act == MLC_ACT_SOFTMAX,
              "unknown activation code %d", act);
  MLC_REQUIRE(prob_dev, "prob_dev is null");
  MLC_REQUIRE(n_img >= 0 && n_cls >= 1 && H >= 1 && W >= 1,
              "bad shape n_img=%lld n_cls=%d H=%d W=%d", (long long)n_img,
              n_cls, H, W);
  ProbArgs a;
  if (int rc = fill_srcset(a.src, src_dev_ptrs, xform, n_src,
                           dtype == MLC_F16 ? 2 : 4, H, W))
    return rc;
  if (n_img == 0) return MLC_OK;
  a.prob = prob_dev;
  a.n_planes = (long long)n_img * n_cls;
  a.plane_px = (long long)H * W;
  a.n_cls = n_cls;
  a.H = H;
  a.W = W;
  a.act = act;
  a.items_per_row = (W + 7) / 8;
  a.items_per_plane = (long long)a.items_per_row * H;
  const long long total = a.n_planes * a.items_per_plane;
  const int grid = (int)min((total + 255) / 256, 32ll * sm_count());
  cudaStream_t st = (cudaStream_t)stream;
#ifdef MLC_BOUNDS_CHECK
  {
    BoundsList bl;
    const size_t eb = dtype == MLC_F16 ? 2 : 4;
    for (int k = 0; k < n_src; ++k)
      bl.add(src_dev_ptrs[k], (size_t)a.n_planes * a.plane_px * eb);
    MLC_CUDA(bounds_commit(bl, st));
  }
#endif
  bool aligned = ((long long)H * W) % 8 == 0 && ((uintptr_t)prob_dev % 16) == 0;
  for (int k = 0; k < n_src; ++k)
    aligned = aligned && ((uintptr_t)src_dev_ptrs[k] % 16) == 0;
  const bool flat = aligned && a.src.any_xf == 0;
  const bool rows = aligned && !flat && (W % 8) == 0 &&
                    !(a.src.any_xf & MLC_XF_TRANSPOSE);
  if (act != MLC_ACT_SOFTMAX && (flat || rows)) {
    const long long items = a.n_planes * (a.plane_px / 8);
    const int vgrid = (int)min((items + 255) / 256, 32ll * sm_count());
    void (*vk)(ProbArgs) =
        dtype == MLC_F16 ? (rows ? seg_prob_vec_kernel<__half, true>
                                 : seg_prob_vec_kernel<__half, false>)
                         : (rows ? seg_prob_vec_kernel<float, true>
                                 : seg_prob_vec_kernel<float, false>);
    vk<<<vgrid, 256, 0, st>>>(a);
  } else if (act == MLC_ACT_SOFTMAX) {
    if (dtype == MLC_F16) seg_prob_kernel<__half, true><<<grid, 256, 0, st>>>(a);
    else seg_prob_kernel<float, true><<<grid, 256, 0, st>>>(a);
  } else {
    if (dtype == MLC_F16) seg_prob_kernel<__half, false><<<grid, 256, 0, st>>>(a);
    else seg_prob_kernel<float, false><<<grid, 256, 0, st>>>(a);
  }
  MLC_LAUNCH_CHECK("seg_prob_kernel");
  return MLC_OK;
}
