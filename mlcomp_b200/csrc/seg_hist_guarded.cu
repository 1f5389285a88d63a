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
#include "seg_hist_guarded.cuh"

namespace mlc {

// ---- the kernel ------------------------------------------------------------
// NS: number of sources when it is one of the instantiated counts (the source
// loop, the group structure and every per-source constant are then resolved
// at compile time: ncu showed 29 of 61 instructions per pixel of the first,
// run-time version were loop / flag / address bookkeeping and spills), 0 =
// any count (run-time loop over groups of four).
// ROWS: flipped sources, rows a whole number of aligned 16-byte items (flips
// are item-index arithmetic); else no transforms, planes are flat arrays.
// BINS: also write every pixel's bin index (mlc_seg_hist_bins).
//
// Per source k the host precomputes (HistArgs::g_*) the item index of output
// item (y, xi) in the source plane as g_b[k] + g_cy[k] * y + g_cx[k] * xi
// (flat mode: 0, 0, 1 with y = 0, xi = item) and whether its eight elements
// come reversed (horizontal flip).
template <int NG>
__device__ __forceinline__ void load_group(const HistArgs &a, int k0,
                                           const __half *const *base, int y,
                                           int xi, uint32_t (&w)[4][4]) {
#pragma unroll
  for (int j = 0; j < NG; ++j) {
    const int idx = a.g_b[k0 + j] + a.g_cy[k0 + j] * y + a.g_cx[k0 + j] * xi;
    const uint4 r = ldg_stream16(base[j] + (size_t)(uint32_t)idx * 8);
    w[j][0] = r.x; w[j][1] = r.y; w[j][2] = r.z; w[j][3] = r.w;
  }
}

// One group of NG sources: reversal masks that occur in practice (none; the
// (id, h, v, hv) order of a 4-flip TTA and its mirror; all) are resolved at
// compile time, any other mask moves the data first.
template <int NG, bool ROWS>
__device__ __forceinline__ void do_group(const HistArgs &a, int k0, int rmask,
                                         uint32_t (&w)[4][4], bool first,
                                         float (&q)[8], uint32_t &dmax) {
  constexpr int kAll = (1 << NG) - 1;
  if (!ROWS || rmask == 0) {
    group_sum<NG, 0>(w, first, q, dmax);
  } else if (NG >= 2 && rmask == (0xA & kAll)) {
    group_sum<NG, 0xA & kAll>(w, first, q, dmax);
  } else if (NG >= 2 && rmask == (0x5 & kAll)) {
    group_sum<NG, 0x5 & kAll>(w, first, q, dmax);
  } else if (rmask == kAll) {
    group_sum<NG, kAll>(w, first, q, dmax);
  } else {
#pragma unroll
    for (int j = 0; j < NG; ++j)
      if ((rmask >> j) & 1) reverse8(w[j]);
    group_sum<NG, 0>(w, first, q, dmax);
  }
}

template <int NS, bool ROWS, bool BINS>
__global__ void __launch_bounds__(kGThreads, MLC_G_MINB)
seg_hist_guarded(const __grid_constant__ HistArgs a) {
  extern __shared__ __align__(16) uint32_t gsm[];
  const int B = a.B, nbins = B + 1;
  // the counters [nbins][32] start on a 16 KiB boundary of the shared window:
  // a counter's address is then (bin << 7) | lane * 4 | base, one LOP3
  const uint32_t gsm_sa = (uint32_t)__cvta_generic_to_shared(gsm);
  const uint32_t cnt_sa = (gsm_sa + 16383u) & ~16383u;
  uint32_t *cnt = gsm + (cnt_sa - gsm_sa) / 4;
  float *thr = reinterpret_cast<float *>(cnt + nbins * 32);  // [B]
  for (int i = threadIdx.x; i < nbins * 32; i += blockDim.x) cnt[i] = 0;
  for (int i = threadIdx.x; i < B; i += blockDim.x) thr[i] = a.thr.v[i];
  __syncthreads();
  const uint32_t slot = cnt_sa | ((threadIdx.x & 31) * 4);
  const long long n_units = a.n_planes * a.units_per_plane;
  const int S = NS ? NS : a.src.n;
  // position on the threshold grid, shifted into [128, 256) (16 fractional
  // bits), once moved down and once up by the error bound: a pixel is
  // decided when both land in the same bin
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
    // this plane of the first sources (the others are added per item)
    constexpr int kHoist = NS > 0 && NS <= 4 ? NS : 1;
    const __half *base[4];
#pragma unroll
    for (int j = 0; j < kHoist; ++j)
      base[j] = static_cast<const __half *>(a.src.ptr[j]) + plane_off;
    for (; it < i1; it += blockDim.x) {
      float q[8];
      uint32_t dmax = 0u;
      uint32_t w[4][4];
      if (NS == 0) {
        for (int k0 = 0; k0 < S; k0 += 4) {
          const int ng = min(4, S - k0);
          const int rm = (a.g_rev >> k0) & 0xF;
          for (int j = 0; j < ng; ++j)
            base[j] = static_cast<const __half *>(a.src.ptr[k0 + j]) +
                      plane_off;
          if (ng == 4) {
            load_group<4>(a, k0, base, y, xi, w);
            do_group<4, ROWS>(a, k0, rm, w, k0 == 0, q, dmax);
          } else if (ng == 3) {
            load_group<3>(a, k0, base, y, xi, w);
            do_group<3, ROWS>(a, k0, rm & 7, w, k0 == 0, q, dmax);
          } else if (ng == 2) {
            load_group<2>(a, k0, base, y, xi, w);
            do_group<2, ROWS>(a, k0, rm & 3, w, k0 == 0, q, dmax);
          } else {
            load_group<1>(a, k0, base, y, xi, w);
            do_group<1, ROWS>(a, k0, rm & 1, w, k0 == 0, q, dmax);
          }
        }
      } else {
#pragma unroll
        for (int k0 = 0; k0 < NS; k0 += 4) {
          constexpr int kFull = NS / 4 * 4;
          if (NS > 4) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              if (k0 + j < NS)
                base[j] = static_cast<const __half *>(a.src.ptr[k0 + j]) +
                          plane_off;
          }
          if (k0 < kFull) {
            load_group<4>(a, k0, base, y, xi, w);
            do_group<4, ROWS>(a, k0, (a.g_rev >> k0) & 0xF, w, k0 == 0, q,
                              dmax);
          } else {
            constexpr int kRest = NS % 4 ? NS % 4 : 1;
            load_group<kRest>(a, k0, base, y, xi, w);
            do_group<kRest, ROWS>(a, k0,
                                  (a.g_rev >> k0) & ((1 << kRest) - 1), w,
                                  k0 == 0, q, dmax);
          }
        }
      }
      finish_item<ROWS, BINS ? 1 : 0>(a, q, dmax, ldg_stream8(gtp + (size_t)it * 8),
                              slot, plane, it,
                              (size_t)plane * (size_t)(a.plane_px >> 3) + it);
      xi += dx;
      if (ROWS) {
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

// ---- the same kernel fed through cp.async ----------------------------------
// ncu on seg_hist_guarded (4-flip TTA): 46 % of the cycles between two issues
// of a warp are spent waiting for its four 128-bit loads (8 warps per
// scheduler, all at the same point of the loop), issue slots 67 % busy.  Here
// every thread copies the NEXT item's 16 bytes per source + 8 truth bytes into
// its own shared-memory slots (cp.async: no registers held while the data is
// in flight) before it works on the current one, then reads them back with
// LDS - a thread only ever touches its own slots, so cp.async.wait_group is
// all the synchronisation needed.
__device__ __forceinline__ void cp_async16(uint32_t dst, const void *src) {
  MLC_CHECK_LOAD(src, 16);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src)
               : "memory");
}
__device__ __forceinline__ void cp_async8(uint32_t dst, const void *src) {
  MLC_CHECK_LOAD(src, 8);
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src)
               : "memory");
}
// next unassigned item of the CTA's unit (see pf_unit)
__shared__ unsigned int s_next_item;

// One unit (a plane or a slice of it) of the cp.async-fed kernel.
// REV >= 0: the reversal mask of the NS sources, known at compile time (the
// flipped sources' pixels are picked by register naming); REV < 0: any mask,
// flipped items are reversed with four PRMT each.
// NT: threads per CTA (compile time: every shared-memory slot of a thread is
// then a constant offset from its first one).
template <int NS, int NT, bool ROWS, bool BINS, int REV>
__device__ __forceinline__ void pf_unit(const HistArgs &a, uint32_t plane,
                                        uint32_t sub, uint32_t slot,
                                        uint32_t my_slot) {
  // a thread's slots of one stage: NS x 16 B of logits at a stride of NT x 16,
  // then 16 B holding the 8 truth bytes; the other stage kStage further on
  constexpr uint32_t kStage = (NS + 1) * NT * 16u;
  const int ipr = a.items_per_row;
  // 32-bit item index from the start of the tensors (the host checked that
  // all items fit): one IMAD.WIDE per source on top of the kernel argument
  const uint32_t ipp = (uint32_t)a.items_per_plane;
  const uint32_t p0 = plane * ipp;
  const int i0 = (int)(sub * (uint32_t)a.items_per_unit);
  const int i1 = (int)min(ipp, (uint32_t)i0 + (uint32_t)a.items_per_unit);
  // Items are handed out to the warps 32 at a time from a counter in shared
  // memory (ncu on the statically split version: 14 % of the warp samples sat
  // at the barrier before the flush - the scheduler lets some warps run ahead
  // and they then idle while the rest finish): a warp fetches its next 32
  // items' data while it works on the current ones, and all warps finish
  // within one round of each other.
  const int lane = threadIdx.x & 31;
  const uint32_t next_sa = (uint32_t)__cvta_generic_to_shared(&s_next_item);
  const int it = i0 + (int)threadIdx.x;  // the first round is static
  auto issue = [&](int it_, uint32_t dst) {
    int y_ = 0, x_ = it_;
    if (ROWS) {  // exact for it_ * ipr < 2^32 (checked by the host)
      y_ = (int)__umulhi((uint32_t)it_, a.g_ipr_magic);
      x_ = it_ - y_ * ipr;
    }
#pragma unroll
    for (int j = 0; j < NS; ++j) {
      const uint32_t idx =
          p0 + (uint32_t)(a.g_b[j] + a.g_cy[j] * y_ + a.g_cx[j] * x_);
      cp_async16(dst + j * NT * 16u,
                 static_cast<const uint4 *>(a.src.ptr[j]) + idx);
    }
    cp_async8(dst + NS * NT * 16u,
              reinterpret_cast<const uint2 *>(a.gt) + (p0 + (uint32_t)it_));
  };
  // The index of the item a stage holds travels with it (third word of the
  // truth slot): nothing but the stage parity is carried around the loop in
  // registers - the loop body needs all 64 of them, and a carried index was
  // spilled and waited for at the end of every round.
  constexpr uint32_t kIdxOff = NS * NT * 16u + 8u;
  bool odd = false;  // which of the two stages holds the current item
  st_shared_u32(my_slot + kIdxOff, (uint32_t)it);
  if (it < i1) issue(it, my_slot);
  asm volatile("cp.async.commit_group;" ::: "memory");
  while (true) {
    const uint32_t cur = odd ? my_slot + kStage : my_slot;
    const int it = (int)lds_u32(cur + kIdxOff);
    if (it - lane >= i1) break;  // the same for all lanes of the warp
    // One lane takes the warp's next 32 items.  The address is offset by a
    // term that is always zero but not provably so (it >= 0): for an atomic
    // on a provably uniform address ptxas builds its warp-aggregation
    // sequence (vote, find-leader, popc, shuffles: ~20 instructions, 11 % of
    // the kernel's stall samples) around this single lane.
    uint32_t nb = 0;
    if (lane == 0)
      asm volatile("atom.shared.add.u32 %0, [%1], 32;"
                   : "=r"(nb)
                   : "r"(next_sa + (((uint32_t)it >> 31) << 2))
                   : "memory");
    nb = __shfl_sync(0xffffffffu, nb, 0);
    {
      const int itn = (int)nb + lane;
      const uint32_t nxt = odd ? my_slot : my_slot + kStage;
      st_shared_u32(nxt + kIdxOff, (uint32_t)itn);
      if (itn < i1) issue(itn, nxt);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 1;" ::: "memory");
    odd = !odd;
    if (it < i1) {
    const uint32_t cur_src = cur;
    float q[8];
    uint32_t dmax = 0u;
    uint32_t w[4][4];
    // NS <= 8: a first group of up to four sources, then the rest
    constexpr int kA = NS < 4 ? NS : 4, kB = NS - kA;
#pragma unroll
    for (int j = 0; j < kA; ++j) {
      const uint4 r = lds_u4(cur_src + j * NT * 16u);
      w[j][0] = r.x; w[j][1] = r.y; w[j][2] = r.z; w[j][3] = r.w;
    }
    if constexpr (REV >= 0)
      group_sum<kA, REV & ((1 << kA) - 1)>(w, true, q, dmax);
    else
      do_group<kA, ROWS>(a, 0, a.g_rev & ((1 << kA) - 1), w, true, q, dmax);
    if constexpr (kB > 0) {
#pragma unroll
      for (int j = 0; j < kB; ++j) {
        const uint4 r = lds_u4(cur_src + (4 + j) * NT * 16u);
        w[j][0] = r.x; w[j][1] = r.y; w[j][2] = r.z; w[j][3] = r.w;
      }
      if constexpr (REV >= 0)
        group_sum<kB, (REV >> 4) & ((1 << kB) - 1)>(w, false, q, dmax);
      else
        do_group<kB, ROWS>(a, 4, (a.g_rev >> 4) & ((1 << kB) - 1), w, false, q,
                           dmax);
    }
    finish_item<ROWS, BINS ? 1 : 0>(a, q, dmax, lds_u2(cur + NS * NT * 16u), slot,
                            blockIdx.x, it, p0 + (uint32_t)it, cur, NT * 16u);
    }
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
}

template <int NS, int NT, bool ROWS, bool BINS>
__global__ void __launch_bounds__(NT, 1024 / NT)
seg_hist_guarded_pf(const __grid_constant__ HistArgs a) {
  extern __shared__ __align__(16) uint32_t gsm[];
  const int B = a.B, nbins = B + 1;
  const uint32_t gsm_sa = (uint32_t)__cvta_generic_to_shared(gsm);
  const uint32_t cnt_sa = (gsm_sa + 16383u) & ~16383u;
  uint32_t *cnt = gsm + (cnt_sa - gsm_sa) / 4;
  float *thr = reinterpret_cast<float *>(cnt + nbins * 32);  // [B]
  for (int i = threadIdx.x; i < nbins * 32; i += NT) cnt[i] = 0;
  for (int i = threadIdx.x; i < B; i += NT) thr[i] = a.thr.v[i];
  if (threadIdx.x == 0)  // the first NT items of the unit are taken statically
    s_next_item = blockIdx.y * (uint32_t)a.items_per_unit + NT;
  __syncthreads();
  const uint32_t slot = cnt_sa | ((threadIdx.x & 31) * 4);
  // two stages behind the threshold table, (NS + 1) x NT x 16 B each
  const uint32_t st_sa = (cnt_sa + nbins * 128 + B * 4 + 15u) & ~15u;
  const uint32_t my_slot = st_sa + threadIdx.x * 16u;
  constexpr int kAll = (1 << NS) - 1;
  const int rev = ROWS ? (int)(a.g_rev & kAll) : 0;
  // one CTA per unit, grid = (planes, slices of a plane): both indices are
  // special registers, so everything derived from them stays in the uniform
  // datapath (a division by units_per_plane would not)
  const uint32_t upp = gridDim.y;
  const uint32_t plane = blockIdx.x, sub = blockIdx.y;
  if (rev == 0)
    pf_unit<NS, NT, ROWS, BINS, 0>(a, plane, sub, slot, my_slot);
  else if (rev == (0xA & kAll))
    pf_unit<NS, NT, ROWS, BINS, 0xA & kAll>(a, plane, sub, slot, my_slot);
  else if (rev == (0x5 & kAll))
    pf_unit<NS, NT, ROWS, BINS, 0x5 & kAll>(a, plane, sub, slot, my_slot);
  else if (rev == kAll)
    pf_unit<NS, NT, ROWS, BINS, kAll>(a, plane, sub, slot, my_slot);
  else
    pf_unit<NS, NT, ROWS, BINS, -1>(a, plane, sub, slot, my_slot);
  uint32_t *out = a.hist + (size_t)plane * 2 * nbins;
  flush_bins(cnt, nbins, out, upp > 1);
}

template <int NS, int NT>
static void (*pick_guarded_pf(bool rows, bool bins))(HistArgs) {
  return bins ? (rows ? seg_hist_guarded_pf<NS, NT, true, true>
                      : seg_hist_guarded_pf<NS, NT, false, true>)
              : (rows ? seg_hist_guarded_pf<NS, NT, true, false>
                      : seg_hist_guarded_pf<NS, NT, false, false>);
}

template <int NS>
static void (*pick_guarded(bool rows, bool bins))(HistArgs) {
  return bins ? (rows ? seg_hist_guarded<NS, true, true>
                      : seg_hist_guarded<NS, false, true>)
              : (rows ? seg_hist_guarded<NS, true, false>
                      : seg_hist_guarded<NS, false, false>);
}

bool seg_hist_guarded_launch(HistArgs &a, bool rows, cudaStream_t st, int *rc) {
  *rc = MLC_OK;
#ifdef MLC_BOUNDS_CHECK
  {
    BoundsList bl;
    for (int k = 0; k < a.src.n; ++k)
      bl.add(a.src.ptr[k], (size_t)a.n_planes * a.plane_px * 2);
    bl.add(a.gt, (size_t)a.n_planes * a.plane_px);
    const cudaError_t e = bounds_commit(bl, st);
    if (e != cudaSuccess) {
      *rc = cuda_fail(e, "bounds_commit");
      return true;
    }
  }
#endif
  const int B = a.B, S = a.src.n;
  if (a.plane_px % 8 != 0 || a.items_per_plane != a.plane_px / 8) return false;
  if (a.plane_px / 8 >= (1ll << 31) - 2 * kGThreads) return false;
  if (!guarded_grid(a)) return false;
  // item index of output item (y, xi) in source k
  const int ipr = a.items_per_row;
  a.g_rev = 0u;
  for (int k = 0; k < S; ++k) {
    const int xf = rows ? a.src.xf[k] : 0;
    const int vy = (xf & MLC_XF_VFLIP) ? 1 : 0, hx = (xf & MLC_XF_HFLIP) ? 1 : 0;
    a.g_cy[k] = rows ? (vy ? -ipr : ipr) : 0;
    a.g_cx[k] = hx ? -1 : 1;
    a.g_b[k] = rows ? (vy ? (a.H - 1) * ipr : 0) + (hx ? ipr - 1 : 0) : 0;
    if (hx) a.g_rev |= 1u << k;
  }
  // 16 KiB of slack to align the counters inside the shared window
  const size_t smem = 16384 + (size_t)(B + 1) * 32 * 4 + (size_t)B * 4 + 16;
  const long long n_units = a.n_planes * a.units_per_plane;
  const int grid = (int)(n_units < (1ll << 22) ? n_units : (1ll << 22));
  const bool bins = a.bins != nullptr;
  // cp.async-fed variant: two stages of S x 16 + 8 bytes per thread; two
  // 512-thread CTAs per SM when they fit the 227 KiB, else one of 1024
  const size_t per_thread = 2 * ((size_t)S + 1) * 16;
  const size_t sm_bytes = 227 * 1024;
  int pf_threads = 0;
#ifndef MLC_G_NO_PF
  // (its 32-bit item index must reach every item of the tensors)
  if (S >= 2 && S <= 5 && a.n_planes * a.items_per_plane < (1ll << 32) &&
      a.n_planes < (1ll << 31) && a.units_per_plane < 65536) {
    if (S <= 4 && 2 * (smem + 512 * per_thread + 1024) <= sm_bytes)
      pf_threads = 512;
    else if (S == 5 && smem + 1024 * per_thread + 1024 <= sm_bytes)
      pf_threads = 1024;
  }
#endif
  // y = item / ipr as __umulhi(item, magic): exact while item * ipr < 2^32
  // (and ipr >= 2: the magic of 1 does not fit 32 bits)
  if (rows && (ipr < 2 || a.items_per_plane * (long long)ipr >= (1ll << 32)))
    pf_threads = 0;
  if (pf_threads) {
    a.g_ipr_magic = rows ? (uint32_t)((1ull << 32) / (unsigned)ipr + 1) : 0u;
    void (*k)(HistArgs) = S == 2   ? pick_guarded_pf<2, 512>(rows, bins)
                          : S == 3 ? pick_guarded_pf<3, 512>(rows, bins)
                          : S == 4 ? pick_guarded_pf<4, 512>(rows, bins)
                                   : pick_guarded_pf<5, 1024>(rows, bins);
    const size_t pf_smem = smem + pf_threads * per_thread;
    const cudaError_t ea = cudaFuncSetAttribute(
        k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pf_smem);
    if (ea != cudaSuccess) {
      *rc = cuda_fail(ea, "cudaFuncSetAttribute(seg_hist_guarded_pf)");
      return true;
    }
    k<<<dim3((unsigned)a.n_planes, (unsigned)a.units_per_plane), pf_threads,
        pf_smem, st>>>(a);
  } else {
    void (*k)(HistArgs) = S == 2   ? pick_guarded<2>(rows, bins)
                          : S == 3 ? pick_guarded<3>(rows, bins)
                          : S == 4 ? pick_guarded<4>(rows, bins)
                          : S == 5 ? pick_guarded<5>(rows, bins)
                          : S == 8 ? pick_guarded<8>(rows, bins)
                                   : pick_guarded<0>(rows, bins);
    k<<<grid, kGThreads, smem, st>>>(a);
  }
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) *rc = cuda_fail(err, "seg_hist_guarded");
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
