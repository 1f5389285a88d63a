// K1, guarded estimate (seg_hist_guarded.cuh), for rows of ANY width and
// alignment (350 x 525 ...): the rows are staged in shared memory by the bulk
// copy engine (TMA unit, 1-D cp.async.bulk) instead of by the threads.
//
// Why: a row of W fp16 values that is no multiple of 16 bytes starts at a
// different offset inside its 16-byte vectors in every row, and flipped
// sources add their own.  The kernel this one replaced (seg_hist_guarded_un:
// every thread copied the two 16-byte vectors that enclose its window, per
// item and source, with cp.async) paid for that with 64-bit address
// arithmetic and a 14-instruction window selection per source and item (70
// instructions per pixel against 49 of the aligned kernel), had room for a
// single stage, and read most 32-byte sectors twice (ncu: 5.5 of 32 bytes
// used per request): 0.43 of the HBM peak on 350 x 525 planes.  Here a
// producer warp asks the copy engine for "the 16-byte vectors that enclose
// these rows of this source" - every sector once - into a ring of slots that
// is many rows deep; the other threads never form a global address.
//
// (A tensor map cannot do the realignment itself: measured on B200, a tiled
// copy whose innermost start coordinate is not a multiple of 16 bytes raises
// "illegal instruction" - tools/_probe/tma_probe.cu - so the copy preserves
// the address modulo 16 and the residual offset, 0..7 elements, uniform per
// row and source, is resolved at the read: three LDS.64, five selects and
// four funnel shifts per source.)
//
// A bulk copy costs the copy engine a fixed ~129 ns per CTA whatever its size
// up to 2 KiB (tools/_probe/bulk_rate.cu: 512 B .. 2 KiB copies all take
// 128.8 ns, 4 KiB 182 ns, from 8 KiB on the HBM rate), and a row of 525 fp16
// values is 1 KiB: one copy per row and source ran at 0.33 of the HBM peak
// with the producer never idle.  Rows of a plane are contiguous, so a copy
// takes a BAND of R consecutive rows of one source (>= 4 KiB) and the rows
// are told apart at the read.
//
// One ring slot = one band of R (a power of two) output rows of the plane:
//   32 bytes per row of the band: one word per source - byte offset, from
//             the slot, of the 8-byte granule in which the window of the
//             row's output item 0 begins | element offset (0..3) of the window
//             in it - and the byte offset of the row's first truth byte.
//             Output item xi begins 16 xi bytes further on or, in a
//             horizontally flipped source, further back (its eight pixels then
//             in reverse order: register naming).
//   then per source: 16 bytes of headroom (a flipped row's partial item begins
//             before the row), the copied vectors, 16 bytes spare
//   then the truth bytes' vectors + 32 bytes spare
// The last item of a row with W % 8 != 0 holds W % 8 real pixels and then
// whatever lies behind (or, flipped, before) the row: not counted
// (finish_item nvalid).
//
// Pipeline: a full / empty mbarrier pair per slot, four to eight slots (as
// many as fit beside the counters).  Lanes 0 .. NS
// of the last warp are the producer (wait for a slot to be empty, write its
// header, issue one copy each: the sources and the truth bytes); the other
// warps take 32 consecutive items of the unit at a time from a shared
// counter; a lane waits for the band its item lies in, reads its words and
// arrives on the band's empty barrier (which counts items).
#include "seg_hist_guarded.cuh"

namespace mlc {

// ring slots: as many as fit beside the counters, at least four (the host
// sizes the bands for four), at most eight.  Depth matters: on 350 x 525
// planes three slots ran at 0.63 of the HBM peak, four at 0.71, five at 0.72
constexpr int kRowsSlots = 4;
constexpr int kRowsMaxSlots = 8;
constexpr int kRowsThreads = 512;
constexpr uint32_t kRowsSmem = 112u * 1024u;  // dynamic, two CTAs per SM
constexpr uint32_t kRowsRing = 78u * 1024u;   // of which the ring may take

struct RowGeom {
  int ipt;              // items per row, the partial last one included
  int nvalid_tail;      // pixels of a row's last item (8: it is a full one)
  uint32_t ipt_magic;   // floor(2^32 / ipt) + 1
  int log_r;            // rows per band = 1 << log_r
  uint32_t pitch;       // bytes of one source's area in a slot
  uint32_t gt_off;      // byte offset of the truth area in a slot
  uint32_t slot_bytes;
  int rows_per_unit;
  uint32_t smem_bytes;  // dynamic shared memory of the launch
};

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive_n(uint32_t bar, uint32_t n) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(n)
               : "memory");
}
// (the time hint lets the hardware park the warp until the phase completes:
// without it the waiting warps polled at full rate - ncu: half of all issued
// instructions - and starved the producer warp of issue slots)
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
      "@!p bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(bar),
      "r"(parity), "r"(100000u)
      : "memory");
}
// `bytes` (a multiple of 16) from 16-byte aligned global memory to shared
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src,
                                         uint32_t bytes, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes "
      "[%0], [%1], %2, [%3];" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar)
      : "memory");
}

struct RowRing {
  uint32_t bar_sa;  // full[s] at + 16 s, empty[s] at + 16 s + 8; behind them
                    // the next unassigned item of the CTA's unit
  uint32_t a_sa;    // slot s at + s * slot_bytes
  int n_slot;
  uint32_t slot_magic;  // floor(2^32 / n_slot) + 1
};

// All 32 lanes of the producer warp take part: lane = 8 * q + j serves copy j
// (j < NS: source j; j == NS: the truth bytes; j > NS: idle) - the lanes with
// q == 0 issue the copies, lane (q, j) writes the header words of the band's
// rows q, q + 4, ...  (The warp shares its scheduler's issue slots with seven
// busy consumer warps: a version whose copy lanes also looped over the rows'
// header words fell behind the consumers.)
template <int NS>
__device__ __forceinline__ void rows_produce(const HistArgs &a,
                                             const RowGeom &g,
                                             const RowRing &ring, uint32_t plane,
                                             int row0, int nrows) {
  const int lane = threadIdx.x & 31, j = lane & 7, q = lane >> 3;
  const int H = a.H, W = a.W, R = 1 << g.log_r;
  const size_t plane_off = (size_t)plane * a.plane_px;
  const bool src = j < NS, used = j <= NS;
  const int xf = src ? a.src.xf[j] : 0;
  const bool vf = xf & MLC_XF_VFLIP;
  const uint32_t esz = src ? 2u : 1u;
  const uintptr_t base =
      src ? reinterpret_cast<uintptr_t>(
                static_cast<const __half *>(a.src.ptr[j]) + plane_off)
          : reinterpret_cast<uintptr_t>(a.gt + plane_off);
  const uint32_t dst_off =
      src ? (32u << g.log_r) + (uint32_t)j * g.pitch + 16u : g.gt_off;
  const uint32_t pos0 = (xf & MLC_XF_HFLIP) ? (uint32_t)(W - 8) : 0u;
  int slot = 0;
  uint32_t round = 0;
  for (int r = 0; r < nrows; r += R) {
    const int nr = min(R, nrows - r), yb = row0 + r;
    const uint32_t slot_sa = ring.a_sa + slot * g.slot_bytes;
    const uint32_t full = ring.bar_sa + slot * 16u;
    // the band's rows of this source are contiguous; the lowest one is
    const int ys_lo = vf ? H - yb - nr : yb;
    const uintptr_t p = base + (size_t)ys_lo * W * esz;
    const uintptr_t from = p & ~(uintptr_t)15;
    const uint32_t bytes =
        used ? (uint32_t)(((p + (uint32_t)(nr * W) * esz + 15u) & ~(uintptr_t)15) -
                          from)
             : 0u;
    uint32_t total = bytes;
#pragma unroll
    for (int d = 1; d <= 4; d <<= 1)
      total += __shfl_xor_sync(0xffffffffu, total, d, 8);
    // (the addresses are ready before the slot is: only the header stores and
    // the copies themselves follow the wait)
    if (round > 0) mbar_wait(full + 8u, (round - 1u) & 1u);
    for (int rr = q; rr < nr; rr += 4) {
      // row rr of the band is the (nr - 1 - rr)-th of the copy when flipped
      const uint32_t at = (uint32_t)((vf ? nr - 1 - rr : rr) * W);
      if (src) {
        const uint32_t pos = (uint32_t)(p & 15) / 2u + pos0 + at;
        st_shared_u32(slot_sa + 32u * rr + 4u * j,
                      (dst_off + (pos >> 2) * 8u) | (pos & 3u));
      } else if (used) {
        st_shared_u32(slot_sa + 32u * rr + 16u, dst_off + (uint32_t)(p & 15) + at);
      }
    }
    __syncwarp();  // the header is written before the barrier is armed
    if (lane == 0) mbar_expect_tx(full, total);
    if (q == 0 && used) {
      MLC_CHECK_LOAD(reinterpret_cast<const void *>(from), bytes);
      bulk_g2s(slot_sa + dst_off, reinterpret_cast<const void *>(from), bytes,
               full);
    }
    if (++slot == ring.n_slot) {
      slot = 0;
      ++round;
    }
  }
}

template <int NS, bool BINS, int REV>
__device__ __forceinline__ void rows_consume(const HistArgs &a,
                                             const RowGeom &g,
                                             const RowRing &ring, uint32_t plane,
                                             int row0, int nrows,
                                             uint32_t slot) {
  const int lane = threadIdx.x & 31;
  const int ipt = g.ipt, W = a.W;
  const int n_items = nrows * ipt;
  const uint32_t next_sa = ring.bar_sa + kRowsMaxSlots * 16u;
  const size_t plane_off = (size_t)plane * a.plane_px;
  int guard = 0;  // always >= 0, not provably so (see pf_unit)
  while (true) {
    // (32 items per counter update: with 64 a band's last items were held
    // longer before their turn and its slot was released later - 0.56
    // against 0.64 of the HBM peak)
    uint32_t nb = 0;
    if (lane == 0)
      asm volatile("atom.shared.add.u32 %0, [%1], 32;"
                   : "=r"(nb)
                   : "r"(next_sa + (((uint32_t)guard >> 31) << 2))
                   : "memory");
    nb = __shfl_sync(0xffffffffu, nb, 0);
    if ((int)nb >= n_items) break;
    const int it = (int)nb + lane;
    guard = it;
    if (it >= n_items) continue;  // (the last chunk of a unit)
    const int r = (int)__umulhi((uint32_t)it, g.ipt_magic);
    const int xi = it - r * ipt;
    const int band = r >> g.log_r, rr = r & ((1 << g.log_r) - 1);
    const uint32_t turn = __umulhi((uint32_t)band, ring.slot_magic);
    const int sl = band - (int)turn * ring.n_slot;
    const uint32_t full = ring.bar_sa + sl * 16u;
    mbar_wait(full, turn & 1u);
    const uint32_t slot_sa = ring.a_sa + sl * g.slot_bytes;
    uint32_t w[4][4];
    const uint32_t row_sa = slot_sa + 32u * rr;
    const uint4 head = lds_u4(row_sa);
    const uint32_t hw[4] = {head.x, head.y, head.z, head.w};
#pragma unroll
    for (int j = 0; j < NS; ++j) {
      const bool rev = REV >= 0 ? ((REV >> j) & 1) : ((a.g_rev >> j) & 1u);
      // three 8-byte loads from the 8-byte granule the window begins in (a
      // lane stride of 16 bytes makes them 4 wavefronts each, 12 against the
      // 8 of two LDS.128 of the enclosing vectors; what is saved is one of
      // two levels of selects: +2.4 %)
      const uint32_t off = hw[j] & 3u;
      const uint32_t q_sa = slot_sa + (hw[j] & ~7u) +
                            (rev ? -(uint32_t)xi * 16u : (uint32_t)xi * 16u);
      const uint2 q0 = lds_u2(q_sa), q1 = lds_u2(q_sa + 8u),
                  q2 = lds_u2(q_sa + 16u);
      const bool b1 = off & 2u;
      const uint32_t b[5] = {b1 ? q0.y : q0.x, b1 ? q1.x : q0.y,
                             b1 ? q1.y : q1.x, b1 ? q2.x : q1.y,
                             b1 ? q2.y : q2.x};
      const uint32_t sh = (off & 1u) * 16u;
#pragma unroll
      for (int k = 0; k < 4; ++k) w[j][k] = __funnelshift_r(b[k], b[k + 1], sh);
      if (REV < 0 && rev) reverse8(w[j]);
    }
    uint2 gt;
    {
      const uint32_t gb = lds_u32(row_sa + 16u) + (uint32_t)xi * 8u;
      const uint32_t ga = slot_sa + (gb & ~3u), sh = (gb & 3u) * 8u;
      const uint32_t t0 = lds_u32(ga), t1 = lds_u32(ga + 4u),
                     t2 = lds_u32(ga + 8u);
      gt = make_uint2(__funnelshift_r(t0, t1, sh), __funnelshift_r(t1, t2, sh));
    }
    // release the band: every lane for its own item, after its own loads (one
    // arrive per warp and band, its count computed by lane 0 behind a
    // __syncwarp, cost 18 issue slots per item against this one: +4.5 %)
    mbar_arrive_n(full + 8u, 1u);
    {
      float q[8];
      uint32_t dmax = 0u;
      if constexpr (REV >= 0)
        group_sum<NS, REV & ((1 << NS) - 1)>(w, true, q, dmax);
      else
        group_sum<NS, 0>(w, true, q, dmax);
      const int y = row0 + r, x0 = xi * 8;
      finish_item<true, BINS ? 2 : 0>(
          a, q, dmax, gt, slot, plane, it, plane_off + (size_t)y * W + (size_t)x0,
          0u, 0u, -y - 1, x0, xi == ipt - 1 ? g.nvalid_tail : 8);
    }
  }
}

template <int NS, bool BINS>
__global__ void __launch_bounds__(kRowsThreads, 1024 / kRowsThreads)
seg_hist_guarded_rows(const __grid_constant__ HistArgs a,
                      const __grid_constant__ RowGeom g) {
  extern __shared__ __align__(128) uint32_t gsm[];
  constexpr int NT = kRowsThreads;
  const int B = a.B, nbins = B + 1;
  const uint32_t gsm_sa = (uint32_t)__cvta_generic_to_shared(gsm);
  const uint32_t end_sa = gsm_sa + g.smem_bytes;
  // counters + thresholds on the last 16 KiB boundary that leaves room for a
  // full window of 128 rows of counters (row 127 takes the pixels beyond a
  // row's end: finish_item); barriers and ring slots before them (the host
  // sized the slots for the 78 KiB that are there at least)
  const uint32_t cnt_sa = (end_sa - 16384u) & ~16383u;
  RowRing ring;
  ring.bar_sa = (gsm_sa + 15u) & ~15u;
  ring.a_sa = ring.bar_sa + kRowsMaxSlots * 16u + 16u;
  ring.n_slot = min(kRowsMaxSlots, (int)((cnt_sa - ring.a_sa) / g.slot_bytes));
  ring.slot_magic = (uint32_t)((1ull << 32) / (unsigned)max(ring.n_slot, 1) + 1);
  uint32_t *cnt = gsm + (cnt_sa - gsm_sa) / 4;
  float *thr = reinterpret_cast<float *>(cnt + nbins * 32);  // [B]
  for (int i = threadIdx.x; i < nbins * 32; i += NT) cnt[i] = 0;
  for (int i = threadIdx.x; i < B; i += NT) thr[i] = a.thr.v[i];
  if (threadIdx.x == 0) {
    st_shared_u32(ring.bar_sa + kRowsMaxSlots * 16u, 0u);
    for (int s = 0; s < ring.n_slot; ++s) {
      mbar_init(ring.bar_sa + s * 16u, 1u);
      mbar_init(ring.bar_sa + s * 16u + 8u, (uint32_t)(g.ipt << g.log_r));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const uint32_t plane = blockIdx.x, sub = blockIdx.y, upp = gridDim.y;
  const int row0 = (int)sub * g.rows_per_unit;
  const int nrows = max(0, min(a.H, row0 + g.rows_per_unit) - row0);
  if (ring.n_slot < 2) __trap();  // (the host sized the slots for four)
  if (threadIdx.x >= NT - 32) {
    rows_produce<NS>(a, g, ring, plane, row0, nrows);
  } else {
    const uint32_t slot = cnt_sa | ((threadIdx.x & 31) * 4);
    constexpr int kAll = (1 << NS) - 1;
    const int rev = (int)(a.g_rev & kAll);
    if (rev == 0)
      rows_consume<NS, BINS, 0>(a, g, ring, plane, row0, nrows, slot);
    else if (rev == (0xA & kAll))
      rows_consume<NS, BINS, 0xA & kAll>(a, g, ring, plane, row0, nrows, slot);
    else if (rev == (0x5 & kAll))
      rows_consume<NS, BINS, 0x5 & kAll>(a, g, ring, plane, row0, nrows, slot);
    else if (rev == kAll)
      rows_consume<NS, BINS, kAll>(a, g, ring, plane, row0, nrows, slot);
    else
      rows_consume<NS, BINS, -1>(a, g, ring, plane, row0, nrows, slot);
  }
  uint32_t *out = a.hist + (size_t)plane * 2 * nbins;
  flush_bins(cnt, nbins, out, upp > 1);
}

// Requires 16-byte aligned source tensors (every vector that is copied then
// contains at least one byte of its tensor), W >= 9, no transpose.
bool seg_hist_guarded_rows_launch(HistArgs &a, cudaStream_t st, int *rc) {
  *rc = MLC_OK;
  const int S = a.src.n, W = a.W, H = a.H;
  if (S < 2 || S > 4) return false;
  if (a.src.any_xf & MLC_XF_TRANSPOSE) return false;
  for (int k = 0; k < S; ++k)
    if (((uintptr_t)a.src.ptr[k] % 16) != 0) return false;
  RowGeom g;
  g.ipt = (W + 7) / 8;
  if (g.ipt < 2) return false;  // (the magic of a division by 1 has 33 bits)
  if ((long long)H * g.ipt * g.ipt >= (1ll << 32) || a.n_planes >= (1ll << 31) ||
      a.units_per_plane >= 65536)
    return false;
  g.nvalid_tail = (W & 7) ? (W & 7) : 8;
  g.ipt_magic = (uint32_t)((1ull << 32) / (unsigned)g.ipt + 1);
  // rows per band: the largest power of two whose four slots fit the ring (a fifth .. eighth slot is used when it fits)
  // (a source's area: 16 bytes of headroom, the vectors that enclose the
  // band's rows, 16 bytes spare for the second vector of the last window)
  g.log_r = -1;
  for (int lr = 0; lr <= 6; ++lr) {
    const uint32_t R = 1u << lr;
    const uint32_t pitch = 16u * ((2u * W * R + 30u + 15u) / 16u + 2u);
    const uint32_t gt_off = 32u * R + (uint32_t)S * pitch;
    const uint32_t slot =
        gt_off + 16u * ((W * R + 15u + 15u) / 16u + 3u);
    if (kRowsSlots * slot > kRowsRing) break;
    g.log_r = lr;
    g.pitch = pitch;
    g.gt_off = gt_off;
    g.slot_bytes = slot;
  }
  if (g.log_r < 0) return false;
  g.smem_bytes = kRowsSmem;
  g.rows_per_unit = (H + a.units_per_plane - 1) / a.units_per_plane;
  if (!guarded_grid(a)) return false;
#ifdef MLC_BOUNDS_CHECK
  {  // the copies may take the 16-byte vectors that enclose a tensor's ends
    BoundsList bl;
    for (int k = 0; k < S; ++k)
      bl.add(a.src.ptr[k], ((size_t)a.n_planes * a.plane_px * 2 + 15) / 16 * 16);
    const uintptr_t g0 = (uintptr_t)a.gt & ~(uintptr_t)15;
    bl.add(reinterpret_cast<const void *>(g0),
           ((uintptr_t)a.gt + (size_t)a.n_planes * a.plane_px + 15) / 16 * 16 - g0);
    const cudaError_t e = bounds_commit(bl, st);
    if (e != cudaSuccess) {
      *rc = cuda_fail(e, "bounds_commit");
      return true;
    }
  }
#endif
  a.g_rev = 0u;
  for (int k = 0; k < S; ++k)
    if (a.src.xf[k] & MLC_XF_HFLIP) a.g_rev |= 1u << k;
  const bool bins = a.bins != nullptr;
  void (*k)(HistArgs, RowGeom) =
      S == 2   ? (bins ? seg_hist_guarded_rows<2, true>
                       : seg_hist_guarded_rows<2, false>)
      : S == 3 ? (bins ? seg_hist_guarded_rows<3, true>
                       : seg_hist_guarded_rows<3, false>)
               : (bins ? seg_hist_guarded_rows<4, true>
                       : seg_hist_guarded_rows<4, false>);
  const cudaError_t ea = cudaFuncSetAttribute(
      k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g.smem_bytes);
  if (ea != cudaSuccess) {
    *rc = cuda_fail(ea, "cudaFuncSetAttribute(seg_hist_guarded_rows)");
    return true;
  }
  k<<<dim3((unsigned)a.n_planes, (unsigned)a.units_per_plane), kRowsThreads,
      g.smem_bytes, st>>>(a, g);
  const cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) *rc = cuda_fail(err, "seg_hist_guarded_rows");
  return true;
}

}  // namespace mlc
