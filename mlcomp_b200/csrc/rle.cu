// K3: threshold (+ min-area reject) -> Kaggle run-length pairs, one CTA per
// plane, one read of the logits.  Contract: include/mlc.h (mlc_rle_encode).
//
// mask2rle (contrib/transform/rle.py:4-13) scans the mask column-major
// (img.T.flatten()).  Phase 1 reads the row-major plane with coalesced row
// segments and transposes in registers: a lane keeps 8 columns and shifts one
// bit per row into 8 words, so after 32 rows it owns 8 finished words of the
// COLUMN-major bit plane, which lives in shared memory as bits[x][Hw]
// (Hw = ceil(H/32) words per column, padding bits zero).  Phase 2 finds run
// starts / ends per word with shifts and popc/ffs, a block-wide scan gives
// every thread its output slot, the CTA reserves its slice of the output with
// one atomic and writes (start, end) then turns end into length in place.
#include "mlc_common.cuh"

namespace mlc {

constexpr size_t kRleMaxSmem = 200 * 1024;

struct RleArgs {
  SrcSet src;
  float thr[MLC_MAX_CLASSES_SEG];
  int min_area[MLC_MAX_CLASSES_SEG];
  int32_t *runs;
  long long capacity;
  long long *off;
  int32_t *cnt;
  int32_t *area;
  const int32_t *known_area;  // or null: the planes' mask areas, known already
  unsigned long long *total;
  long long n_planes;
  int C, H, W, Hw;
  int act;
  int vec_ok;  // fp16 sources 16-byte aligned, rows a multiple of 8 px
};

__device__ __forceinline__ int block_excl_scan(int v, int *warp_tot,
                                               int *total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) warp_tot[w] = inc;
  __syncthreads();
  int add = 0, all = 0;
  const int n_warps = blockDim.x >> 5;
  for (int k = 0; k < n_warps; ++k) {
    const int t = warp_tot[k];
    if (k < w) add += t;
    all += t;
  }
  __syncthreads();
  *total = all;
  return add + inc - v;
}

// fp16 code -> 16-bit key whose unsigned order is the order of the values
// (-NaN < -inf < ... < -0 < +0 < ... < +inf < +NaN).
__device__ __forceinline__ uint32_t half_key(uint32_t code) {
  return code ^ ((code & 0x8000u) ? 0xFFFFu : 0x8000u);
}
constexpr uint32_t kKeyNegInf = 0x03FFu, kKeyPosInf = 0xFC00u;

// 8 columns x `rows` rows of one fp16 plane -> 8 column words (bit r = row r
// passes).  RAW: compare the raw codes, else their order-preserving keys.
// Same rows through one packed fp16 compare per pixel pair (HSET2.GE against
// the threshold VALUE x'_c, the code of key_c): act(x) >= thr  <=>  x >= x'_c
// for the non-decreasing activation, NaN compares false, -0 and +0 share
// their activation so no bisection ever separates them.  5 instructions per
// pair (an integer range test on order-preserving keys needed 9: 0.535 ->
// 0.487 ms on 512 Severstal images, C4 0.99 -> 0.95 ms).
template <bool FULL>  // FULL: all 32 rows of the band exist - no row tests
__device__ __forceinline__ void code_rows_t(const __half *prow, int W, int rows,
                                            __half2 t2, uint32_t word[8]) {
  for (int r0 = 0; r0 < (FULL ? 32 : rows); r0 += 4) {
    uint4 v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      v[q] = (FULL || r0 + q < rows) ? ldg_stream16(prow)
                                     : make_uint4(0, 0, 0, 0);
      prow += W;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t w4[4] = {v[q].x, v[q].y, v[q].z, v[q].w};
      const uint32_t bit = (FULL || r0 + q < rows) ? 1u << (r0 + q) : 0u;
#pragma unroll
      for (int h = 0; h < 4; ++h) {
        const uint32_t m =
            __hge2_mask(*reinterpret_cast<const __half2 *>(&w4[h]), t2);
        word[2 * h] |= prmt32(m, 0u, 0x0000u) & bit;
        word[2 * h + 1] |= prmt32(m, 0u, 0x2222u) & bit;
      }
    }
  }
}
__device__ __forceinline__ void code_rows_hset(const __half *row0, int W,
                                               int rows, uint32_t t2bits,
                                               uint32_t word[8]) {
  const __half2 t2 = *reinterpret_cast<const __half2 *>(&t2bits);
  if (rows == 32) code_rows_t<true>(row0, W, rows, t2, word);
  else code_rows_t<false>(row0, W, rows, t2, word);
}


// Conflict-free store of a lane's 8 column words: with an odd column pitch
// the bank of (x, yb) is (x * Hp + yb) % 32, and lane l writing column
// 8l + ((s + l/4) & 7) in step s makes the 32 banks of a step distinct
// (8*Hp*l spreads l%4 over the multiples of 8, Hp*j spreads j over the
// residues mod 8).  The register array is rotated by l/4 with three
// conditional-swap stages instead of indexed dynamically.
__device__ __forceinline__ void store_columns(uint32_t *bits, int Hp, int x0,
                                              int yb, int nv, int lane,
                                              uint32_t word[8]) {
  const int rot = (lane >> 2) & 7;
  if (rot & 4) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t t = word[i];
      word[i] = word[i + 4];
      word[i + 4] = t;
    }
  }
  if (rot & 2) {
    const uint32_t t0 = word[0], t1 = word[1];
#pragma unroll
    for (int i = 0; i < 6; ++i) word[i] = word[i + 2];
    word[6] = t0;
    word[7] = t1;
  }
  if (rot & 1) {
    const uint32_t t0 = word[0];
#pragma unroll
    for (int i = 0; i < 7; ++i) word[i] = word[i + 1];
    word[7] = t0;
  }
#pragma unroll
  for (int s2 = 0; s2 < 8; ++s2) {  // word[s2] is column (s2 + rot) & 7
    const int j = (s2 + rot) & 7;
    if (j < nv) bits[(x0 + j) * Hp + yb] = word[s2];
  }
}

// BYTES (MODE 4): one untransformed uint8 source, no activation - masks, or
// the per-pixel bin indices mlc_seg_hist_bins keeps (mask = bin >= k).  Eight
// rows of 8 bytes in flight per lane; value >= k for four bytes at once:
//   t = ((w & 0x7F7F7F7F) | 0x80808080) - (k & 0x7F) * 0x01010101
// has bit 7 of a byte set iff its low 7 bits are >= those of k (no borrow
// crosses bytes), and with the top bit of k known per class
//   k < 128 : b >= k  <=>  top(b) |  low-part test
//   k >= 128: b >= k  <=>  top(b) &  low-part test.
template <bool FULL, bool KHI>  // FULL: 32 rows; KHI: the top bit of k
__device__ __forceinline__ void byte_rows_t(const uint8_t *prow, int W,
                                            int rows, uint32_t k7,
                                            uint32_t word[8]) {
  for (int r0 = 0; r0 < (FULL ? 32 : rows); r0 += 8) {
    uint2 v[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      v[q] = (FULL || r0 + q < rows) ? ldg_stream8(prow) : make_uint2(0u, 0u);
      prow += W;
    }
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const uint32_t bit = (FULL || r0 + q < rows) ? 1u << (r0 + q) : 0u;
      const uint32_t w2[2] = {v[q].x, v[q].y};
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const uint32_t t = ((w2[h] & 0x7F7F7F7Fu) | 0x80808080u) - k7;
        const uint32_t ge = KHI ? (t & w2[h]) : (t | w2[h]);
#pragma unroll
        for (int b = 0; b < 4; ++b)
          if (ge & (0x80u << (8 * b))) word[4 * h + b] |= bit;
      }
    }
  }
}
__device__ __forceinline__ void byte_rows(const uint8_t *row0, int W, int rows,
                                          uint32_t k, uint32_t word[8]) {
  const uint32_t k7 = (k & 0x7Fu) * 0x01010101u;
  const bool khi = (k & 0x80u) != 0, full = rows == 32;
  if (full) {
    if (khi) byte_rows_t<true, true>(row0, W, rows, k7, word);
    else byte_rows_t<true, false>(row0, W, rows, k7, word);
  } else {
    if (khi) byte_rows_t<false, true>(row0, W, rows, k7, word);
    else byte_rows_t<false, false>(row0, W, rows, k7, word);
  }
}

// CODE: one untransformed fp16 source.  act(x) >= thr is then a statement
// about the 16-bit code alone, and because the activation evaluated here is
// non-decreasing over the fp16 values (all 63 488 finite codes, asserted by
// tests/test_gpu_parity.py::test_sigmoid_is_monotone_over_fp16) it equals
// key(code) in [key_c, key(+inf)], with key_c found once per CTA and class by
// a 16-step bisection with the exact activation.  No MUFU work per pixel.
__device__ __forceinline__ int block_excl_max(int v, int *warp_tot) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc = max(inc, t);
  }
  if (lane == 31) warp_tot[w] = inc;
  __syncthreads();
  int before = 0;
  for (int k = 0; k < w; ++k) before = max(before, warp_tot[k]);
  __syncthreads();
  const int up = __shfl_up_sync(0xffffffffu, inc, 1);
  return max(before, lane ? up : 0);
}

// BOUND (MODE 2): several fp16 sources and / or flips (TTA, ensembles), an
// element-wise activation.  The mean of S activations is monotone in each of
// them (fp32 add, multiply and divide round monotonically), so with x'_c the
// smallest fp16 value for which S copies of act(x') average to >= thr
// (bisection with the very arithmetic of fused_prob8):
//   every source >= x'_c  =>  mean >= mean(act(x'_c) x S) >= thr      -> 1
//   every source <  x'_c  =>  mean <= mean(act(pred x'_c) x S) < thr  -> 0
// (NaN compares false: a pixel whose sources all fail is 0, as NaN >= thr
// is.)  One packed fp16 compare per pixel pair and source decides the pixels
// on which the variants agree; the others - the rims of the masks - are
// collected as bits and evaluated exactly, one pixel at a time, afterwards.
enum { kRleArith = 0, kRleCode = 1, kRleBound = 2, kRleBoundVec = 3,
       kRleBytes = 4 };

// One source's share of a BOUND tile: bit r of wk[i] <- source pixel under
// output pixel (y0 + r, x0 + i) >= x'_c, for the rows of `fastmask` (the
// others stay 0).  Four rows are loaded before any is compared.  VEC: rows
// are 16-byte aligned (W % 8 == 0, aligned tensors) - one 128-bit load per
// row; otherwise five aligned words per row, shifted into place, for any
// 2-byte alignment.  The 8-element window of an hflipped or partial item may
// reach up to 8 elements into the neighbouring row: the caller keeps the
// outermost rows of the tensor out of `fastmask`.
// FULL: every row of the tile is compared (fastmask all ones): no per-row
// tests.  REV: the source is h-flipped - the pairs are compared in source
// order and the results land in the mirrored columns (pair j holds pixels
// 7-2j and 6-2j), so the data is never permuted.
template <bool VEC, bool FULL, bool REV>
__device__ __forceinline__ void bound_rows_t(const __half *prow, long long step,
                                             uint32_t fastmask, __half2 t2,
                                             uint32_t wk[8]) {
  const int top = FULL ? 32 : 32 - __clz(fastmask);
  for (int r0 = 0; r0 < top; r0 += 4) {
    uint32_t w[4][4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int r = r0 + q;
      const bool on = FULL || ((fastmask >> r) & 1u);
      const __half *p = prow;
      prow += step;
      if (VEC) {
        const uint4 v = on ? ldg_stream16(p) : make_uint4(0, 0, 0, 0);
        w[q][0] = v.x; w[q][1] = v.y; w[q][2] = v.z; w[q][3] = v.w;
      } else {
        const uintptr_t a = reinterpret_cast<uintptr_t>(p);
        const uint32_t *pw = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
        const uint32_t sh = (uint32_t)(a & 2) * 8;
        uint32_t t[5];
#pragma unroll
        for (int j = 0; j < 5; ++j) t[j] = on ? ldg_nc32(pw + j) : 0u;
#pragma unroll
        for (int j = 0; j < 4; ++j)
          w[q][j] = __funnelshift_r(t[j], t[j + 1], sh);
      }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int r = r0 + q;
      const uint32_t bit = FULL ? 1u << r : ((fastmask >> r) & 1u) << r;
#pragma unroll
      for (int h = 0; h < 4; ++h) {
        const uint32_t m =
            __hge2_mask(*reinterpret_cast<const __half2 *>(&w[q][h]), t2);
        wk[REV ? 7 - 2 * h : 2 * h] |= prmt32(m, 0u, 0x0000u) & bit;
        wk[REV ? 6 - 2 * h : 2 * h + 1] |= prmt32(m, 0u, 0x2222u) & bit;
      }
    }
  }
}
template <bool VEC>
__device__ __forceinline__ void bound_rows(const __half *plane, int xf, int H,
                                           int W, int y0, int x0,
                                           uint32_t fastmask, uint32_t t2bits,
                                           uint32_t wk[8]) {
  const __half2 t2 = *reinterpret_cast<const __half2 *>(&t2bits);
  const bool rev = (xf & MLC_XF_HFLIP) != 0, vf = (xf & MLC_XF_VFLIP) != 0;
  const int xs = rev ? W - 8 - x0 : x0;
  // the row pointer advances by +-W elements: all address arithmetic that
  // does not depend on the row stays out of the loop
  const long long step = vf ? -(long long)W : (long long)W;
  const __half *prow = plane + ((long long)(vf ? H - 1 - y0 : y0) * W + xs);
  if (fastmask == 0xFFFFFFFFu) {
    if (rev) bound_rows_t<VEC, true, true>(prow, step, fastmask, t2, wk);
    else bound_rows_t<VEC, true, false>(prow, step, fastmask, t2, wk);
  } else {
    if (rev) bound_rows_t<VEC, false, true>(prow, step, fastmask, t2, wk);
    else bound_rows_t<VEC, false, false>(prow, step, fastmask, t2, wk);
  }
}
template <bool VEC, bool FULL, bool REV>
__device__ __forceinline__ void bound_rows_t(const float *prow, long long step,
                                             uint32_t fastmask, float t,
                                             uint32_t wk[8]) {
  const int top = FULL ? 32 : 32 - __clz(fastmask);
  for (int r0 = 0; r0 < top; r0 += 2) {
    float u[2][8];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int r = r0 + q;
      const bool on = FULL || ((fastmask >> r) & 1u);
      const float *p = prow;
      prow += step;
      if (VEC) {
        const uint4 a = on ? ldg_stream16(p) : make_uint4(0, 0, 0, 0);
        const uint4 b = on ? ldg_stream16(p + 4) : make_uint4(0, 0, 0, 0);
        u[q][0] = __uint_as_float(a.x); u[q][1] = __uint_as_float(a.y);
        u[q][2] = __uint_as_float(a.z); u[q][3] = __uint_as_float(a.w);
        u[q][4] = __uint_as_float(b.x); u[q][5] = __uint_as_float(b.y);
        u[q][6] = __uint_as_float(b.z); u[q][7] = __uint_as_float(b.w);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          u[q][j] = on ? __uint_as_float(ldg_nc32(p + j)) : 0.f;
      }
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const int r = r0 + q;
      const uint32_t bit = FULL ? 1u << r : ((fastmask >> r) & 1u) << r;
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (u[q][REV ? 7 - i : i] >= t) wk[i] |= bit;
    }
  }
}
template <bool VEC>
__device__ __forceinline__ void bound_rows(const float *plane, int xf, int H,
                                           int W, int y0, int x0,
                                           uint32_t fastmask, uint32_t t_bits,
                                           uint32_t wk[8]) {
  const float t = __uint_as_float(t_bits);
  const bool rev = (xf & MLC_XF_HFLIP) != 0, vf = (xf & MLC_XF_VFLIP) != 0;
  const int xs = rev ? W - 8 - x0 : x0;
  const long long step = vf ? -(long long)W : (long long)W;
  const float *prow = plane + ((long long)(vf ? H - 1 - y0 : y0) * W + xs);
  if (fastmask == 0xFFFFFFFFu) {
    if (rev) bound_rows_t<VEC, true, true>(prow, step, fastmask, t, wk);
    else bound_rows_t<VEC, true, false>(prow, step, fastmask, t, wk);
  } else {
    if (rev) bound_rows_t<VEC, false, true>(prow, step, fastmask, t, wk);
    else bound_rows_t<VEC, false, false>(prow, step, fastmask, t, wk);
  }
}
template <bool VEC>
__device__ __forceinline__ void bound_rows(const uint8_t *, int, int, int, int,
                                           int, uint32_t, uint32_t,
                                           uint32_t[8]) {}  // never selected

// exact fused probability of ONE pixel: the scalar twin of fused_prob8 for
// sources without transposes
template <typename T>
__device__ __forceinline__ float fused_prob1(const SrcSet &s, size_t plane_off,
                                             int H, int W, int y, int x,
                                             int act) {
  float p = 0.f;
  for (int k0 = 0; k0 < s.n; k0 += 4) {
    float v[4];  // up to four sources' loads in flight together
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (k0 + j < s.n) {
        const int xf = s.xf[k0 + j];
        const int ys = (xf & MLC_XF_VFLIP) ? H - 1 - y : y;
        const int xs = (xf & MLC_XF_HFLIP) ? W - 1 - x : x;
        v[j] = to_f32<T>(static_cast<const T *>(
            s.ptr[k0 + j])[plane_off + (size_t)ys * W + xs]);
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (k0 + j < s.n) {
        float t = v[j];
        if (act == MLC_ACT_SIGMOID)
          t = t >= -86.0f ? sigmoid_in_range(t) : activate(t, act);
        p = k0 + j == 0 ? t : p + t;
      }
    }
  }
  float q[8] = {p, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  mean8(q, s.n);
  return q[0];
}

// NT threads: the integer paths (MODE >= 1) need <= 64 registers and run as
// ONE 1024-thread CTA per SM when the bit plane is large, else as TWO
// 512-thread CTAs whose phases overlap (one streams while the other scans).
template <typename T, int MODE, bool SM, int NT>
__global__ void __launch_bounds__(NT, (MODE >= 1 && NT == 512) ? 2 : 1)
rle_kernel(const __grid_constant__ RleArgs a) {
  // the code path is light on registers: twice the threads, twice the loads
  // in flight per SM (one CTA per SM when the bit plane is 128 KiB)
  constexpr bool CODE = MODE == kRleCode;
  constexpr bool BOUND = MODE == kRleBound || MODE == kRleBoundVec;
  constexpr bool BYTES = MODE == kRleBytes;
  constexpr int kRleThreads = NT;
  extern __shared__ uint32_t bits[];  // [W][Hw]
  __shared__ int warp_tot[kRleThreads / 32];
  __shared__ long long s_base;
  __shared__ uint32_t key_thr[MLC_MAX_CLASSES_SEG];
  if (BYTES) {
    // smallest byte value that passes (float)b >= thr: 0 .. 256 (none)
    if ((int)threadIdx.x < a.C) {
      const float thr = a.thr[threadIdx.x];
      key_thr[threadIdx.x] = thr <= 0.f ? 0u
                             : thr > 255.f ? 256u
                                           : (uint32_t)ceilf(thr);
    }
    __syncthreads();
  }
  if (CODE || BOUND) {
    if ((int)threadIdx.x < a.C) {
      const float thr = a.thr[threadIdx.x];
      // as S equal sources: does x pass?
      auto passes = [&](float x) {
        const float v = activate(x, a.act);
        float q[8] = {v, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int k = 1; k < a.src.n; ++k) q[0] = q[0] + v;
        mean8(q, a.src.n);
        return q[0] >= thr;
      };
      if (sizeof(T) == 4) {
        // fp32 sources: same bisection over the 32-bit order-preserving keys
        // (the device sigmoid is non-decreasing over ALL fp32 values:
        // tools/check_sigmoid_monotone_f32.py sweeps the 2^32 inputs).
        // key_thr holds the BITS of the smallest passing value; a NaN
        // pattern when nothing passes (x >= NaN is false for every x).
        const uint32_t k_ninf = 0x007FFFFFu, k_pinf = 0xFF800000u;
        uint32_t lo = k_ninf, hi = k_pinf + 1;  // keys of -inf, one past +inf
        while (lo < hi) {
          const uint32_t mid = lo + ((hi - lo) >> 1);
          const uint32_t b = mid ^ ((mid & 0x80000000u) ? 0x80000000u
                                                        : 0xFFFFFFFFu);
          if (passes(__uint_as_float(b))) hi = mid;
          else lo = mid + 1;
        }
        key_thr[threadIdx.x] =
            lo > k_pinf ? 0x7FC00000u
                        : lo ^ ((lo & 0x80000000u) ? 0x80000000u : 0xFFFFFFFFu);
      } else {
        // smallest key in [key(-inf), key(+inf)] whose value passes; one past
        // +inf if none does
        uint32_t lo = kKeyNegInf, hi = kKeyPosInf + 1;
        while (lo < hi) {
          const uint32_t mid = (lo + hi) >> 1;
          const uint32_t code = mid ^ ((mid & 0x8000u) ? 0x8000u : 0xFFFFu);
          if (passes(__half2float(__ushort_as_half((unsigned short)code))))
            hi = mid;
          else
            lo = mid + 1;
        }
        key_thr[threadIdx.x] = lo;
      }
    }
    __syncthreads();
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n_warps = kRleThreads / 32;
  const int H = a.H, W = a.W, Hw = a.Hw;
  const int Hp = Hw | 1;  // odd column pitch: see the phase-1 store below
  const int nW = W * Hw;
  const int hl = (H - 1) & 31;  // top valid bit of a column's last word
  const int xchunks = (W + 255) / 256;
  const int n_tiles = Hw * xchunks;

  for (long long plane = blockIdx.x; plane < a.n_planes; plane += gridDim.x) {
    const int c = (int)(plane % a.C);
    const float thr = a.thr[c];
    const size_t plane_off = (size_t)plane * H * W;
    const bool inner = plane > 0 && plane + 1 < a.n_planes;
    if (a.known_area) {
      // the caller knows this mask's area (from the histograms): an empty or
      // rejected mask (dice.py:13-15) has no runs - nothing to read
      const int ka = a.known_area[plane];
      if (ka == 0 || (a.min_area[c] > 0 && ka < a.min_area[c])) {
        if (threadIdx.x == 0) {
          a.off[plane] = 0;
          a.cnt[plane] = 0;
          if (a.area) a.area[plane] = ka;
        }
        continue;
      }
    }

    // ---- phase 1: column-major bit plane ---------------------------------
    for (int tile = warp; tile < n_tiles; tile += n_warps) {
      const int yb = tile / xchunks;
      const int x0 = (tile - yb * xchunks) * 256 + lane * 8;
      const int nv = min(8, W - x0);
      uint32_t word[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      if (nv > 0) {
        const int rows = min(32, H - yb * 32);
        if (CODE && nv == 8 && (W & 7) == 0 && a.vec_ok) {
          // rows are 16-byte aligned: one 128-bit load per row, one packed
          // compare per pixel pair against x'_c (the value of key_c)
          const uint32_t kc = key_thr[c];
          if (kc <= kKeyPosInf) {
            const uint32_t xc = kc ^ ((kc & 0x8000u) ? 0x8000u : 0xFFFFu);
            const __half *row0 = static_cast<const __half *>(a.src.ptr[0]) +
                                 plane_off + (size_t)(yb * 32) * W + x0;
            code_rows_hset(row0, W, rows, xc | (xc << 16), word);
          }
        } else if (BYTES) {
          const uint32_t k = key_thr[c];
          const uint8_t *row0 = static_cast<const uint8_t *>(a.src.ptr[0]) +
                                plane_off + (size_t)(yb * 32) * W + x0;
          if (k == 0) {
            const uint32_t valid = rows == 32 ? ~0u : (1u << rows) - 1u;
#pragma unroll
            for (int i = 0; i < 8; ++i) word[i] = valid;
          } else if (k < 256) {
            byte_rows(row0, W, rows, k, word);
          }
        } else if (BOUND) {
          const uint32_t kc = key_thr[c];
          // fp16: key -> code, both halves of a packed pair; fp32: the bits
          const uint32_t xc = kc ^ ((kc & 0x8000u) ? 0x8000u : 0xFFFFu);
          const uint32_t t2 = sizeof(T) == 4 ? kc : (xc | (xc << 16));
          if (sizeof(T) == 4 || kc <= kKeyPosInf) {
            // rows whose windows stay inside the tensor for every flip
            const uint32_t valid = rows == 32 ? ~0u : (1u << rows) - 1u;
            uint32_t fastmask = valid;
            if (!inner) {
              if (yb == 0) fastmask &= ~1u;
              if (yb == Hw - 1) fastmask &= ~(1u << hl);
            }
            uint32_t any[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
            for (int i = 0; i < 8; ++i) word[i] = fastmask;
            if (fastmask) {
              for (int k = 0; k < a.src.n; ++k) {
                uint32_t wk[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                const T *pl = static_cast<const T *>(a.src.ptr[k]) + plane_off;
                bound_rows<MODE == kRleBoundVec>(pl, a.src.xf[k], H, W, yb * 32,
                                                 x0, fastmask, t2, wk);
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                  word[i] &= wk[i];
                  any[i] |= wk[i];
                }
              }
            }
            // undecided: the sources disagree, or the row was not compared
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              any[i] = i < nv ? (any[i] | (valid & ~fastmask)) : 0u;
              if (i >= nv) word[i] = 0u;
            }
            // Undecided pixels (sources disagree, or not compared): exact,
            // one pixel per lane and turn - ONE copy of the evaluation, each
            // lane walking its own list whatever the column.
#pragma unroll
            for (int i = 0; i < 8; ++i) any[i] &= ~word[i];
            while (true) {
              int i = -1;
              uint32_t m = 0;
#pragma unroll
              for (int j = 7; j >= 0; --j)
                if (any[j]) {
                  i = j;
                  m = any[j];
                }
              if (i < 0) break;
              const int r = __ffs(m) - 1;
              const bool pass = fused_prob1<T>(a.src, plane_off, H, W,
                                            yb * 32 + r, x0 + i, a.act) >= thr;
#pragma unroll
              for (int j = 0; j < 8; ++j)
                if (j == i) {
                  any[j] = m & (m - 1);
                  if (pass) word[j] |= 1u << r;
                }
            }
          }
        } else {
          for (int r = 0; r < rows; ++r) {
            float p[8];
            const PlaneCtx cx = {plane, a.n_planes, (long long)H * W, a.C};
            fused_prob8<T, SM>(a.src, plane_off, cx, H, W, yb * 32 + r, x0, nv,
                           inner, a.act, p);
#pragma unroll
            for (int i = 0; i < 8; ++i)
              word[i] |= (uint32_t)(i < nv && p[i] >= thr) << r;
          }
        }
        store_columns(bits, Hp, x0, yb, nv, lane, word);
      }
    }
    __syncthreads();

    // ---- phase 2: runs -----------------------------------------------------
    // Logical word q = x * Hw + yb lives at bits[x * Hp + yb].  Every thread
    // owns `per` consecutive logical words (odd: neighbouring threads start
    // in different banks).
    const int per = ((nW + kRleThreads - 1) / kRleThreads) | 1;
    const int q0 = min(nW, (int)threadIdx.x * per);
    const int q1 = min(nW, q0 + per);
    const int xq0 = q0 / Hw, ybq0 = q0 - xq0 * Hw;
    int n_starts = 0, n_px = 0, last_start = 0;
    // Sliding window over the chunk: one LDS per word; the edge bit of the
    // previous logical word (its top valid bit) is carried in a register.
    uint32_t pbit0 = 0;  // edge bit of the logical word before q0
    if (q0 > 0 && q0 < nW) {
      const int xp = ybq0 > 0 ? xq0 : xq0 - 1;
      const int yp = ybq0 > 0 ? ybq0 - 1 : Hw - 1;
      pbit0 = (bits[xp * Hp + yp] >> (yp == Hw - 1 ? hl : 31)) & 1u;
    }
    uint32_t open0 = 0;  // a run is open across the start of my chunk
    {
      int x = xq0, yb = ybq0;
      uint32_t pb = pbit0;
      uint32_t w = q0 < q1 ? bits[x * Hp + yb] : 0u;
      open0 = pb & w & 1u;
      for (int q = q0; q < q1; ++q) {
        const int top = yb == Hw - 1 ? hl : 31;
        const uint32_t starts = w & ~((w << 1) | pb);
        n_starts += __popc(starts);
        n_px += __popc(w);
        if (starts) last_start = x * H + yb * 32 + 32 - __clz(starts);  // 1-based
        pb = (w >> top) & 1u;
        if (++yb == Hw) { yb = 0; ++x; }
        w = q + 1 < q1 ? bits[x * Hp + yb] : 0u;
      }
    }
    int tot_runs, tot_px;
    const int my_first = block_excl_scan(n_starts, warp_tot, &tot_runs);
    (void)block_excl_scan(n_px, warp_tot, &tot_px);
    // start of the run that is open where my chunk begins = the latest start
    // anywhere before it
    const int carry = block_excl_max(last_start, warp_tot);
    const int min_area = a.min_area[c];
    if (min_area > 0 && tot_px < min_area) tot_runs = 0;  // dice.py:13-15 rule
    if (threadIdx.x == 0) {
      long long base = (long long)atomicAdd(a.total, 2ull * tot_runs);
      if (base + 2ll * tot_runs > a.capacity) base = -1;
      s_base = base;
      a.off[plane] = base;
      a.cnt[plane] = tot_runs;
      if (a.area) a.area[plane] = tot_px;
    }
    __syncthreads();
    const long long base = s_base;
    if (tot_runs > 0 && base >= 0) {
      // complete (start, length) pairs, one 8-byte store each, emitted by the
      // thread in whose chunk the run ENDS (rle.py:11-12)
      int2 *out = reinterpret_cast<int2 *>(a.runs + base);
      int ei = my_first - (int)open0;  // runs ended before my chunk
      int cur = carry;                 // start of the currently open run
      int x = xq0, yb = ybq0;
      uint32_t pb = pbit0;
      uint32_t w = q0 < q1 ? bits[x * Hp + yb] : 0u;
      for (int q = q0; q < q1; ++q) {
        const int top = yb == Hw - 1 ? hl : 31;
        const int pos0 = x * H + yb * 32 + 1;  // 1-based, column-major
        if (++yb == Hw) { yb = 0; ++x; }
        // next logical word (possibly the neighbour thread's first one)
        const uint32_t wn = q + 1 < nW ? bits[x * Hp + yb] : 0u;
        const uint32_t starts = w & ~((w << 1) | pb);
        uint32_t ends = w & ~((w >> 1) | ((wn & 1u) << top));
        while (ends) {
          const int e = __ffs(ends) - 1;
          ends &= ends - 1;
          const uint32_t s_le = starts & (0xFFFFFFFFu >> (31 - e));
          const int st = s_le ? pos0 + 31 - __clz(s_le) : cur;
          out[ei++] = make_int2(st, pos0 + e - st + 1);
        }
        if (starts) cur = pos0 + 31 - __clz(starts);
        pb = (w >> top) & 1u;
        w = wn;
      }
    }
    __syncthreads();
  }
}

// ---- planes whose bit plane does not fit in shared memory -------------------
// (more than ~1.6 M pixels.)  No transpose: a thread owns whole columns
// (column x -> thread x mod 512, so a warp reads 32 neighbouring columns of a
// row together), walks them twice - once to count run starts, once to emit -
// and only the per-column start counts and latest start positions live in
// shared memory (8 bytes per column).  The plane is read twice and the
// activation evaluated per pixel: a fallback for correctness, not a fast path.
constexpr int kBigThreads = 512;

template <typename T, bool SM>
__global__ void __launch_bounds__(kBigThreads)
rle_big_kernel(const __grid_constant__ RleArgs a) {
  extern __shared__ uint32_t bits[];  // reused: nst[W] then lst[W]
  __shared__ int warp_tot[kBigThreads / 32];
  __shared__ long long s_base;
  const int H = a.H, W = a.W;
  int *nst = reinterpret_cast<int *>(bits);  // starts in column x -> exclusive
  int *lst = nst + W;                        // latest start <= column x
  for (long long plane = blockIdx.x; plane < a.n_planes; plane += gridDim.x) {
    const int c = (int)(plane % a.C);
    const float thr = a.thr[c];
    const size_t plane_off = (size_t)plane * H * W;
    const PlaneCtx cx = {plane, a.n_planes, (long long)H * W, a.C};
    if (a.known_area) {
      const int ka = a.known_area[plane];
      if (ka == 0 || (a.min_area[c] > 0 && ka < a.min_area[c])) {
        if (threadIdx.x == 0) {
          a.off[plane] = 0;
          a.cnt[plane] = 0;
          if (a.area) a.area[plane] = ka;
        }
        continue;
      }
    }
    auto bit = [&](int y, int x) -> bool {
      float p[8];
      fused_prob8<T, SM>(a.src, plane_off, cx, H, W, y, x, 1, false, a.act, p);
      return p[0] >= thr;
    };
    // ---- pass 1: starts per column -----------------------------------------
    int my_px = 0;
    for (int x = threadIdx.x; x < W; x += kBigThreads) {
      bool prev = x > 0 ? bit(H - 1, x - 1) : false;
      int n = 0, last = 0;
      for (int y = 0; y < H; ++y) {
        const bool b = bit(y, x);
        if (b && !prev) {
          ++n;
          last = x * H + y + 1;  // 1-based, column-major
        }
        my_px += b;
        prev = b;
      }
      nst[x] = n;
      lst[x] = last;
    }
    __syncthreads();
    // ---- scans over the columns (thread = a contiguous chunk of columns) ----
    const int per = (W + kBigThreads - 1) / kBigThreads;
    const int c0 = min(W, (int)threadIdx.x * per), c1 = min(W, c0 + per);
    int sum = 0, mx = 0;
    for (int x = c0; x < c1; ++x) {
      sum += nst[x];
      mx = max(mx, lst[x]);
    }
    int tot_runs, tot_px;
    int running = block_excl_scan(sum, warp_tot, &tot_runs);
    (void)block_excl_scan(my_px, warp_tot, &tot_px);
    int runmax = block_excl_max(mx, warp_tot);
    for (int x = c0; x < c1; ++x) {
      const int n = nst[x], m = lst[x];
      nst[x] = running;  // starts before column x
      lst[x] = runmax;   // start of the run open at the top of column x
      running += n;
      runmax = max(runmax, m);
    }
    const int min_area = a.min_area[c];
    if (min_area > 0 && tot_px < min_area) tot_runs = 0;  // dice.py:13-15 rule
    if (threadIdx.x == 0) {
      long long base = (long long)atomicAdd(a.total, 2ull * tot_runs);
      if (base + 2ll * tot_runs > a.capacity) base = -1;
      s_base = base;
      a.off[plane] = base;
      a.cnt[plane] = tot_runs;
      if (a.area) a.area[plane] = tot_px;
    }
    __syncthreads();
    const long long base = s_base;
    // ---- pass 2: one (start, length) pair per run, written where it ends ----
    if (tot_runs > 0 && base >= 0) {
      int2 *out = reinterpret_cast<int2 *>(a.runs + base);
      for (int x = threadIdx.x; x < W; x += kBigThreads) {
        bool prev = x > 0 ? bit(H - 1, x - 1) : false;
        int idx = nst[x] - 1;  // index of the run open at the top, if any
        int cur = lst[x];
        bool b = bit(0, x);
        for (int y = 0; y < H; ++y) {
          const bool nb = y + 1 < H ? bit(y + 1, x)
                                    : (x + 1 < W ? bit(0, x + 1) : false);
          const int pos = x * H + y + 1;
          if (b && !prev) {
            ++idx;
            cur = pos;
          }
          if (b && !nb) out[idx] = make_int2(cur, pos - cur + 1);
          prev = b;
          b = nb;
        }
      }
    }
    __syncthreads();
  }
}

}  // namespace mlc

using namespace mlc;

// Developer builds (-DMLC_DEV_KNOBS) only: MLC_RLE_ARITH=1 keeps multi-source
// planes on the all-arithmetic path, MLC_RLE_GRID=1/0 forces one CTA per plane
// / a persistent grid (A/B measurements; results are identical).  The shipped
// library reads no environment.
static bool arith_only() {
#ifdef MLC_DEV_KNOBS
  static const bool v = [] {
    const char *e = getenv("MLC_RLE_ARITH");
    return e && e[0] == '1';
  }();
  return v;
#else
  return false;
#endif
}

static int rle_encode_impl(const void *const *src_dev_ptrs,
                           const int32_t *xform, int n_src, int dtype,
                           int64_t n_img, int n_cls, int H, int W,
                           const float *thr_host,
                           const int32_t *min_area_host, int act,
                           int32_t *runs_dev, int64_t capacity,
                           int64_t *off_dev, int32_t *cnt_dev,
                           int32_t *area_dev, int64_t *total_dev,
                           const int32_t *known_area_dev, void *stream) {
  MLC_REQUIRE(dtype == MLC_F16 || dtype == MLC_F32 || dtype == MLC_U8,
              "dtype %d is not MLC_F16/MLC_F32/MLC_U8", dtype);
  MLC_REQUIRE(act == MLC_ACT_NONE || act == MLC_ACT_SIGMOID ||
                  act == MLC_ACT_SOFTMAX,
              "unknown activation code %d", act);
  MLC_REQUIRE(n_img >= 0 && n_cls >= 1 && n_cls <= MLC_MAX_CLASSES_SEG &&
                  H >= 1 && W >= 1,
              "bad shape n_img=%lld n_cls=%d H=%d W=%d", (long long)n_img,
              n_cls, H, W);
  MLC_REQUIRE((long long)H * W < (1ll << 31) - 2, "plane of %d x %d too large",
              H, W);
  MLC_REQUIRE(thr_host && min_area_host, "thr_host / min_area_host is null");
  MLC_REQUIRE(off_dev && cnt_dev && total_dev, "off/cnt/total is null");
  MLC_REQUIRE(capacity >= 0 && (runs_dev || capacity == 0),
              "runs_dev is null with capacity %lld", (long long)capacity);
  MLC_REQUIRE(((uintptr_t)runs_dev % 8) == 0,
              "runs_dev must be 8-byte aligned (pairs are stored whole)");
  RleArgs a;
  const int eb = dtype == MLC_F16 ? 2 : dtype == MLC_F32 ? 4 : 1;
  if (int rc = fill_srcset(a.src, src_dev_ptrs, xform, n_src, eb, H, W))
    return rc;
  cudaStream_t st = (cudaStream_t)stream;
  MLC_CUDA(cudaMemsetAsync(total_dev, 0, 8, st));
  if (n_img == 0) return MLC_OK;
  for (int c = 0; c < n_cls; ++c) {
    MLC_REQUIRE(thr_host[c] == thr_host[c], "thr[%d] is NaN", c);
    MLC_REQUIRE(min_area_host[c] >= 0, "min_area[%d] is negative", c);
    a.thr[c] = thr_host[c];
    a.min_area[c] = min_area_host[c];
  }
  a.runs = runs_dev;
  a.capacity = capacity;
  a.off = reinterpret_cast<long long *>(off_dev);
  a.cnt = cnt_dev;
  a.area = area_dev;
  a.known_area = known_area_dev;
  a.total = reinterpret_cast<unsigned long long *>(total_dev);
  a.n_planes = (long long)n_img * n_cls;
  a.C = n_cls;
  a.H = H;
  a.W = W;
  a.Hw = (H + 31) / 32;
  a.act = act;
#ifdef MLC_BOUNDS_CHECK
  {
    BoundsList bl;
    for (int k = 0; k < n_src; ++k)
      bl.add(src_dev_ptrs[k], (size_t)a.n_planes * H * W * eb);
    MLC_CUDA(bounds_commit(bl, st));
  }
#endif
  const size_t smem = (size_t)W * (a.Hw | 1) * 4;
  if (smem > kRleMaxSmem) {
    // the column-walking fallback: 8 bytes of shared memory per column
    const size_t big = (size_t)W * 8;
    if (big > kRleMaxSmem) {
      set_error("plane %d x %d: neither its bit plane (%zu bytes) nor its "
                "column table (%zu bytes) fits the %zu bytes of shared memory "
                "this build uses",
                H, W, smem, big, kRleMaxSmem);
      return MLC_ERR_UNSUPPORTED;
    }
    const bool sm_ = act == MLC_ACT_SOFTMAX;
    void (*bk)(RleArgs) =
        dtype == MLC_F16 ? (sm_ ? rle_big_kernel<__half, true>
                                : rle_big_kernel<__half, false>)
        : dtype == MLC_F32 ? (sm_ ? rle_big_kernel<float, true>
                                  : rle_big_kernel<float, false>)
                           : rle_big_kernel<uint8_t, false>;
    MLC_CUDA(cudaFuncSetAttribute(
        bk, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kRleMaxSmem));
    const int bgrid = (int)min(a.n_planes, 2ll * sm_count());
    bk<<<bgrid, kBigThreads, big, st>>>(a);
    MLC_LAUNCH_CHECK("rle_big_kernel");
    return MLC_OK;
  }
  const bool code = dtype == MLC_F16 && n_src == 1 && a.src.any_xf == 0 &&
                    act != MLC_ACT_SOFTMAX;
  // rows a whole number of 16-byte vectors and 16-byte aligned tensors
  bool aligned = dtype == MLC_F16 ? (W % 8) == 0 : (W % 4) == 0;
  for (int k = 0; k < n_src; ++k)
    aligned = aligned && ((uintptr_t)src_dev_ptrs[k] % 16) == 0;
  a.vec_ok = dtype != MLC_U8 && aligned;
  const bool sm = act == MLC_ACT_SOFTMAX;
  // several sources / flips without a transpose, and every fp32 job:
  // bounded by packed compares
  const bool bound = (dtype == MLC_F16 || dtype == MLC_F32) && !code && !sm &&
                     W >= 8 && !(a.src.any_xf & MLC_XF_TRANSPOSE) &&
                     !arith_only();
  // uint8 masks / bin planes with whole 8-byte row segments: packed compares
  const bool bytes = dtype == MLC_U8 && n_src == 1 && a.src.any_xf == 0 &&
                     act == MLC_ACT_NONE && (W % 8) == 0 &&
                     ((uintptr_t)src_dev_ptrs[0] % 8) == 0 && !arith_only();
  // CODE: two 512-thread CTAs per SM when two bit planes (+ slack) fit, so
  // that one streams while the other scans.  BOUND: one 1024-thread CTA
  // (measured: the exact pass of a 512-thread pair costs more than the
  // overlap gains - 2.60 vs 2.36 ms on 512 Severstal 4-flip images).
  const bool wide = bound || ((code || bytes) && smem > 100 * 1024);
  const int nt = wide ? 1024 : 512;
  void (*kern)(RleArgs) =
      dtype == MLC_F16
          ? (code ? (wide ? rle_kernel<__half, kRleCode, false, 1024>
                          : rle_kernel<__half, kRleCode, false, 512>)
             : bound
                 ? (a.vec_ok
                        ? rle_kernel<__half, kRleBoundVec, false, 1024>
                        : rle_kernel<__half, kRleBound, false, 1024>)
                 : sm ? rle_kernel<__half, kRleArith, true, 512>
                      : rle_kernel<__half, kRleArith, false, 512>)
      : dtype == MLC_F32
          ? (bound ? (a.vec_ok ? rle_kernel<float, kRleBoundVec, false, 1024>
                               : rle_kernel<float, kRleBound, false, 1024>)
             : sm ? rle_kernel<float, kRleArith, true, 512>
                  : rle_kernel<float, kRleArith, false, 512>)
      : bytes ? (wide ? rle_kernel<uint8_t, kRleBytes, false, 1024>
                      : rle_kernel<uint8_t, kRleBytes, false, 512>)
              : rle_kernel<uint8_t, kRleArith, false, 512>;
  MLC_CUDA(cudaFuncSetAttribute(kern,
                                cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)kRleMaxSmem));
  // persistent grid = exactly the CTAs that are resident at once (registers,
  // threads and shared memory all count): a larger grid would run its last
  // CTAs as a second, half-empty wave
  int per_sm = 1;
  MLC_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, nt,
                                                         smem));
  per_sm = max(1, per_sm);
  // One CTA per plane, balanced by the hardware scheduler, for the light
  // 512-thread kernels (Severstal single source 0.474 -> 0.463 ms, bin planes
  // 0.487 -> 0.451 ms) and for BOUND, whose planes differ in cost with their
  // rims (60 % empty planes, disagreement along the rims: 1.69 -> 1.61 ms;
  // heavy disagreement 4.75 -> 3.28 ms; uniform synthetic planes pay 2 % for
  // the per-CTA bisection).  The single-source kernel on 135 KB bit planes
  // stays a persistent grid (C4 0.955 vs 0.98 ms).
  // (with known areas most planes end at once: let the scheduler balance)
  bool per_plane = bound || (!wide && (code || bytes)) || known_area_dev;
#ifdef MLC_DEV_KNOBS
  if (const char *e_grid = getenv("MLC_RLE_GRID")) per_plane = e_grid[0] == '1';
#endif
  const int grid = per_plane
                       ? (int)min(a.n_planes, (long long)1 << 22)
                       : (int)min(a.n_planes, (long long)per_sm * sm_count());
  kern<<<grid, nt, smem, st>>>(a);
  MLC_LAUNCH_CHECK("rle_kernel");
  return MLC_OK;
}

extern "C" int mlc_rle_encode(const void *const *src_dev_ptrs,
                              const int32_t *xform, int n_src, int dtype,
                              int64_t n_img, int n_cls, int H, int W,
                              const float *thr_host,
                              const int32_t *min_area_host, int act,
                              int32_t *runs_dev, int64_t capacity,
                              int64_t *off_dev, int32_t *cnt_dev,
                              int32_t *area_dev, int64_t *total_dev,
                              void *stream) {
  return rle_encode_impl(src_dev_ptrs, xform, n_src, dtype, n_img, n_cls, H, W,
                         thr_host, min_area_host, act, runs_dev, capacity,
                         off_dev, cnt_dev, area_dev, total_dev, nullptr,
                         stream);
}

extern "C" int mlc_rle_encode_known(const void *const *src_dev_ptrs,
                                    const int32_t *xform, int n_src, int dtype,
                                    int64_t n_img, int n_cls, int H, int W,
                                    const float *thr_host,
                                    const int32_t *min_area_host, int act,
                                    const int32_t *known_area_dev,
                                    int32_t *runs_dev, int64_t capacity,
                                    int64_t *off_dev, int32_t *cnt_dev,
                                    int32_t *area_dev, int64_t *total_dev,
                                    void *stream) {
  MLC_REQUIRE(known_area_dev, "known_area_dev is null");
  return rle_encode_impl(src_dev_ptrs, xform, n_src, dtype, n_img, n_cls, H, W,
                         thr_host, min_area_host, act, runs_dev, capacity,
                         off_dev, cnt_dev, area_dev, total_dev, known_area_dev,
                         stream);
}
