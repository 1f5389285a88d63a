// Shared pieces of the histogram kernels (seg_hist.cu, seg_hist_guarded.cu).
#pragma once

#include "mlc_common.cuh"

namespace mlc {

constexpr int kHistThreads = 512;
constexpr int kMaxUnitPx = 1 << 20;  // 32 lanes x 32 768 px < 65 536 per half

struct HistArgs {
  SrcSet src;
  ThrTable thr;
  const uint8_t *gt;
  uint32_t *hist;
  const uint8_t *lut;
  uint8_t *bins;  // optional [plane][H][W] bin index of every pixel (or null)
  long long n_planes;
  long long plane_px;  // H*W
  int n_cls;
  int H, W;            // geometry used for fetches (flat mode: 1, H*W)
  int B;
  int act;
  int units_per_plane;
  long long items_per_unit;  // items of 8 px
  long long items_per_plane;
  long long full_items;  // row mode: H * (W / 8)
  int items_per_row;     // row mode: W / 8 full items per row
  int flat;
  float scale, bias;
  // seg_hist_guarded (see seg_hist_guarded.cu): grid position of the mean of
  // the S sigmoids, f = q * g_kq + bias with q their sum, evaluated at
  // f + 128 -/+ the error bound (16 fractional bits in [128, 256)); per-source
  // item-index constants of the flips
  float g_kq, g_fb_lo, g_fb_hi;
  int g_b[MLC_MAX_SOURCES], g_cy[MLC_MAX_SOURCES], g_cx[MLC_MAX_SOURCES];
  uint32_t g_rev;  // bit k: source k holds its items reversed (h-flip)
  uint32_t g_ipr_magic;  // floor(2^32 / items_per_row) + 1
};

__device__ __forceinline__ void flush_bins(uint32_t *cnt, int nbins,
                                           uint32_t *out, bool atomic) {
  __syncthreads();
  for (int j = threadIdx.x; j < nbins; j += blockDim.x) {
    uint32_t lo = 0, hi = 0;
#pragma unroll 8
    for (int i = 0; i < 32; ++i) {
      const int col = (j + i) & 31;  // rotate: conflict-free column walk
      const uint32_t w = cnt[j * 32 + col];
      cnt[j * 32 + col] = 0;
      lo += w & 0xffffu;
      hi += w >> 16;
    }
    if (atomic) {
      if (lo) atomicAdd(out + j, lo);
      if (hi) atomicAdd(out + nbins + j, hi);
    } else {
      out[j] = lo;
      out[nbins + j] = hi;
    }
  }
  __syncthreads();
}

// Optional second output of the multi-source kernels: the bin index
// cnt(p) of every pixel as one byte (B <= 255), so that a later pass - RLE at
// a threshold the search has yet to choose - reads 1 byte per pixel instead of
// all the sources again: (p >= thr[k])  <=>  (bin > k).
__device__ __forceinline__ void store_bins8(uint8_t *dst, const int c[8],
                                            int nv) {
  if (nv == 8 && (reinterpret_cast<uintptr_t>(dst) & 7) == 0) {
    uint2 v;
    v.x = (uint32_t)c[0] | ((uint32_t)c[1] << 8) | ((uint32_t)c[2] << 16) |
          ((uint32_t)c[3] << 24);
    v.y = (uint32_t)c[4] | ((uint32_t)c[5] << 8) | ((uint32_t)c[6] << 16) |
          ((uint32_t)c[7] << 24);
    *reinterpret_cast<uint2 *>(dst) = v;
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (i < nv) dst[i] = (uint8_t)c[i];
  }
}

// seg_hist_guarded.cu: fp16 sources under a sigmoid, aligned layouts.
// Returns true when the job was launched (or failed: *rc), false when the
// threshold table / layout is outside what the kernel serves.
bool seg_hist_guarded_launch(HistArgs &a, bool rows, cudaStream_t st, int *rc);
// the same for rows of any width / alignment, staged by the TMA unit
// (seg_hist_guarded_rows.cu; a.H, a.W as for the fetches; W >= 9, 16-byte
// aligned source tensors, no transpose)
bool seg_hist_guarded_rows_launch(HistArgs &a, cudaStream_t st, int *rc);

}  // namespace mlc
