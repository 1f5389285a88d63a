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
  long long n_planes;
  long long plane_px;  // H*W
  int H, W;            // geometry used for fetches (flat mode: 1, H*W)
  int B;
  int act;
  int units_per_plane;
  long long items_per_unit;  // items of 8 px
  long long items_per_plane;
  int items_per_row;
  int flat;
  float scale, bias;
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

// ---- generic kernel: any number of sources, flips, f16/f32 ---------------
template <typename T>
__global__ void __launch_bounds__(kHistThreads)
seg_hist_generic(const __grid_constant__ HistArgs a) {
  extern __shared__ uint32_t smem[];
  const int nbins = a.B + 1;
  uint32_t *cnt = smem;                                   // [nbins][32]
  float *thr = reinterpret_cast<float *>(smem + nbins * 32);  // [B]
  for (int i = threadIdx.x; i < nbins * 32; i += blockDim.x) cnt[i] = 0;
  for (int i = threadIdx.x; i < a.B; i += blockDim.x) thr[i] = a.thr.v[i];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const long long n_units = a.n_planes * a.units_per_plane;

  for (long long unit = blockIdx.x; unit < n_units; unit += gridDim.x) {
    const long long plane = unit / a.units_per_plane;
    const int sub = (int)(unit - plane * a.units_per_plane);
    const size_t plane_off = (size_t)plane * a.plane_px;
    const uint8_t *gtp = a.gt + plane_off;
    const int shift = a.flat ? (int)(plane_off & 7) : 0;
    const long long n_items =
        a.flat ? (a.plane_px + shift + 7) / 8 : a.items_per_plane;
    const long long i0 = (long long)sub * a.items_per_unit;
    const long long i1 = min(n_items, i0 + a.items_per_unit);

    for (long long it = i0 + threadIdx.x; it < i1; it += blockDim.x) {
      int y, x0, nv;
      if (a.flat) {
        y = 0;
        x0 = it == 0 ? 0 : (int)(it * 8 - shift);
        nv = it == 0 ? min(8 - shift, a.W) : min(8, a.W - x0);
      } else {
        y = (int)(it / a.items_per_row);
        x0 = (int)(it - (long long)y * a.items_per_row) * 8;
        nv = min(8, a.W - x0);
      }
      float p[8];
      fused_prob8<T>(a.src, plane_off, a.H, a.W, y, x0, nv, a.act, p);
      float g[8];
      fetch8<uint8_t>(gtp, 0, a.H, a.W, y, x0, nv, g);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (i < nv) {
          const int c = count_ge(p[i], thr, a.B, a.scale, a.bias);
          atomicAdd(&cnt[c * 32 + lane], g[i] != 0.f ? 0x10000u : 1u);
        }
      }
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

__global__ void __launch_bounds__(kHistThreads, 2)
seg_hist_lut(const __grid_constant__ HistArgs a) {
  extern __shared__ uint32_t smem[];
  const int nbins = a.B + 1;
  uint8_t *lut = reinterpret_cast<uint8_t *>(smem);   // 64 KiB
  uint32_t *cnt = smem + MLC_LUT_BYTES / 4;           // [nbins][32]
  {
    const uint4 *g = reinterpret_cast<const uint4 *>(a.lut);
    uint4 *s = reinterpret_cast<uint4 *>(lut);
    for (int i = threadIdx.x; i < MLC_LUT_BYTES / 16; i += blockDim.x)
      s[i] = g[i];
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

// ---- fused probabilities only --------------------------------------------
struct ProbArgs {
  SrcSet src;
  float *prob;
  long long n_planes;
  long long plane_px;
  int H, W;
  int act;
  int items_per_row;
  long long items_per_plane;
};

template <typename T>
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
    fused_prob8<T>(a.src, plane_off, a.H, a.W, y, x0, nv, a.act, p);
    float *o = a.prob + plane_off + (size_t)y * a.W + x0;
    for (int i = 0; i < nv; ++i) o[i] = p[i];
  }
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
  return MLC_OK;
}

extern "C" int mlc_seg_hist(const void *const *src_dev_ptrs,
                            const int32_t *xform, int n_src, int dtype,
                            const uint8_t *gt_dev, int64_t n_img, int n_cls,
                            int H, int W, const float *thr_host, int n_thr,
                            int act, const uint8_t *lut_dev,
                            uint32_t *hist_dev, void *stream) {
  if (int rc = check_thr(thr_host, n_thr)) return rc;
  MLC_REQUIRE(dtype == MLC_F16 || dtype == MLC_F32,
              "dtype %d is not MLC_F16/MLC_F32", dtype);
  MLC_REQUIRE(act == MLC_ACT_NONE || act == MLC_ACT_SIGMOID,
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
  a.n_planes = (long long)n_img * n_cls;
  a.plane_px = (long long)H * W;
  a.B = n_thr;
  a.act = act;
  a.flat = a.src.any_xf == 0;
  a.H = a.flat ? 1 : H;
  a.W = a.flat ? (int)a.plane_px : W;
  a.items_per_row = (a.W + 7) / 8;
  // linear first guess for count_ge: k ~ (p - thr[0]) * (B-1)/(thr[B-1]-thr[0])
  const float span = thr_host[n_thr - 1] - thr_host[0];
  a.scale = (n_thr > 1 && span > 0) ? (float)(n_thr - 1) / span : 0.f;
  a.bias = 1.0f - thr_host[0] * a.scale;

  const int sms = sm_count();
  const bool table = lut_dev && n_src == 1 && dtype == MLC_F16 && a.flat &&
                     (a.plane_px % 8) == 0 &&
                     ((uintptr_t)src_dev_ptrs[0] % 16) == 0 &&
                     ((uintptr_t)gt_dev % 8) == 0;
  // flat mode may prepend one partial item to reach 16-byte alignment
  a.items_per_plane = a.flat ? (a.plane_px + 7 + 7) / 8
                             : (long long)a.items_per_row * H;
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

  if (table) {
    const size_t smem = MLC_LUT_BYTES + (size_t)(n_thr + 1) * 32 * 4;
    // per-device attribute, cheap: set on every call rather than cache
    MLC_CUDA(cudaFuncSetAttribute(seg_hist_lut,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  MLC_LUT_BYTES + (MLC_MAX_THR + 1) * 128));
    const int grid = (int)min(n_units, 2ll * sms);
    seg_hist_lut<<<grid, kHistThreads, smem, st>>>(a);
    MLC_LAUNCH_CHECK("seg_hist_lut");
  } else {
    const size_t smem = (size_t)(n_thr + 1) * 32 * 4 + (size_t)n_thr * 4;
    const int grid = (int)min(n_units, 4ll * sms);
    if (dtype == MLC_F16)
      seg_hist_generic<__half><<<grid, kHistThreads, smem, st>>>(a);
    else
      seg_hist_generic<float><<<grid, kHistThreads, smem, st>>>(a);
    MLC_LAUNCH_CHECK("seg_hist_generic");
  }
  return MLC_OK;
}

extern "C" int mlc_seg_prob(const void *const *src_dev_ptrs,
                            const int32_t *xform, int n_src, int dtype,
                            int64_t n_img, int n_cls, int H, int W, int act,
                            float *prob_dev, void *stream) {
  MLC_REQUIRE(dtype == MLC_F16 || dtype == MLC_F32,
              "dtype %d is not MLC_F16/MLC_F32", dtype);
  MLC_REQUIRE(act == MLC_ACT_NONE || act == MLC_ACT_SIGMOID,
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
  a.H = H;
  a.W = W;
  a.act = act;
  a.items_per_row = (W + 7) / 8;
  a.items_per_plane = (long long)a.items_per_row * H;
  const long long total = a.n_planes * a.items_per_plane;
  const int grid = (int)min((total + 255) / 256, 32ll * sm_count());
  if (dtype == MLC_F16)
    seg_prob_kernel<__half><<<grid, 256, 0, (cudaStream_t)stream>>>(a);
  else
    seg_prob_kernel<float><<<grid, 256, 0, (cudaStream_t)stream>>>(a);
  MLC_LAUNCH_CHECK("seg_prob_kernel");
  return MLC_OK;
}
