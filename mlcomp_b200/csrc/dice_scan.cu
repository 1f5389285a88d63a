// K1b: per-plane histograms -> suffix scan -> P(t), I(t), T -> dice for the
// whole (threshold x min-area) grid, summed over images in a fixed order.
// Contract: include/mlc.h (mlc_dice_search, mlc_plane_dice).
#include "mlc_common.cuh"

namespace mlc {

struct AreaTable {
  int v[MLC_MAX_CLASSES_SEG * MLC_MAX_AREAS];
};

// dice.py:13-23 from exact integers.  t, p, i < 2^31.
__device__ __forceinline__ double dice_point(unsigned t, unsigned p, unsigned i,
                                             int min_area, int empty_one,
                                             double eps) {
  if (min_area > 0 && p < (unsigned)min_area) {  // dice.py:13-15
    p = 0;
    i = 0;
  }
  if (empty_one && t == 0 && p == 0) return 1.0;  // dice.py:17-18
  // dice.py:20-23: 2*I (exact) / (T + P + eps) in fp64
  return (double)(2ull * i) / ((double)((unsigned long long)t + p) + eps);
}

// Inclusive suffix sum over j (bins) of packed (total << 32 | positives)
// across the CTA; returns S[j] = sum_{i >= j} v[i].
__device__ __forceinline__ unsigned long long block_suffix_sum(
    unsigned long long v, unsigned long long *warp_tot) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int n_warps = blockDim.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned long long t = __shfl_down_sync(0xffffffffu, v, o);
    if (lane + o < 32) v += t;
  }
  if (lane == 0) warp_tot[w] = v;
  __syncthreads();
  unsigned long long add = 0;
  for (int k = w + 1; k < n_warps; ++k) add += warp_tot[k];
  __syncthreads();
  return v + add;
}

struct SearchArgs {
  const uint32_t *hist;
  double *partial;               // [n_chunks][C][B][A]
  double *partial_iou;           // same shape, or null
  unsigned long long *micro;     // [C][1+2B] or null
  AreaTable areas;               // [C][A]
  long long n_img;
  int C, B, A;
  int imgs_per_chunk;
  int empty_one;
  double eps;
};

// grid = (C, n_chunks); thread j owns bin j (j <= B) and threshold k = j.
__global__ void __launch_bounds__(256)
dice_search_kernel(const __grid_constant__ SearchArgs a) {
  __shared__ unsigned long long warp_tot[8];
  __shared__ unsigned long long S[MLC_MAX_THR + 2];
  const int c = blockIdx.x, chunk = blockIdx.y;
  const int j = threadIdx.x;
  const int nb = a.B + 1;
  const long long n0 = (long long)chunk * a.imgs_per_chunk;
  const long long n1 = min(a.n_img, n0 + a.imgs_per_chunk);
  double acc[MLC_MAX_AREAS], acc_iou[MLC_MAX_AREAS];
#pragma unroll
  for (int m = 0; m < MLC_MAX_AREAS; ++m) acc[m] = acc_iou[m] = 0.0;
  unsigned long long sumP = 0, sumI = 0, sumT = 0;
  int area[MLC_MAX_AREAS];
#pragma unroll
  for (int m = 0; m < MLC_MAX_AREAS; ++m)
    area[m] = m < a.A ? a.areas.v[c * a.A + m] : 0;

  // the loop is a chain of dependent global loads (one image after the
  // other): fetch image n+1 while image n is scanned
  unsigned nh0 = 0, nh1 = 0;
  if (j < nb && n0 < n1) {
    const uint32_t *h = a.hist + ((size_t)n0 * a.C + c) * 2 * nb;
    nh0 = h[j];
    nh1 = h[nb + j];
  }
  for (long long n = n0; n < n1; ++n) {
    const unsigned h0 = nh0, h1 = nh1;
    if (j < nb && n + 1 < n1) {
      const uint32_t *h = a.hist + ((size_t)(n + 1) * a.C + c) * 2 * nb;
      nh0 = h[j];
      nh1 = h[nb + j];
    }
    unsigned long long v = 0;
    if (j < nb) v = ((unsigned long long)(h0 + h1) << 32) | h1;
    const unsigned long long s = block_suffix_sum(v, warp_tot);
    if (j < nb) S[j] = s;
    if (j == 0) S[nb] = 0ull;
    __syncthreads();
    if (j < a.B) {
      const unsigned long long up = S[j + 1];  // bins above threshold j
      const unsigned p = (unsigned)(up >> 32), i = (unsigned)up;
      const unsigned t = (unsigned)S[0];
      // One divide per (plane, threshold): the min-area rule only switches
      // between the full dice and the dice of an emptied prediction
      // (dice.py:13-18), so the candidates differ by a select.
      const double d_full = dice_point(t, p, i, 0, a.empty_one, a.eps);
      const double d_zero = (a.empty_one && t == 0) ? 1.0 : 0.0;
#pragma unroll
      for (int m = 0; m < MLC_MAX_AREAS; ++m)
        if (m < a.A)
          acc[m] += (area[m] > 0 && p < (unsigned)area[m]) ? d_zero : d_full;
      if (a.partial_iou) {
        // IoU of the same masks under the same two rules: I / (T + P - I +
        // eps); an emptied prediction has I = P = 0 (builder-defined: the
        // reference has no IoU)
        const double u_full =
            (a.empty_one && t == 0 && p == 0)
                ? 1.0
                : (double)i / ((double)((unsigned long long)t + p - i) + a.eps);
#pragma unroll
        for (int m = 0; m < MLC_MAX_AREAS; ++m)
          if (m < a.A)
            acc_iou[m] +=
                (area[m] > 0 && p < (unsigned)area[m]) ? d_zero : u_full;
      }
      sumP += p;
      sumI += i;
      if (j == 0) sumT += t;
    }
    __syncthreads();
  }
  if (j < a.B) {
    double *out = a.partial +
                  (((size_t)chunk * a.C + c) * a.B + j) * a.A;
#pragma unroll
    for (int m = 0; m < MLC_MAX_AREAS; ++m)
      if (m < a.A) out[m] = acc[m];
    if (a.partial_iou) {
      double *o2 = a.partial_iou +
                   (((size_t)chunk * a.C + c) * a.B + j) * a.A;
#pragma unroll
      for (int m = 0; m < MLC_MAX_AREAS; ++m)
        if (m < a.A) o2[m] = acc_iou[m];
    }
    if (a.micro) {
      unsigned long long *mo = a.micro + (size_t)c * (1 + 2 * a.B);
      if (j == 0 && sumT) atomicAdd(mo, sumT);
      if (sumP) atomicAdd(mo + 1 + j, sumP);
      if (sumI) atomicAdd(mo + 1 + a.B + j, sumI);
    }
  }
}

// Sum the per-chunk partials in chunk order (fixed order => reproducible).
__global__ void dice_reduce_kernel(const double *partial, double *out,
                                   int n_chunks, long long cells) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= cells) return;
  double s = 0.0;
  for (int k = 0; k < n_chunks; ++k) s += partial[(size_t)k * cells + i];
  out[i] = s;
}

// Image chunks per class.  A function of the SHAPE alone, on purpose: the
// workspace size (mlc_dice_search_workspace) must not depend on which device
// is current, and the chunk count fixes the fp64 summation order, so results
// are bit-identical on any sm_100 part.  4 x 148 CTAs fill a B200; a part
// with another SM count only runs a slightly different number of waves.
static int chunks_for(int64_t n_img, int n_cls) {
  long long per_cls = (4ll * 148 + n_cls - 1) / n_cls;
  if (per_cls < 1) per_cls = 1;
  return (int)(n_img < per_cls ? (n_img < 1 ? 1 : n_img) : per_cls);
}

struct PlaneArgs {
  const uint32_t *hist;
  double *dice;        // or null
  int32_t *area_out;   // or null: P(k), the mask area before the min-area rule
  long long n_planes;
  int C, B;
  int empty_one;
  double eps;
  int k[MLC_MAX_CLASSES_SEG];
  int area[MLC_MAX_CLASSES_SEG];
};

// one warp per plane
__global__ void __launch_bounds__(256)
plane_dice_kernel(const __grid_constant__ PlaneArgs a) {
  const int lane = threadIdx.x & 31;
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (warp >= a.n_planes) return;
  const int c = (int)(warp % a.C);
  const int nb = a.B + 1;
  const uint32_t *h = a.hist + (size_t)warp * 2 * nb;
  const int k = a.k[c];
  unsigned t = 0, p = 0, i = 0;
  for (int j = lane; j < nb; j += 32) {
    const unsigned h0 = h[j], h1 = h[nb + j];
    t += h1;
    if (j > k) {
      p += h0 + h1;
      i += h1;
    }
  }
  for (int o = 16; o; o >>= 1) {
    t += __shfl_xor_sync(0xffffffffu, t, o);
    p += __shfl_xor_sync(0xffffffffu, p, o);
    i += __shfl_xor_sync(0xffffffffu, i, o);
  }
  if (lane == 0) {
    if (a.dice)
      a.dice[warp] = dice_point(t, p, i, a.area[c], a.empty_one, a.eps);
    if (a.area_out) a.area_out[warp] = (int32_t)p;
  }
}

}  // namespace mlc

using namespace mlc;

extern "C" size_t mlc_dice_search_workspace(int64_t n_img, int n_cls, int n_thr,
                                            int n_area) {
  if (n_img < 0 || n_cls < 1 || n_thr < 1 || n_area < 1) return 0;
  return (size_t)chunks_for(n_img, n_cls) * n_cls * n_thr * n_area *
         sizeof(double);
}

static int dice_search_impl(const uint32_t *hist_dev, int64_t n_img,
                            int n_cls, int n_thr,
                            const int32_t *min_area_host, int n_area,
                            int empty_one, double eps, double *dice_sum_dev,
                            double *iou_sum_dev, int64_t *micro_dev,
                            void *work_dev, size_t work_bytes, void *stream) {
  MLC_REQUIRE(hist_dev && dice_sum_dev && min_area_host,
              "hist/dice_sum/min_area pointer is null");
  MLC_REQUIRE(n_img >= 0 && n_cls >= 1 && n_cls <= MLC_MAX_CLASSES_SEG,
              "bad n_img=%lld / n_cls=%d (max %d classes)", (long long)n_img,
              n_cls, MLC_MAX_CLASSES_SEG);
  MLC_REQUIRE(n_thr >= 1 && n_thr <= MLC_MAX_THR, "n_thr=%d outside [1, %d]",
              n_thr, MLC_MAX_THR);
  MLC_REQUIRE(n_area >= 1 && n_area <= MLC_MAX_AREAS,
              "n_area=%d outside [1, %d]", n_area, MLC_MAX_AREAS);
  const size_t one = mlc_dice_search_workspace(n_img, n_cls, n_thr, n_area);
  const size_t need = iou_sum_dev ? 2 * one : one;
  if (!work_dev || work_bytes < need) {
    set_error("workspace of %zu bytes needed, %zu given", need, work_bytes);
    return MLC_ERR_CAPACITY;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const long long cells = (long long)n_cls * n_thr * n_area;
  if (micro_dev)
    MLC_CUDA(cudaMemsetAsync(micro_dev, 0,
                             (size_t)n_cls * (1 + 2 * n_thr) * 8, st));
  if (n_img == 0) {
    MLC_CUDA(cudaMemsetAsync(dice_sum_dev, 0, (size_t)cells * 8, st));
    if (iou_sum_dev)
      MLC_CUDA(cudaMemsetAsync(iou_sum_dev, 0, (size_t)cells * 8, st));
    return MLC_OK;
  }
  SearchArgs a;
  a.hist = hist_dev;
  a.partial = static_cast<double *>(work_dev);
  a.partial_iou = iou_sum_dev
                      ? reinterpret_cast<double *>(
                            static_cast<char *>(work_dev) + one)
                      : nullptr;
  a.micro = reinterpret_cast<unsigned long long *>(micro_dev);
  for (int i = 0; i < n_cls * n_area; ++i) {
    MLC_REQUIRE(min_area_host[i] >= 0, "min_area[%d] is negative", i);
    a.areas.v[i] = min_area_host[i];
  }
  a.n_img = n_img;
  a.C = n_cls;
  a.B = n_thr;
  a.A = n_area;
  const int n_chunks = chunks_for(n_img, n_cls);
  a.imgs_per_chunk = (int)((n_img + n_chunks - 1) / n_chunks);
  a.empty_one = empty_one;
  a.eps = eps;
  const int threads = ((n_thr + 1) + 31) / 32 * 32;  // one per bin
  dice_search_kernel<<<dim3(n_cls, n_chunks), threads, 0, st>>>(a);
  MLC_LAUNCH_CHECK("dice_search_kernel");
  dice_reduce_kernel<<<(unsigned)((cells + 127) / 128), 128, 0, st>>>(
      a.partial, dice_sum_dev, n_chunks, cells);
  MLC_LAUNCH_CHECK("dice_reduce_kernel");
  if (iou_sum_dev) {
    dice_reduce_kernel<<<(unsigned)((cells + 127) / 128), 128, 0, st>>>(
        a.partial_iou, iou_sum_dev, n_chunks, cells);
    MLC_LAUNCH_CHECK("dice_reduce_kernel");
  }
  return MLC_OK;
}

extern "C" int mlc_dice_search(const uint32_t *hist_dev, int64_t n_img,
                               int n_cls, int n_thr,
                               const int32_t *min_area_host, int n_area,
                               int empty_one, double eps, double *dice_sum_dev,
                               int64_t *micro_dev, void *work_dev,
                               size_t work_bytes, void *stream) {
  return dice_search_impl(hist_dev, n_img, n_cls, n_thr, min_area_host, n_area,
                          empty_one, eps, dice_sum_dev, nullptr, micro_dev,
                          work_dev, work_bytes, stream);
}

extern "C" int mlc_dice_iou_search(const uint32_t *hist_dev, int64_t n_img,
                                   int n_cls, int n_thr,
                                   const int32_t *min_area_host, int n_area,
                                   int empty_one, double eps,
                                   double *dice_sum_dev, double *iou_sum_dev,
                                   int64_t *micro_dev, void *work_dev,
                                   size_t work_bytes, void *stream) {
  MLC_REQUIRE(iou_sum_dev, "iou_sum_dev is null");
  return dice_search_impl(hist_dev, n_img, n_cls, n_thr, min_area_host, n_area,
                          empty_one, eps, dice_sum_dev, iou_sum_dev, micro_dev,
                          work_dev, work_bytes, stream);
}

extern "C" int mlc_plane_dice(const uint32_t *hist_dev, int64_t n_img,
                              int n_cls, int n_thr, const int32_t *k_host,
                              const int32_t *min_area_host, int empty_one,
                              double eps, double *dice_dev, void *stream) {
  MLC_REQUIRE(hist_dev && dice_dev && k_host && min_area_host,
              "hist/dice/k/min_area pointer is null");
  MLC_REQUIRE(n_img >= 0 && n_cls >= 1 && n_cls <= MLC_MAX_CLASSES_SEG,
              "bad n_img=%lld / n_cls=%d (max %d classes)", (long long)n_img,
              n_cls, MLC_MAX_CLASSES_SEG);
  MLC_REQUIRE(n_thr >= 1 && n_thr <= MLC_MAX_THR, "n_thr=%d outside [1, %d]",
              n_thr, MLC_MAX_THR);
  if (n_img == 0) return MLC_OK;
  PlaneArgs a;
  a.hist = hist_dev;
  a.dice = dice_dev;
  a.area_out = nullptr;
  a.n_planes = (long long)n_img * n_cls;
  a.C = n_cls;
  a.B = n_thr;
  a.empty_one = empty_one;
  a.eps = eps;
  for (int c = 0; c < n_cls; ++c) {
    MLC_REQUIRE(k_host[c] >= 0 && k_host[c] < n_thr,
                "threshold index %d of class %d outside [0, %d)", k_host[c], c,
                n_thr);
    MLC_REQUIRE(min_area_host[c] >= 0, "min_area of class %d is negative", c);
    a.k[c] = k_host[c];
    a.area[c] = min_area_host[c];
  }
  const long long blocks = (a.n_planes + 7) / 8;
  plane_dice_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(a);
  MLC_LAUNCH_CHECK("plane_dice_kernel");
  return MLC_OK;
}

extern "C" int mlc_plane_area(const uint32_t *hist_dev, int64_t n_img,
                              int n_cls, int n_thr, const int32_t *k_host,
                              int32_t *area_dev, void *stream) {
  MLC_REQUIRE(hist_dev && area_dev && k_host, "hist/area/k pointer is null");
  MLC_REQUIRE(n_img >= 0 && n_cls >= 1 && n_cls <= MLC_MAX_CLASSES_SEG,
              "bad n_img=%lld / n_cls=%d (max %d classes)", (long long)n_img,
              n_cls, MLC_MAX_CLASSES_SEG);
  MLC_REQUIRE(n_thr >= 1 && n_thr <= MLC_MAX_THR, "n_thr=%d outside [1, %d]",
              n_thr, MLC_MAX_THR);
  if (n_img == 0) return MLC_OK;
  PlaneArgs a;
  a.hist = hist_dev;
  a.dice = nullptr;
  a.area_out = area_dev;
  a.n_planes = (long long)n_img * n_cls;
  a.C = n_cls;
  a.B = n_thr;
  a.empty_one = 0;
  a.eps = 0.0;
  for (int c = 0; c < n_cls; ++c) {
    MLC_REQUIRE(k_host[c] >= 0 && k_host[c] < n_thr,
                "threshold index %d of class %d outside [0, %d)", k_host[c], c,
                n_thr);
    a.k[c] = k_host[c];
    a.area[c] = 0;
  }
  const long long blocks = (a.n_planes + 7) / 8;
  plane_dice_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(a);
  MLC_LAUNCH_CHECK("plane_dice_kernel");
  return MLC_OK;
}

// ---------------------------------------------------------------------------
// mlc_dice_search_static: the same search for callers that own no workspace.
//
// mlc_dice_search needs a scratch buffer for its per-chunk partial sums and
// expects the caller to allocate it.  Two kinds of caller of this library
// cannot: a scoring loop that runs beside an inference engine whose memory
// pool was sized to the whole card at start-up (any later cudaMalloc fails or
// is forbidden by the deployment), and a thin binding from the reference's
// worker process (ctypes + NumPy), which receives its input
// and output buffers from somebody else and has no device allocator at all.
// For them the scratch is this fixed array inside the library image: every
// GPU that loads libmlcomp_b200.so carries these 4 MiB until the process
// exits, whether or not the entry point is ever called.
//
// The array's name is only meaningful to device code; host code must not
// pass it to a kernel.  The device address is looked up once per device with
// cudaGetSymbolAddress (that device current) and cached.
// ---------------------------------------------------------------------------
namespace mlc {
constexpr size_t kStaticPartialDoubles = 512 * 1024;  // 4 MiB
static __device__ __align__(256) double g_dice_partials[kStaticPartialDoubles];

// Device address of g_dice_partials on the CURRENT device, or nullptr with
// *err set.  First use on a device performs the symbol look-up.
static double *static_partials(cudaError_t *err) {
  // written without a lock: two threads racing on a device's first call both
  // look the symbol up and store the same address (idempotent)
  static double *cache[64] = {nullptr};
  int dev = 0;
  *err = cudaGetDevice(&dev);
  if (*err != cudaSuccess) return nullptr;
  if (dev < 0 || dev >= 64) {
    *err = cudaErrorInvalidDevice;
    return nullptr;
  }
  if (cache[dev] == nullptr) {
    void *p = nullptr;
    *err = cudaGetSymbolAddress(&p, g_dice_partials);
    if (*err != cudaSuccess) return nullptr;
    cache[dev] = static_cast<double *>(p);
  }
  return cache[dev];
}
}  // namespace mlc

extern "C" int mlc_dice_search_static(const uint32_t *hist_dev, int64_t n_img,
                                      int n_cls, int n_thr,
                                      const int32_t *min_area_host, int n_area,
                                      int empty_one, double eps,
                                      double *dice_sum_dev, int64_t *micro_dev,
                                      void *stream) {
  const size_t cap = kStaticPartialDoubles * sizeof(double);
  // shape errors are reported by mlc_dice_search below; here only the size
  const size_t need = mlc_dice_search_workspace(n_img, n_cls, n_thr, n_area);
  if (need > cap) {
    set_error("mlc_dice_search_static: this search needs %zu bytes of scratch "
              "but the built-in array holds %zu (n_cls*n_thr*n_area = %lld "
              "cells); use mlc_dice_search with a caller-provided workspace",
              need, cap, (long long)n_cls * n_thr * n_area);
    return MLC_ERR_CAPACITY;
  }
  cudaError_t e = cudaSuccess;
  double *scratch = static_partials(&e);
  if (scratch == nullptr)
    return cuda_fail(e, "cudaGetSymbolAddress(g_dice_partials)");
  // Nothing in the array has to start at zero: dice_search_kernel overwrites
  // every partial that dice_reduce_kernel reads back (empty chunks write
  // zeros), and micro_dev is cleared by mlc_dice_search itself.
  return mlc_dice_search(hist_dev, n_img, n_cls, n_thr, min_area_host, n_area,
                         empty_one, eps, dice_sum_dev, micro_dev, scratch, cap,
                         stream);
}
