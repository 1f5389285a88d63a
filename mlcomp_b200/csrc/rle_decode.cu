// Inverse of K3: (start, length) pairs -> uint8 masks on the device.
// rle2mask (contrib/transform/rle.py:16-31) builds ground truth from the
// competition's RLE strings with a Python loop over the runs; a validation
// set of tens of thousands of planes is decoded here in one launch.
// Contract: include/mlc.h (mlc_rle_decode).
//
// One CTA per plane.  The runs are painted into a COLUMN-major bit plane in
// shared memory (bits[x][Hp], Hp odd: the transposed read below is
// conflict-free), boundary words with atomicOr, interior words with plain
// stores; the CTA then writes the row-major bytes with coalesced stores.
// Planes too large for shared memory are painted pixel by pixel in global
// memory.
#include "mlc_common.cuh"

namespace mlc {

constexpr int kDecThreads = 512;
constexpr size_t kDecMaxSmem = 200 * 1024;

struct DecArgs {
  const int32_t *runs;
  const long long *off;
  const int32_t *cnt;
  uint8_t *mask;
  long long n_planes;
  int H, W, Hw, Hp;
  int in_smem;
};

__global__ void __launch_bounds__(kDecThreads)
rle_decode_kernel(const __grid_constant__ DecArgs a) {
  extern __shared__ uint32_t bits[];  // [W][Hp]
  const int H = a.H, W = a.W, Hp = a.Hp;
  const long long px = (long long)H * W;
  for (long long plane = blockIdx.x; plane < a.n_planes; plane += gridDim.x) {
    const long long src = a.off[plane];
    const int n_runs = src < 0 ? 0 : a.cnt[plane];
    const int2 *runs = reinterpret_cast<const int2 *>(a.runs + (src < 0 ? 0 : src));
    uint8_t *out = a.mask + plane * px;
    if (!a.in_smem) {  // big plane: zero, then paint run by run
      for (long long i = threadIdx.x; i < px; i += kDecThreads) out[i] = 0;
      __syncthreads();
      for (int r = threadIdx.x; r < n_runs; r += kDecThreads) {
        const int2 run = runs[r];
        // clipped to the plane on both sides: a malformed row (start < 1,
        // negative length, run past the end) must not write outside it
        long long p = (long long)run.x - 1, e = min(px, p + run.y);
        p = max(p, 0ll);
        for (; p < e; ++p) out[(p % H) * W + p / H] = 1;
      }
      __syncthreads();
      continue;
    }
    for (int i = threadIdx.x; i < W * Hp; i += kDecThreads) bits[i] = 0;
    __syncthreads();
    for (int r = threadIdx.x; r < n_runs; r += kDecThreads) {
      const int2 run = runs[r];
      long long p = (long long)run.x - 1;               // 0-based, col-major
      long long e = min(px, p + (long long)run.y);      // exclusive
      p = max(p, 0ll);  // start < 1 in a malformed row: clip, do not write
      while (p < e) {
        const int x = (int)(p / H), y = (int)(p - (long long)x * H);
        const int in_col = (int)min((long long)(H - y), e - p);  // this column
        int yy = y, left = in_col;
        while (left > 0) {
          const int b = yy & 31, n = min(32 - b, left);
          const uint32_t m = (n == 32 ? 0xFFFFFFFFu : ((1u << n) - 1u)) << b;
          uint32_t *wd = &bits[x * Hp + (yy >> 5)];
          if (n == 32) *wd = m;  // interior word: no other run touches it
          else atomicOr(wd, m);
          yy += n;
          left -= n;
        }
        p += in_col;
      }
    }
    __syncthreads();
    // row-major bytes: a thread takes 4 columns x 32 rows - four column
    // words, then one 4-byte store per row (a warp's stores of a row are 128
    // contiguous bytes); ~3 instructions per pixel
    const int w4 = (W + 3) >> 2;
    const int Hw = a.Hw;
    for (int i = threadIdx.x; i < Hw * w4; i += kDecThreads) {
      const int yb = i / w4, x0 = (i - yb * w4) * 4;
      uint32_t c0 = bits[x0 * Hp + yb];
      uint32_t c1 = x0 + 1 < W ? bits[(x0 + 1) * Hp + yb] : 0u;
      uint32_t c2 = x0 + 2 < W ? bits[(x0 + 2) * Hp + yb] : 0u;
      uint32_t c3 = x0 + 3 < W ? bits[(x0 + 3) * Hp + yb] : 0u;
      const int rows = min(32, H - yb * 32);
      uint8_t *o = out + (long long)(yb * 32) * W + x0;
      const bool vec = x0 + 4 <= W && (W & 3) == 0 &&
                       (reinterpret_cast<uintptr_t>(o) & 3) == 0;
      for (int r = 0; r < rows; ++r) {
        const uint32_t v = (c0 & 1u) | ((c1 & 1u) << 8) | ((c2 & 1u) << 16) |
                           ((c3 & 1u) << 24);
        c0 >>= 1; c1 >>= 1; c2 >>= 1; c3 >>= 1;
        if (vec) {
          *reinterpret_cast<uint32_t *>(o) = v;
        } else {
          for (int k = 0; k < 4 && x0 + k < W; ++k)
            o[k] = (uint8_t)(v >> (8 * k));
        }
        o += W;
      }
    }
    __syncthreads();
  }
}

}  // namespace mlc

using namespace mlc;

extern "C" int mlc_rle_decode(const int32_t *runs_dev, const int64_t *off_dev,
                              const int32_t *cnt_dev, int64_t n_planes, int H,
                              int W, uint8_t *mask_dev, void *stream) {
  MLC_REQUIRE(n_planes >= 0 && H >= 1 && W >= 1,
              "bad shape n_planes=%lld H=%d W=%d", (long long)n_planes, H, W);
  MLC_REQUIRE((long long)H * W < (1ll << 31) - 2, "plane of %d x %d too large",
              H, W);
  MLC_REQUIRE(off_dev && cnt_dev && mask_dev, "off/cnt/mask pointer is null");
  MLC_REQUIRE(((uintptr_t)runs_dev % 8) == 0, "runs_dev must be 8-byte aligned");
  if (n_planes == 0) return MLC_OK;
  DecArgs a;
  a.runs = runs_dev;
  a.off = reinterpret_cast<const long long *>(off_dev);
  a.cnt = cnt_dev;
  a.mask = mask_dev;
  a.n_planes = n_planes;
  a.H = H;
  a.W = W;
  a.Hw = (H + 31) / 32;
  a.Hp = a.Hw | 1;
  const size_t smem = (size_t)W * a.Hp * 4;
  a.in_smem = smem <= kDecMaxSmem;
  MLC_CUDA(cudaFuncSetAttribute(rle_decode_kernel,
                                cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)kDecMaxSmem));
  const int per_sm = a.in_smem
                         ? (int)max((size_t)1, min((size_t)4, (220 * 1024) / (smem + 1024)))
                         : 4;
  const int grid = (int)min((long long)n_planes, (long long)per_sm * sm_count());
  rle_decode_kernel<<<grid, kDecThreads, a.in_smem ? smem : 0,
                      (cudaStream_t)stream>>>(a);
  MLC_LAUNCH_CHECK("rle_decode_kernel");
  return MLC_OK;
}
