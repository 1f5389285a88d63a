// Report tiles on the device: the colour composite of predicted masks and the
// class outlines over an image that SegmentationReportBuilder draws with NumPy
// and OpenCV on the host (mlcomp/worker/reports/segmenation.py:71-96).
// Contract: include/mlc.h (mlc_mask_composite, mlc_mask_outline).
//
// Both are plain streaming kernels, one thread per pixel (four pixels per
// thread where alignment allows): masks are read once (the outline reads a
// 13-pixel diamond per class, served by L1), images are written once.
#include "mlc_common.cuh"

namespace mlc {

struct Palette {
  uint8_t rgb[MLC_MAX_CLASSES_SEG][3];
};

// encode_pred (segmenation.py:71-80): res += (plane * color).astype(uint8)
// per class, everything in uint8 - the product and the running sum both wrap
// modulo 256 (a mask value of 1 simply adds the colour).
template <int PX>
__global__ void __launch_bounds__(256)
mask_composite_kernel(const uint8_t *__restrict__ mask,
                      uint8_t *__restrict__ out, long long n_img, int C,
                      long long plane_px, const __grid_constant__ Palette pal) {
  const long long items = plane_px / PX;
  const long long total = n_img * items;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
       t < total; t += (long long)gridDim.x * blockDim.x) {
    const long long img = t / items;
    const long long p0 = (t - img * items) * PX;
    uint32_t acc[PX][3];
#pragma unroll
    for (int i = 0; i < PX; ++i) acc[i][0] = acc[i][1] = acc[i][2] = 0;
    for (int c = 0; c < C; ++c) {
      const uint8_t *src = mask + (img * C + c) * plane_px + p0;
      uint32_t m[PX];
      if (PX == 4) {
        const uint32_t w = *reinterpret_cast<const uint32_t *>(src);
#pragma unroll
        for (int i = 0; i < PX; ++i) m[i] = (w >> (8 * i)) & 0xFFu;
      } else {
        m[0] = src[0];
      }
#pragma unroll
      for (int i = 0; i < PX; ++i)
#pragma unroll
        for (int ch = 0; ch < 3; ++ch)
          acc[i][ch] += (m[i] * pal.rgb[c][ch]) & 0xFFu;  // uint8 product
    }
    uint8_t *dst = out + (img * plane_px + p0) * 3;
    if (PX == 4) {
      uint32_t b[12];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) b[i * 3 + ch] = acc[i][ch] & 0xFFu;
      uint32_t *d4 = reinterpret_cast<uint32_t *>(dst);
#pragma unroll
      for (int j = 0; j < 3; ++j)
        d4[j] = b[4 * j] | (b[4 * j + 1] << 8) | (b[4 * j + 2] << 16) |
                (b[4 * j + 3] << 24);
    } else {
#pragma unroll
      for (int ch = 0; ch < 3; ++ch) dst[ch] = (uint8_t)(acc[0][ch] & 0xFFu);
    }
  }
}

// plot_mask (segmenation.py:82-96).  The reference traces every border with
// cv2.findContours(RETR_LIST, CHAIN_APPROX_NONE) and then hands the POINT
// array of each contour to cv2.polylines(..., thickness 2): one degenerate
// poly-line per contour point.  Which pixels that paints (established
// against OpenCV 4.13, fixtures in tests/golden/composite.npz):
//   * contour points = mask pixels with a zero 4-neighbour (frame = zero);
//   * a thickness-2 line from a point to itself = that pixel + its four
//     4-neighbours, clipped to the image;
//   * classes in order, later over earlier.
// So pixel p takes the colour of the LAST class that has a contour point in
// the plus-shaped neighbourhood of p.
__global__ void __launch_bounds__(256)
mask_outline_kernel(uint8_t *__restrict__ img, const uint8_t *__restrict__ mask,
                    long long n_img, int C, int H, int W,
                    const __grid_constant__ Palette pal) {
  const long long plane_px = (long long)H * W;
  const long long total = n_img * plane_px;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x;
       t < total; t += (long long)gridDim.x * blockDim.x) {
    const long long im = t / plane_px;
    const int p = (int)(t - im * plane_px);
    const int y = p / W, x = p - y * W;
    int hit = -1;
    for (int c = 0; c < C; ++c) {
      const uint8_t *m = mask + (im * C + c) * plane_px;
      auto at = [&](int yy, int xx) -> bool {
        return yy >= 0 && yy < H && xx >= 0 && xx < W &&
               m[(long long)yy * W + xx] != 0;
      };
      auto border = [&](int yy, int xx) -> bool {
        return at(yy, xx) && !(at(yy - 1, xx) && at(yy + 1, xx) &&
                               at(yy, xx - 1) && at(yy, xx + 1));
      };
      if (border(y, x) || border(y - 1, x) || border(y + 1, x) ||
          border(y, x - 1) || border(y, x + 1))
        hit = c;
    }
    if (hit >= 0) {
      uint8_t *d = img + t * 3;
      d[0] = pal.rgb[hit][0];
      d[1] = pal.rgb[hit][1];
      d[2] = pal.rgb[hit][2];
    }
  }
}

}  // namespace mlc

using namespace mlc;

static int fill_palette(Palette &pal, const uint8_t *colors_host, int n_cls,
                        uint8_t r, uint8_t g, uint8_t b) {
  for (int c = 0; c < n_cls; ++c) {
    pal.rgb[c][0] = colors_host ? colors_host[3 * c] : r;
    pal.rgb[c][1] = colors_host ? colors_host[3 * c + 1] : g;
    pal.rgb[c][2] = colors_host ? colors_host[3 * c + 2] : b;
  }
  return MLC_OK;
}

extern "C" int mlc_mask_composite(const uint8_t *mask_dev, int64_t n_img,
                                  int n_cls, int H, int W,
                                  const uint8_t *colors_host, uint8_t *out_dev,
                                  void *stream) {
  MLC_REQUIRE(n_img >= 0 && n_cls >= 1 && n_cls <= MLC_MAX_CLASSES_SEG &&
                  H >= 1 && W >= 1,
              "bad shape n_img=%lld n_cls=%d H=%d W=%d", (long long)n_img,
              n_cls, H, W);
  MLC_REQUIRE((long long)H * W < (1ll << 31), "plane of %d x %d too large", H,
              W);
  if (n_img == 0) return MLC_OK;
  MLC_REQUIRE(mask_dev && out_dev, "mask_dev / out_dev is null");
  Palette pal;
  fill_palette(pal, colors_host, n_cls, 255, 255, 255);  // :75-77 default
  const long long plane_px = (long long)H * W;
  cudaStream_t st = (cudaStream_t)stream;
  const bool vec = (plane_px % 4) == 0 && ((uintptr_t)mask_dev % 4) == 0 &&
                   ((uintptr_t)out_dev % 4) == 0;
  const long long threads = n_img * (vec ? plane_px / 4 : plane_px);
  const int grid =
      (int)max(1ll, min((threads + 255) / 256, (long long)sm_count() * 32));
  if (vec)
    mask_composite_kernel<4><<<grid, 256, 0, st>>>(mask_dev, out_dev, n_img,
                                                   n_cls, plane_px, pal);
  else
    mask_composite_kernel<1><<<grid, 256, 0, st>>>(mask_dev, out_dev, n_img,
                                                   n_cls, plane_px, pal);
  MLC_LAUNCH_CHECK("mask_composite_kernel");
  return MLC_OK;
}

extern "C" int mlc_mask_outline(uint8_t *img_dev, const uint8_t *mask_dev,
                                int64_t n_img, int n_cls, int H, int W,
                                const uint8_t *colors_host, void *stream) {
  MLC_REQUIRE(n_img >= 0 && n_cls >= 1 && n_cls <= MLC_MAX_CLASSES_SEG &&
                  H >= 1 && W >= 1,
              "bad shape n_img=%lld n_cls=%d H=%d W=%d", (long long)n_img,
              n_cls, H, W);
  MLC_REQUIRE((long long)H * W < (1ll << 31), "plane of %d x %d too large", H,
              W);
  if (n_img == 0) return MLC_OK;
  MLC_REQUIRE(img_dev && mask_dev, "img_dev / mask_dev is null");
  Palette pal;
  fill_palette(pal, colors_host, n_cls, 0, 255, 0);  // :92 default
  const long long total = n_img * (long long)H * W;
  const int grid =
      (int)max(1ll, min((total + 255) / 256, (long long)sm_count() * 32));
  mask_outline_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
      img_dev, mask_dev, n_img, n_cls, H, W, pal);
  MLC_LAUNCH_CHECK("mask_outline_kernel");
  return MLC_OK;
}
