// F1: the text of the run-length strings on the device.
// mask2rle ends with ' '.join(str(x) for x in runs) (contrib/transform/
// rle.py:13); for a submission of thousands of masks that join is the last
// per-pixel-scale Python loop left.  Here every plane's (start, length) pairs
// become ASCII decimal text separated by single blanks: digit counts ->
// block scan -> digits into a shared-memory tile -> byte-coalesced store.
// Contract: include/mlc.h (mlc_rle_format).
#include "mlc_common.cuh"

namespace mlc {

constexpr int kFmtThreads = 256;
constexpr int kFmtPerThread = 8;
constexpr int kFmtTile = kFmtThreads * kFmtPerThread;   // values per tile
constexpr int kFmtTileChars = kFmtTile * 11;            // <= 10 digits + blank

struct FmtArgs {
  const int32_t *runs;
  const long long *off;
  const int32_t *cnt;
  char *chars;
  long long capacity;
  long long *str_off;
  int32_t *str_len;
  unsigned long long *total;
  long long n_planes;
};

__device__ __forceinline__ int n_digits(uint32_t v) {
  int n = 1;
  if (v >= 10u) n = 2;
  if (v >= 100u) n = 3;
  if (v >= 1000u) n = 4;
  if (v >= 10000u) n = 5;
  if (v >= 100000u) n = 6;
  if (v >= 1000000u) n = 7;
  if (v >= 10000000u) n = 8;
  if (v >= 100000000u) n = 9;
  if (v >= 1000000000u) n = 10;
  return n;
}

__device__ __forceinline__ int block_scan_excl(int v, int *warp_tot,
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
  for (int k = 0; k < kFmtThreads / 32; ++k) {
    const int t = warp_tot[k];
    if (k < w) add += t;
    all += t;
  }
  __syncthreads();
  *total = all;
  return add + inc - v;
}

__global__ void __launch_bounds__(kFmtThreads)
rle_format_kernel(const __grid_constant__ FmtArgs a) {
  __shared__ char tile[kFmtTileChars];
  __shared__ int warp_tot[kFmtThreads / 32];
  __shared__ long long s_base;
  for (long long plane = blockIdx.x; plane < a.n_planes;
       plane += gridDim.x) {
    const long long src = a.off[plane];
    const int n_val = src < 0 ? 0 : 2 * a.cnt[plane];
    const int32_t *vals = a.runs + (src < 0 ? 0 : src);
    // pass 1: characters of the whole string (each value + one blank, minus
    // the blank after the last value)
    int mine = 0;
    for (int i = threadIdx.x; i < n_val; i += kFmtThreads)
      mine += n_digits((uint32_t)vals[i]) + 1;
    int all;
    (void)block_scan_excl(mine, warp_tot, &all);
    const int len = n_val ? all - 1 : 0;
    if (threadIdx.x == 0) {
      long long base = (long long)atomicAdd(a.total, (unsigned long long)len);
      if (base + len > a.capacity) base = -1;
      s_base = base;
      a.str_off[plane] = base;
      a.str_len[plane] = len;
    }
    __syncthreads();
    long long out = s_base;
    if (out < 0 || len == 0) continue;
    // pass 2: tile by tile
    for (int t0 = 0; t0 < n_val; t0 += kFmtTile) {
      uint32_t v[kFmtPerThread];
      int nd[kFmtPerThread], chars = 0;
      const int first = t0 + threadIdx.x * kFmtPerThread;
#pragma unroll
      for (int j = 0; j < kFmtPerThread; ++j) {
        const int i = first + j;
        v[j] = i < n_val ? (uint32_t)vals[i] : 0u;
        nd[j] = i < n_val ? n_digits(v[j]) : 0;
        chars += i < n_val ? nd[j] + 1 : 0;
      }
      int tile_chars;
      int pos = block_scan_excl(chars, warp_tot, &tile_chars);
#pragma unroll
      for (int j = 0; j < kFmtPerThread; ++j) {
        if (first + j < n_val) {
          uint32_t x = v[j];
          for (int d = nd[j] - 1; d >= 0; --d) {
            tile[pos + d] = (char)('0' + x % 10u);
            x /= 10u;
          }
          tile[pos + nd[j]] = ' ';
          pos += nd[j] + 1;
        }
      }
      __syncthreads();
      // the very last blank of the string is not written
      const int n_out = (t0 + kFmtTile >= n_val) ? tile_chars - 1 : tile_chars;
      for (int i = threadIdx.x; i < n_out; i += kFmtThreads)
        a.chars[out + i] = tile[i];
      out += tile_chars;
      __syncthreads();
    }
  }
}

}  // namespace mlc

using namespace mlc;

extern "C" int mlc_rle_format(const int32_t *runs_dev, const int64_t *off_dev,
                              const int32_t *cnt_dev, int64_t n_planes,
                              char *chars_dev, int64_t capacity,
                              int64_t *str_off_dev, int32_t *str_len_dev,
                              int64_t *total_dev, void *stream) {
  MLC_REQUIRE(n_planes >= 0, "n_planes=%lld is negative", (long long)n_planes);
  MLC_REQUIRE(off_dev && cnt_dev && str_off_dev && str_len_dev && total_dev,
              "off/cnt/str_off/str_len/total pointer is null");
  MLC_REQUIRE(capacity >= 0 && (chars_dev || capacity == 0),
              "chars_dev is null with capacity %lld", (long long)capacity);
  cudaStream_t st = (cudaStream_t)stream;
  MLC_CUDA(cudaMemsetAsync(total_dev, 0, 8, st));
  if (n_planes == 0) return MLC_OK;
  FmtArgs a;
  a.runs = runs_dev;
  a.off = reinterpret_cast<const long long *>(off_dev);
  a.cnt = cnt_dev;
  a.chars = chars_dev;
  a.capacity = capacity;
  a.str_off = reinterpret_cast<long long *>(str_off_dev);
  a.str_len = str_len_dev;
  a.total = reinterpret_cast<unsigned long long *>(total_dev);
  a.n_planes = n_planes;
  const int grid = (int)(n_planes < 8ll * sm_count() ? n_planes
                                                     : 8ll * sm_count());
  rle_format_kernel<<<grid, kFmtThreads, 0, st>>>(a);
  MLC_LAUNCH_CHECK("rle_format_kernel");
  return MLC_OK;
}
