// Error plumbing, device query and argument marshalling shared by all entry
// points of libmlcomp_b200.
#include "mlc_common.cuh"

#include <cstring>
#include <string>

namespace mlc {

static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char *what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return MLC_ERR_CUDA;
}

int sm_count() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) !=
            cudaSuccess ||
        n <= 0)
      n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

int fill_srcset(SrcSet &s, const void *const *ptrs, const int32_t *xform,
                int n_src, int elem_bytes, int H, int W) {
  MLC_REQUIRE(ptrs != nullptr, "src_dev_ptrs is null");
  MLC_REQUIRE(n_src >= 1 && n_src <= MLC_MAX_SOURCES,
              "n_src=%d outside [1, %d]", n_src, MLC_MAX_SOURCES);
  s.n = n_src;
  s.any_xf = 0;
  for (int i = 0; i < MLC_MAX_SOURCES; ++i) {
    s.ptr[i] = nullptr;
    s.xf[i] = 0;
  }
  for (int i = 0; i < n_src; ++i) {
    MLC_REQUIRE(ptrs[i] != nullptr, "source %d is a null pointer", i);
    MLC_REQUIRE(((uintptr_t)ptrs[i] % elem_bytes) == 0,
                "source %d is not aligned to its element size", i);
    const int xf = xform ? xform[i] : 0;
    MLC_REQUIRE(xf >= 0 && xf < 8, "source %d: xform %d is not an MLC_XF_* mask",
                i, xf);
    s.ptr[i] = ptrs[i];
    s.xf[i] = (uint8_t)xf;
    s.any_xf |= xf;
  }
  (void)H;
  (void)W;
  return MLC_OK;
}

}  // namespace mlc

extern "C" int mlc_version(void) { return 100; }  // 0.1.0

extern "C" const char *mlc_last_error(void) { return mlc::g_err; }
