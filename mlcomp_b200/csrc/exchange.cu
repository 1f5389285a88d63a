// The exchange step of the sharded search: every rank
// contributes ONE packed int64 buffer - the bit patterns of its fp64 dice
// partial sums, then its integer counts - the buffers are all-gathered (NCCL),
// and this kernel folds the gathered rows: fp64 values are added in rank order
// (the same order on every rank and in every rerun: bit-identical scores),
// integers are summed exactly.  One launch instead of a chain of small tensor
// ops.  Contract: include/mlc.h (mlc_combine_ranks).
#include "mlc_common.cuh"

namespace mlc {

__global__ void combine_ranks_kernel(const long long *rows, int world,
                                     long long n_f64, long long n_total,
                                     long long *out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_total) return;
  if (i < n_f64) {
    double s = __longlong_as_double(rows[i]);
    for (int r = 1; r < world; ++r)
      s += __longlong_as_double(rows[(size_t)r * n_total + i]);
    out[i] = __double_as_longlong(s);
  } else {
    long long s = 0;
    for (int r = 0; r < world; ++r) s += rows[(size_t)r * n_total + i];
    out[i] = s;
  }
}

}  // namespace mlc

using namespace mlc;

extern "C" int mlc_combine_ranks(const int64_t *gathered_dev, int world,
                                 int64_t n_f64, int64_t n_i64,
                                 int64_t *out_dev, void *stream) {
  MLC_REQUIRE(gathered_dev && out_dev, "gathered/out pointer is null");
  MLC_REQUIRE(world >= 1 && n_f64 >= 0 && n_i64 >= 0,
              "bad sizes world=%d n_f64=%lld n_i64=%lld", world,
              (long long)n_f64, (long long)n_i64);
  const long long total = n_f64 + n_i64;
  if (total == 0) return MLC_OK;
  combine_ranks_kernel<<<(unsigned)((total + 255) / 256), 256, 0,
                         (cudaStream_t)stream>>>(
      reinterpret_cast<const long long *>(gathered_dev), world, n_f64, total,
      reinterpret_cast<long long *>(out_dev));
  MLC_LAUNCH_CHECK("combine_ranks_kernel");
  return MLC_OK;
}
