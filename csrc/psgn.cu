// K-PSGN (linear layers): ghost norms  out[i] = |a_i|^2 * |d_i|^2 (+ |d_i|^2 for the bias).
// One CTA per sample; both rows are streamed once.
#include "common.cuh"

namespace nfb {

template <typename A, typename D>
__global__ void __launch_bounds__(256)
    linear_psgn_kernel(const A* __restrict__ act, const D* __restrict__ dout, int64_t fin,
                       int64_t fout, int bias, float* __restrict__ out) {
  __shared__ float red[33];
  const int64_t i = blockIdx.x;
  float a2 = 0.f, d2 = 0.f;
  for (int64_t j = threadIdx.x; j < fin; j += blockDim.x) {
    const float v = to_f32(act[i * fin + j]);
    a2 = fmaf(v, v, a2);
  }
  for (int64_t j = threadIdx.x; j < fout; j += blockDim.x) {
    const float v = to_f32(dout[i * fout + j]);
    d2 = fmaf(v, v, d2);
  }
  a2 = block_sum(a2, red);
  d2 = block_sum(d2, red);
  if (threadIdx.x == 0) out[i] = a2 * d2 + (bias ? d2 : 0.f);
}

}  // namespace nfb

extern "C" int nfb_linear_psgn(const void* act, int act_dtype, const void* dout, int dout_dtype,
                               int64_t B, int64_t in_features, int64_t out_features, int bias,
                               float* out, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(act && dout && out, NFB_EINVAL, "linear_psgn: null pointer");
  NFB_REQUIRE(B >= 1 && B <= 2147483647LL, NFB_ERANGE, "linear_psgn: bad batch");
  cudaStream_t st = as_stream(stream);
  return dispatch_dtype(act_dtype, [&](auto ta) {
    using A = decltype(ta);
    return dispatch_dtype(dout_dtype, [&](auto td) {
      using D = decltype(td);
      linear_psgn_kernel<A, D><<<static_cast<unsigned>(B), 256, 0, st>>>(
          static_cast<const A*>(act), static_cast<const D*>(dout), in_features, out_features,
          bias, out);
      return launch_status("linear_psgn");
    });
  });
}
