// K-PSGN (linear layers): ghost norms  out[i] = |a_i|^2 * |d_i|^2 (+ |d_i|^2 for the bias).
// One CTA per sample; both rows are streamed once.  Also the batch form used by the grouped
// engine's DP-SGD mode (batch_ghost_kernel below).
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

// Ghost norm of a BATCH weight gradient:  |sum_b d_b (x) x_b|_F^2 = sum_{b,b'} (d_b . d_b') (x_b . x_b').
// gram_x [C][B][B] holds x_b . x_b' (made by the grouped tcgen05 GEMM with split-K atomics; it is
// consumed here and handed back zeroed), d is [C][B][J] bf16.  out[c] = that norm + add_a[c] + add_b[c]
// (squared norms of the client's other gradients), i.e. the complete squared gradient norm a DP-SGD
// clip needs - without the J x I weight gradient ever being written.
__global__ void __launch_bounds__(256)
    batch_ghost_kernel(float* __restrict__ gram_x, const __nv_bfloat16* __restrict__ d, int B, int J,
                       const float* __restrict__ add_a, const float* __restrict__ add_b,
                       float* __restrict__ out) {
  extern __shared__ __nv_bfloat16 sd[];   // [B][J]
  __shared__ float red[33];
  const int c = blockIdx.x;
  const __nv_bfloat16* dc = d + static_cast<int64_t>(c) * B * J;
  for (int i = threadIdx.x; i < B * J; i += blockDim.x) sd[i] = dc[i];
  __syncthreads();
  float* gx = gram_x + static_cast<int64_t>(c) * B * B;
  float acc = 0.f;
  for (int pr = threadIdx.x; pr < B * B; pr += blockDim.x) {
    const int b0 = pr / B, b1 = pr - b0 * B;
    float gd = 0.f;
    for (int j = 0; j < J; j += 2) {
      const float2 u = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(sd + b0 * J + j));
      const float2 v = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(sd + b1 * J + j));
      gd = fmaf(u.x, v.x, fmaf(u.y, v.y, gd));
    }
    acc = fmaf(gd, gx[pr], acc);
    gx[pr] = 0.f;
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0)
    out[c] = fmaxf(acc, 0.f) + (add_a ? add_a[c] : 0.f) + (add_b ? add_b[c] : 0.f);
}

}  // namespace nfb

extern "C" int nfb_batch_ghost_norm(float* gram_x, const void* d, int clients, int batch, int J,
                                    const float* add_a, const float* add_b, float* out,
                                    void* stream) {
  using namespace nfb;
  NFB_REQUIRE(gram_x && d && out, NFB_EINVAL, "batch_ghost_norm: null pointer");
  NFB_REQUIRE(clients >= 1 && batch >= 1 && J >= 2 && (J & 1) == 0 &&
                  static_cast<size_t>(batch) * J * 2 <= 48 * 1024,
              NFB_ERANGE, "batch_ghost_norm: batch x J bf16 values must fit 48 KB");
  batch_ghost_kernel<<<clients, 256, static_cast<size_t>(batch) * J * 2, as_stream(stream)>>>(
      gram_x, static_cast<const __nv_bfloat16*>(d), batch, J, add_a, add_b, out);
  return launch_status("batch_ghost_norm");
}

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
