// K-XENT: log-softmax + NLL + accuracy + dlogits for G groups of [B, C] logits.
// One warp per sample row (C is small: 10 for MNIST/CIFAR); one CTA per group slice.
// Loss/accuracy are reduced per group deterministically: a group is handled by ONE CTA,
// which loops over its B rows, so no cross-CTA atomics are needed.
#include "common.cuh"

namespace nfb {

template <typename T, typename D>
__global__ void __launch_bounds__(256)
    xent_kernel(const T* __restrict__ logits, const void* __restrict__ targets, int targets_i64,
                int B, int C, D* __restrict__ dlogits, float grad_scale,
                float* __restrict__ loss_sum, float* __restrict__ correct,
                float* __restrict__ stats) {
  __shared__ float s_loss[8], s_hit[8];
  const int64_t g = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const T* lg = logits + g * static_cast<int64_t>(B) * C;
  D* dl = dlogits + g * static_cast<int64_t>(B) * C;
  float w_loss = 0.f, w_hit = 0.f;
  for (int b = warp; b < B; b += nwarps) {
    const T* row = lg + static_cast<int64_t>(b) * C;
    const int64_t t = targets_i64 ? static_cast<const int64_t*>(targets)[g * B + b]
                                  : static_cast<const int*>(targets)[g * B + b];
    // pass 1: max and argmax (first index wins ties, like torch.argmax)
    float mx = -INFINITY;
    int arg = 0x7fffffff;
    for (int c = lane; c < C; c += 32) {
      const float v = to_f32(row[c]);
      if (v > mx) { mx = v; arg = c; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (om > mx || (om == mx && oa < arg)) { mx = om; arg = oa; }
    }
    // pass 2: sum exp
    float se = 0.f;
    for (int c = lane; c < C; c += 32) se += __expf(to_f32(row[c]) - mx);
    se = warp_sum(se);
    const float lse = mx + __logf(se);
    // pass 3: gradient
    const float inv = 1.f / se;
    for (int c = lane; c < C; c += 32) {
      const float pr = __expf(to_f32(row[c]) - mx) * inv;
      dl[static_cast<int64_t>(b) * C + c] = from_f32<D>((pr - (c == t ? 1.f : 0.f)) * grad_scale);
    }
    if (lane == 0) {
      const float zt = (t >= 0 && t < C) ? to_f32(row[t]) : 0.f;
      w_loss += lse - zt;
      w_hit += (arg == t) ? 1.f : 0.f;
    }
  }
  if (lane == 0) { s_loss[warp] = w_loss; s_hit[warp] = w_hit; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float l = 0.f, h = 0.f;
    for (int w = 0; w < nwarps; ++w) { l += s_loss[w]; h += s_hit[w]; }
    if (loss_sum) loss_sum[g] = l;
    if (correct) correct[g] = h;
    if (stats) {
      // One CTA owns group g: plain read-modify-write, stream order serialises launches.
      stats[3 * g + 0] += l;
      stats[3 * g + 1] += h;
      stats[3 * g + 2] += static_cast<float>(B);
    }
  }
}

}  // namespace nfb

extern "C" int nfb_xent_fused(const void* logits, int dtype, const void* targets,
                              int targets_i64, int64_t G, int B, int C, void* dlogits,
                              int dlogits_dtype, float grad_scale, float* loss_sum,
                              float* correct, float* stats, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(logits && targets && dlogits, NFB_EINVAL, "xent_fused: null pointer");
  NFB_REQUIRE(G >= 1 && G <= 2147483647LL && B >= 1 && C >= 1, NFB_ERANGE,
              "xent_fused: bad G/B/C");
  cudaStream_t st = as_stream(stream);
  const unsigned grid = static_cast<unsigned>(G);
  return dispatch_dtype(dtype, [&](auto tin) {
    using T = decltype(tin);
    return dispatch_dtype(dlogits_dtype, [&](auto tout) {
      using D = decltype(tout);
      xent_kernel<T, D><<<grid, 256, 0, st>>>(static_cast<const T*>(logits), targets,
                                              targets_i64, B, C, static_cast<D*>(dlogits),
                                              grad_scale, loss_sum, correct, stats);
      return launch_status("xent_fused");
    });
  });
}
