// K-RNG: fill a buffer from the Philox stream; K-MASK: additive masking.
#include "common.cuh"
#include "philox.cuh"

namespace nfb {

// One thread = one Philox group = 4 consecutive outputs.
template <typename T, int KIND>
__global__ void __launch_bounds__(256) philox_fill_kernel(T* __restrict__ out, int64_t n,
                                                          float scale, uint64_t seed,
                                                          uint64_t group0) {
  const int64_t groups = (n + 3) >> 2;
  for (int64_t g = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; g < groups;
       g += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    float4 z;
    if (KIND == 0) z = Philox::normal4(group0 + g, seed);
    else if (KIND == 1) z = Philox::laplace4(group0 + g, seed);
    else z = Philox::uniform4(group0 + g, seed);
    const float v[4] = {z.x * scale, z.y * scale, z.z * scale, z.w * scale};
    const int64_t base = g << 2;
    if (base + 4 <= n && sizeof(T) == 4) {
      *reinterpret_cast<float4*>(reinterpret_cast<float*>(out) + base) =
          make_float4(v[0], v[1], v[2], v[3]);
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if (base + i < n) out[base + i] = from_f32<T>(v[i]);
    }
  }
}

__global__ void __launch_bounds__(256) mask_add_kernel(float* __restrict__ x,
                                                       float* __restrict__ cum, int64_t n,
                                                       uint64_t seed, uint64_t group0) {
  const int64_t groups = (n + 3) >> 2;
  for (int64_t g = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; g < groups;
       g += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const float4 m = Philox::uniform4(group0 + g, seed);
    const float v[4] = {m.x, m.y, m.z, m.w};
    const int64_t base = g << 2;
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (base + i < n) {
        x[base + i] += v[i];
        cum[base + i] += v[i];
      }
  }
}

static int grid_for(int64_t work_items, int block) {
  int64_t blocks = (work_items + block - 1) / block;
  const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

}  // namespace nfb

extern "C" int nfb_philox_fill(void* out, int out_dtype, int64_t n, int kind, float scale,
                               uint64_t seed, uint64_t offset, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(out != nullptr || n == 0, NFB_EINVAL, "philox_fill: null output");
  NFB_REQUIRE(n >= 0, NFB_EINVAL, "philox_fill: negative length");
  NFB_REQUIRE((offset & 3u) == 0, NFB_EINVAL, "philox_fill: offset must be a multiple of 4");
  NFB_REQUIRE(kind >= 0 && kind <= 2, NFB_EINVAL, "philox_fill: kind must be 0,1,2");
  if (n == 0) return 0;
  NFB_REQUIRE(out_dtype != NFB_DT_F32 || aligned16(out), NFB_EINVAL,
              "philox_fill: fp32 output must be 16-byte aligned");
  const int grid = grid_for((n + 3) / 4, 256);
  const uint64_t g0 = offset >> 2;
  cudaStream_t s = as_stream(stream);
  return dispatch_dtype(out_dtype, [&](auto tag) {
    using T = decltype(tag);
    T* o = static_cast<T*>(out);
    if (kind == 0) philox_fill_kernel<T, 0><<<grid, 256, 0, s>>>(o, n, scale, seed, g0);
    else if (kind == 1) philox_fill_kernel<T, 1><<<grid, 256, 0, s>>>(o, n, scale, seed, g0);
    else philox_fill_kernel<T, 2><<<grid, 256, 0, s>>>(o, n, scale, seed, g0);
    return launch_status("philox_fill");
  });
}

extern "C" int nfb_mask_add(float* x, float* cumulative, int64_t n, uint64_t seed,
                            uint64_t offset, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(x && cumulative, NFB_EINVAL, "mask_add: null pointer");
  NFB_REQUIRE((offset & 3u) == 0, NFB_EINVAL, "mask_add: offset must be a multiple of 4");
  if (n <= 0) return 0;
  mask_add_kernel<<<grid_for((n + 3) / 4, 256), 256, 0, as_stream(stream)>>>(
      x, cumulative, n, seed, offset >> 2);
  return launch_status("mask_add");
}
