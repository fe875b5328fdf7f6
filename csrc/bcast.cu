// K-BCAST (single-GPU form): global model -> every client slot, fused with the
// optimiser-state reset.  Reads the N-float master once per 16-byte group and writes it
// to K fp32 rows, K bf16 rows and zeroes K momentum rows (each optional).  Write-bound:
// K*N*(4+2+4) bytes.  The multi-GPU form (fedavg_allreduce.cu) performs the same writes
// through peer/multicast pointers directly from the aggregation kernel.
#include "common.cuh"

namespace nfb {

__global__ void __launch_bounds__(256)
    bcast_kernel(const float* __restrict__ master, int64_t N, int K, float* __restrict__ slots,
                 int64_t slots_stride, __nv_bfloat16* __restrict__ lowp, int64_t lowp_stride,
                 float* __restrict__ mom, int64_t mom_stride) {
  // blockIdx.y walks client rows in strides so the grid stays bounded for large K.
  const int64_t groups = N >> 2;
  for (int64_t g = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; g < groups;
       g += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const float4 v = *reinterpret_cast<const float4*>(master + 4 * g);
    const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y);
    const __nv_bfloat162 hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 packed;
    packed.x = *reinterpret_cast<const uint32_t*>(&lo);
    packed.y = *reinterpret_cast<const uint32_t*>(&hi);
    for (int k = blockIdx.y; k < K; k += gridDim.y) {
      if (slots) *reinterpret_cast<float4*>(slots + k * slots_stride + 4 * g) = v;
      if (lowp) *reinterpret_cast<uint2*>(lowp + k * lowp_stride + 4 * g) = packed;
      if (mom) *reinterpret_cast<float4*>(mom + k * mom_stride + 4 * g) = make_float4(0, 0, 0, 0);
    }
  }
}

}  // namespace nfb

extern "C" int nfb_bcast_global(const float* master, int64_t N, int64_t K, float* slots,
                                int64_t slots_stride, void* slots_bf16, int64_t lowp_stride,
                                float* momentum, int64_t momentum_stride, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(master, NFB_EINVAL, "bcast_global: null master");
  NFB_REQUIRE((N & 3) == 0, NFB_EINVAL, "bcast_global: N must be a multiple of 4");
  NFB_REQUIRE(K >= 0 && K <= (1 << 20), NFB_ERANGE, "bcast_global: K out of range");
  NFB_REQUIRE(aligned16(master) && (!slots || aligned16(slots)) &&
                  (!momentum || aligned16(momentum)) &&
                  (!slots_bf16 || (reinterpret_cast<uintptr_t>(slots_bf16) & 7u) == 0),
              NFB_EINVAL, "bcast_global: misaligned buffer");
  NFB_REQUIRE((slots_stride & 3) == 0 && (lowp_stride & 3) == 0 && (momentum_stride & 3) == 0,
              NFB_EINVAL, "bcast_global: strides must be multiples of 4 elements");
  if (N == 0 || K == 0 || (!slots && !slots_bf16 && !momentum)) return 0;
  const int64_t groups = N >> 2;
  int64_t bx = (groups + 255) / 256;
  const int64_t cap = static_cast<int64_t>(sm_count()) * 4;
  if (bx > cap) bx = cap;
  int by = static_cast<int>(K < 8 ? K : 8);
  dim3 grid(static_cast<unsigned>(bx), static_cast<unsigned>(by));
  bcast_kernel<<<grid, 256, 0, as_stream(stream)>>>(
      master, N, static_cast<int>(K), slots, slots_stride,
      static_cast<__nv_bfloat16*>(slots_bf16), lowp_stride, momentum, momentum_stride);
  return launch_status("bcast_global");
}
