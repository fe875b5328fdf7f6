// K-AGG (single-GPU form): fused sample-weighted FedAvg reduction.
//
//   out[j] (+)= sum_k w_k c_k x[k,j] + sigma' z[j] m[j] ;  out_lowp[j] = cast(out[j])
//
// One launch covers what the reference does as upload + list->tensor rebuild +
// K*|keys| mul/add_ pairs + K clip passes + K*|keys| noise tensors (see the header).
// The per-client scalars (weights from raw sample counts and the participation mask,
// DP clip factors from the K-STATS sums of squares, the folded noise std) are computed
// by every CTA in a short prologue from device memory, so no host value — and no host
// sync — sits between local training and aggregation.
//
// Bandwidth shape: reads K*N*sizeof(T) once, writes N*4 (+N*2).  Each thread owns one
// 16-byte column group and walks the K rows with UNROLL independent streaming loads in
// flight (rows are far apart in memory, so the loads hit different DRAM pages/channels).
// The multi-GPU form (csrc/fedavg_allreduce.cu) reuses `column_reduce` for its local
// phase and then reduces the per-GPU partials over NVLink from inside the same kernel.
#include "common.cuh"
#include "fedavg_device.cuh"
#include "philox.cuh"

namespace nfb {

struct AggEpilogue {
  float noise_std;            // base std; multiplied by sqrt(sum w^2) in-kernel
  const uint8_t* noise_mask;  // optional [N]
  const float* unit_noise;    // optional [N] (oracle-injected z)
  uint64_t seed, group0;
  float* out;
  void* out_lowp;
  int lowp_dtype;
  bool accumulate;
  nfb_bcast_target bc;        // rows > 0: K-BCAST fused - replicate the result into the client rows
};

// Noise + accumulate + store for V consecutive outputs starting at col (col % 4 == 0).
template <int V>
__device__ __forceinline__ void agg_store(const AggEpilogue& ep, float sigma, int64_t col,
                                          float* acc) {
  if (sigma > 0.f) {
#pragma unroll
    for (int q = 0; q < V / 4; ++q) {
      const int64_t c4 = col + 4 * q;
      float z[4];
      if (ep.unit_noise) {
        const float4 u = *reinterpret_cast<const float4*>(ep.unit_noise + c4);
        z[0] = u.x; z[1] = u.y; z[2] = u.z; z[3] = u.w;
      } else {
        const float4 u = Philox::normal4(ep.group0 + (c4 >> 2), ep.seed);
        z[0] = u.x; z[1] = u.y; z[2] = u.z; z[3] = u.w;
      }
      if (ep.noise_mask) {
        const uchar4 m = *reinterpret_cast<const uchar4*>(ep.noise_mask + c4);
        z[0] *= m.x ? 1.f : 0.f; z[1] *= m.y ? 1.f : 0.f;
        z[2] *= m.z ? 1.f : 0.f; z[3] *= m.w ? 1.f : 0.f;
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[4 * q + e] = fmaf(sigma, z[e], acc[4 * q + e]);
    }
  }
#pragma unroll
  for (int q = 0; q < V / 4; ++q) {
    float4* o = reinterpret_cast<float4*>(ep.out + col + 4 * q);
    if (ep.accumulate) {
      const float4 prev = *o;
      acc[4 * q + 0] += prev.x; acc[4 * q + 1] += prev.y;
      acc[4 * q + 2] += prev.z; acc[4 * q + 3] += prev.w;
    }
    *o = make_float4(acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]);
  }
  if (ep.bc.rows > 0) {
    // K-BCAST in the same pass: the finished columns go straight into every client's fp32 row and
    // bf16 working copy, and the optimiser state of the next round starts at zero.  The rows may
    // be the very rows this thread has just reduced (same column, same thread: no hazard).
#pragma unroll
    for (int q = 0; q < V / 4; ++q) {
      const float4 v = make_float4(acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]);
      const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
      uint2 pk;
      pk.x = *reinterpret_cast<const uint32_t*>(&lo);
      pk.y = *reinterpret_cast<const uint32_t*>(&hi);
      const int64_t c4 = col + 4 * q;
      for (int64_t k = 0; k < ep.bc.rows; ++k) {
        if (ep.bc.slots) *reinterpret_cast<float4*>(ep.bc.slots + k * ep.bc.slots_stride + c4) = v;
        if (ep.bc.lowp)
          *reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(ep.bc.lowp) + k * ep.bc.lowp_stride + c4) = pk;
        if (ep.bc.momentum)
          *reinterpret_cast<float4*>(ep.bc.momentum + k * ep.bc.momentum_stride + c4) =
              make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
  }
  if (ep.out_lowp) {
    if (ep.lowp_dtype == NFB_DT_BF16) {
      __nv_bfloat16* o = static_cast<__nv_bfloat16*>(ep.out_lowp) + col;
#pragma unroll
      for (int e = 0; e < V; e += 2)
        *reinterpret_cast<__nv_bfloat162*>(o + e) = __floats2bfloat162_rn(acc[e], acc[e + 1]);
    } else {
      __half* o = static_cast<__half*>(ep.out_lowp) + col;
#pragma unroll
      for (int e = 0; e < V; e += 2)
        *reinterpret_cast<__half2*>(o + e) = __floats2half2_rn(acc[e], acc[e + 1]);
    }
  }
}

template <typename T, bool BCAST>
__global__ void __launch_bounds__(kAggThreads)
    fedavg_reduce_kernel(const T* x, int K, int64_t N, int64_t row_stride,
                         const float* __restrict__ counts, const float* __restrict__ sumsq,
                         const float* __restrict__ mask, float clip_norm, bool normalise,
                         AggEpilogue ep) {
  extern __shared__ float coef[];  // [K]
  __shared__ float red[33];
  constexpr int V = Vec16<T>::N;
  float total = 0.f;
  const float w2 = agg_prologue(coef, K, counts, sumsq, mask, clip_norm, normalise, red, &total);
  // every client masked out (dropped, screened as non-finite, all stragglers): there is nothing
  // to average - keep the previous global model instead of overwriting it with zeros (with the
  // fused broadcast the unchanged model still has to reach the client rows)
  const bool keep = normalise && !(total > 0.f);
  if (keep && !BCAST) return;
  const float sigma = (ep.noise_std > 0.f && !keep) ? ep.noise_std * sqrtf(w2) : 0.f;

  const int64_t groups = N / V;  // N % V == 0 enforced by the host wrapper
  for (int64_t g = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; g < groups;
       g += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    float acc[V];
    if (BCAST && keep) {
#pragma unroll
      for (int e = 0; e < V; ++e) acc[e] = ep.out[g * V + e];
    } else {
      column_reduce<T, !BCAST>(x, row_stride, K, coef, g * V, acc);
    }
    agg_store<V>(ep, sigma, g * V, acc);
  }
}

}  // namespace nfb

extern "C" int nfb_fedavg_reduce(const void* x, int x_dtype, int64_t K, int64_t N,
                                 int64_t row_stride, const float* counts, const float* sumsq,
                                 const float* mask, float clip_norm, float noise_std,
                                 const uint8_t* noise_mask, const float* unit_noise,
                                 uint64_t seed, uint64_t offset, float* out, void* out_lowp,
                                 int lowp_dtype, int flags, void* reserved, void* stream) {
  using namespace nfb;
  const nfb_bcast_target* bc = static_cast<const nfb_bcast_target*>(reserved);
  NFB_REQUIRE(x && counts && out, NFB_EINVAL, "fedavg_reduce: null pointer");
  NFB_REQUIRE(K >= 1 && K <= 12000, NFB_ERANGE, "fedavg_reduce: K=%lld outside [1, 12000]",
              static_cast<long long>(K));
  NFB_REQUIRE(N >= 0 && row_stride >= N, NFB_EINVAL, "fedavg_reduce: bad N/row_stride");
  NFB_REQUIRE((N & 7) == 0 && (row_stride & 7) == 0, NFB_EINVAL,
              "fedavg_reduce: N and row_stride must be multiples of 8 elements (FlatLayout pads "
              "to 64)");
  NFB_REQUIRE(aligned16(x) && aligned16(out) && (!out_lowp || aligned16(out_lowp)) &&
                  (!unit_noise || aligned16(unit_noise)) &&
                  (!noise_mask || (reinterpret_cast<uintptr_t>(noise_mask) & 3u) == 0),
              NFB_EINVAL, "fedavg_reduce: buffers must be 16-byte aligned");
  NFB_REQUIRE((offset & 3u) == 0, NFB_EINVAL, "fedavg_reduce: offset must be a multiple of 4");
  NFB_REQUIRE(!out_lowp || lowp_dtype == NFB_DT_BF16 || lowp_dtype == NFB_DT_F16, NFB_EINVAL,
              "fedavg_reduce: low-precision copy must be bf16 or fp16");
  if (N == 0) return 0;

  AggEpilogue ep;
  ep.noise_std = noise_std;
  ep.noise_mask = noise_mask;
  ep.unit_noise = unit_noise;
  ep.seed = seed;
  ep.group0 = offset >> 2;
  ep.out = out;
  ep.out_lowp = out_lowp;
  ep.lowp_dtype = lowp_dtype;
  ep.accumulate = (flags & 2) != 0;
  ep.bc = nfb_bcast_target{};
  if (bc && bc->rows > 0) {
    NFB_REQUIRE(!ep.accumulate, NFB_EINVAL, "fedavg_reduce: a partial sum cannot be broadcast");
    NFB_REQUIRE((!bc->slots || aligned16(bc->slots)) && (!bc->momentum || aligned16(bc->momentum)) &&
                    (!bc->lowp || (reinterpret_cast<uintptr_t>(bc->lowp) & 7u) == 0) &&
                    (bc->slots_stride & 7) == 0 && (bc->lowp_stride & 7) == 0 &&
                    (bc->momentum_stride & 7) == 0,
                NFB_EINVAL, "fedavg_reduce: misaligned broadcast target");
    ep.bc = *bc;
  }
  const bool do_bcast = ep.bc.rows > 0;
  const bool normalise = (flags & 1) != 0;
  cudaStream_t st = as_stream(stream);
  const size_t smem = static_cast<size_t>(K) * sizeof(float);

  return dispatch_dtype(x_dtype, [&](auto tag) {
    using T = decltype(tag);
    constexpr int V = Vec16<T>::N;
    const int64_t groups = N / V;
    int64_t blocks = (groups + kAggThreads - 1) / kAggThreads;
    const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
    if (blocks > cap) blocks = cap;
    if (do_bcast)
      fedavg_reduce_kernel<T, true><<<static_cast<int>(blocks), kAggThreads, smem, st>>>(
          static_cast<const T*>(x), static_cast<int>(K), N, row_stride, counts, sumsq, mask,
          clip_norm, normalise, ep);
    else
      fedavg_reduce_kernel<T, false><<<static_cast<int>(blocks), kAggThreads, smem, st>>>(
          static_cast<const T*>(x), static_cast<int>(K), N, row_stride, counts, sumsq, mask,
          clip_norm, normalise, ep);
    return launch_status("fedavg_reduce");
  });
}
