// Device helpers shared by the single-GPU (fedavg_reduce.cu) and multi-GPU
// (fedavg_allreduce.cu) forms of K-AGG.
#pragma once
#include "common.cuh"

namespace nfb {

constexpr int kAggThreads = 256;
constexpr int kAggUnroll = 8;

// Prologue: coef[k] = w_k * c_k into shared memory; returns sum_k w_k^2 (all threads).
__device__ __forceinline__ float agg_prologue(float* __restrict__ coef, int K,
                                              const float* __restrict__ counts,
                                              const float* __restrict__ sumsq,
                                              const float* __restrict__ mask, float clip_norm,
                                              bool normalise, float* red,
                                              float* total_out = nullptr) {
  float part = 0.f;
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    const float e = counts[k] * (mask ? mask[k] : 1.f);
    coef[k] = e;
    part += e;
  }
  const float total = block_sum(part, red);
  if (total_out) *total_out = total;
  const float inv = normalise ? (total > 0.f ? 1.f / total : 0.f) : 1.f;
  float w2 = 0.f;
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    const float w = coef[k] * inv;
    float c = 1.f;
    if (clip_norm > 0.f && sumsq) c = fminf(clip_norm / (sqrtf(sumsq[k]) + 1e-6f), 1.f);
    coef[k] = w * c;
    w2 += w * w;
  }
  return block_sum(w2, red);  // block_sum ends with __syncthreads(): coef[] is visible
}

// acc[0..V) = sum_k coef[k] * x[k, col .. col+V)   (col = 16-byte column group)
// NC = false: plain loads instead of the non-coherent path, for launches that write the rows they
// read (K-AGG fused with K-BCAST replicates the result into the client rows in place).
template <typename T, bool NC = true>
__device__ __forceinline__ void column_reduce(const T* x, int64_t row_stride,
                                              int K, const float* __restrict__ coef,
                                              int64_t col, float* acc) {
  constexpr int V = Vec16<T>::N;
#pragma unroll
  for (int e = 0; e < V; ++e) acc[e] = 0.f;
  const T* p = x + col;
  int k = 0;
  for (; k + kAggUnroll <= K; k += kAggUnroll) {
    uint4 r[kAggUnroll];
#pragma unroll
    for (int u = 0; u < kAggUnroll; ++u) {
      // A zero coefficient means "client absent": do not touch its row at all.
      r[u] = (coef[k + u] != 0.f) ? (NC ? ld_stream(p + (k + u) * row_stride)
                                        : ld_plain(p + (k + u) * row_stride))
                                  : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < kAggUnroll; ++u) {
      float f[V];
      unpack<T>(r[u], f);
      const float c = coef[k + u];
#pragma unroll
      for (int e = 0; e < V; ++e) acc[e] = fmaf(c, f[e], acc[e]);
    }
  }
  for (; k < K; ++k) {
    const float c = coef[k];
    if (c == 0.f) continue;
    float f[V];
    unpack<T>(NC ? ld_stream(p + k * row_stride) : ld_plain(p + k * row_stride), f);
#pragma unroll
    for (int e = 0; e < V; ++e) acc[e] = fmaf(c, f[e], acc[e]);
  }
}

}  // namespace nfb
