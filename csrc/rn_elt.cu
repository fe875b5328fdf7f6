// K-ELT for the ResNet cohort engine: everything around the tcgen05 convolutions of
// csrc/rn_conv.cu, each as ONE launch for all clients of the GPU (client = blockIdx.y):
//
//   im2col3      stem input [3, 32, 32] fp32 -> [32*32, 64] bf16 patch rows (27 taps + zero pad)
//   bn_apply     y = relu(gamma * (x - mean) * rstd + beta [+ residual | + bn2(residual)]) from the
//                batch sums the conv epilogue accumulated; also the running-statistics update
//                (momentum 0.1, unbiased variance, num_batches_tracked += 1) in the client's row
//   bn_bwd_red   dbeta = sum g, dgamma = sum g * xhat with g = dout * (out > 0)  (ReLU mask), for
//                the main branch and optionally the projection shortcut's BatchNorm at once
//   bn_bwd_apply dx = gamma * rstd * (g - dbeta / n - xhat * dgamma / n)
//   head         global average pool -> fc -> log-softmax / NLL / accuracy -> dlogits -> fc
//                gradients -> gradient of the pooled features broadcast back over the 4x4 grid and
//                masked by the ReLU of the feature map (gradient w.r.t. the pre-ReLU block output)
//   wm_convert   conv weights between the arena's torch layout [co][ci][kh][kw] and the engine's
//                tap-major fp32 master + bf16 working copy [tap][co][ci] (once per round each way)
//   tap_sgd      SGD (momentum, weight decay) on tap-major ranges whose gradient the wgrad kernel
//                accumulated with atomics; the other convolutions are updated in the wgrad epilogue
//   small_sgd    SGD for BatchNorm weight/bias and the classifier; also re-zeroes the batch sums
//
// No reference counterpart: these are the torch library calls (BatchNorm2d, ReLU, AdaptiveAvgPool2d,
// Linear, CrossEntropyLoss, SGD.step) that a ResNet runs inside nanofed/trainer/base.py:134-160.
#include "common.cuh"
#include "philox.cuh"

namespace nfb {
namespace rn {

struct BnRef {               // one BatchNorm layer of one launch (offsets into a client's fp32 row)
  const float* stats;        // [clients][2][C]: sum, sum of squares of the conv output
  long long gamma, beta, rmean, rvar, nbt;
};

__device__ __forceinline__ void bn_coeffs(const BnRef& bn, const float* row, int g, int C, int c,
                                          float inv_n, float eps, float& mean, float& rstd) {
  const float s = bn.stats[(static_cast<long long>(g) * 2) * C + c];
  const float q = bn.stats[(static_cast<long long>(g) * 2 + 1) * C + c];
  mean = s * inv_n;
  const float var = fmaxf(q * inv_n - mean * mean, 0.f);
  rstd = rsqrtf(var + eps);
}

// ---- stem im2col ------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    im2col3_kernel(const float* __restrict__ x, long long x_gs, int B, __nv_bfloat16* __restrict__ out) {
  pdl_trigger();
  pdl_wait();      // programmatic dependent launch: see common.cuh
  // one thread = one position x 8 patch entries; entry k = (kh*3 + kw)*3 + ci.  A thread keeps its
  // entry octet for the whole launch (32 positions per block trip), so the tap / channel of its 8
  // entries is computed once; octets 4..7 are the zero padding of K = 27 -> 64.
  const int g = blockIdx.y;
  const int oct = threadIdx.x & 7;
  int coff[8], dh[8], dw[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int k = oct * 8 + e;
    const int ci = k % 3, tap = k / 3;
    coff[e] = k < 27 ? ci * 1024 : -1;
    dh[e] = tap / 3 - 1;
    dw[e] = tap % 3 - 1;
  }
  const long long positions = static_cast<long long>(B) * 1024;
  for (long long pos = blockIdx.x * 32ll + (threadIdx.x >> 3); pos < positions; pos += gridDim.x * 32ll) {
    const int b = static_cast<int>(pos >> 10), h = static_cast<int>((pos >> 5) & 31),
              w = static_cast<int>(pos & 31);
    const float* img = x + g * x_gs + static_cast<long long>(b) * 3072;
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int hh = h + dh[e], ww = w + dw[e];
      const bool in = coff[e] >= 0 && hh >= 0 && hh < 32 && ww >= 0 && ww < 32;
      v[e] = in ? __ldg(img + coff[e] + hh * 32 + ww) : 0.f;
    }
    st_vec(out + ((static_cast<long long>(g) * positions + pos) * 64 + oct * 8), pack<__nv_bfloat16>(v));
  }
}

// ---- BatchNorm forward / backward ------------------------------------------------------------
// Thread layout shared by the three kernels: a thread owns ONE channel octet (16 bytes) for the
// whole launch, threads of a block cover 256 / (C / 8) consecutive positions; every per-channel
// coefficient therefore lives in registers, the loop body is loads -> FMAs -> store with no index
// arithmetic beyond one pointer increment, and U positions are in flight per thread (the kernels
// are latency-bound otherwise: 3 loads per 64 bytes of traffic).
constexpr int BN_U = 4;

struct BnApplyArgs {
  const __nv_bfloat16* x;     // raw conv output [clients][P][C]
  const __nv_bfloat16* res;   // optional residual (already activated input, or a raw shortcut conv)
  __nv_bfloat16* y;
  float* slots;               // client rows (parameters + running statistics)
  long long row_stride;
  BnRef bn, bn_res;           // bn_res.stats == null: residual is added as is
  int P, C, relu;
  float eps, momentum;
};

template <bool RES, bool RES_BN>
__global__ void __launch_bounds__(256) bn_apply_kernel(const BnApplyArgs a) {
  pdl_trigger();
  pdl_wait();      // programmatic dependent launch: see common.cuh
  const int g = blockIdx.y, C = a.C, octs = C >> 3;
  const int oct = threadIdx.x % octs, r = threadIdx.x / octs, R = 256 / octs, c0 = oct * 8;
  float* row = a.slots + g * a.row_stride;
  const float inv_n = 1.f / static_cast<float>(a.P);
  float sc[8], sh[8], sc2[8], sh2[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int c = c0 + e;
    float mean, rstd;
    bn_coeffs(a.bn, row, g, C, c, inv_n, a.eps, mean, rstd);
    sc[e] = row[a.bn.gamma + c] * rstd;
    sh[e] = row[a.bn.beta + c] - mean * sc[e];
    float mean2 = 0.f, rstd2 = 1.f;
    sc2[e] = 1.f;
    sh2[e] = 0.f;
    if constexpr (RES_BN) {
      bn_coeffs(a.bn_res, row, g, C, c, inv_n, a.eps, mean2, rstd2);
      sc2[e] = row[a.bn_res.gamma + c] * rstd2;
      sh2[e] = row[a.bn_res.beta + c] - mean2 * sc2[e];
    }
    if (blockIdx.x == 0 && r == 0) {
      // torch.nn.BatchNorm2d in training mode: running = (1 - m) * running + m * batch, with the
      // unbiased batch variance
      const float unb = static_cast<float>(a.P) / static_cast<float>(a.P - 1);
      const float var = 1.f / (rstd * rstd) - a.eps;
      row[a.bn.rmean + c] = (1.f - a.momentum) * row[a.bn.rmean + c] + a.momentum * mean;
      row[a.bn.rvar + c] = (1.f - a.momentum) * row[a.bn.rvar + c] + a.momentum * var * unb;
      if (c == 0) row[a.bn.nbt] += 1.f;
      if constexpr (RES_BN) {
        const float var2 = 1.f / (rstd2 * rstd2) - a.eps;
        row[a.bn_res.rmean + c] = (1.f - a.momentum) * row[a.bn_res.rmean + c] + a.momentum * mean2;
        row[a.bn_res.rvar + c] = (1.f - a.momentum) * row[a.bn_res.rvar + c] + a.momentum * var2 * unb;
        if (c == 0) row[a.bn_res.nbt] += 1.f;
      }
    }
  }
  const long long base = static_cast<long long>(g) * a.P * C + c0;
  for (int p0 = blockIdx.x * R * BN_U; p0 < a.P; p0 += gridDim.x * R * BN_U) {
    uint4 xv[BN_U], rv[BN_U];
#pragma unroll
    for (int u = 0; u < BN_U; ++u) {
      const int p = p0 + u * R + r;
      if (p < a.P) {
        xv[u] = ld_stream(a.x + base + static_cast<long long>(p) * C);
        if constexpr (RES) rv[u] = ld_stream(a.res + base + static_cast<long long>(p) * C);
      }
    }
#pragma unroll
    for (int u = 0; u < BN_U; ++u) {
      const int p = p0 + u * R + r;
      if (p >= a.P) continue;
      float v[8], rr[8];
      unpack<__nv_bfloat16>(xv[u], v);
      if constexpr (RES) unpack<__nv_bfloat16>(rv[u], rr);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        float o = fmaf(v[e], sc[e], sh[e]);
        if constexpr (RES) o += RES_BN ? fmaf(rr[e], sc2[e], sh2[e]) : rr[e];
        v[e] = a.relu ? fmaxf(o, 0.f) : o;
      }
      st_vec(a.y + base + static_cast<long long>(p) * C, pack<__nv_bfloat16>(v));
    }
  }
}

struct BnBwdArgs {
  const __nv_bfloat16* dout;   // gradient w.r.t. the activated output [clients][P][C]
  const __nv_bfloat16* out;    // activated output (ReLU mask), or null
  const __nv_bfloat16* x;      // raw conv output of the main branch
  const __nv_bfloat16* x2;     // raw conv output of the projection shortcut, or null
  __nv_bfloat16* dx;           // gradient w.r.t. x   (apply pass)
  __nv_bfloat16* dx2;          // gradient w.r.t. x2  (apply pass)
  const float* slots;
  long long row_stride;
  float* sgrad;                // compact gradient rows [clients][sgrad_stride]
  long long sgrad_stride;
  long long dgamma, dbeta, dgamma2, dbeta2;   // offsets into a sgrad row
  BnRef bn, bn2;
  int P, C;
  float eps;
  // apply pass without a reduce pass: the dgrad kernel that produced dout left sum(g), sum(g * x)
  // per channel here ([clients][2][C]); dgamma = rstd * (sum gx - mean * sum g), dbeta = sum g are
  // derived on the fly and written to the gradient row by the first block
  const float* raw_sums;
};

// g = dout where the activated output is positive (bf16 > 0  <=>  sign clear and magnitude != 0)
__device__ __forceinline__ void relu_mask(uint4& d, const uint4& o) {
  auto m = [](uint32_t dv, uint32_t ov) {
    const uint32_t lo = ((ov & 0x8000u) == 0u && (ov & 0x7fffu) != 0u) ? 0x0000ffffu : 0u;
    const uint32_t hi = ((ov & 0x80000000u) == 0u && (ov & 0x7fff0000u) != 0u) ? 0xffff0000u : 0u;
    return dv & (lo | hi);
  };
  d.x = m(d.x, o.x); d.y = m(d.y, o.y); d.z = m(d.z, o.z); d.w = m(d.w, o.w);
}

template <bool TWO>
__global__ void __launch_bounds__(256) bn_bwd_reduce_kernel(const BnBwdArgs a) {
  pdl_trigger();
  pdl_wait();      // programmatic dependent launch: see common.cuh
  __shared__ float scratch[256 * 8];
  const int g = blockIdx.y, C = a.C, octs = C >> 3;
  const int oct = threadIdx.x % octs, r = threadIdx.x / octs, R = 256 / octs, c0 = oct * 8;
  const float* row = a.slots + g * a.row_stride;
  const float inv_n = 1.f / static_cast<float>(a.P);
  float rs[8], mrs[8], rs2[8], mrs2[8];        // xhat = x * rs - mrs
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    float mean, rstd;
    bn_coeffs(a.bn, row, g, C, c0 + e, inv_n, a.eps, mean, rstd);
    rs[e] = rstd; mrs[e] = mean * rstd;
    rs2[e] = mrs2[e] = 0.f;
    if constexpr (TWO) {
      bn_coeffs(a.bn2, row, g, C, c0 + e, inv_n, a.eps, mean, rstd);
      rs2[e] = rstd; mrs2[e] = mean * rstd;
    }
  }
  float sg[8], sgx[8], sgx2[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) sg[e] = sgx[e] = sgx2[e] = 0.f;
  const long long base = static_cast<long long>(g) * a.P * C + c0;
  for (int p0 = blockIdx.x * R * BN_U; p0 < a.P; p0 += gridDim.x * R * BN_U) {
    uint4 dv[BN_U], xv[BN_U], x2v[BN_U];
#pragma unroll
    for (int u = 0; u < BN_U; ++u) {
      const int p = p0 + u * R + r;
      dv[u] = xv[u] = x2v[u] = make_uint4(0, 0, 0, 0);
      if (p < a.P) {
        const long long off = base + static_cast<long long>(p) * C;
        dv[u] = ld_stream(a.dout + off);
        if (a.out) relu_mask(dv[u], ld_stream(a.out + off));
        xv[u] = ld_stream(a.x + off);
        if constexpr (TWO) x2v[u] = ld_stream(a.x2 + off);
      }
    }
#pragma unroll
    for (int u = 0; u < BN_U; ++u) {
      float d[8], xx[8];
      unpack<__nv_bfloat16>(dv[u], d);      // rows past the end hold zeros: they add nothing
      unpack<__nv_bfloat16>(xv[u], xx);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        sg[e] += d[e];
        sgx[e] = fmaf(d[e], fmaf(xx[e], rs[e], -mrs[e]), sgx[e]);
      }
      if constexpr (TWO) {
        unpack<__nv_bfloat16>(x2v[u], xx);
#pragma unroll
        for (int e = 0; e < 8; ++e) sgx2[e] = fmaf(d[e], fmaf(xx[e], rs2[e], -mrs2[e]), sgx2[e]);
      }
    }
  }
  // reduce over the R threads that share a channel octet, one fp32 atomic per channel and block
  float* grow = a.sgrad + g * a.sgrad_stride;
  for (int pass = 0; pass < (TWO ? 3 : 2); ++pass) {
    const float* src = pass == 0 ? sg : pass == 1 ? sgx : sgx2;
    __syncthreads();
#pragma unroll
    for (int e = 0; e < 8; ++e) scratch[threadIdx.x * 8 + e] = src[e];
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += 256) {
      float t = 0.f;
      for (int rr = 0; rr < R; ++rr) t += scratch[(rr * octs + (c >> 3)) * 8 + (c & 7)];
      if (pass == 0) {
        atomicAdd(grow + a.dbeta + c, t);
        if (TWO) atomicAdd(grow + a.dbeta2 + c, t);
      } else if (pass == 1) {
        atomicAdd(grow + a.dgamma + c, t);
      } else {
        atomicAdd(grow + a.dgamma2 + c, t);
      }
    }
  }
}

template <bool TWO>
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const BnBwdArgs a) {
  pdl_trigger();
  pdl_wait();      // programmatic dependent launch: see common.cuh
  const int g = blockIdx.y, C = a.C, octs = C >> 3;
  const int oct = threadIdx.x % octs, r = threadIdx.x / octs, R = 256 / octs, c0 = oct * 8;
  const float* row = a.slots + g * a.row_stride;
  const float* grow = a.sgrad + g * a.sgrad_stride;
  const float inv_n = 1.f / static_cast<float>(a.P);
  // dx = s * (g - dbeta / n - xhat * dgamma / n) = ka * g + kb * x + kc,  s = gamma * rstd
  float ka[8], kb[8], kc[8], ka2[8], kb2[8], kc2[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int c = c0 + e;
    float mean, rstd;
    bn_coeffs(a.bn, row, g, C, c, inv_n, a.eps, mean, rstd);
    const float s = row[a.bn.gamma + c] * rstd;
    float dgamma, dbeta;
    if (a.raw_sums) {
      dbeta = a.raw_sums[(static_cast<long long>(g) * 2) * C + c];
      dgamma = rstd * (a.raw_sums[(static_cast<long long>(g) * 2 + 1) * C + c] - mean * dbeta);
      if (blockIdx.x == 0 && r == 0) {
        a.sgrad[g * a.sgrad_stride + a.dgamma + c] = dgamma;
        a.sgrad[g * a.sgrad_stride + a.dbeta + c] = dbeta;
      }
    } else {
      dgamma = grow[a.dgamma + c];
      dbeta = grow[a.dbeta + c];
    }
    const float t = s * dgamma * inv_n * rstd;
    ka[e] = s; kb[e] = -t; kc[e] = t * mean - s * dbeta * inv_n;
    ka2[e] = kb2[e] = kc2[e] = 0.f;
    if constexpr (TWO) {
      bn_coeffs(a.bn2, row, g, C, c, inv_n, a.eps, mean, rstd);
      const float s2 = row[a.bn2.gamma + c] * rstd;
      const float t2 = s2 * grow[a.dgamma2 + c] * inv_n * rstd;
      ka2[e] = s2; kb2[e] = -t2; kc2[e] = t2 * mean - s2 * grow[a.dbeta2 + c] * inv_n;
    }
  }
  const long long base = static_cast<long long>(g) * a.P * C + c0;
  for (int p0 = blockIdx.x * R * BN_U; p0 < a.P; p0 += gridDim.x * R * BN_U) {
    uint4 dv[BN_U], xv[BN_U], x2v[BN_U];
#pragma unroll
    for (int u = 0; u < BN_U; ++u) {
      const int p = p0 + u * R + r;
      if (p < a.P) {
        const long long off = base + static_cast<long long>(p) * C;
        dv[u] = ld_stream(a.dout + off);
        if (a.out) relu_mask(dv[u], ld_stream(a.out + off));
        xv[u] = ld_stream(a.x + off);
        if constexpr (TWO) x2v[u] = ld_stream(a.x2 + off);
      }
    }
#pragma unroll
    for (int u = 0; u < BN_U; ++u) {
      const int p = p0 + u * R + r;
      if (p >= a.P) continue;
      const long long off = base + static_cast<long long>(p) * C;
      float d[8], xx[8], o[8];
      unpack<__nv_bfloat16>(dv[u], d);
      unpack<__nv_bfloat16>(xv[u], xx);
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = fmaf(ka[e], d[e], fmaf(kb[e], xx[e], kc[e]));
      st_vec(a.dx + off, pack<__nv_bfloat16>(o));
      if constexpr (TWO) {
        unpack<__nv_bfloat16>(x2v[u], xx);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = fmaf(ka2[e], d[e], fmaf(kb2[e], xx[e], kc2[e]));
        st_vec(a.dx2 + off, pack<__nv_bfloat16>(o));
      }
    }
  }
}

// ---- classifier head ------------------------------------------------------------------------------
// feat [clients][B][pp positions][F] bf16 -> average pool -> logits = W [NC][F] pooled + b.
// grid (B / 8, clients), 512 threads (one per feature; F == 512), 8 samples per block.
constexpr int HEAD_NC = 10;
__global__ void __launch_bounds__(512)
    head_kernel(const __nv_bfloat16* __restrict__ feat, int B, int pp, const int64_t* __restrict__ labels,
                long long label_gs, const float* __restrict__ slots, long long row_stride,
                long long w_off, long long b_off, float* __restrict__ sgrad, long long sgrad_stride,
                long long dw_off, long long db_off, __nv_bfloat16* __restrict__ dfeat,
                float* __restrict__ stats) {
  pdl_trigger();
  pdl_wait();      // programmatic dependent launch: see common.cuh
  constexpr int F = 512, SB = 8, NC = HEAD_NC;
  __shared__ float part[16][SB * NC];     // per-warp partial logits
  __shared__ float logit[SB * NC];
  __shared__ float dl[SB * NC];
  const int g = blockIdx.y, b0 = blockIdx.x * SB, f = threadIdx.x;
  const int warp = f >> 5, lane = f & 31;
  const float* row = slots + g * row_stride;
  float w[NC];
#pragma unroll
  for (int j = 0; j < NC; ++j) w[j] = row[w_off + j * F + f];
  float pooled[SB];
  const float inv_pp = 1.f / static_cast<float>(pp);
#pragma unroll
  for (int s = 0; s < SB; ++s) {
    const __nv_bfloat16* src = feat + ((static_cast<long long>(g) * B + b0 + s) * pp) * F + f;
    float t = 0.f;
    for (int p = 0; p < pp; ++p) t += __bfloat162float(src[static_cast<long long>(p) * F]);
    pooled[s] = t * inv_pp;
  }
#pragma unroll
  for (int s = 0; s < SB; ++s)
#pragma unroll
    for (int j = 0; j < NC; ++j) {
      const float t = warp_sum(pooled[s] * w[j]);
      if (lane == 0) part[warp][s * NC + j] = t;
    }
  __syncthreads();
  if (f < SB * NC) {
    float t = row[b_off + f % NC];
    for (int wq = 0; wq < 16; ++wq) t += part[wq][f];
    logit[f] = t;
  }
  __syncthreads();
  if (f < SB) {
    const int s = f;
    const int y = static_cast<int>(labels[g * label_gs + b0 + s]);
    float mx = -3.4e38f;
    int arg = 0;
    for (int j = 0; j < NC; ++j)
      if (logit[s * NC + j] > mx) { mx = logit[s * NC + j]; arg = j; }
    float z = 0.f;
    for (int j = 0; j < NC; ++j) z += __expf(logit[s * NC + j] - mx);
    const float lse = mx + __logf(z);
    const float invb = 1.f / static_cast<float>(B);
    for (int j = 0; j < NC; ++j)
      dl[s * NC + j] = (__expf(logit[s * NC + j] - lse) - (j == y ? 1.f : 0.f)) * invb;
    atomicAdd(stats + g * 3 + 0, lse - logit[s * NC + y]);
    atomicAdd(stats + g * 3 + 1, arg == y ? 1.f : 0.f);
    atomicAdd(stats + g * 3 + 2, 1.f);
  }
  __syncthreads();
  float* grow = sgrad + g * sgrad_stride;
  float dwj[NC];
#pragma unroll
  for (int j = 0; j < NC; ++j) dwj[j] = 0.f;
#pragma unroll
  for (int s = 0; s < SB; ++s) {
    float df = 0.f;
#pragma unroll
    for (int j = 0; j < NC; ++j) {
      df = fmaf(dl[s * NC + j], w[j], df);
      dwj[j] = fmaf(dl[s * NC + j], pooled[s], dwj[j]);
    }
    // gradient w.r.t. the PRE-ReLU block output: zero where the feature was clipped by the ReLU
    const __nv_bfloat16 o = __float2bfloat16_rn(df * inv_pp), zero = __float2bfloat16_rn(0.f);
    const long long base = ((static_cast<long long>(g) * B + b0 + s) * pp) * F + f;
    for (int p = 0; p < pp; ++p)
      dfeat[base + static_cast<long long>(p) * F] =
          __bfloat162float(feat[base + static_cast<long long>(p) * F]) > 0.f ? o : zero;
  }
#pragma unroll
  for (int j = 0; j < NC; ++j) atomicAdd(grow + dw_off + j * F + f, dwj[j]);
  if (f < NC) {
    float t = 0.f;
    for (int s = 0; s < SB; ++s) t += dl[s * NC + f];
    atomicAdd(grow + db_off + f, t);
  }
}

// ---- optimizer ----------------------------------------------------------------------------------
struct ConvEntry {            // one convolution in the conv_sgd table (device memory)
  long long w_off;            // fp32 master [co][ci][taps] inside a client row
  long long t_off;            // tap-major [taps][co][ci] inside a client's dwt / wt row
  long long col_start;        // first (co, ci) column of this layer in the flattened index space
  int cout, cin, taps, stem;  // stem: cin == 64 patch entries, only 27 real (k = tap*3 + ci)
};

// dir 0: client rows (torch layout [co][ci][kh][kw]) -> fp32 tap-major master wm + bf16 copy wt
// dir 1: wm -> client rows.  Once per round each; local steps work on the tap-major copies only.
__global__ void __launch_bounds__(256)
    wm_convert_kernel(const ConvEntry* __restrict__ table, int n_layers, long long total_cols,
                      float* __restrict__ slots, long long row_stride, float* __restrict__ wm,
                      __nv_bfloat16* __restrict__ wt, long long t_stride, int dir) {
  pdl_trigger();
  pdl_wait();      // programmatic dependent launch: see common.cuh
  const int g = blockIdx.y;
  float* row = slots + g * row_stride;
  float* mrow = wm + g * t_stride;
  __nv_bfloat16* trow = wt + g * t_stride;
  for (long long col = blockIdx.x * 256ll + threadIdx.x; col < total_cols; col += gridDim.x * 256ll) {
    int l = 0;
    while (l + 1 < n_layers && col >= table[l + 1].col_start) ++l;
    const ConvEntry e = table[l];
    const long long local = col - e.col_start;
    const int co = static_cast<int>(local / e.cin), ci = static_cast<int>(local % e.cin);
    if (e.stem) {
      const long long ti = e.t_off + static_cast<long long>(co) * e.cin + ci;
      if (ci >= 27) {
        if (dir == 0) { mrow[ti] = 0.f; trow[ti] = __float2bfloat16_rn(0.f); }
        continue;
      }
      const long long wi = e.w_off + (static_cast<long long>(co) * 3 + ci % 3) * 9 + ci / 3;
      if (dir == 0) { const float w = row[wi]; mrow[ti] = w; trow[ti] = __float2bfloat16_rn(w); }
      else row[wi] = mrow[ti];
      continue;
    }
    const long long wi0 = e.w_off + local * e.taps;
    for (int t = 0; t < e.taps; ++t) {
      const long long ti = e.t_off + (static_cast<long long>(t) * e.cout + co) * e.cin + ci;
      if (dir == 0) { const float w = row[wi0 + t]; mrow[ti] = w; trow[ti] = __float2bfloat16_rn(w); }
      else row[wi0 + t] = mrow[ti];
    }
  }
}

// SGD on tap-major ranges: w -= lr * (mu * v + g + wd * w), wt = bf16(w), g = 0.  Used for the
// convolutions whose weight gradient was accumulated with atomics (split-K) or when momentum is on;
// the others are updated inside the wgrad epilogue (csrc/rn_conv.cu).
struct TapRange { long long t_off, len; };

// Client-side DP-SGD (reference: nanofed/trainer/private.py:54-86 - clip the batch gradient of the
// client to C, add N(0, (sigma C)^2) to every parameter's gradient) inside the two update kernels.
// A client's gradient lives in two buffers, the tap-major conv gradients [0, t_stride) and the
// compact BatchNorm / classifier row; idx_base places a kernel's buffer in that index space so the
// two kernels draw from disjoint Philox counters.  Stream position = (steps done before this local
// epoch (device counter) + step) x all clients of the job.
struct RnDp {
  const float* row_sumsq;              // [clients of this launch] squared gradient norm; null: no DP
  float clip_norm, noise_std;
  unsigned long long seed;
  const unsigned long long* ctr;       // device: local steps of earlier rounds
  int step;                            // step inside this local epoch
  int client0, clients_total;          // global index of the launch's first client / of the job
  long long row_groups;                // Philox groups (4 values) per client and step
  long long idx_base;
};
__device__ __forceinline__ float rn_dp_scale(const RnDp& dp, int g) {
  return fminf(dp.clip_norm / (sqrtf(dp.row_sumsq[g]) + 1e-6f), 1.f);
}
__device__ __forceinline__ unsigned long long rn_dp_group0(const RnDp& dp, int g) {
  const unsigned long long t = *dp.ctr + static_cast<unsigned long long>(dp.step);
  return (t * dp.clients_total + dp.client0 + g) * static_cast<unsigned long long>(dp.row_groups);
}

__global__ void __launch_bounds__(256)
    tap_sgd_kernel(const TapRange* __restrict__ ranges, float* __restrict__ wm,
                   float* __restrict__ dwt, float* __restrict__ mom, __nv_bfloat16* __restrict__ wt,
                   long long t_stride, float lr, float mu, float wd, const RnDp dp) {
  pdl_trigger();
  pdl_wait();      // programmatic dependent launch: see common.cuh
  const TapRange rg = ranges[blockIdx.y];
  const long long base = blockIdx.z * t_stride + rg.t_off;
  const float gs = dp.row_sumsq ? rn_dp_scale(dp, blockIdx.z) : 1.f;
  const unsigned long long group0 =
      dp.row_sumsq ? rn_dp_group0(dp, blockIdx.z) + ((dp.idx_base + rg.t_off) >> 2) : 0ull;
  for (long long i = (blockIdx.x * 256ll + threadIdx.x) * 4; i < rg.len; i += gridDim.x * 1024ll) {
    float4 w = *reinterpret_cast<float4*>(wm + base + i);
    float4 gr = *reinterpret_cast<const float4*>(dwt + base + i);
    if (dp.row_sumsq) {
      const float4 z = Philox::normal4(group0 + (i >> 2), dp.seed);
      gr.x = fmaf(dp.noise_std, z.x, gr.x * gs); gr.y = fmaf(dp.noise_std, z.y, gr.y * gs);
      gr.z = fmaf(dp.noise_std, z.z, gr.z * gs); gr.w = fmaf(dp.noise_std, z.w, gr.w * gs);
    }
    float d[4] = {fmaf(wd, w.x, gr.x), fmaf(wd, w.y, gr.y), fmaf(wd, w.z, gr.z), fmaf(wd, w.w, gr.w)};
    if (mom) {
      float4 v = *reinterpret_cast<float4*>(mom + base + i);
      d[0] = fmaf(mu, v.x, d[0]); d[1] = fmaf(mu, v.y, d[1]);
      d[2] = fmaf(mu, v.z, d[2]); d[3] = fmaf(mu, v.w, d[3]);
      *reinterpret_cast<float4*>(mom + base + i) = make_float4(d[0], d[1], d[2], d[3]);
    }
    w.x = fmaf(-lr, d[0], w.x); w.y = fmaf(-lr, d[1], w.y);
    w.z = fmaf(-lr, d[2], w.z); w.w = fmaf(-lr, d[3], w.w);
    *reinterpret_cast<float4*>(wm + base + i) = w;
    *reinterpret_cast<float4*>(dwt + base + i) = make_float4(0.f, 0.f, 0.f, 0.f);
    __nv_bfloat162 lo = __floats2bfloat162_rn(w.x, w.y), hi = __floats2bfloat162_rn(w.z, w.w);
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&lo);
    pk.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(wt + base + i) = pk;
  }
}

struct SmallEntry { long long p_off, g_off, len; };

__global__ void __launch_bounds__(256)
    small_sgd_kernel(const SmallEntry* __restrict__ table, float* __restrict__ slots,
                     long long row_stride, float* __restrict__ momentum, float* __restrict__ sgrad,
                     long long sgrad_stride, float lr, float mu, float wd,
                     float* __restrict__ zero_ptr, long long zero_len, const RnDp dp) {
  pdl_trigger();
  pdl_wait();      // programmatic dependent launch: see common.cuh
  const SmallEntry e = table[blockIdx.x];
  const int g = blockIdx.y;
  const float gs = dp.row_sumsq ? rn_dp_scale(dp, g) : 1.f;
  const unsigned long long group0 = dp.row_sumsq ? rn_dp_group0(dp, g) : 0ull;
  // the BatchNorm batch sums of this step have been consumed: hand them back zeroed
  for (long long i = (static_cast<long long>(g) * gridDim.x + blockIdx.x) * 256 + threadIdx.x;
       i < zero_len; i += 256ll * gridDim.x * gridDim.y)
    zero_ptr[i] = 0.f;
  float* p = slots + g * row_stride + e.p_off;
  float* m = momentum ? momentum + g * row_stride + e.p_off : nullptr;
  float* gr = sgrad + g * sgrad_stride + e.g_off;
  for (long long i = threadIdx.x; i < e.len; i += 256) {
    float w = p[i];
    float gi = gr[i];
    if (dp.row_sumsq) {
      const long long idx = dp.idx_base + e.g_off + i;
      const float4 z = Philox::normal4(group0 + (idx >> 2), dp.seed);
      const int lane = static_cast<int>(idx & 3);
      gi = fmaf(dp.noise_std, lane == 0 ? z.x : lane == 1 ? z.y : lane == 2 ? z.z : z.w, gi * gs);
    }
    float d = fmaf(wd, w, gi);
    if (m) { d = fmaf(mu, m[i], d); m[i] = d; }
    p[i] = fmaf(-lr, d, w);
    gr[i] = 0.f;
  }
}

static RnDp make_dp(const nfb_rn_dp* dp) {
  RnDp d{};
  if (dp && dp->row_sumsq) {
    d.row_sumsq = dp->row_sumsq; d.clip_norm = dp->clip_norm; d.noise_std = dp->noise_std;
    d.seed = dp->seed; d.ctr = dp->step_base; d.step = dp->step; d.client0 = dp->client0;
    d.clients_total = dp->clients_total; d.row_groups = dp->row_groups; d.idx_base = dp->idx_base;
  }
  return d;
}

static int blocks_for(long long work_items, int clients) {
  long long want = (work_items + 255) / 256;
  const long long cap = (8ll * sm_count() + clients - 1) / clients;
  if (want > cap) want = cap;
  return static_cast<int>(want < 1 ? 1 : want);
}

}  // namespace rn
}  // namespace nfb

extern "C" {

int nfb_rn_im2col3(const float* x, long long x_gs, int clients, int batch, void* out, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(x && out && clients >= 1 && batch >= 1, NFB_EINVAL, "rn_im2col3: bad argument");
  dim3 grid(rn::blocks_for(static_cast<long long>(batch) * 8192, clients), clients);
  launch_pdl(rn::im2col3_kernel, dim3(grid), dim3(256), 0, as_stream(stream), x, x_gs, batch,
                                                          static_cast<__nv_bfloat16*>(out));
  return launch_status("rn_im2col3");
}

static nfb::rn::BnRef bn_ref(const float* stats, const long long* offs) {
  return nfb::rn::BnRef{stats, offs[0], offs[1], offs[2], offs[3], offs[4]};
}

int nfb_rn_bn_apply(const void* x, const void* res, void* y, float* slots, long long row_stride,
                    const float* stats, const long long* bn_offs, const float* res_stats,
                    const long long* res_offs, int clients, int P, int C, int relu, float eps,
                    float momentum, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(x && y && slots && stats && bn_offs, NFB_EINVAL, "rn_bn_apply: null pointer");
  NFB_REQUIRE(C % 8 == 0 && C <= 512 && P > 1, NFB_EINVAL, "rn_bn_apply: bad shape");
  rn::BnApplyArgs a{};
  a.x = static_cast<const __nv_bfloat16*>(x);
  a.res = static_cast<const __nv_bfloat16*>(res);
  a.y = static_cast<__nv_bfloat16*>(y);
  a.slots = slots; a.row_stride = row_stride;
  a.bn = bn_ref(stats, bn_offs);
  if (res_stats) {
    NFB_REQUIRE(res && res_offs, NFB_EINVAL, "rn_bn_apply: shortcut BatchNorm without a residual");
    a.bn_res = bn_ref(res_stats, res_offs);
  }
  a.P = P; a.C = C; a.relu = relu; a.eps = eps; a.momentum = momentum;
  NFB_REQUIRE(256 % (C / 8) == 0, NFB_EINVAL, "rn_bn_apply: C / 8 must divide 256");
  // four trips of the load loop per block: with one, the deep layers (4096 or 1024 positions per
  // client) ran 500-1000 blocks that were mostly prologue (32 coefficient loads per thread)
  dim3 grid(rn::blocks_for(static_cast<long long>(P) * C / 8 / (4 * rn::BN_U), clients), clients);
  cudaStream_t st = as_stream(stream);
  if (!res) launch_pdl(rn::bn_apply_kernel<false, false>, dim3(grid), dim3(256), 0, st, a);
  else if (!res_stats) launch_pdl(rn::bn_apply_kernel<true, false>, dim3(grid), dim3(256), 0, st, a);
  else launch_pdl(rn::bn_apply_kernel<true, true>, dim3(grid), dim3(256), 0, st, a);
  return launch_status("rn_bn_apply");
}

int nfb_rn_bn_bwd(int apply, const void* dout, const void* out, const void* x, const void* x2,
                  void* dx, void* dx2, const float* slots, long long row_stride, float* sgrad,
                  long long sgrad_stride, const long long* grad_offs /* dgamma dbeta dgamma2 dbeta2 */,
                  const float* stats, const long long* bn_offs, const float* stats2,
                  const long long* bn2_offs, int clients, int P, int C, float eps,
                  const float* raw_sums, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(dout && x && slots && sgrad && grad_offs && stats && bn_offs, NFB_EINVAL,
              "rn_bn_bwd: null pointer");
  NFB_REQUIRE(C % 8 == 0 && C <= 512 && 256 % (C / 8) == 0, NFB_EINVAL, "rn_bn_bwd: bad channels");
  NFB_REQUIRE(!x2 || (stats2 && bn2_offs), NFB_EINVAL, "rn_bn_bwd: second branch incomplete");
  NFB_REQUIRE(!apply || (dx && (!x2 || dx2)), NFB_EINVAL, "rn_bn_bwd: apply needs outputs");
  rn::BnBwdArgs a{};
  a.dout = static_cast<const __nv_bfloat16*>(dout);
  a.out = static_cast<const __nv_bfloat16*>(out);
  a.x = static_cast<const __nv_bfloat16*>(x);
  a.x2 = static_cast<const __nv_bfloat16*>(x2);
  a.dx = static_cast<__nv_bfloat16*>(dx);
  a.dx2 = static_cast<__nv_bfloat16*>(dx2);
  a.slots = slots; a.row_stride = row_stride; a.sgrad = sgrad; a.sgrad_stride = sgrad_stride;
  a.dgamma = grad_offs[0]; a.dbeta = grad_offs[1]; a.dgamma2 = grad_offs[2]; a.dbeta2 = grad_offs[3];
  a.bn = bn_ref(stats, bn_offs);
  if (x2) a.bn2 = bn_ref(stats2, bn2_offs);
  a.P = P; a.C = C; a.eps = eps;
  NFB_REQUIRE(!raw_sums || (apply && !x2), NFB_EINVAL,
              "rn_bn_bwd: raw sums replace the reduce pass of a single-branch apply");
  a.raw_sums = raw_sums;
  cudaStream_t st = as_stream(stream);
  if (!apply) {
    // fewer, longer blocks: every block ends with C atomics per sum
    dim3 grid(rn::blocks_for(static_cast<long long>(P) * C / 8 / (4 * rn::BN_U), clients), clients);
    if (x2) launch_pdl(rn::bn_bwd_reduce_kernel<true>, dim3(grid), dim3(256), 0, st, a);
    else launch_pdl(rn::bn_bwd_reduce_kernel<false>, dim3(grid), dim3(256), 0, st, a);
    return launch_status("rn_bn_bwd_reduce");
  }
  dim3 grid(rn::blocks_for(static_cast<long long>(P) * C / 8 / (4 * rn::BN_U), clients), clients);
  if (x2) launch_pdl(rn::bn_bwd_apply_kernel<true>, dim3(grid), dim3(256), 0, st, a);
  else launch_pdl(rn::bn_bwd_apply_kernel<false>, dim3(grid), dim3(256), 0, st, a);
  return launch_status("rn_bn_bwd_apply");
}

int nfb_rn_head(const void* feat, int clients, int batch, int positions, const int64_t* labels,
                long long label_gs, const float* slots, long long row_stride, long long w_off,
                long long b_off, float* sgrad, long long sgrad_stride, long long dw_off,
                long long db_off, void* dfeat, float* stats, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(feat && labels && slots && sgrad && dfeat && stats, NFB_EINVAL, "rn_head: null pointer");
  NFB_REQUIRE(batch % 8 == 0 && positions >= 1, NFB_EINVAL, "rn_head: batch must be a multiple of 8");
  dim3 grid(batch / 8, clients);
  launch_pdl(rn::head_kernel, dim3(grid), dim3(512), 0, as_stream(stream), 
      static_cast<const __nv_bfloat16*>(feat), batch, positions, labels, label_gs, slots, row_stride,
      w_off, b_off, sgrad, sgrad_stride, dw_off, db_off, static_cast<__nv_bfloat16*>(dfeat), stats);
  return launch_status("rn_head");
}

int nfb_rn_wm_convert(const void* table, int n_layers, long long total_cols, float* slots,
                      long long row_stride, float* wm, void* wt, long long t_stride, int clients,
                      int to_slots, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(table && slots && wm && wt && n_layers >= 1, NFB_EINVAL, "rn_wm_convert: null pointer");
  dim3 grid(rn::blocks_for(total_cols, clients), clients);
  launch_pdl(rn::wm_convert_kernel, dim3(grid), dim3(256), 0, as_stream(stream), 
      static_cast<const rn::ConvEntry*>(table), n_layers, total_cols, slots, row_stride, wm,
      static_cast<__nv_bfloat16*>(wt), t_stride, to_slots);
  return launch_status("rn_wm_convert");
}

int nfb_rn_tap_sgd(const void* ranges, int n_ranges, long long max_len, float* wm, float* dwt,
                   float* momentum, void* wt, long long t_stride, int clients, float lr, float mu,
                   float wd, const nfb_rn_dp* dp, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(ranges && wm && dwt && wt && n_ranges >= 1, NFB_EINVAL, "rn_tap_sgd: null pointer");
  const int per = (4 * sm_count() + n_ranges * clients - 1) / (n_ranges * clients);
  long long want = (max_len / 4 + 255) / 256;
  if (want > per) want = per;
  dim3 grid(static_cast<unsigned>(want < 1 ? 1 : want), n_ranges, clients);
  launch_pdl(rn::tap_sgd_kernel, dim3(grid), dim3(256), 0, as_stream(stream), 
      static_cast<const rn::TapRange*>(ranges), wm, dwt, momentum, static_cast<__nv_bfloat16*>(wt),
      t_stride, lr, mu, wd, rn::make_dp(dp));
  return launch_status("rn_tap_sgd");
}

int nfb_rn_small_sgd(const void* table, int n_entries, float* slots, long long row_stride,
                     float* momentum, float* sgrad, long long sgrad_stride, int clients, float lr,
                     float mu, float wd, float* zero_ptr, long long zero_len, const nfb_rn_dp* dp,
                     void* stream) {
  using namespace nfb;
  NFB_REQUIRE(table && slots && sgrad && n_entries >= 1, NFB_EINVAL, "rn_small_sgd: null pointer");
  NFB_REQUIRE(!dp || !dp->row_sumsq || (dp->step_base && dp->clients_total >= 1), NFB_EINVAL,
              "rn_small_sgd: DP needs the device step counter and the job's client count");
  dim3 grid(n_entries, clients);
  launch_pdl(rn::small_sgd_kernel, dim3(grid), dim3(256), 0, as_stream(stream), 
      static_cast<const rn::SmallEntry*>(table), slots, row_stride, momentum, sgrad, sgrad_stride,
      lr, mu, wd, zero_ptr, zero_ptr ? zero_len : 0, rn::make_dp(dp));
  return launch_status("rn_small_sgd");
}

}  // extern "C"
