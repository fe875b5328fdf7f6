// K-OPT: fused SGD / Adam(W) on flat fp32 buffers, optional DP-SGD front end
// (device-side clip factor + Philox Gaussian noise) and bf16 working-copy write-back.
// One pass: reads p, g (+ state), writes p (+ state, + bf16 copy).  HBM-bound.
#include "common.cuh"
#include "philox.cuh"

namespace nfb {

struct SgdArgs {
  float lr, momentum, dampening, weight_decay, noise_std;
  int nesterov, first_step;
  const float* grad_scale;
  uint64_t seed, group0;
  int64_t row_len;      // > 0: the flat buffer is [rows][row_len] ...
  int64_t zero_prefix;  // ... and g[row][0 : zero_prefix) is zeroed after use (atomic accumulators)
  int64_t skip_lo, skip_len;  // columns [skip_lo, skip_lo + skip_len) of every row are left alone
  int64_t zero_from;          // > 0: g[row][zero_from : row_len) is zeroed after use as well
  // Optional permuted bf16 side output for ONE parameter stored [rows][inner][taps] at columns
  // [perm_lo, perm_lo + perm_len): written as [rows][taps][inner] into perm_out + row * perm_gs
  // (the tap-major conv weights the implicit-GEMM kernels read; saves a permutation kernel per step).
  __nv_bfloat16* perm_out;
  int64_t perm_gs, perm_lo, perm_len;
  int perm_inner, perm_taps;
  // DP-SGD per row (= per client): g' = g * min(1, clip_norm / ||g_row||) + noise_std * z, with the
  // row's squared gradient norm from K-STATS (reference: clip_grad_norm_ + per-parameter noise of
  // nanofed/trainer/private.py:54-86).  noise_mask limits the noise to real parameter entries;
  // step_ctr (device) advances the noise stream from step to step inside a replayed CUDA graph.
  const float* row_sumsq;
  float clip_norm;
  const uint8_t* noise_mask;
  const unsigned long long* step_ctr;
  // K-STATS side output (optional, [rows]): += sum of the squares of the UPDATED parameters of each
  // row, over the columns this launch touches - with the matching output of the fused weight-
  // gradient kernel for the skipped range this is the per-client sum(x^2) that central DP clips by,
  // so the last local step makes the separate pass over the client rows unnecessary.
  float* sumsq_out;
};

// SUMSQ: instantiation with the sum-of-squares side output (a separate one: its shared-memory
// array and ten more registers cost the plain kernel 3 us per step when it was a run-time switch)
template <typename G, bool SUMSQ>
__global__ void __launch_bounds__(256)
    sgd_kernel(float* __restrict__ p, G* __restrict__ g, float* __restrict__ mom,
               __nv_bfloat16* __restrict__ lowp, int64_t n, SgdArgs a) {
  __shared__ float s_rowsq[SUMSQ ? 256 : 1];   // per-row partial sums of this block (rows <= 256)
  const int64_t n_rows = SUMSQ ? n / a.row_len : 0;
  if constexpr (SUMSQ) {
    s_rowsq[threadIdx.x] = 0.f;
    __syncthreads();
  }
  const float gs0 = a.grad_scale ? *a.grad_scale : 1.f;
  const uint64_t group0 = a.group0 + (a.step_ctr ? *a.step_ctr * static_cast<uint64_t>((n + 3) >> 2) : 0ull);
  // with a skipped column range the thread index runs over the remaining columns only
  const int64_t eff_row = a.row_len - a.skip_len;
  const int64_t groups = a.skip_len > 0 ? (n / a.row_len) * (eff_row >> 2) : (n + 3) >> 2;
  for (int64_t q = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; q < groups;
       q += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    int64_t base = q << 2;
    if (a.skip_len > 0) {
      const int64_t row = base / eff_row;
      const int64_t col = base - row * eff_row;
      base = row * a.row_len + (col < a.skip_lo ? col : col + a.skip_len);
    }
    const int cnt = (n - base) >= 4 ? 4 : static_cast<int>(n - base);
    float gs = gs0;
    // row / column of this group for the per-row clip factor and the noise mask: 32-bit division
    // where the buffer allows it (a 64-bit divide per group cost 10 us of a 124 us pass)
    int64_t dp_row = 0, dp_col = 0;
    if (a.row_sumsq || a.noise_mask || SUMSQ) {
      if (n < (1ll << 31)) {
        const uint32_t b32 = static_cast<uint32_t>(base), rl = static_cast<uint32_t>(a.row_len);
        const uint32_t r32 = b32 / rl;
        dp_row = r32;
        dp_col = b32 - r32 * rl;
      } else {
        dp_row = base / a.row_len;
        dp_col = base - dp_row * a.row_len;
      }
    }
    if (a.row_sumsq) gs *= fminf(a.clip_norm / (sqrtf(a.row_sumsq[dp_row]) + 1e-6f), 1.f);
    float pv[4], gv[4], mv[4] = {0, 0, 0, 0};
    if (cnt == 4) {
      const float4 t = *reinterpret_cast<const float4*>(p + base);
      pv[0] = t.x; pv[1] = t.y; pv[2] = t.z; pv[3] = t.w;
    } else {
      for (int e = 0; e < 4; ++e) pv[e] = e < cnt ? p[base + e] : 0.f;
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) gv[e] = e < cnt ? to_f32(g[base + e]) * gs : 0.f;
    if (a.zero_prefix > 0 || a.zero_from > 0) {
      // gradient entries that the next step accumulates into with atomics: hand them back zeroed
      const int64_t col = base % a.row_len;
      if (col < a.zero_prefix || (a.zero_from > 0 && col + 3 >= a.zero_from)) {
        for (int e = 0; e < cnt; ++e) {
          const int64_t ce = (base + e) % a.row_len;
          if (ce < a.zero_prefix || (a.zero_from > 0 && ce >= a.zero_from))
            g[base + e] = from_f32<G>(0.f);
        }
      }
    }
    if (a.noise_std > 0.f) {
      const float4 z = Philox::normal4(group0 + (base >> 2), a.seed);
      float zz[4] = {z.x, z.y, z.z, z.w};
      if (a.noise_mask) {
        if (cnt == 4 && (dp_col & 3) == 0) {          // one 4-byte load per group
          const uchar4 m = *reinterpret_cast<const uchar4*>(a.noise_mask + dp_col);
          if (!m.x) zz[0] = 0.f;
          if (!m.y) zz[1] = 0.f;
          if (!m.z) zz[2] = 0.f;
          if (!m.w) zz[3] = 0.f;
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (e >= cnt || !a.noise_mask[dp_col + e]) zz[e] = 0.f;
        }
      }
      gv[0] = fmaf(a.noise_std, zz[0], gv[0]); gv[1] = fmaf(a.noise_std, zz[1], gv[1]);
      gv[2] = fmaf(a.noise_std, zz[2], gv[2]); gv[3] = fmaf(a.noise_std, zz[3], gv[3]);
    }
    if (a.weight_decay != 0.f) {
#pragma unroll
      for (int e = 0; e < 4; ++e) gv[e] = fmaf(a.weight_decay, pv[e], gv[e]);
    }
    if (mom) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        if (e < cnt) {
          mv[e] = a.first_step ? gv[e]
                               : fmaf(a.momentum, mom[base + e], (1.f - a.dampening) * gv[e]);
          mom[base + e] = mv[e];
          gv[e] = a.nesterov ? fmaf(a.momentum, mv[e], gv[e]) : mv[e];
        }
      }
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) pv[e] = fmaf(-a.lr, gv[e], pv[e]);
    if (cnt == 4) {
      *reinterpret_cast<float4*>(p + base) = make_float4(pv[0], pv[1], pv[2], pv[3]);
    } else {
      for (int e = 0; e < cnt; ++e) p[base + e] = pv[e];
    }
    if (lowp) {
      for (int e = 0; e < cnt; ++e) lowp[base + e] = __float2bfloat16_rn(pv[e]);
    }
    if constexpr (SUMSQ) {                 // row_len % 4 == 0: a group never straddles two rows
      float t = 0.f;
      for (int e = 0; e < cnt; ++e) t = fmaf(pv[e], pv[e], t);
      if (n_rows <= 256) atomicAdd(&s_rowsq[dp_row], t);
      else atomicAdd(a.sumsq_out + dp_row, t);
    }
    if (a.perm_out) {
      const int64_t row = base / a.row_len;
      const int64_t col0 = base - row * a.row_len;
      if (col0 + 3 >= a.perm_lo && col0 < a.perm_lo + a.perm_len) {
        const int block = a.perm_inner * a.perm_taps;
        for (int e = 0; e < cnt; ++e) {
          const int64_t idx = col0 + e - a.perm_lo;
          if (idx < 0 || idx >= a.perm_len) continue;
          const int r = static_cast<int>(idx / block), rem = static_cast<int>(idx % block);
          const int ci = rem / a.perm_taps, t = rem % a.perm_taps;
          a.perm_out[row * a.perm_gs + static_cast<int64_t>(r) * block + t * a.perm_inner + ci] =
              __float2bfloat16_rn(pv[e]);
        }
      }
    }
  }
  if (SUMSQ && n_rows <= 256) {
    __syncthreads();
    if (threadIdx.x < n_rows && s_rowsq[threadIdx.x] != 0.f)
      atomicAdd(a.sumsq_out + threadIdx.x, s_rowsq[threadIdx.x]);
  }
}

struct AdamArgs {
  float lr, beta1, beta2, eps, weight_decay, bc1, bc2;
  int decoupled;
  const float* grad_scale;
};

template <typename G>
__global__ void __launch_bounds__(256)
    adam_kernel(float* __restrict__ p, const G* __restrict__ g, float* __restrict__ m,
                float* __restrict__ v, __nv_bfloat16* __restrict__ lowp, int64_t n, AdamArgs a) {
  const float gs = a.grad_scale ? *a.grad_scale : 1.f;
  const float inv_sqrt_bc2 = rsqrtf(a.bc2);
  const float step = a.lr / a.bc1;
  for (int64_t i = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    float pv = p[i];
    float gv = to_f32(g[i]) * gs;
    if (a.weight_decay != 0.f) {
      if (a.decoupled) pv *= (1.f - a.lr * a.weight_decay);
      else gv = fmaf(a.weight_decay, pv, gv);
    }
    const float mv = fmaf(a.beta1, m[i], (1.f - a.beta1) * gv);
    const float vv = fmaf(a.beta2, v[i], (1.f - a.beta2) * gv * gv);
    m[i] = mv;
    v[i] = vv;
    pv -= step * mv / (sqrtf(vv) * inv_sqrt_bc2 + a.eps);
    p[i] = pv;
    if (lowp) lowp[i] = __float2bfloat16_rn(pv);
  }
}

static int flat_grid(int64_t items) {
  int64_t b = (items + 255) / 256;
  const int64_t cap = static_cast<int64_t>(sm_count()) * 8;
  if (b > cap) b = cap;
  return static_cast<int>(b < 1 ? 1 : b);
}

}  // namespace nfb

static int fused_sgd_impl(float* p, void* g, int g_dtype, float* momentum_buf, void* lowp_out,
                          int64_t n, float lr, float momentum, float dampening,
                          float weight_decay, int nesterov, int first_step,
                          const float* grad_scale, float noise_std, uint64_t seed,
                          uint64_t offset, int64_t row_len, int64_t zero_prefix, int64_t skip_lo,
                          int64_t skip_hi, int64_t zero_from, void* stream,
                          void* perm_out = nullptr, int64_t perm_gs = 0, int64_t perm_lo = 0,
                          int64_t perm_rows = 0, int perm_inner = 0, int perm_taps = 0,
                          const float* row_sumsq = nullptr, float clip_norm = 0.f,
                          const uint8_t* noise_mask = nullptr,
                          const unsigned long long* step_ctr = nullptr,
                          float* sumsq_out = nullptr) {
  using namespace nfb;
  NFB_REQUIRE(p && g, NFB_EINVAL, "fused_sgd: null pointer");
  NFB_REQUIRE(!sumsq_out || (row_len > 0 && n % row_len == 0 && (row_len & 3) == 0), NFB_EINVAL,
              "fused_sgd: the sum-of-squares output needs rows of a multiple of 4 elements");
  NFB_REQUIRE(aligned16(p), NFB_EINVAL, "fused_sgd: parameters must be 16-byte aligned");
  NFB_REQUIRE((offset & 3u) == 0, NFB_EINVAL, "fused_sgd: offset must be a multiple of 4");
  if (n <= 0) return 0;
  NFB_REQUIRE(zero_prefix == 0 || (row_len > 0 && zero_prefix <= row_len && n % row_len == 0),
              NFB_EINVAL, "fused_sgd: bad row_len/zero_prefix");
  NFB_REQUIRE(zero_from == 0 || (row_len > 0 && zero_from > 0 && zero_from <= row_len &&
                                 n % row_len == 0),
              NFB_EINVAL, "fused_sgd: bad zero_from");
  const int64_t skip_len = skip_hi - skip_lo;
  NFB_REQUIRE(skip_len == 0 ||
                  (row_len > 0 && n % row_len == 0 && skip_lo >= 0 && skip_len > 0 &&
                   skip_hi <= row_len && ((row_len | skip_lo | skip_len) & 3) == 0),
              NFB_EINVAL, "fused_sgd: skipped range must lie in a row, multiples of 4");
  const int64_t perm_len = perm_rows * perm_inner * perm_taps;
  NFB_REQUIRE(!perm_out || (row_len > 0 && n % row_len == 0 && perm_len > 0 && perm_lo >= 0 &&
                            perm_lo + perm_len <= row_len && perm_gs >= perm_len &&
                            (skip_len == 0 || perm_lo + perm_len <= skip_lo || perm_lo >= skip_hi)),
              NFB_EINVAL, "fused_sgd: bad permuted-output range");
  SgdArgs a{lr, momentum, dampening, weight_decay, noise_std, nesterov, first_step,
            grad_scale, seed, offset >> 2, row_len > 0 ? row_len : n, zero_prefix, skip_lo,
            skip_len, zero_from, static_cast<__nv_bfloat16*>(perm_out), perm_gs, perm_lo,
            perm_out ? perm_len : 0, perm_inner, perm_taps, row_sumsq, clip_norm, noise_mask,
            step_ctr, sumsq_out};
  NFB_REQUIRE(!(row_sumsq || noise_mask) || (row_len > 0 && n % row_len == 0), NFB_EINVAL,
              "fused_sgd: per-row clipping / noise mask need row_len");
  float* mom = (momentum != 0.f) ? momentum_buf : nullptr;
  cudaStream_t st = as_stream(stream);
  const int64_t work = skip_len > 0 ? (n / row_len) * (row_len - skip_len) : n;
  if (work <= 0) return 0;
  return dispatch_dtype(g_dtype, [&](auto tag) {
    using G = decltype(tag);
    if (sumsq_out)
      sgd_kernel<G, true><<<flat_grid((work + 3) / 4), 256, 0, st>>>(
          p, static_cast<G*>(g), mom, static_cast<__nv_bfloat16*>(lowp_out), n, a);
    else
      sgd_kernel<G, false><<<flat_grid((work + 3) / 4), 256, 0, st>>>(
          p, static_cast<G*>(g), mom, static_cast<__nv_bfloat16*>(lowp_out), n, a);
    return launch_status("fused_sgd");
  });
}

extern "C" int nfb_fused_sgd(float* p, void* g, int g_dtype, float* momentum_buf,
                             void* lowp_out, int64_t n, float lr, float momentum,
                             float dampening, float weight_decay, int nesterov, int first_step,
                             const float* grad_scale, float noise_std, uint64_t seed,
                             uint64_t offset, int64_t row_len, int64_t zero_prefix, void* stream) {
  return fused_sgd_impl(p, g, g_dtype, momentum_buf, lowp_out, n, lr, momentum, dampening,
                        weight_decay, nesterov, first_step, grad_scale, noise_std, seed, offset,
                        row_len, zero_prefix, 0, 0, 0, stream);
}

// Same update, except that columns [skip_lo, skip_hi) of every row are not touched (neither
// read nor written): they belong to a parameter whose update was already applied by the kernel
// that produced its gradient (nfb_tc_linear_wgrad_sgd).  zero_from > 0 additionally zeroes
// g[row][zero_from : row_len) after use (a second block of atomically accumulated entries).
extern "C" int nfb_fused_sgd_skip(float* p, void* g, int g_dtype, float* momentum_buf,
                                  void* lowp_out, int64_t n, float lr, float momentum,
                                  float dampening, float weight_decay, int nesterov,
                                  int first_step, const float* grad_scale, float noise_std,
                                  uint64_t seed, uint64_t offset, int64_t row_len,
                                  int64_t zero_prefix, int64_t skip_lo, int64_t skip_hi,
                                  int64_t zero_from, void* stream) {
  return fused_sgd_impl(p, g, g_dtype, momentum_buf, lowp_out, n, lr, momentum, dampening,
                        weight_decay, nesterov, first_step, grad_scale, noise_std, seed, offset,
                        row_len, zero_prefix, skip_lo, skip_hi, zero_from, stream);
}

// nfb_fused_sgd_skip plus a permuted bf16 side output: the parameter stored as
// [perm_rows][perm_inner][perm_taps] at columns perm_lo.. of every row is also written as
// [perm_rows][perm_taps][perm_inner] (bf16) to perm_out + row * perm_group_stride.
extern "C" int nfb_fused_sgd_perm(float* p, void* g, int g_dtype, float* momentum_buf,
                                  void* lowp_out, int64_t n, float lr, float momentum,
                                  float dampening, float weight_decay, int nesterov,
                                  int first_step, const float* grad_scale, float noise_std,
                                  uint64_t seed, uint64_t offset, int64_t row_len,
                                  int64_t zero_prefix, int64_t skip_lo, int64_t skip_hi,
                                  int64_t zero_from, void* perm_out, int64_t perm_group_stride,
                                  int64_t perm_lo, int64_t perm_rows, int perm_inner,
                                  int perm_taps, void* stream) {
  NFB_REQUIRE(perm_out, NFB_EINVAL, "fused_sgd_perm: null permuted output");
  return fused_sgd_impl(p, g, g_dtype, momentum_buf, lowp_out, n, lr, momentum, dampening,
                        weight_decay, nesterov, first_step, grad_scale, noise_std, seed, offset,
                        row_len, zero_prefix, skip_lo, skip_hi, zero_from, stream, perm_out,
                        perm_group_stride, perm_lo, perm_rows, perm_inner, perm_taps);
}

// DP-SGD for a cohort in one launch: every row (client) is clipped to clip_norm by its own gradient
// norm (row_sumsq from K-STATS) and gets Gaussian noise of std noise_std on the entries noise_mask
// marks; step_ctr (device counter, may be null) makes the noise stream advance inside a replayed
// graph.  perm_out may be null.  sumsq_out (optional, [rows]): += sum of the squares of the updated
// parameters per row.  With row_sumsq == NULL, noise_std == 0 this is plain SGD with that side
// output.  Other arguments as nfb_fused_sgd_perm.
extern "C" int nfb_fused_sgd_dp(float* p, void* g, int g_dtype, float* momentum_buf,
                                void* lowp_out, int64_t n, float lr, float momentum,
                                float dampening, float weight_decay, int nesterov, int first_step,
                                float noise_std, uint64_t seed, uint64_t offset, int64_t row_len,
                                int64_t zero_prefix, int64_t skip_lo, int64_t skip_hi,
                                int64_t zero_from, const float* row_sumsq,
                                float clip_norm, const uint8_t* noise_mask,
                                const unsigned long long* step_ctr, void* perm_out,
                                int64_t perm_group_stride, int64_t perm_lo, int64_t perm_rows,
                                int perm_inner, int perm_taps, float* sumsq_out, void* stream) {
  return fused_sgd_impl(p, g, g_dtype, momentum_buf, lowp_out, n, lr, momentum, dampening,
                        weight_decay, nesterov, first_step, nullptr, noise_std, seed, offset,
                        row_len, zero_prefix, skip_lo, skip_hi, zero_from, stream, perm_out,
                        perm_group_stride,
                        perm_lo, perm_rows, perm_inner, perm_taps, row_sumsq, clip_norm, noise_mask,
                        step_ctr, sumsq_out);
}

extern "C" int nfb_fused_adam(float* p, const void* g, int g_dtype, float* exp_avg,
                              float* exp_avg_sq, void* lowp_out, int64_t n, float lr,
                              float beta1, float beta2, float eps, float weight_decay,
                              float bias_corr1, float bias_corr2, int decoupled,
                              const float* grad_scale, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(p && g && exp_avg && exp_avg_sq, NFB_EINVAL, "fused_adam: null pointer");
  NFB_REQUIRE(bias_corr1 > 0.f && bias_corr2 > 0.f, NFB_EINVAL,
              "fused_adam: bias corrections must be positive (step >= 1)");
  if (n <= 0) return 0;
  AdamArgs a{lr, beta1, beta2, eps, weight_decay, bias_corr1, bias_corr2, decoupled, grad_scale};
  cudaStream_t st = as_stream(stream);
  return dispatch_dtype(g_dtype, [&](auto tag) {
    using G = decltype(tag);
    adam_kernel<G><<<flat_grid(n), 256, 0, st>>>(p, static_cast<const G*>(g), exp_avg,
                                                 exp_avg_sq,
                                                 static_cast<__nv_bfloat16*>(lowp_out), n, a);
    return launch_status("fused_adam");
  });
}
