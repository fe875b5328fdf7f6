// GroupNorm(8) for the ResNet cohort engine (model "resnet18_gn": the BatchNorm-free variant that
// DP training uses, since per-sample gradients do not exist through batch statistics).
//
//   gn_stats      per image and group: sum, sum of squares of the raw conv output
//   gn_apply      y = relu(gamma_c * (x - mean_ig) * rstd_ig + beta_c [+ residual | + gn2(residual)])
//   gn_bwd_reduce per image and group: s1 = sum gamma_c g, s2 = sum gamma_c g xhat; per client and
//                 channel: dgamma_c += sum g xhat, dbeta_c += sum g  (main branch and, with x2, the
//                 projection shortcut's GroupNorm at once)
//   gn_bwd_apply  dx = rstd * (gamma_c g - s1 / m - xhat * s2 / m),  m = H * W * C / 8
//
// Activations are [images][H * W][C] bf16; a thread owns one channel octet (16 bytes, inside one
// group because C / 8 >= 8) of one image, grid = (position splits, images).  Statistics are per
// IMAGE, so unlike BatchNorm they cannot ride in the convolution epilogue's per-client column sums:
// a 256-position tile covers 1/4 .. 16 images depending on the layer.
//
// No reference counterpart: these are torch.nn.GroupNorm / ReLU calls of a model trained by
// nanofed/trainer/base.py:134-160 (and per sample by nanofed/trainer/private.py:54-86).
#include "common.cuh"

namespace nfb {
namespace rn {

constexpr int GN_GROUPS = 8;
constexpr int GN_U = 4;

struct GnRef {
  const float* stats;        // [images][8][2]: sum, sum of squares
  long long gamma, beta;     // offsets into a client's fp32 row
};

__device__ __forceinline__ void gn_moments(const float* stats, long long img, int grp, float inv_m,
                                           float eps, float& mean, float& rstd) {
  const float s = stats[(img * GN_GROUPS + grp) * 2];
  const float q = stats[(img * GN_GROUPS + grp) * 2 + 1];
  mean = s * inv_m;
  rstd = rsqrtf(fmaxf(q * inv_m - mean * mean, 0.f) + eps);
}

__global__ void __launch_bounds__(256)
    gn_stats_kernel(const __nv_bfloat16* __restrict__ x, int HW, int C, float* __restrict__ stats) {
  pdl_trigger();
  pdl_wait();      // programmatic dependent launch: see common.cuh
  __shared__ float red[GN_GROUPS * 2];
  const long long img = blockIdx.y;
  const int octs = C >> 3, oct = threadIdx.x % octs, r = threadIdx.x / octs, R = 256 / octs;
  const int grp = oct * 64 / C;
  if (threadIdx.x < GN_GROUPS * 2) red[threadIdx.x] = 0.f;
  __syncthreads();
  float s = 0.f, q = 0.f;
  const long long base = img * HW * C + oct * 8;
  for (int p0 = blockIdx.x * R * GN_U; p0 < HW; p0 += gridDim.x * R * GN_U) {
    uint4 v[GN_U];
#pragma unroll
    for (int u = 0; u < GN_U; ++u) {
      const int p = p0 + u * R + r;
      v[u] = p < HW ? ld_stream(x + base + static_cast<long long>(p) * C) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < GN_U; ++u) {
      float f[8];
      unpack<__nv_bfloat16>(v[u], f);
#pragma unroll
      for (int e = 0; e < 8; ++e) { s += f[e]; q = fmaf(f[e], f[e], q); }
    }
  }
  atomicAdd(&red[grp * 2], s);
  atomicAdd(&red[grp * 2 + 1], q);
  __syncthreads();
  if (threadIdx.x < GN_GROUPS * 2)
    atomicAdd(stats + img * GN_GROUPS * 2 + threadIdx.x, red[threadIdx.x]);
}

struct GnApplyArgs {
  const __nv_bfloat16* x;
  const __nv_bfloat16* res;
  __nv_bfloat16* y;
  const float* slots;
  long long row_stride;
  GnRef gn, gn_res;          // gn_res.stats == null: the residual is added as is
  int HW, C, imgs_per_client, relu;
  float eps;
};

template <bool RES, bool RES_GN>
__global__ void __launch_bounds__(256) gn_apply_kernel(const GnApplyArgs a) {
  pdl_trigger();
  pdl_wait();      // programmatic dependent launch: see common.cuh
  const long long img = blockIdx.y;
  const int C = a.C, octs = C >> 3, oct = threadIdx.x % octs, r = threadIdx.x / octs, R = 256 / octs;
  const int grp = oct * 64 / C, c0 = oct * 8;
  const float* row = a.slots + (img / a.imgs_per_client) * a.row_stride;
  const float inv_m = 1.f / (static_cast<float>(a.HW) * static_cast<float>(C / GN_GROUPS));
  float mean, rstd, mean2 = 0.f, rstd2 = 1.f;
  gn_moments(a.gn.stats, img, grp, inv_m, a.eps, mean, rstd);
  if constexpr (RES_GN) gn_moments(a.gn_res.stats, img, grp, inv_m, a.eps, mean2, rstd2);
  float sc[8], sh[8], sc2[8], sh2[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    sc[e] = row[a.gn.gamma + c0 + e] * rstd;
    sh[e] = row[a.gn.beta + c0 + e] - mean * sc[e];
    sc2[e] = 1.f; sh2[e] = 0.f;
    if constexpr (RES_GN) {
      sc2[e] = row[a.gn_res.gamma + c0 + e] * rstd2;
      sh2[e] = row[a.gn_res.beta + c0 + e] - mean2 * sc2[e];
    }
  }
  const long long base = img * a.HW * C + c0;
  for (int p0 = blockIdx.x * R * GN_U; p0 < a.HW; p0 += gridDim.x * R * GN_U) {
    uint4 xv[GN_U], rv[GN_U];
#pragma unroll
    for (int u = 0; u < GN_U; ++u) {
      const int p = p0 + u * R + r;
      if (p < a.HW) {
        xv[u] = ld_stream(a.x + base + static_cast<long long>(p) * C);
        if constexpr (RES) rv[u] = ld_stream(a.res + base + static_cast<long long>(p) * C);
      }
    }
#pragma unroll
    for (int u = 0; u < GN_U; ++u) {
      const int p = p0 + u * R + r;
      if (p >= a.HW) continue;
      float v[8], rr[8];
      unpack<__nv_bfloat16>(xv[u], v);
      if constexpr (RES) unpack<__nv_bfloat16>(rv[u], rr);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        float o = fmaf(v[e], sc[e], sh[e]);
        if constexpr (RES) o += RES_GN ? fmaf(rr[e], sc2[e], sh2[e]) : rr[e];
        v[e] = a.relu ? fmaxf(o, 0.f) : o;
      }
      st_vec(a.y + base + static_cast<long long>(p) * C, pack<__nv_bfloat16>(v));
    }
  }
}

struct GnBwdArgs {
  const __nv_bfloat16* dout;   // gradient w.r.t. the pre-ReLU block output [images][HW][C]
  const __nv_bfloat16* x;      // raw conv output of the main branch
  const __nv_bfloat16* x2;     // raw conv output of the projection shortcut, or null
  __nv_bfloat16* dx;
  __nv_bfloat16* dx2;
  const float* slots;
  long long row_stride;
  float* sgrad;                // compact gradient rows [clients][sgrad_stride]
  long long sgrad_stride;
  long long dgamma, dbeta, dgamma2, dbeta2;
  GnRef gn, gn2;
  float* bsums;                // [images][8][2]: s1, s2 of the main branch
  float* bsums2;               // same for the shortcut branch
  int HW, C, imgs_per_client;
  float eps;
};

template <bool TWO>
__global__ void __launch_bounds__(256) gn_bwd_reduce_kernel(const GnBwdArgs a) {
  pdl_trigger();
  pdl_wait();      // programmatic dependent launch: see common.cuh
  __shared__ float scratch[256 * 8];
  __shared__ float red[GN_GROUPS * 4];
  const long long img = blockIdx.y;
  const int C = a.C, octs = C >> 3, oct = threadIdx.x % octs, r = threadIdx.x / octs, R = 256 / octs;
  const int grp = oct * 64 / C, c0 = oct * 8;
  const long long client = img / a.imgs_per_client;
  const float* row = a.slots + client * a.row_stride;
  const float inv_m = 1.f / (static_cast<float>(a.HW) * static_cast<float>(C / GN_GROUPS));
  float mean, rstd, mean2 = 0.f, rstd2 = 0.f;
  gn_moments(a.gn.stats, img, grp, inv_m, a.eps, mean, rstd);
  if constexpr (TWO) gn_moments(a.gn2.stats, img, grp, inv_m, a.eps, mean2, rstd2);
  const float mrs = mean * rstd, mrs2 = mean2 * rstd2;
  if (threadIdx.x < GN_GROUPS * 4) red[threadIdx.x] = 0.f;
  float sg[8], sgx[8], sgx2[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) sg[e] = sgx[e] = sgx2[e] = 0.f;
  const long long base = img * a.HW * C + c0;
  for (int p0 = blockIdx.x * R * GN_U; p0 < a.HW; p0 += gridDim.x * R * GN_U) {
    uint4 dv[GN_U], xv[GN_U], x2v[GN_U];
#pragma unroll
    for (int u = 0; u < GN_U; ++u) {
      const int p = p0 + u * R + r;
      dv[u] = xv[u] = x2v[u] = make_uint4(0, 0, 0, 0);
      if (p < a.HW) {
        const long long off = base + static_cast<long long>(p) * C;
        dv[u] = ld_stream(a.dout + off);
        xv[u] = ld_stream(a.x + off);
        if constexpr (TWO) x2v[u] = ld_stream(a.x2 + off);
      }
    }
#pragma unroll
    for (int u = 0; u < GN_U; ++u) {
      float d[8], xx[8];
      unpack<__nv_bfloat16>(dv[u], d);      // rows past the end hold zeros: they add nothing
      unpack<__nv_bfloat16>(xv[u], xx);
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        sg[e] += d[e];
        sgx[e] = fmaf(d[e], fmaf(xx[e], rstd, -mrs), sgx[e]);
      }
      if constexpr (TWO) {
        unpack<__nv_bfloat16>(x2v[u], xx);
#pragma unroll
        for (int e = 0; e < 8; ++e) sgx2[e] = fmaf(d[e], fmaf(xx[e], rstd2, -mrs2), sgx2[e]);
      }
    }
  }
  // group sums of this image: s1 = sum gamma g, s2 = sum gamma g xhat
  float s1 = 0.f, s2 = 0.f, t1 = 0.f, t2 = 0.f;
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const float gm = row[a.gn.gamma + c0 + e];
    s1 = fmaf(gm, sg[e], s1);
    s2 = fmaf(gm, sgx[e], s2);
    if constexpr (TWO) {
      const float gm2 = row[a.gn2.gamma + c0 + e];
      t1 = fmaf(gm2, sg[e], t1);
      t2 = fmaf(gm2, sgx2[e], t2);
    }
  }
  __syncthreads();
  atomicAdd(&red[grp * 4], s1);
  atomicAdd(&red[grp * 4 + 1], s2);
  if constexpr (TWO) {
    atomicAdd(&red[grp * 4 + 2], t1);
    atomicAdd(&red[grp * 4 + 3], t2);
  }
  __syncthreads();
  if (threadIdx.x < GN_GROUPS * 4) {
    const int gq = threadIdx.x >> 2, which = threadIdx.x & 3;
    if (which < 2) atomicAdd(a.bsums + (img * GN_GROUPS + gq) * 2 + which, red[threadIdx.x]);
    else if (TWO) atomicAdd(a.bsums2 + (img * GN_GROUPS + gq) * 2 + (which - 2), red[threadIdx.x]);
  }
  // parameter gradients: reduce over the R threads that share a channel octet, then one fp32
  // atomic per channel and block into the client's gradient row
  float* grow = a.sgrad + client * a.sgrad_stride;
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
__global__ void __launch_bounds__(256) gn_bwd_apply_kernel(const GnBwdArgs a) {
  pdl_trigger();
  pdl_wait();      // programmatic dependent launch: see common.cuh
  const long long img = blockIdx.y;
  const int C = a.C, octs = C >> 3, oct = threadIdx.x % octs, r = threadIdx.x / octs, R = 256 / octs;
  const int grp = oct * 64 / C, c0 = oct * 8;
  const float* row = a.slots + (img / a.imgs_per_client) * a.row_stride;
  const float inv_m = 1.f / (static_cast<float>(a.HW) * static_cast<float>(C / GN_GROUPS));
  // dx = rstd * (gamma g - s1 / m - (x - mean) rstd s2 / m) = ka[c] * g + kb * x + kc
  float mean, rstd;
  gn_moments(a.gn.stats, img, grp, inv_m, a.eps, mean, rstd);
  const float s1 = a.bsums[(img * GN_GROUPS + grp) * 2], s2 = a.bsums[(img * GN_GROUPS + grp) * 2 + 1];
  const float kb = -rstd * rstd * s2 * inv_m, kc = -kb * mean - rstd * s1 * inv_m;
  float ka[8], ka2[8], kb2 = 0.f, kc2 = 0.f;
#pragma unroll
  for (int e = 0; e < 8; ++e) { ka[e] = rstd * row[a.gn.gamma + c0 + e]; ka2[e] = 0.f; }
  if constexpr (TWO) {
    float mean2, rstd2;
    gn_moments(a.gn2.stats, img, grp, inv_m, a.eps, mean2, rstd2);
    const float t1 = a.bsums2[(img * GN_GROUPS + grp) * 2], t2 = a.bsums2[(img * GN_GROUPS + grp) * 2 + 1];
    kb2 = -rstd2 * rstd2 * t2 * inv_m;
    kc2 = -kb2 * mean2 - rstd2 * t1 * inv_m;
#pragma unroll
    for (int e = 0; e < 8; ++e) ka2[e] = rstd2 * row[a.gn2.gamma + c0 + e];
  }
  const long long base = img * a.HW * C + c0;
  for (int p0 = blockIdx.x * R * GN_U; p0 < a.HW; p0 += gridDim.x * R * GN_U) {
    uint4 dv[GN_U], xv[GN_U], x2v[GN_U];
#pragma unroll
    for (int u = 0; u < GN_U; ++u) {
      const int p = p0 + u * R + r;
      if (p < a.HW) {
        const long long off = base + static_cast<long long>(p) * C;
        dv[u] = ld_stream(a.dout + off);
        xv[u] = ld_stream(a.x + off);
        if constexpr (TWO) x2v[u] = ld_stream(a.x2 + off);
      }
    }
#pragma unroll
    for (int u = 0; u < GN_U; ++u) {
      const int p = p0 + u * R + r;
      if (p >= a.HW) continue;
      const long long off = base + static_cast<long long>(p) * C;
      float d[8], xx[8], o[8];
      unpack<__nv_bfloat16>(dv[u], d);
      unpack<__nv_bfloat16>(xv[u], xx);
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = fmaf(ka[e], d[e], fmaf(kb, xx[e], kc));
      st_vec(a.dx + off, pack<__nv_bfloat16>(o));
      if constexpr (TWO) {
        unpack<__nv_bfloat16>(x2v[u], xx);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = fmaf(ka2[e], d[e], fmaf(kb2, xx[e], kc2));
        st_vec(a.dx2 + off, pack<__nv_bfloat16>(o));
      }
    }
  }
}

// position splits per image: enough blocks to fill the machine on the wide early layers, one block
// per image where an image is a few kilobytes
static int gn_splits(int HW, int C, int iters = 1) {
  const int rows_per_block = (256 / (C / 8)) * GN_U * iters;
  int s = (HW + rows_per_block - 1) / rows_per_block;
  return s < 1 ? 1 : (s > 8 ? 8 : s);
}

}  // namespace rn
}  // namespace nfb

extern "C" {

int nfb_rn_gn_stats(const void* x, int imgs, int HW, int C, float* stats, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(x && stats && imgs >= 1 && HW >= 1, NFB_EINVAL, "rn_gn_stats: bad argument");
  NFB_REQUIRE(C % 64 == 0 && C <= 512 && 256 % (C / 8) == 0, NFB_EINVAL, "rn_gn_stats: bad channels");
  NFB_REQUIRE(imgs <= 65535, NFB_ERANGE, "rn_gn_stats: too many images for one launch");
  launch_pdl(rn::gn_stats_kernel, dim3(dim3(rn::gn_splits(HW, C, 4), imgs)), dim3(256), 0, as_stream(stream), 
      static_cast<const __nv_bfloat16*>(x), HW, C, stats);
  return launch_status("rn_gn_stats");
}

int nfb_rn_gn_apply(const void* x, const void* res, void* y, const float* slots,
                    long long row_stride, const float* stats, const long long* gn_offs,
                    const float* res_stats, const long long* res_offs, int imgs,
                    int imgs_per_client, int HW, int C, int relu, float eps, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(x && y && slots && stats && gn_offs, NFB_EINVAL, "rn_gn_apply: null pointer");
  NFB_REQUIRE(C % 64 == 0 && C <= 512 && 256 % (C / 8) == 0, NFB_EINVAL, "rn_gn_apply: bad channels");
  NFB_REQUIRE(imgs >= 1 && imgs <= 65535 && imgs_per_client >= 1, NFB_ERANGE, "rn_gn_apply: images");
  NFB_REQUIRE(!res_stats || (res && res_offs), NFB_EINVAL, "rn_gn_apply: residual branch incomplete");
  rn::GnApplyArgs a{};
  a.x = static_cast<const __nv_bfloat16*>(x);
  a.res = static_cast<const __nv_bfloat16*>(res);
  a.y = static_cast<__nv_bfloat16*>(y);
  a.slots = slots; a.row_stride = row_stride;
  a.gn = rn::GnRef{stats, gn_offs[0], gn_offs[1]};
  if (res_stats) a.gn_res = rn::GnRef{res_stats, res_offs[0], res_offs[1]};
  a.HW = HW; a.C = C; a.imgs_per_client = imgs_per_client; a.relu = relu; a.eps = eps;
  const dim3 grid(rn::gn_splits(HW, C, 2), imgs);
  cudaStream_t st = as_stream(stream);
  if (!res) launch_pdl(rn::gn_apply_kernel<false, false>, dim3(grid), dim3(256), 0, st, a);
  else if (!res_stats) launch_pdl(rn::gn_apply_kernel<true, false>, dim3(grid), dim3(256), 0, st, a);
  else launch_pdl(rn::gn_apply_kernel<true, true>, dim3(grid), dim3(256), 0, st, a);
  return launch_status("rn_gn_apply");
}

int nfb_rn_gn_bwd(int apply, const void* dout, const void* x, const void* x2, void* dx, void* dx2,
                  const float* slots, long long row_stride, float* sgrad, long long sgrad_stride,
                  const long long* grad_offs, const float* stats, const long long* gn_offs,
                  const float* stats2, const long long* gn2_offs, float* bsums, float* bsums2,
                  int imgs, int imgs_per_client, int HW, int C, float eps, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(dout && x && slots && sgrad && grad_offs && stats && gn_offs && bsums, NFB_EINVAL,
              "rn_gn_bwd: null pointer");
  NFB_REQUIRE(C % 64 == 0 && C <= 512 && 256 % (C / 8) == 0, NFB_EINVAL, "rn_gn_bwd: bad channels");
  NFB_REQUIRE(imgs >= 1 && imgs <= 65535 && imgs_per_client >= 1, NFB_ERANGE, "rn_gn_bwd: images");
  NFB_REQUIRE(!x2 || (stats2 && gn2_offs && bsums2), NFB_EINVAL, "rn_gn_bwd: second branch incomplete");
  NFB_REQUIRE(!apply || (dx && (!x2 || dx2)), NFB_EINVAL, "rn_gn_bwd: apply needs outputs");
  rn::GnBwdArgs a{};
  a.dout = static_cast<const __nv_bfloat16*>(dout);
  a.x = static_cast<const __nv_bfloat16*>(x);
  a.x2 = static_cast<const __nv_bfloat16*>(x2);
  a.dx = static_cast<__nv_bfloat16*>(dx);
  a.dx2 = static_cast<__nv_bfloat16*>(dx2);
  a.slots = slots; a.row_stride = row_stride; a.sgrad = sgrad; a.sgrad_stride = sgrad_stride;
  a.dgamma = grad_offs[0]; a.dbeta = grad_offs[1]; a.dgamma2 = grad_offs[2]; a.dbeta2 = grad_offs[3];
  a.gn = rn::GnRef{stats, gn_offs[0], gn_offs[1]};
  if (x2) a.gn2 = rn::GnRef{stats2, gn2_offs[0], gn2_offs[1]};
  a.bsums = bsums; a.bsums2 = bsums2;
  a.HW = HW; a.C = C; a.imgs_per_client = imgs_per_client; a.eps = eps;
  const dim3 grid(rn::gn_splits(HW, C, 2), imgs);
  cudaStream_t st = as_stream(stream);
  if (!apply) {
    // the reduce pass has a fixed cost per block (moments, two shared-memory reductions, C
    // atomics): four trips of the load loop per block instead of one
    const dim3 rgrid(rn::gn_splits(HW, C, 4), imgs);
    if (x2) launch_pdl(rn::gn_bwd_reduce_kernel<true>, rgrid, dim3(256), 0, st, a);
    else launch_pdl(rn::gn_bwd_reduce_kernel<false>, rgrid, dim3(256), 0, st, a);
  } else {
    if (x2) launch_pdl(rn::gn_bwd_apply_kernel<true>, dim3(grid), dim3(256), 0, st, a);
    else launch_pdl(rn::gn_bwd_apply_kernel<false>, dim3(grid), dim3(256), 0, st, a);
  }
  return launch_status(apply ? "rn_gn_bwd_apply" : "rn_gn_bwd_reduce");
}

}  // extern "C"
