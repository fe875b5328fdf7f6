// K-ELT / K-XENT for the flagship MNIST-CNN cohort step: every non-GEMM op of
// nanofed/models/mnist.py:16-28 and its backward, written for C clients at once.
//
// Layout conventions (shared with ops/tc.py and models/mnist_cohort.py):
//   x     [C, B, 28, 28]        fp32 input images
//   a1    [C, B*676, 32]  bf16  conv1+ReLU output, NHWC on a 26x26 grid (row = b*676+oh*26+ow)
//   a2    [C, B*676, 64]  bf16  conv2+bias+ReLU on the SAME grid (valid for oh,ow < 24)
//   p     [C, B, 9216]    bf16  maxpool+dropout output in PyTorch flatten order ch*144+ph*12+pw
//   pidx  [C, B, 144, 64] u8    bits 0-1 argmax in the 2x2 window, bit 2 max>0, bit 3 dropout keep
//   hpre  [C, B, 128]     fp32  fc1 pre-activation (without bias)
//   dh    [C, B, 128]     bf16  gradient w.r.t. hpre
//   dp    [C, B, 9216]    bf16  gradient w.r.t. p
//   da2   [C, B*676, 64]  bf16  gradient w.r.t. conv2 pre-activation, ZERO outside oh,ow < 24
//   da1   [C, B*676, 32]  bf16  gradient w.r.t. conv1 pre-activation (ReLU mask applied)
// Weights/gradients are addressed inside the flat arena rows through element offsets.
#include "common.cuh"
#include "philox.cuh"

namespace nfb {

constexpr int kG = 26;          // grid width/height after conv1
constexpr int kGrid = kG * kG;  // 676 positions per image
constexpr int kPool = 12;

// -------------------------------------------------------------------------------------
// conv2 weight permutation: arena fp32 [co][ci][t] -> bf16 [co][t*32+ci]
// -------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    prep_conv2_weight_kernel(const float* __restrict__ slots, int64_t row_stride, int64_t w_off,
                             __nv_bfloat16* __restrict__ w2p) {
  const int c = blockIdx.y;
  const float* w = slots + c * row_stride + w_off;
  __nv_bfloat16* o = w2p + static_cast<int64_t>(c) * 64 * 288;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 64 * 288; i += gridDim.x * blockDim.x) {
    const int co = i / 288, rem = i % 288, t = rem / 32, ci = rem % 32;
    o[i] = __float2bfloat16_rn(w[co * 288 + ci * 9 + t]);
  }
}

// -------------------------------------------------------------------------------------
// conv1 (1 -> 32, 3x3) + bias + ReLU, direct on CUDA cores (K = 9 is too thin for UMMA)
// one thread = one grid position x 8 output channels; one block = 256 threads
// -------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    conv1_fwd_kernel(const float* __restrict__ x, int64_t x_cstride,
                     const float* __restrict__ slots, int64_t row_stride, int64_t w_off,
                     int64_t b_off, int B, __nv_bfloat16* __restrict__ a1) {
  // item = (image row-pair slot: b, oh, ow-pair (13 per row), channel octet): two horizontally
  // adjacent outputs share their 3x4 input window and every weight read (the kernel is
  // LSU-bound: ncu showed l1tex at 88 % with one output per item).
  __shared__ __align__(16) float swt[9 * 32];  // [tap][co]
  __shared__ float sb[32];
  const int c = blockIdx.y;
  const float* w = slots + c * row_stride + w_off;
  for (int i = threadIdx.x; i < 288; i += blockDim.x) swt[(i % 9) * 32 + i / 9] = w[i];
  if (threadIdx.x < 32) sb[threadIdx.x] = slots[c * row_stride + b_off + threadIdx.x];
  __syncthreads();
  const int64_t items = static_cast<int64_t>(B) * kG * 13 * 4;
  const float* xc = x + static_cast<int64_t>(c) * x_cstride;
  __nv_bfloat16* oc = a1 + static_cast<int64_t>(c) * B * kGrid * 32;
  for (int64_t it = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; it < items;
       it += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int oct = static_cast<int>(it & 3);
    const int64_t pr = it >> 2;                       // (b, oh, pair)
    const int pair = static_cast<int>(pr % 13);
    const int64_t rowi = pr / 13;                     // b * 26 + oh
    const int b = static_cast<int>(rowi / kG), oh = static_cast<int>(rowi % kG);
    const int ow = pair * 2;
    const float* px = xc + b * 784 + oh * 28 + ow;
    float in[3][4];
#pragma unroll
    for (int kh = 0; kh < 3; ++kh) {
      const float2 lo = *reinterpret_cast<const float2*>(px + kh * 28);      // ow even => 8-B aligned
      const float2 hi = *reinterpret_cast<const float2*>(px + kh * 28 + 2);
      in[kh][0] = lo.x; in[kh][1] = lo.y; in[kh][2] = hi.x; in[kh][3] = hi.y;
    }
    float o0[8], o1[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o0[j] = o1[j] = sb[oct * 8 + j];
#pragma unroll
    for (int kh = 0; kh < 3; ++kh)
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const int t = kh * 3 + kw;
        const float4 w0 = *reinterpret_cast<const float4*>(swt + t * 32 + oct * 8);
        const float4 w1 = *reinterpret_cast<const float4*>(swt + t * 32 + oct * 8 + 4);
        const float a = in[kh][kw], bb = in[kh][kw + 1];
        o0[0] = fmaf(a, w0.x, o0[0]); o0[1] = fmaf(a, w0.y, o0[1]);
        o0[2] = fmaf(a, w0.z, o0[2]); o0[3] = fmaf(a, w0.w, o0[3]);
        o0[4] = fmaf(a, w1.x, o0[4]); o0[5] = fmaf(a, w1.y, o0[5]);
        o0[6] = fmaf(a, w1.z, o0[6]); o0[7] = fmaf(a, w1.w, o0[7]);
        o1[0] = fmaf(bb, w0.x, o1[0]); o1[1] = fmaf(bb, w0.y, o1[1]);
        o1[2] = fmaf(bb, w0.z, o1[2]); o1[3] = fmaf(bb, w0.w, o1[3]);
        o1[4] = fmaf(bb, w1.x, o1[4]); o1[5] = fmaf(bb, w1.y, o1[5]);
        o1[6] = fmaf(bb, w1.z, o1[6]); o1[7] = fmaf(bb, w1.w, o1[7]);
      }
#pragma unroll
    for (int j = 0; j < 8; ++j) { o0[j] = fmaxf(o0[j], 0.f); o1[j] = fmaxf(o1[j], 0.f); }
    const int64_t pos = rowi * kG + ow;               // == b*676 + oh*26 + ow
    st_vec(oc + pos * 32 + oct * 8, pack<__nv_bfloat16>(o0));
    st_vec(oc + (pos + 1) * 32 + oct * 8, pack<__nv_bfloat16>(o1));
  }
}

// -------------------------------------------------------------------------------------
// 2x2 max-pool + dropout(p1) forward; one block per image, staged through shared memory so
// that both the NHWC read and the PyTorch-order (ch*144 + pos) write are coalesced
// -------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    pool_fwd_kernel(const __nv_bfloat16* __restrict__ a2, int B, float drop_p,
                    const unsigned long long* __restrict__ rng_state, uint64_t seed,
                    __nv_bfloat16* __restrict__ p, uint8_t* __restrict__ pidx) {
  __shared__ __nv_bfloat16 sp[144 * 66];  // [pos][ch], row padded to 66 to avoid bank conflicts
  const int b = blockIdx.x, c = blockIdx.y;
  const int64_t img = static_cast<int64_t>(c) * B + b;
  const __nv_bfloat16* src = a2 + img * kGrid * 64;
  uint8_t* idx = pidx + img * 144 * 64;
  const uint64_t step = rng_state ? *rng_state : 0ull;
  const float keep_scale = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  // item = (pooled position, channel octet): 144 * 8 items
  for (int it = threadIdx.x; it < 144 * 8; it += blockDim.x) {
    const int oct = it & 7, pos = it >> 3;
    const int ph = pos / kPool, pw = pos % kPool;
    const int q = (2 * ph) * kG + 2 * pw;
    float v[4][8];
    unpack<__nv_bfloat16>(ld_plain(src + (q)*64 + oct * 8), v[0]);
    unpack<__nv_bfloat16>(ld_plain(src + (q + 1) * 64 + oct * 8), v[1]);
    unpack<__nv_bfloat16>(ld_plain(src + (q + kG) * 64 + oct * 8), v[2]);
    unpack<__nv_bfloat16>(ld_plain(src + (q + kG + 1) * 64 + oct * 8), v[3]);
    // dropout keep bits: one Philox group (4 uniforms) per 4 channels
    const uint64_t elem = (img * 144 + pos) * 64 + oct * 8;
    float u[8];
    if (drop_p > 0.f) {
      const float4 u0 = Philox::uniform4((step << 34) + (elem >> 2), seed);
      const float4 u1 = Philox::uniform4((step << 34) + (elem >> 2) + 1, seed);
      u[0] = u0.x; u[1] = u0.y; u[2] = u0.z; u[3] = u0.w;
      u[4] = u1.x; u[5] = u1.y; u[6] = u1.z; u[7] = u1.w;
    }
    uint8_t code[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float m = v[0][j];
      int a = 0;
#pragma unroll
      for (int w = 1; w < 4; ++w)
        if (v[w][j] > m) { m = v[w][j]; a = w; }
      const bool keep = drop_p > 0.f ? (u[j] >= drop_p) : true;
      code[j] = static_cast<uint8_t>(a | (m > 0.f ? 4 : 0) | (keep ? 8 : 0));
      sp[pos * 66 + oct * 8 + j] = __float2bfloat16_rn(keep ? m * keep_scale : 0.f);
    }
    *reinterpret_cast<uint2*>(idx + pos * 64 + oct * 8) = *reinterpret_cast<const uint2*>(code);
  }
  __syncthreads();
  // write out in PyTorch flatten order: out[ch*144 + pos]
  __nv_bfloat16* dst = p + img * 9216;
  for (int i8 = threadIdx.x; i8 < 9216 / 8; i8 += blockDim.x) {   // 144 % 8 == 0
    const int i = i8 * 8;
    const int ch = i / 144, pos = i % 144;
    __align__(16) __nv_bfloat16 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = sp[(pos + j) * 66 + ch];
    *reinterpret_cast<uint4*>(dst + i) = *reinterpret_cast<const uint4*>(v);
  }
}

// -------------------------------------------------------------------------------------
// backward of dropout + max-pool + ReLU: dp -> da2 on the 26-grid (zeros elsewhere),
// plus conv2 bias gradient (column sums, atomics into the gradient slab)
// -------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    pool_bwd_kernel(const __nv_bfloat16* __restrict__ dp, const uint8_t* __restrict__ pidx, int B,
                    float drop_p, __nv_bfloat16* __restrict__ da2, float* __restrict__ grads,
                    int64_t row_stride, int64_t b2_off) {
  // The first version was issue-bound (ncu: 74 % issue slots, 57.6 M warp instructions): per
  // output vector it tested 8 code bytes one by one, read 8 two-byte values and re-rounded 8
  // values for the bias sum.  Now the 8 codes are compared with two byte-SIMD instructions, the
  // values come as one 16-byte shared load and are masked as bf16 pairs, and the bias gradient is
  // summed over the 144 pooled positions (each contributes once) instead of the 676 grid ones.
  //
  // sd = [pos][ch] gradient w.r.t. the pooled pre-dropout value, rows of 72 bf16 (16-byte
  // aligned); the channel octet is XOR-swizzled with the position octet so that the transposing
  // two-byte stores of phase 1 spread over banks.
  __shared__ __align__(16) __nv_bfloat16 sd[144 * 72];
  __shared__ float sbias[64];
  const int b = blockIdx.x, c = blockIdx.y;
  const int64_t img = static_cast<int64_t>(c) * B + b;
  const __nv_bfloat16* src = dp + img * 9216;
  const uint8_t* idx = pidx + img * 144 * 64;
  const float keep_scale = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  const bool scale = keep_scale != 1.f;   // false when the producer of dp already applied it
  if (threadIdx.x < 64) sbias[threadIdx.x] = 0.f;
  for (int i8 = threadIdx.x; i8 < 9216 / 8; i8 += blockDim.x) {   // 16-byte loads, 144 % 8 == 0
    const int i = i8 * 8;
    const int ch = i / 144, pos = i % 144;
    const uint4 raw = ld_stream(src + i);
    const __nv_bfloat16* v = reinterpret_cast<const __nv_bfloat16*>(&raw);
    const int col = (((ch >> 3) ^ ((pos >> 3) & 7)) << 3) + (ch & 7);
#pragma unroll
    for (int j = 0; j < 8; ++j) sd[(pos + j) * 72 + col] = v[j];
  }
  __syncthreads();
  // 8 channels of pooled position `pos`, masked by an 8-byte compare result (0xff = keep)
  auto masked = [&](int pos, int oct, uint32_t m0, uint32_t m1) {
    uint4 v = *reinterpret_cast<const uint4*>(sd + pos * 72 + ((oct ^ ((pos >> 3) & 7)) << 3));
    if (scale) {
      float f[8];
      unpack<__nv_bfloat16>(v, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] *= keep_scale;
      v = pack<__nv_bfloat16>(f);
    }
    v.x &= __byte_perm(m0, 0, 0x1100);
    v.y &= __byte_perm(m0, 0, 0x3322);
    v.z &= __byte_perm(m1, 0, 0x1100);
    v.w &= __byte_perm(m1, 0, 0x3322);
    return v;
  };
  // conv2 bias gradient: every pooled element with ReLU and keep bits set reaches exactly one
  // grid position.  blockDim.x % 8 == 0, so a thread keeps its channel octet.
  float bsum[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  for (int it = threadIdx.x; it < 144 * 8; it += blockDim.x) {
    const int oct = it & 7, pos = it >> 3;
    const uint2 codes = *reinterpret_cast<const uint2*>(idx + pos * 64 + oct * 8);
    const uint4 v = masked(pos, oct, __vcmpeq4(codes.x & 0x0c0c0c0cu, 0x0c0c0c0cu),
                           __vcmpeq4(codes.y & 0x0c0c0c0cu, 0x0c0c0c0cu));
    float f[8];
    unpack<__nv_bfloat16>(v, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) bsum[j] += f[j];
  }
  __nv_bfloat16* dst = da2 + img * kGrid * 64;
  // item = (grid position, channel octet): 676 * 8 items, every position is written; four items
  // per trip with their code loads issued together
  for (int it0 = threadIdx.x; it0 < kGrid * 8; it0 += 4 * blockDim.x) {
    uint2 codes[4];
    int pos[4];
    uint32_t want[4];
    bool live[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int it = it0 + u * blockDim.x;
      live[u] = it < kGrid * 8;
      const int q = it >> 3, oh = q / kG, ow = q % kG;
      const bool inside = live[u] && oh < 24 && ow < 24;
      pos[u] = inside ? (oh >> 1) * kPool + (ow >> 1) : 0;
      // code byte: bits 0-1 argmax slot, bit 2 ReLU passed, bit 3 kept by dropout
      want[u] = 0x0c0c0c0cu | (static_cast<uint32_t>((oh & 1) * 2 + (ow & 1)) * 0x01010101u);
      codes[u] = make_uint2(0u, 0u);                 // never equals `want`: zeros outside
      if (inside) codes[u] = *reinterpret_cast<const uint2*>(idx + pos[u] * 64 + (it & 7) * 8);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (!live[u]) continue;
      const int it = it0 + u * blockDim.x;
      const uint4 v = masked(pos[u], it & 7, __vcmpeq4(codes[u].x & 0x0f0f0f0fu, want[u]),
                             __vcmpeq4(codes[u].y & 0x0f0f0f0fu, want[u]));
      st_vec(dst + (it >> 3) * 64 + (it & 7) * 8, v);
    }
  }
  {
    const int oct = threadIdx.x & 7;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      // lanes with equal (lane & 7) share an octet: reduce over lane bits 3,4 first
      float v = bsum[j];
      v += __shfl_xor_sync(0xffffffffu, v, 8);
      v += __shfl_xor_sync(0xffffffffu, v, 16);
      if ((threadIdx.x & 31) < 8 && v != 0.f) atomicAdd(&sbias[oct * 8 + j], v);
    }
  }
  __syncthreads();
  if (threadIdx.x < 64 && sbias[threadIdx.x] != 0.f)
    atomicAdd(grads + c * row_stride + b2_off + threadIdx.x, sbias[threadIdx.x]);
}

// -------------------------------------------------------------------------------------
// The classifier head, one block per client:
//   h = dropout(relu(hpre + b_fc1)) ; logits = h Wfc2^T + b_fc2 ; log-softmax / NLL / accuracy
//   dlogits = (softmax - onehot)/B ; dWfc2, dbfc2 ; dh = (dlogits Wfc2) * relu'(.) * dropout ;
//   dbfc1 = colsum(dh).   Accumulates (loss_sum, correct, B) into stats[c] — the device-side
// counters that replace the reference's two .item() syncs per batch — and bumps the
// per-step dropout counter.
// -------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    fc_head_kernel(float* __restrict__ hpre, const int64_t* __restrict__ labels,
                   int64_t label_stride, int B, const float* __restrict__ slots,
                   float* __restrict__ grads, int64_t row_stride, int64_t b1_off, int64_t w2_off,
                   int64_t b2_off, float drop_p, unsigned long long* __restrict__ rng_state,
                   int advance_rng, uint64_t seed, __nv_bfloat16* __restrict__ dh,
                   float* __restrict__ stats) {
  // grid (clients, splits): block (c, s) owns batch rows [r0, r1) of client c.  With one block
  // per client (the first version) 32 of 148 SMs worked for 34 us; parameter-gradient partial
  // sums of the row blocks meet in fp32 atomics (targets are zeroed by K-OPT after use).
  extern __shared__ float smem[];
  const int splits = gridDim.y;
  const int rows_per = (B + splits - 1) / splits;
  const int r0 = blockIdx.y * rows_per;
  const int nb = max(min(B, r0 + rows_per) - r0, 0);
  float* h = smem;                 // [rows_per][128]   (later reused for dh in fp32)
  float* w2 = h + rows_per * 128;  // [10][128]
  float* lg = w2 + 1280;           // [rows_per][10] logits -> dlogits
  float* red = lg + rows_per * 10; // [64]
  const int c = blockIdx.x;
  const float* row = slots + c * row_stride;
  float* grow = grads + c * row_stride;
  const uint64_t step = rng_state ? rng_state[0] : 0ull;
  const float keep_scale = drop_p > 0.f ? 1.f / (1.f - drop_p) : 1.f;
  const bool shared_targets = splits > 1;

  for (int i = threadIdx.x; i < 1280; i += blockDim.x) w2[i] = row[w2_off + i];
  float* hp = hpre + (static_cast<int64_t>(c) * B + r0) * 128;
  for (int i4 = threadIdx.x; i4 < nb * 32; i4 += blockDim.x) {   // 4 elements per Philox draw
    const int i = i4 * 4;
    const float4 hv = *reinterpret_cast<const float4*>(hp + i);
    // hpre is a split-K atomic accumulator: leave it zeroed for the next step's fc1 forward
    *reinterpret_cast<float4*>(hp + i) = make_float4(0.f, 0.f, 0.f, 0.f);
    const float4 bv = *reinterpret_cast<const float4*>(row + b1_off + (i & 127));
    float v[4] = {fmaxf(hv.x + bv.x, 0.f), fmaxf(hv.y + bv.y, 0.f), fmaxf(hv.z + bv.z, 0.f),
                  fmaxf(hv.w + bv.w, 0.f)};
    if (drop_p > 0.f) {
      const uint64_t grp = (static_cast<uint64_t>(c) * B + r0) * 32 + i4;
      const float4 u = Philox::uniform4((step << 34) + (1ull << 33) + grp, seed);
      v[0] = u.x >= drop_p ? v[0] * keep_scale : 0.f;
      v[1] = u.y >= drop_p ? v[1] * keep_scale : 0.f;
      v[2] = u.z >= drop_p ? v[2] * keep_scale : 0.f;
      v[3] = u.w >= drop_p ? v[3] * keep_scale : 0.f;
    }
    *reinterpret_cast<float4*>(h + i) = make_float4(v[0], v[1], v[2], v[3]);
  }
  __syncthreads();
  // logits
  // one warp per (row, class) dot product: lane l takes elements 4l..4l+3 of both vectors (16-byte
  // conflict-free reads).  One THREAD per dot product read h[b][j] for 16 different b at the same
  // j - a 16-way bank conflict on every one of its 128 loads.
  for (int i = threadIdx.x >> 5; i < nb * 10; i += blockDim.x >> 5) {
    const int b = i / 10, k = i % 10, l4 = (threadIdx.x & 31) * 4;
    const float4 hv = *reinterpret_cast<const float4*>(h + b * 128 + l4);
    const float4 wv = *reinterpret_cast<const float4*>(w2 + k * 128 + l4);
    float acc = fmaf(hv.x, wv.x, fmaf(hv.y, wv.y, fmaf(hv.z, wv.z, hv.w * wv.w)));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) lg[i] = acc + row[b2_off + k];
  }
  __syncthreads();
  // softmax / loss / accuracy / dlogits : one thread per sample
  float my_loss = 0.f, my_hit = 0.f;
  for (int b = threadIdx.x; b < nb; b += blockDim.x) {
    float* l = lg + b * 10;
    const int t = static_cast<int>(labels[c * label_stride + r0 + b]);
    float mx = l[0];
    int arg = 0;
    for (int k = 1; k < 10; ++k)
      if (l[k] > mx) { mx = l[k]; arg = k; }
    float se = 0.f;
    for (int k = 0; k < 10; ++k) se += __expf(l[k] - mx);
    const float lse = mx + __logf(se);
    my_loss += lse - l[t];
    my_hit += (arg == t) ? 1.f : 0.f;
    const float inv = 1.f / se, sc = 1.f / static_cast<float>(B);
    for (int k = 0; k < 10; ++k) l[k] = (__expf(l[k] - mx) * inv - (k == t ? 1.f : 0.f)) * sc;
  }
  my_loss = block_sum(my_loss, red);
  my_hit = block_sum(my_hit, red);
  if (threadIdx.x == 0 && nb > 0) {
    if (shared_targets) {
      atomicAdd(&stats[3 * c + 0], my_loss);
      atomicAdd(&stats[3 * c + 1], my_hit);
      atomicAdd(&stats[3 * c + 2], static_cast<float>(nb));
    } else {
      stats[3 * c + 0] += my_loss;
      stats[3 * c + 1] += my_hit;
      stats[3 * c + 2] += static_cast<float>(nb);
    }
  }
  // fc2 gradients over this block's rows
  for (int i = threadIdx.x; i < 1280; i += blockDim.x) {
    const int k = i >> 7, j = i & 127;
    float acc = 0.f;
    for (int b = 0; b < nb; ++b) acc = fmaf(lg[b * 10 + k], h[b * 128 + j], acc);
    if (shared_targets) atomicAdd(&grow[w2_off + i], acc);
    else grow[w2_off + i] = acc;
  }
  if (threadIdx.x < 10) {
    float acc = 0.f;
    for (int b = 0; b < nb; ++b) acc += lg[b * 10 + threadIdx.x];
    if (shared_targets) atomicAdd(&grow[b2_off + threadIdx.x], acc);
    else grow[b2_off + threadIdx.x] = acc;
  }
  __syncthreads();
  // dh (gradient w.r.t. hpre), written in place over h
  __nv_bfloat16* dhc = dh + (static_cast<int64_t>(c) * B + r0) * 128;
  for (int i = threadIdx.x; i < nb * 128; i += blockDim.x) {
    const int b = i >> 7, j = i & 127;
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 10; ++k) acc = fmaf(lg[b * 10 + k], w2[k * 128 + j], acc);
    const float g = h[i] > 0.f ? acc * keep_scale : 0.f;  // relu' and dropout share h > 0
    const __nv_bfloat16 gb = __float2bfloat16_rn(g);
    dhc[i] = gb;
    h[i] = __bfloat162float(gb);
  }
  __syncthreads();
  if (threadIdx.x < 128) {
    float acc = 0.f;
    for (int b = 0; b < nb; ++b) acc += h[b * 128 + threadIdx.x];
    if (shared_targets) atomicAdd(&grow[b1_off + threadIdx.x], acc);
    else grow[b1_off + threadIdx.x] = acc;
  }
  // dropout step counter: bumped by whichever block finishes LAST (rng_state[1] counts arrivals
  // and is left at zero), so no block can read the new value as its own step
  if (advance_rng && rng_state && threadIdx.x == 0) {
    __threadfence();
    const unsigned long long total = static_cast<unsigned long long>(gridDim.x) * gridDim.y;
    if (atomicAdd(&rng_state[1], 1ull) == total - 1) {
      rng_state[1] = 0ull;
      rng_state[0] = step + 1;
    }
  }
}

// -------------------------------------------------------------------------------------
// conv1 weight/bias gradient, one block per image:
//   dW1[co][t] += sum_q da1[q][co] * x[q + s_t] ;  db1[co] += sum_q da1[q][co]
// -------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
    conv1_wgrad_kernel(const __nv_bfloat16* __restrict__ da1, const float* __restrict__ x,
                       int64_t x_cstride, int B, float* __restrict__ grads, int64_t row_stride,
                       int64_t w_off, int64_t b_off) {
  // half-warp = one position (16 lanes x 2 channels, one 4-byte load each: a full 64-byte NHWC
  // row); every x-window value read from shared memory feeds two channels (the one-channel
  // version was LSU-bound: 9 LDS per 10 FMA, l1tex 73 % in ncu).  16 position slots per block.
  __shared__ float sx[784];
  __shared__ float part[16][32][10];
  const int b = blockIdx.x, c = blockIdx.y;
  const int64_t img = static_cast<int64_t>(c) * B + b;
  for (int i = threadIdx.x; i < 784; i += blockDim.x)
    sx[i] = x[c * x_cstride + static_cast<int64_t>(b) * 784 + i];
  __syncthreads();
  const int cp = threadIdx.x & 15, slot = threadIdx.x >> 4;     // channels 2cp, 2cp+1
  const __nv_bfloat162* d = reinterpret_cast<const __nv_bfloat162*>(da1 + img * kGrid * 32) + cp;
  float a0[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, a1[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
  auto tap = [&](int q, __nv_bfloat162 gv) {
    const float2 g = __bfloat1622float2(gv);
    const int oh = q / kG, ow = q % kG;
    const float* px = sx + oh * 28 + ow;
#pragma unroll
    for (int kh = 0; kh < 3; ++kh)
#pragma unroll
      for (int kw = 0; kw < 3; ++kw) {
        const float xv = px[kh * 28 + kw];
        a0[kh * 3 + kw] = fmaf(g.x, xv, a0[kh * 3 + kw]);
        a1[kh * 3 + kw] = fmaf(g.y, xv, a1[kh * 3 + kw]);
      }
    a0[9] += g.x;
    a1[9] += g.y;
  };
  int q = slot;
  for (; q + 48 < kGrid; q += 64) {                 // 4 positions in flight per thread
    const __nv_bfloat162 g0 = d[(q)*16], g1 = d[(q + 16) * 16], g2 = d[(q + 32) * 16],
                         g3 = d[(q + 48) * 16];
    tap(q, g0); tap(q + 16, g1); tap(q + 32, g2); tap(q + 48, g3);
  }
  for (; q < kGrid; q += 16) tap(q, d[q * 16]);
#pragma unroll
  for (int t = 0; t < 10; ++t) { part[slot][2 * cp][t] = a0[t]; part[slot][2 * cp + 1][t] = a1[t]; }
  __syncthreads();
  for (int i = threadIdx.x; i < 320; i += blockDim.x) {
    const int cc = i / 10, t = i % 10;
    float s = 0.f;
#pragma unroll
    for (int sl = 0; sl < 16; ++sl) s += part[sl][cc][t];
    float* dst = grads + c * row_stride + (t < 9 ? w_off + cc * 9 + t : b_off + cc);
    atomicAdd(dst, s);
  }
}

}  // namespace nfb

// =====================================================================================
// C entry points
// =====================================================================================
extern "C" {

int nfb_mnist_prep_conv2(const float* slots, int64_t row_stride, int64_t w_off, int C,
                         void* w2p, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(slots && w2p && C >= 1, NFB_EINVAL, "mnist_prep_conv2: bad argument");
  prep_conv2_weight_kernel<<<dim3(18, C), 256, 0, as_stream(stream)>>>(
      slots, row_stride, w_off, static_cast<__nv_bfloat16*>(w2p));
  return launch_status("mnist_prep_conv2");
}

int nfb_mnist_conv1_fwd(const float* x, int64_t x_cstride, const float* slots, int64_t row_stride,
                        int64_t w_off, int64_t b_off, int C, int B, void* a1, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(x && slots && a1 && C >= 1 && B >= 1, NFB_EINVAL, "mnist_conv1_fwd: bad argument");
  const int64_t items = static_cast<int64_t>(B) * kG * 13 * 4;
  // ~8 items per thread: the per-block weight staging is amortised, and 148*8/C blocks per
  // client still fill the machine
  int bx = static_cast<int>((items + 2047) / 2048);
  if (bx > 1024) bx = 1024;
  if (bx < 1) bx = 1;
  conv1_fwd_kernel<<<dim3(bx, C), 256, 0, as_stream(stream)>>>(
      x, x_cstride, slots, row_stride, w_off, b_off, B, static_cast<__nv_bfloat16*>(a1));
  return launch_status("mnist_conv1_fwd");
}

int nfb_mnist_pool_fwd(const void* a2, int C, int B, float drop_p, const void* rng_state,
                       uint64_t seed, void* p, uint8_t* pidx, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(a2 && p && pidx && C >= 1 && B >= 1, NFB_EINVAL, "mnist_pool_fwd: bad argument");
  pool_fwd_kernel<<<dim3(B, C), 256, 0, as_stream(stream)>>>(
      static_cast<const __nv_bfloat16*>(a2), B, drop_p,
      static_cast<const unsigned long long*>(rng_state), seed, static_cast<__nv_bfloat16*>(p),
      pidx);
  return launch_status("mnist_pool_fwd");
}

int nfb_mnist_pool_bwd(const void* dp, const uint8_t* pidx, int C, int B, float drop_p, void* da2,
                       float* grads, int64_t row_stride, int64_t b2_off, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(dp && pidx && da2 && grads, NFB_EINVAL, "mnist_pool_bwd: null pointer");
  pool_bwd_kernel<<<dim3(B, C), 256, 0, as_stream(stream)>>>(
      static_cast<const __nv_bfloat16*>(dp), pidx, B, drop_p, static_cast<__nv_bfloat16*>(da2),
      grads, row_stride, b2_off);
  return launch_status("mnist_pool_bwd");
}

int nfb_mnist_fc_head(float* hpre, const int64_t* labels, int64_t label_stride, int C,
                      int B, const float* slots, float* grads, int64_t row_stride,
                      int64_t b1_off, int64_t w2_off, int64_t b2_off, float drop_p,
                      void* rng_state, int advance_rng, uint64_t seed, void* dh, float* stats,
                      void* stream) {
  using namespace nfb;
  NFB_REQUIRE(hpre && labels && slots && grads && dh && stats, NFB_EINVAL,
              "mnist_fc_head: null pointer");
  NFB_REQUIRE(B >= 1 && B <= 256, NFB_ERANGE, "mnist_fc_head: batch %d outside [1, 256]", B);
  const int splits = B >= 64 ? 8 : (B >= 32 ? 4 : 1);   // 256 blocks for 32 clients x batch 64
  const int rows_per = (B + splits - 1) / splits;
  const size_t smem = (static_cast<size_t>(rows_per) * 128 + 1280 +
                       static_cast<size_t>(rows_per) * 10 + 64) * sizeof(float);
  static bool attr = false;
  if (!attr) {
    cudaFuncSetAttribute(fc_head_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    attr = true;
  }
  fc_head_kernel<<<dim3(C, splits), 256, smem, as_stream(stream)>>>(
      hpre, labels, label_stride, B, slots, grads, row_stride, b1_off, w2_off, b2_off, drop_p,
      static_cast<unsigned long long*>(rng_state), advance_rng, seed,
      static_cast<__nv_bfloat16*>(dh), stats);
  return launch_status("mnist_fc_head");
}

int nfb_mnist_conv1_wgrad(const void* da1, const float* x, int64_t x_cstride, int C, int B,
                          float* grads,
                          int64_t row_stride, int64_t w_off, int64_t b_off, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(da1 && x && grads, NFB_EINVAL, "mnist_conv1_wgrad: null pointer");
  conv1_wgrad_kernel<<<dim3(B, C), 256, 0, as_stream(stream)>>>(
      static_cast<const __nv_bfloat16*>(da1), x, x_cstride, B, grads, row_stride, w_off, b_off);
  return launch_status("mnist_conv1_wgrad");
}

}  // extern "C"
