// K-GEMM / K-CONV: one warp-specialised tcgen05 kernel family for every GEMM-shaped op of
// the cohort training step, with the CLIENT as the group dimension (grid.z) so that C
// clients' identical-shape, different-weight problems run as one launch.
//
//   D_t[g][m, n] = sum over k-blocks of  A[g][m + a_mn(t), k + a_k(t)] * B[g][n + b_mn(t), k + b_k(t)]
//
// * operands arrive by TMA (cp.async.bulk.tensor, 3-D maps: element, row, client) into a
//   STAGES-deep shared-memory ring with 64/128-byte hardware swizzle; either operand may be
//   K-major ([rows = M or N][K contiguous]) or MN-major ([rows = K][M or N contiguous]), so
//   forward, dgrad and wgrad all read the SAME activation/weight tensors without transposes;
// * one elected thread issues tcgen05.mma (UMMA 128 x BN x 16, bf16 -> fp32) into TMEM;
// * "taps" shift the TMA coordinates per k-block: a 3x3 convolution on a row-flattened
//   NHWC image grid is 9 shifted GEMMs accumulated in TMEM (implicit GEMM without im2col);
//   for weight gradients the 9 taps own 9 TMEM accumulators that share one A tile;
// * 4 epilogue warps read TMEM with tcgen05.ld and apply bias / ReLU / ReLU-mask, then
//   store (fp32 or bf16, arbitrary element strides => free transposes/permutations) or
//   atomically accumulate (split-K, weight gradients).
//
// Warp roles (256 threads): 0 = TMA producer, 1 = MMA issuer, 2 = TMEM allocator,
// 4..7 = epilogue (warp w owns TMEM lanes 32*(w-4) .. +31).
#include <cstdlib>

#include "common.cuh"
#include "philox.cuh"
#include "tc_common.cuh"

namespace nfb {
namespace tc {

constexpr int kMaxTaps = 9;
constexpr int BM = 128;       // UMMA M (TMEM lanes)
constexpr int UMMA_K = 16;    // bf16

struct Tap {
  int a_k, a_mn, b_k, b_mn;
};

struct GemmArgs {
  int taps;          // 1..9
  int inner;         // k-blocks per tap (TPS == 1) or total k-blocks (TPS > 1)
  int k_split;       // K range split over this many CTAs (needs atomic epilogue if > 1)
  Tap tap[kMaxTaps];
  // epilogue
  void* out;
  int out_bf16;      // 0: fp32, 1: bf16
  int atomic;        // fp32 atomicAdd instead of store
  long long out_gs, out_ms, out_ns, out_ts;  // element strides: group, m, n, tap
  const float* bias; // optional
  long long bias_gs;
  int bias_on_m;     // bias index = m (1) or n (0)
  int relu;
  const __nv_bfloat16* mask_src;  // optional: out = 0 where mask_src[same index] <= 0
  int m_valid, n_valid;
  int vec_store;     // n contiguous (out_ns == 1), non-atomic, 16-byte aligned rows
  int tma_store;     // stage 32-row x 128-byte chunks in smem and write them with TMA (map_o)
  float sgd_lr, sgd_wd;  // EPI == 1: w <- w - lr * (acc + wd * w), w addressed by map_o
  // EPI == 1 with DP-SGD: g' = acc * min(1, clip / sqrt(dp_ss[group])) + dp_noise * z, where z is
  // the SAME Philox stream K-OPT uses for the other parameters of the row (element index inside
  // the [groups][dp_row_len] slab = group * dp_row_len + dp_w_off + m * n_valid + n)
  const float* dp_ss;
  float dp_clip, dp_noise;
  unsigned long long dp_seed, dp_group0, dp_groups_total;
  const unsigned long long* dp_step;
  long long dp_row_len, dp_w_off;
  // EPI == 1: optional [groups] accumulator of sum(w_new^2) over this weight (K-STATS side output:
  // central DP needs every client's squared update norm, and this weight never passes through K-OPT)
  float* sumsq_out;
  int* status;       // set to 1 if a pipeline wait timed out
};

// EPI: 0 = store / accumulate the product, 1 = the product is a weight gradient that is applied
// to the fp32 master weights in the epilogue (SGD) and never written to memory, 2 = the product is
// one sample's weight gradient and only the sum of its squares is accumulated into out[group].
template <bool A_MN, int A_SWZ, bool B_MN, int B_SWZ, int BK, int BN, int TPS, int STAGES,
          int EPI = 0>
struct Cfg {
  static constexpr int A_TILE = BM * BK * 2;                 // bytes
  static constexpr int B_TILE = BN * BK * 2;
  static constexpr int STAGE = A_TILE + TPS * B_TILE;
  static constexpr int A_ATOM = A_SWZ / 2;                   // MN elements per swizzle atom
  static constexpr int B_ATOM = B_SWZ / 2;
  static constexpr int TMEM_NEED = BN * TPS;
  static constexpr int TMEM_COLS = TMEM_NEED <= 32 ? 32 : TMEM_NEED <= 64 ? 64
                                   : TMEM_NEED <= 128 ? 128 : TMEM_NEED <= 256 ? 256 : 512;
  // 4 epilogue warps x [32 rows x 128 B]; EPI 1: WBUF weight buffers per warp plus a
  // [32 rows x 64 B] bf16 tile per warp.  The fused update is a streaming pass over the fp32 master
  // (fc1 of the flagship: one k-block of MMAs per 64 KB of weights), so what counts is bytes in
  // flight: with a single operand stage (K = one k-block) the second stage's 32 KB become two more
  // buffers per warp and ALL four 32-column chunks of a 128-column tile are requested before the
  // accumulator is even waited for.
  static constexpr int WBUF = (EPI == 1 && STAGES == 1) ? 4 : 2;
  static constexpr int STAGING = EPI == 1 ? 4 * (WBUF * 4096 + 2048) : 4 * 4096;
  static constexpr size_t SMEM = static_cast<size_t>(STAGES) * STAGE + STAGING + 1024 /*align*/ + 256;
  static_assert(A_MN || A_SWZ == BK * 2, "K-major: swizzle span must equal BK bytes");
  static_assert(B_MN || B_SWZ == BK * 2, "K-major: swizzle span must equal BK bytes");
  static_assert(A_TILE % 1024 == 0 && B_TILE % 1024 == 0, "tiles must keep 1024-B alignment");
  static_assert(BN % 16 == 0 && BN >= 16 && BN <= 256, "UMMA N");
  static_assert(TMEM_NEED <= 512, "TMEM columns");
  static_assert(EPI == 0 || (TPS == 1 && BN % 32 == 0), "SGD epilogue: one accumulator, 32-col chunks");
};

template <bool MN, int SWZ, int BK_, int ROWS>
__device__ __forceinline__ void load_operand(uint8_t* dst, const CUtensorMap* map, uint64_t* bar,
                                             int k0, int mn0, int g) {
  if constexpr (!MN) {
    tma_load_3d(dst, map, bar, k0, mn0, g);  // box [BK, ROWS, 1]
  } else {
    constexpr int ATOM = SWZ / 2;
#pragma unroll
    for (int a = 0; a < ROWS / ATOM; ++a)    // box [ATOM, BK, 1] per swizzle atom
      tma_load_3d(dst + a * (BK_ * SWZ), map, bar, mn0 + a * ATOM, k0, g);
  }
}

template <bool MN, int SWZ, int BK_>
__device__ __forceinline__ uint64_t operand_desc(uint32_t tile_addr, int kk) {
  if constexpr (!MN) {
    // K-major: 8-row groups are 8*SWZ bytes apart; one atom along K; step K by 32 bytes
    return make_smem_desc(tile_addr + kk * (UMMA_K * 2), 0, 8 * SWZ, swizzle_code_for_bytes(SWZ));
  } else {
    // MN-major: SBO = 8 K-rows, LBO = next MN atom (BK rows further); step K by 16 rows
    return make_smem_desc(tile_addr + kk * (UMMA_K * SWZ), BK_ * SWZ, 8 * SWZ,
                          swizzle_code_for_bytes(SWZ));
  }
}

template <bool A_MN, int A_SWZ, bool B_MN, int B_SWZ, int BK, int BN, int TPS, int STAGES,
          int MINB = 1, int EPI = 0>
__global__ void __launch_bounds__(256, MINB)
    tc_gemm_kernel(const __grid_constant__ CUtensorMap map_a,
                   const __grid_constant__ CUtensorMap map_b,
                   const __grid_constant__ CUtensorMap map_o,
                   const __grid_constant__ CUtensorMap map_l,   // EPI 1: bf16 copy of the weights
                   const GemmArgs args) {
  using C = Cfg<A_MN, A_SWZ, B_MN, B_SWZ, BK, BN, TPS, STAGES, EPI>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* staging = smem + STAGES * C::STAGE;          // 1024-aligned (stage sizes are)
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + C::STAGING);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* acc_bar = empty_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 1);
  uint64_t* w_bar = acc_bar + 2;                        // EPI 1: [4 warps][WBUF buffers]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int g = blockIdx.z / args.k_split;
  const int split = blockIdx.z % args.k_split;

  // iteration space: TPS == 1 -> (tap, k-block) pairs ; TPS > 1 -> k-blocks (all taps each)
  const int iters_total = (TPS == 1) ? args.taps * args.inner : args.inner;
  const int per = (iters_total + args.k_split - 1) / args.k_split;
  const int it_lo = split * per;
  const int it_hi = min(it_lo + per, iters_total);
  const int n_iters = max(it_hi - it_lo, 0);

  if (threadIdx.x == 0) {
    prefetch_tmap(&map_a);
    prefetch_tmap(&map_b);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(acc_bar, 1);
    if constexpr (EPI == 1) {
      prefetch_tmap(&map_o);
      prefetch_tmap(&map_l);
      for (int i = 0; i < 4 * C::WBUF; ++i) mbar_init(&w_bar[i], 1);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<C::TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      for (int i = 0; i < n_iters; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        if (!mbar_wait(&empty_bar[s], ph ^ 1)) { *args.status = 1; break; }
        uint8_t* stage = smem + s * C::STAGE;
        mbar_expect_tx(&full_bar[s], C::STAGE);
        const int it = it_lo + i;
        if constexpr (TPS == 1) {
          const int t = it / args.inner, c = it % args.inner;
          const Tap tp = args.tap[t];
          load_operand<A_MN, A_SWZ, BK, BM>(stage, &map_a, &full_bar[s], tp.a_k + c * BK,
                                            m0 + tp.a_mn, g);
          load_operand<B_MN, B_SWZ, BK, BN>(stage + C::A_TILE, &map_b, &full_bar[s],
                                            tp.b_k + c * BK, n0 + tp.b_mn, g);
        } else {
          const Tap t0 = args.tap[0];
          load_operand<A_MN, A_SWZ, BK, BM>(stage, &map_a, &full_bar[s], t0.a_k + it * BK,
                                            m0 + t0.a_mn, g);
          for (int t = 0; t < TPS; ++t) {
            const Tap tp = args.tap[t];
            load_operand<B_MN, B_SWZ, BK, BN>(stage + C::A_TILE + t * C::B_TILE, &map_b,
                                              &full_bar[s], tp.b_k + it * BK, n0 + tp.b_mn, g);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN ? 1u : 0u, B_MN ? 1u : 0u);
      bool ok = true;
      for (int i = 0; i < n_iters && ok; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        if (!mbar_wait(&full_bar[s], ph)) { *args.status = 1; ok = false; break; }
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + s * C::STAGE);
#pragma unroll
        for (int t = 0; t < TPS; ++t) {
          const uint32_t b_addr = a_addr + C::A_TILE + t * C::B_TILE;
#pragma unroll
          for (int kk = 0; kk < BK / UMMA_K; ++kk) {
            const uint64_t da = operand_desc<A_MN, A_SWZ, BK>(a_addr, kk);
            const uint64_t db = operand_desc<B_MN, B_SWZ, BK>(b_addr, kk);
            umma_f16(tmem_base + t * BN, da, db, idesc, (i > 0 || kk > 0) ? 1u : 0u);
          }
        }
        umma_commit(&empty_bar[s]);  // frees the smem slot when these MMAs retire
      }
      umma_commit(acc_bar);          // accumulators complete
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int sub = warp - 4;                 // TMEM sub-partition == warp % 4
    const int m = m0 + sub * 32 + lane;       // this thread's output row
    bool ok = true;
    [[maybe_unused]] int sgd_chunks = 0;
    if constexpr (EPI == 1) {
      // the weight tiles do not depend on the MMA: start fetching the first WBUF 32-column
      // chunks of this warp's 32 rows before waiting for the accumulator
      sgd_chunks = min(BN, args.n_valid - n0) / 32;
      if (lane == 0) {
        for (int c = 0; c < C::WBUF && c < sgd_chunks; ++c) {
          mbar_expect_tx(&w_bar[sub * C::WBUF + c], 4096);
          tma_load_3d(staging + sub * (C::WBUF * 4096) + c * 4096, &map_o,
                      &w_bar[sub * C::WBUF + c], n0 + c * 32, m0 + sub * 32, g);
        }
      }
    }
    if (n_iters > 0) ok = mbar_wait(acc_bar, 0);
    if (!ok && lane == 0) *args.status = 1;
    tc_fence_after();
    const bool row_ok = ok && n_iters > 0 && m < args.m_valid;
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(sub * 32) << 16);
    if constexpr (EPI == 2) {
      // ---- K-PSGN: the product is ONE SAMPLE's weight gradient; only its squared Frobenius norm
      // leaves the kernel (rows / columns beyond the operands were zero-filled by TMA, so they add
      // nothing and need no masking) ----
      float acc = 0.f;
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 16) {
        uint32_t r[16];
        tmem_ld_32x32b_x16(lane_addr + c0, r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) acc = fmaf(__uint_as_float(r[j]), __uint_as_float(r[j]), acc);
      }
      acc = warp_sum(acc);
      if (lane == 0 && ok && n_iters > 0) atomicAdd(static_cast<float*>(args.out) + g, acc);
    } else if constexpr (EPI == 1) {
      // ---- weight gradient applied in place: w <- w - lr * (dW + wd * w) ----
      // Per 32-column chunk: the fp32 weights arrive by TMA in a 128-byte-swizzled [32 x 128 B]
      // buffer, each thread updates its own row in place, and the same buffer is the source of
      // the TMA store back; the bf16 working copy goes out through a 64-byte-swizzled tile.
      // The gradient itself never reaches memory.
      uint8_t* lstage = staging + 4 * (C::WBUF * 4096) + sub * 2048;
      const float lr = args.sgd_lr, wd = args.sgd_wd;
      const bool live = ok && n_iters > 0;
      const float cs = args.dp_ss ? fminf(args.dp_clip / (sqrtf(args.dp_ss[g]) + 1e-6f), 1.f) : 1.f;
      const unsigned long long dp_g0 =
          args.dp_group0 + (args.dp_step ? *args.dp_step * args.dp_groups_total : 0ull);
      const long long dp_row0 = static_cast<long long>(g) * args.dp_row_len + args.dp_w_off +
                                static_cast<long long>(m) * args.n_valid + n0;
      float wsq = 0.f;
#pragma unroll 1
      for (int c = 0; c < sgd_chunks; ++c) {
        const int buf = c % C::WBUF;
        uint8_t* wbuf = staging + sub * (C::WBUF * 4096) + buf * 4096;
        uint32_t r[32];
        __syncwarp();                         // lane 0 may come late from the refill below
        tmem_ld_32x32b_x16(lane_addr + c * 32, r);
        tmem_ld_32x32b_x16(lane_addr + c * 32 + 16, r + 16);
        tmem_ld_wait();
        if (!mbar_wait(&w_bar[sub * C::WBUF + buf], (c / C::WBUF) & 1)) { *args.status = 1; break; }
        // bf16 tile of the previous chunk must have been read by its store before we overwrite it
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
        float nw[32];
#pragma unroll
        for (int p = 0; p < 8; ++p) {
          uint8_t* addr = wbuf + lane * 128 + ((p ^ (lane & 7)) * 16);
          float w4[4];
          unpack<float>(*reinterpret_cast<const uint4*>(addr), w4);
          float zn[4] = {0.f, 0.f, 0.f, 0.f};
          if (args.dp_noise > 0.f) {
            const float4 z = Philox::normal4(dp_g0 + ((dp_row0 + c * 32 + 4 * p) >> 2), args.dp_seed);
            zn[0] = args.dp_noise * z.x; zn[1] = args.dp_noise * z.y;
            zn[2] = args.dp_noise * z.z; zn[3] = args.dp_noise * z.w;
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float gr = fmaf(wd, w4[e], fmaf(cs, __uint_as_float(r[4 * p + e]), zn[e]));
            nw[4 * p + e] = fmaf(-lr, gr, w4[e]);
          }
          st_vec(addr, pack<float>(nw + 4 * p));
        }
#pragma unroll
        for (int q = 0; q < 4; ++q)
          st_vec(lstage + lane * 64 + ((q ^ ((lane >> 1) & 3)) * 16), pack<__nv_bfloat16>(nw + 8 * q));
        fence_proxy_async();
        __syncwarp();
        if (lane == 0 && live) {
          asm volatile(
              "cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                  &map_o),
              "r"(smem_u32(wbuf)), "r"(n0 + c * 32), "r"(m0 + sub * 32), "r"(g)
              : "memory");
          asm volatile(
              "cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                  &map_l),
              "r"(smem_u32(lstage)), "r"(n0 + c * 32), "r"(m0 + sub * 32), "r"(g)
              : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          if (c + C::WBUF < sgd_chunks) {
            // refill this buffer with chunk c + WBUF once the store above has read it
            asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            mbar_expect_tx(&w_bar[sub * C::WBUF + buf], 4096);
            tma_load_3d(wbuf, &map_o, &w_bar[sub * C::WBUF + buf], n0 + (c + C::WBUF) * 32,
                        m0 + sub * 32, g);
          }
        }
        if (args.sumsq_out) {   // off the store's critical path: the tile is on its way already
#pragma unroll
          for (int e = 0; e < 32; ++e) wsq = fmaf(nw[e], nw[e], wsq);
        }
      }
      if (args.sumsq_out) {   // rows / columns past the weight's edge hold zeros (TMA fill)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) wsq += __shfl_xor_sync(0xffffffffu, wsq, o);
        if (lane == 0 && live) atomicAdd(args.sumsq_out + g, wsq);
      }
      if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    } else if (args.tma_store) {
      // Row-per-thread global stores touch 32 different lines per instruction; instead each
      // warp stages [32 rows x 128 B] chunks in swizzled shared memory (conflict-free 16-byte
      // stores) and a bulk tensor store writes them as whole lines.
      uint8_t* my_stage = staging + sub * 4096;
      const int cols = args.out_bf16 ? 64 : 32;           // columns per 128-byte chunk
      const bool live = ok && n_iters > 0;
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += cols) {
        if (n0 + c0 >= args.n_valid) break;
        uint32_t r[64];
#pragma unroll
        for (int q = 0; q < 4; ++q)
          if (q * 16 < cols) tmem_ld_32x32b_x16(lane_addr + c0 + q * 16, r + q * 16);
        tmem_ld_wait();
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
        float v[64];
#pragma unroll
        for (int j = 0; j < 64; ++j) {
          if (j < cols) {
            v[j] = __uint_as_float(r[j]);
            if (args.bias && split == 0)
              v[j] += args.bias[static_cast<long long>(g) * args.bias_gs +
                                (args.bias_on_m ? m : (n0 + c0 + j))];
            if (args.relu) v[j] = fmaxf(v[j], 0.f);
          }
        }
#pragma unroll
        for (int p = 0; p < 8; ++p) {                    // eight 16-byte pieces per row
          uint4 piece;
          if (args.out_bf16) piece = pack<__nv_bfloat16>(v + 8 * p);
          else piece = pack<float>(v + 4 * p);
          st_vec(my_stage + lane * 128 + ((p ^ (lane & 7)) * 16), piece);
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0 && live) {
          asm volatile(
              "cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                  &map_o),
              "r"(smem_u32(my_stage)), "r"(n0 + c0), "r"(m0 + sub * 32), "r"(g)
              : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
      if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    } else
#pragma unroll 1
    for (int t = 0; t < TPS; ++t) {
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 16) {
        uint32_t r[16];
        tmem_ld_32x32b_x16(lane_addr + t * BN + c0, r);   // warp-collective: all lanes execute
        tmem_ld_wait();
        if (!row_ok) continue;
        const long long row_off = static_cast<long long>(g) * args.out_gs +
                                  static_cast<long long>(m) * args.out_ms +
                                  static_cast<long long>(t) * args.out_ts;
        float v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          v[j] = __uint_as_float(r[j]);
          if (args.bias && split == 0)
            v[j] += args.bias[static_cast<long long>(g) * args.bias_gs +
                              (args.bias_on_m ? m : (n0 + c0 + j))];
          if (args.relu) v[j] = fmaxf(v[j], 0.f);
        }
        const long long idx0 = row_off + static_cast<long long>(n0 + c0) * args.out_ns;
        if (args.vec_store && n0 + c0 + 16 <= args.n_valid) {
          // n is the contiguous output dimension and everything is 16-byte aligned
          if (args.mask_src) {
            const uint4 mk0 = *reinterpret_cast<const uint4*>(args.mask_src + idx0);
            const uint4 mk1 = *reinterpret_cast<const uint4*>(args.mask_src + idx0 + 8);
            float mf[16];
            unpack<__nv_bfloat16>(mk0, mf);
            unpack<__nv_bfloat16>(mk1, mf + 8);
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (!(mf[j] > 0.f)) v[j] = 0.f;
          }
          if (args.out_bf16) {
            __nv_bfloat16* o = static_cast<__nv_bfloat16*>(args.out) + idx0;
            st_vec(o, pack<__nv_bfloat16>(v));
            st_vec(o + 8, pack<__nv_bfloat16>(v + 8));
          } else {
            float* o = static_cast<float*>(args.out) + idx0;
#pragma unroll
            for (int q = 0; q < 4; ++q)
              *reinterpret_cast<float4*>(o + 4 * q) =
                  make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
          }
          continue;
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int n = n0 + c0 + j;
          if (n >= args.n_valid) break;
          const long long idx = idx0 + static_cast<long long>(j) * args.out_ns;
          float o = v[j];
          if (args.mask_src && !(__bfloat162float(args.mask_src[idx]) > 0.f)) o = 0.f;
          if (args.atomic) {
            atomicAdd(static_cast<float*>(args.out) + idx, o);
          } else if (args.out_bf16) {
            static_cast<__nv_bfloat16*>(args.out)[idx] = __float2bfloat16_rn(o);
          } else {
            static_cast<float*>(args.out)[idx] = o;
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<C::TMEM_COLS>(tmem_base);
}

}  // namespace tc
}  // namespace nfb

// =====================================================================================
// Host side: tensor-map cache, variant table, C entry point
// =====================================================================================
#include <map>
#include <mutex>
#include <tuple>

namespace nfb {
namespace tc {

using EncodeFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                              const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                              const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                              CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeFn encode_fn() {
  static EncodeFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeFn>(p);
  });
  return fn;
}


using MapKey = std::tuple<const void*, long long, long long, long long, long long, long long,
                          int, int, int>;
static std::map<MapKey, CUtensorMap> g_maps;
static std::mutex g_maps_mu;

int get_map(const Operand& op, int box_inner, int box_rows, int swz_bytes,
                   CUtensorMap* out) {
  MapKey key{op.ptr, op.inner, op.rows, op.groups, op.row_stride, op.group_stride,
             box_inner, box_rows, swz_bytes};
  std::lock_guard<std::mutex> lock(g_maps_mu);
  auto it = g_maps.find(key);
  if (it != g_maps.end()) { *out = it->second; return 0; }
  EncodeFn enc = encode_fn();
  if (!enc) return fail(NFB_ESTATE, "tc_gemm: cuTensorMapEncodeTiled unavailable");
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(op.inner), static_cast<cuuint64_t>(op.rows),
                        static_cast<cuuint64_t>(op.groups)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(op.row_stride) * 2,
                           static_cast<cuuint64_t>(op.group_stride) * 2};
  cuuint32_t box[3] = {static_cast<cuuint32_t>(box_inner), static_cast<cuuint32_t>(box_rows), 1};
  cuuint32_t estr[3] = {1, 1, 1};
  if (op.groups == 1) strides[1] = static_cast<cuuint64_t>(op.rows) * strides[0];
  if ((strides[0] & 15) || (strides[1] & 15) || (reinterpret_cast<uintptr_t>(op.ptr) & 15))
    return fail(NFB_EINVAL, "tc_gemm: operand base/strides must be 16-byte aligned");
  const CUtensorMapSwizzle sw = swz_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                : swz_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                                                  : CU_TENSOR_MAP_SWIZZLE_32B;
  CUtensorMap m;
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(op.ptr), dims,
                   strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(NFB_EINVAL, "tc_gemm: cuTensorMapEncodeTiled failed (%d)",
                                     static_cast<int>(r));
  if (g_maps.size() > 4096) g_maps.clear();
  g_maps[key] = m;
  *out = m;
  return 0;
}

// Output map for the TMA-store epilogue: [groups][m rows][n cols], n contiguous.
int get_out_map(void* ptr, long long n, long long m, long long groups, long long ms,
                       long long gs, bool bf16, CUtensorMap* out) {
  const int esz = bf16 ? 2 : 4;
  MapKey key{ptr, n, m, groups, ms, gs, bf16 ? -64 : -32, 32, 128};
  std::lock_guard<std::mutex> lock(g_maps_mu);
  auto it = g_maps.find(key);
  if (it != g_maps.end()) { *out = it->second; return 0; }
  EncodeFn enc = encode_fn();
  if (!enc) return fail(NFB_ESTATE, "tc_gemm: cuTensorMapEncodeTiled unavailable");
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(n), static_cast<cuuint64_t>(m),
                        static_cast<cuuint64_t>(groups)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(ms) * esz, static_cast<cuuint64_t>(gs) * esz};
  if (groups == 1) strides[1] = static_cast<cuuint64_t>(m) * strides[0];
  cuuint32_t box[3] = {static_cast<cuuint32_t>(bf16 ? 64 : 32), 32, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUtensorMap mo;
  CUresult r = enc(&mo, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32,
                   3, ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(NFB_EINVAL, "tc_gemm: output tensor map rejected (%d)",
                                     static_cast<int>(r));
  g_maps[key] = mo;
  *out = mo;
  return 0;
}

// bf16 [groups][m rows][n cols] map with box [32 cols, 32 rows] and 64-byte swizzle: the working
// copy of the weights written by the SGD epilogue.
int get_lowp_map(void* ptr, long long n, long long m, long long groups, long long ms,
                        long long gs, CUtensorMap* out) {
  MapKey key{ptr, n, m, groups, ms, gs, -16, 32, 64};
  std::lock_guard<std::mutex> lock(g_maps_mu);
  auto it = g_maps.find(key);
  if (it != g_maps.end()) { *out = it->second; return 0; }
  EncodeFn enc = encode_fn();
  if (!enc) return fail(NFB_ESTATE, "tc_gemm: cuTensorMapEncodeTiled unavailable");
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(n), static_cast<cuuint64_t>(m),
                        static_cast<cuuint64_t>(groups)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(ms) * 2, static_cast<cuuint64_t>(gs) * 2};
  if (groups == 1) strides[1] = static_cast<cuuint64_t>(m) * strides[0];
  if ((strides[0] & 15) || (strides[1] & 15) || (reinterpret_cast<uintptr_t>(ptr) & 15))
    return fail(NFB_EINVAL, "tc_gemm: bf16 weight copy must be 16-byte aligned");
  cuuint32_t box[3] = {32, 32, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUtensorMap ml;
  CUresult r = enc(&ml, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, ptr, dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B,
                   CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(NFB_EINVAL, "tc_gemm: bf16 weight map rejected (%d)",
                                     static_cast<int>(r));
  g_maps[key] = ml;
  *out = ml;
  return 0;
}

template <bool A_MN, int A_SWZ, bool B_MN, int B_SWZ, int BK, int BN, int TPS, int STAGES,
          int MINB = 1, int EPI = 0>
static int launch(const Operand& a, const Operand& b, int m_tiles, int n_tiles, int groups,
                  GemmArgs args, cudaStream_t st, void* lowp = nullptr, long long lowp_gs = 0) {
  using C = Cfg<A_MN, A_SWZ, B_MN, B_SWZ, BK, BN, TPS, STAGES, EPI>;
  CUtensorMap ma, mb, mo, ml;
  int rc = A_MN ? get_map(a, C::A_ATOM, BK, A_SWZ, &ma) : get_map(a, BK, BM, A_SWZ, &ma);
  if (rc) return rc;
  rc = B_MN ? get_map(b, C::B_ATOM, BK, B_SWZ, &mb) : get_map(b, BK, BN, B_SWZ, &mb);
  if (rc) return rc;
  auto kern = tc_gemm_kernel<A_MN, A_SWZ, B_MN, B_SWZ, BK, BN, TPS, STAGES, MINB, EPI>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(C::SMEM));
    if (e != cudaSuccess) return fail(static_cast<int>(e), "tc_gemm: smem attribute: %s",
                                      cudaGetErrorString(e));
    attr_set = true;
  }
  // TMA-store epilogue when the output is a plain row-major [m][n] tile with 128-byte chunks
  const int chunk_cols = args.out_bf16 ? 64 : 32;
  args.tma_store = (args.vec_store && TPS == 1 && args.k_split == 1 && !args.mask_src &&
                    BN % chunk_cols == 0 && getenv("NFB_GEMM_NO_TMA_STORE") == nullptr) ? 1 : 0;
  mo = ma;
  ml = ma;
  if constexpr (EPI == 1) {
    if (!args.tma_store || args.out_bf16 || !lowp || args.n_valid % 32)
      return fail(NFB_EINVAL, "tc_gemm: SGD epilogue needs a dense fp32 [m][n] weight tile, "
                              "n a multiple of 32, and a bf16 copy");
    rc = get_lowp_map(lowp, args.n_valid, args.m_valid, groups, args.out_ms, lowp_gs, &ml);
    if (rc) return rc;
  }
  if (args.tma_store) {
    rc = get_out_map(args.out, args.n_valid, args.m_valid, groups, args.out_ms, args.out_gs,
                     args.out_bf16 != 0, &mo);
    if (rc) return rc;
  }
  dim3 grid(m_tiles, n_tiles, groups * args.k_split);
  kern<<<grid, 256, C::SMEM, st>>>(ma, mb, mo, ml, args);
  return launch_status("tc_gemm");
}

}  // namespace tc
}  // namespace nfb

extern "C" {

// nfb_tc_operand / nfb_tc_epilogue: include/nanofed_b200.h

// variant: 0 conv_fwd, 1 conv_dgrad, 2 conv_wgrad, 3 fc_fwd, 4 fc_dgrad, 5 fc_wgrad,
//          6 fc_dgrad with M = batch (row-major bf16 output, TMA-store epilogue)
int nfb_tc_gemm(int variant, const nfb_tc_operand* A, const nfb_tc_operand* B, int m_tiles,
                int n_tiles, int groups, int taps, int inner, int k_split, const int* tap_table,
                const nfb_tc_epilogue* epi, int* status, void* stream) {
  using namespace nfb;
  using namespace nfb::tc;
  NFB_REQUIRE(A && B && epi && status && epi->out, NFB_EINVAL, "tc_gemm: null pointer");
  NFB_REQUIRE(taps >= 1 && taps <= kMaxTaps && inner >= 1 && k_split >= 1, NFB_EINVAL,
              "tc_gemm: bad taps/inner/k_split");
  NFB_REQUIRE(m_tiles >= 1 && n_tiles >= 1 && groups >= 1 &&
                  static_cast<long long>(groups) * k_split <= 65535, NFB_ERANGE,
              "tc_gemm: bad grid");
  NFB_REQUIRE(k_split == 1 || epi->atomic, NFB_EINVAL, "tc_gemm: split-K needs atomic epilogue");
  GemmArgs g{};
  g.taps = taps; g.inner = inner; g.k_split = k_split;
  for (int t = 0; t < taps; ++t) {
    if (tap_table) g.tap[t] = Tap{tap_table[4 * t], tap_table[4 * t + 1], tap_table[4 * t + 2],
                                  tap_table[4 * t + 3]};
    else g.tap[t] = Tap{0, 0, 0, 0};
  }
  g.out = epi->out; g.out_bf16 = epi->out_bf16; g.atomic = epi->atomic;
  g.out_gs = epi->out_gs; g.out_ms = epi->out_ms; g.out_ns = epi->out_ns; g.out_ts = epi->out_ts;
  g.bias = epi->bias; g.bias_gs = epi->bias_gs; g.bias_on_m = epi->bias_on_m;
  g.relu = epi->relu;
  g.mask_src = static_cast<const __nv_bfloat16*>(epi->mask_src);
  g.m_valid = epi->m_valid; g.n_valid = epi->n_valid;
  g.status = status;
  const int esz = epi->out_bf16 ? 2 : 4;
  g.vec_store = (!epi->atomic && epi->out_ns == 1 &&
                 (reinterpret_cast<uintptr_t>(epi->out) & 15) == 0 &&
                 ((epi->out_gs * esz) & 15) == 0 && ((epi->out_ms * esz) & 15) == 0 &&
                 ((epi->out_ts * esz) & 15) == 0 &&
                 (!epi->mask_src || (reinterpret_cast<uintptr_t>(epi->mask_src) & 15) == 0))
                    ? 1 : 0;
  Operand a{A->ptr, A->inner, A->rows, A->groups, A->row_stride, A->group_stride};
  Operand b{B->ptr, B->inner, B->rows, B->groups, B->row_stride, B->group_stride};
  cudaStream_t st = as_stream(stream);
  switch (variant) {
    case 0: return launch<false, 64, false, 64, 32, 64, 1, 6>(a, b, m_tiles, n_tiles, groups, g, st);
    case 1: return launch<false, 128, true, 64, 64, 32, 1, 4>(a, b, m_tiles, n_tiles, groups, g, st);
    case 2:
      NFB_REQUIRE(taps == 9, NFB_EINVAL, "tc_gemm: conv_wgrad needs 9 taps");
      return launch<true, 128, true, 64, 64, 32, 9, 4>(a, b, m_tiles, n_tiles, groups, g, st);
    case 3: return launch<false, 128, false, 128, 64, 64, 1, 3, 2>(a, b, m_tiles, n_tiles, groups, g, st);
    case 4: return launch<true, 128, false, 128, 64, 64, 1, 4>(a, b, m_tiles, n_tiles, groups, g, st);
    // variants 5/6 have only 1-2 k-blocks per tile: every CTA is one latency-bound chain
    // (setup -> TMA -> MMA -> epilogue), so two CTAs per SM (2 stages, <=128 registers) are what
    // overlaps them; measured 107 us at 1 CTA/SM for the 128x9216 weight gradient.
    case 5: return launch<true, 128, true, 128, 64, 128, 1, 2, 2>(a, b, m_tiles, n_tiles, groups, g, st);
    case 6: return launch<false, 128, true, 128, 64, 128, 1, 2, 2>(a, b, m_tiles, n_tiles, groups, g, st);
    default: return fail(NFB_EINVAL, "tc_gemm: unknown variant %d", variant);
  }
}


// fc weight gradient fused with the optimizer step (momentum-free SGD):
//   w[g][j][i] <- w - lr * (sum_b dy[g][b][j] * x[g][b][i] + wd * w),  w_lowp <- bf16(w)
// sumsq_out (optional, [groups]): += sum of the squares of the NEW weights of each group.
// A = dy [groups][batch][j] and B = x [groups][batch][i] (both MN-major), w fp32 rows of length i
// `w_gs` / `lowp_gs` elements apart per group.  The gradient is never written.  The caller must
// have finished every read of the OLD weights (e.g. the data gradient) on this stream.
int nfb_tc_linear_wgrad_sgd(const nfb_tc_operand* A, const nfb_tc_operand* B, int groups, int j,
                            int i, int inner, float* w, long long w_gs, void* w_lowp,
                            long long lowp_gs, float lr, float weight_decay, float* sumsq_out,
                            int* status, void* stream) {
  using namespace nfb;
  using namespace nfb::tc;
  NFB_REQUIRE(A && B && w && w_lowp && status, NFB_EINVAL, "linear_wgrad_sgd: null pointer");
  NFB_REQUIRE(groups >= 1 && groups <= 65535 && j >= 1 && i >= 32 && i % 32 == 0 && inner >= 1,
              NFB_EINVAL, "linear_wgrad_sgd: bad shape");
  NFB_REQUIRE((reinterpret_cast<uintptr_t>(w) & 15) == 0 && (w_gs & 3) == 0, NFB_EINVAL,
              "linear_wgrad_sgd: weights must be 16-byte aligned");
  GemmArgs g{};
  g.taps = 1; g.inner = inner; g.k_split = 1;
  g.tap[0] = Tap{0, 0, 0, 0};
  g.out = w; g.out_bf16 = 0; g.atomic = 0;
  g.out_gs = w_gs; g.out_ms = i; g.out_ns = 1; g.out_ts = 0;
  g.m_valid = j; g.n_valid = i;
  g.vec_store = 1;
  g.sgd_lr = lr; g.sgd_wd = weight_decay;
  g.sumsq_out = sumsq_out;
  g.status = status;
  Operand a{A->ptr, A->inner, A->rows, A->groups, A->row_stride, A->group_stride};
  Operand b{B->ptr, B->inner, B->rows, B->groups, B->row_stride, B->group_stride};
  if (inner == 1)   // K = one k-block: nothing to pipeline, the smem goes to the weight buffers
    return launch<true, 128, true, 128, 64, 128, 1, 1, 2, 1>(a, b, (j + BM - 1) / BM,
                                                           (i + 127) / 128, groups, g,
                                                           as_stream(stream), w_lowp, lowp_gs);
  return launch<true, 128, true, 128, 64, 128, 1, 2, 2, 1>(a, b, (j + BM - 1) / BM, (i + 127) / 128,
                                                         groups, g, as_stream(stream), w_lowp,
                                                         lowp_gs);
}

// nfb_tc_linear_wgrad_sgd with the DP-SGD front end in the epilogue: the gradient tile in TMEM is
// scaled by the client's clip factor min(1, clip / sqrt(row_sumsq[client])) and gets Gaussian noise
// from the Philox stream at the weight's position inside the [groups][row_len] parameter slab
// (w_off = offset of this weight in a row), i.e. the same stream nfb_fused_sgd_dp uses for the
// rest of the row.  row_sumsq must hold the COMPLETE squared gradient norm of each client (this
// weight's part is obtained without materialising it: ghost norm, nfb_mnist_dp_norm).
int nfb_tc_linear_wgrad_sgd_dp(const nfb_tc_operand* A, const nfb_tc_operand* B, int groups, int j,
                               int i, int inner, float* w, long long w_gs, void* w_lowp,
                               long long lowp_gs, float lr, float weight_decay,
                               const float* row_sumsq, float clip_norm, float noise_std,
                               uint64_t seed, uint64_t offset, const unsigned long long* step_ctr,
                               long long row_len, long long w_off, int* status, void* stream) {
  using namespace nfb;
  using namespace nfb::tc;
  NFB_REQUIRE(A && B && w && w_lowp && status && row_sumsq, NFB_EINVAL, "linear_wgrad_sgd_dp: null pointer");
  NFB_REQUIRE(groups >= 1 && groups <= 65535 && j >= 1 && i >= 32 && i % 32 == 0 && inner >= 1,
              NFB_EINVAL, "linear_wgrad_sgd_dp: bad shape");
  NFB_REQUIRE((row_len & 3) == 0 && (w_off & 3) == 0 && (offset & 3u) == 0, NFB_EINVAL,
              "linear_wgrad_sgd_dp: row length / offsets must be multiples of 4");
  GemmArgs g{};
  g.taps = 1; g.inner = inner; g.k_split = 1;
  g.tap[0] = Tap{0, 0, 0, 0};
  g.out = w; g.out_bf16 = 0; g.atomic = 0;
  g.out_gs = w_gs; g.out_ms = i; g.out_ns = 1; g.out_ts = 0;
  g.m_valid = j; g.n_valid = i;
  g.vec_store = 1;
  g.sgd_lr = lr; g.sgd_wd = weight_decay;
  g.dp_ss = row_sumsq; g.dp_clip = clip_norm; g.dp_noise = noise_std; g.dp_seed = seed;
  g.dp_group0 = offset >> 2;
  g.dp_groups_total = static_cast<unsigned long long>((static_cast<long long>(groups) * row_len + 3) >> 2);
  g.dp_step = step_ctr; g.dp_row_len = row_len; g.dp_w_off = w_off;
  g.status = status;
  Operand a{A->ptr, A->inner, A->rows, A->groups, A->row_stride, A->group_stride};
  Operand b{B->ptr, B->inner, B->rows, B->groups, B->row_stride, B->group_stride};
  if (inner == 1)   // K = one k-block: nothing to pipeline, the smem goes to the weight buffers
    return launch<true, 128, true, 128, 64, 128, 1, 1, 2, 1>(a, b, (j + BM - 1) / BM,
                                                           (i + 127) / 128, groups, g,
                                                           as_stream(stream), w_lowp, lowp_gs);
  return launch<true, 128, true, 128, 64, 128, 1, 2, 2, 1>(a, b, (j + BM - 1) / BM, (i + 127) / 128,
                                                         groups, g, as_stream(stream), w_lowp,
                                                         lowp_gs);
}

// K-PSGN for convolutions (and any layer whose per-sample weight gradient is delta_i^T U_i):
//   out[i] += || sum_p dy[i][p][:]^T u[i][p][:] ||_F^2     for every sample i (= GEMM group),
// on tensor cores, both operands MN-major ([P rows][channels contiguous], bf16); the per-sample
// gradient lives only in TMEM.  out must be zero on entry (CTAs of one sample add atomically).
// Replaces the clip_grad_norm_ of nanofed/trainer/private.py:62-63 for TRUE per-sample clipping.
int nfb_tc_psgn(const nfb_tc_operand* dy, const nfb_tc_operand* u, int samples, int cout, int kdim,
                int inner, float* out, int* status, void* stream) {
  using namespace nfb;
  using namespace nfb::tc;
  NFB_REQUIRE(dy && u && out && status, NFB_EINVAL, "tc_psgn: null pointer");
  NFB_REQUIRE(samples >= 1 && samples <= 65535 && cout >= 1 && kdim >= 1 && inner >= 1, NFB_EINVAL,
              "tc_psgn: bad shape");
  GemmArgs g{};
  g.taps = 1; g.inner = inner; g.k_split = 1;
  g.tap[0] = Tap{0, 0, 0, 0};
  g.out = out; g.m_valid = cout; g.n_valid = kdim;
  g.status = status;
  Operand a{dy->ptr, dy->inner, dy->rows, dy->groups, dy->row_stride, dy->group_stride};
  Operand b{u->ptr, u->inner, u->rows, u->groups, u->row_stride, u->group_stride};
  return launch<true, 128, true, 128, 64, 128, 1, 2, 2, 2>(a, b, (cout + BM - 1) / BM,
                                                          (kdim + 127) / 128, samples, g,
                                                          as_stream(stream));
}

}  // extern "C"
