// K-CONV (generic): tcgen05 implicit-GEMM convolutions for NHWC bf16 activations with the CLIENT
// as the group dimension - forward, data gradient and weight gradient of 3x3 (stride 1/2, pad 1)
// and 1x1 (stride 1/2) convolutions for any channel count that is a multiple of 64 and any
// power-of-two image grid from 4x4 to 32x32.  This is what the ResNet-18 cohort engine
// (nanofed_b200/models/resnet_cohort.py) runs on; no reference counterpart (the reference's
// convolutions are torch library calls inside nanofed/trainer/base.py:134-160).
//
// No im2col and no padded storage: a tile of 128 (or 32 / 64) output positions is ONE 5-D TMA box
//   [64 channels, bw columns, 1, bh rows, bn images]        (bw * bh * bn = positions per box)
// of the activation tensor, which lands in shared memory as [positions x 128 B] with the
// 128-byte swizzle, i.e. directly as a K-major (forward / dgrad) or MN-major (wgrad) UMMA
// operand.  A filter tap (kh, kw) is the same box with its start coordinate moved by
// (kh - 1, kw - 1); positions outside the image are filled with zeros by the TMA unit
// (out-of-bounds fill), which IS the padding.  Stride 2 reads through a "parity view" of the
// same memory, [2C, W/2, 2, H/2, N]: input pixel (2*ho + kh - 1, 2*wo + kw - 1) is parity
// ((kh+1)&1, (kw+1)&1) at half-resolution coordinate (ho - (kh==0), wo - (kw==0)).
//
//   forward    D[pos, co]      = sum_tap sum_ci X[pos + tap, ci] * Wt[tap][co][ci]
//              A = X boxes (K-major), B = Wt rows (K-major); epilogue: bf16 store through a TMA
//              box + per-channel sum / sum of squares of the stored values (BatchNorm batch
//              statistics) reduced in shared memory, one fp32 atomic per channel and CTA.
//   dgrad      D[pos, ci]      = sum_tap sum_co dY[pos - tap, co] * Wt[tap][co][ci]
//              A = dY boxes (K-major), B = the SAME Wt read as an MN-major operand; a stride-2
//              convolution's data gradient is four launches, one per output parity class, that
//              store through the parity view; a second (A, B) pair can be appended to the
//              k-loop (the 1x1 projection shortcut); epilogue can add a residual gradient
//              masked by a ReLU output.
//   wgrad      D[(tap, ci), co] = sum_pos X[pos + tap, ci] * dY[pos, co]
//              both operands MN-major: the M rows are two 64-channel atoms (tap, ci-chunk), the
//              K dimension runs over positions (split over CTAs), fp32 result stored or
//              atomically added in tap-major layout dWt[tap][co][ci].
//
// Warp roles (256 threads): 0 = TMA producer, 1 = MMA issuer, 2 = TMEM allocator, 4..7 = epilogue.
#include <map>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "tc_common.cuh"

namespace nfb {
namespace rn {

using namespace nfb::tc;

constexpr int kMaxTaps = 10;
constexpr int UMMA_K = 16;

struct Tap {
  int dh, dw, ph, pw, wrow, src;
  int cls;   // output parity class this tap belongs to (-1: every class); see GemmArgs::n_cls
};

__device__ __forceinline__ void tma_load_5d(void* smem_dst, const CUtensorMap* map, uint64_t* bar,
                                            int c0, int c1, int c2, int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}
__device__ __forceinline__ void tma_store_5d(const CUtensorMap* map, const void* smem_src, int c0,
                                             int c1, int c2, int c3, int c4) {
  asm volatile(
      "cp.async.bulk.tensor.5d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];" ::"l"(
          map),
      "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
      : "memory");
}

// first image / first grid row of position box `pb` (boxes of POS positions, counted over all
// images of all clients) on an H x W grid
__device__ __forceinline__ void box_origin(int pb, int pos, int H, int W, int& n0, int& h0) {
  const int hw = H * W;
  if (hw >= pos) {
    const int per = hw / pos;
    n0 = pb / per;
    h0 = (pb - n0 * per) * (pos / W);
  } else {
    n0 = pb * (pos / hw);
    h0 = 0;
  }
}

// ---------------------------------------------------------------------------------------------
// forward / data gradient
// ---------------------------------------------------------------------------------------------
struct GemmArgs {
  int H, W;              // position grid of the M side (output positions)
  int P;                 // positions per client
  int n_taps;            // total taps (primary first, then secondary)
  int kc[2];             // 64-channel k-chunks per tap for source 0 / 1
  int cpar[2];           // channel count of A tensor 0 / 1 (parity views add pw * cpar)
  Tap tap[kMaxTaps];
  int n_out;             // output channels
  int out_ph, out_pw, out_c;  // output parity class (parity view) and channel count of the output
  int clients, n_cls;         // n_cls = 4: grid.z also runs over the parity classes (taps carry cls)
  float* stats;          // [clients][2][n_out] or null
  const __nv_bfloat16* residual;  // [clients * P][n_out] added to the result (dgrad) or null
  const __nv_bfloat16* res_mask;  // residual counts only where res_mask > 0 (or null)
  // ReLU backward of the layer this gradient flows INTO, fused: the stored value is zeroed where
  // out_mask (dense tensor of the output's shape, e.g. the activation whose gradient this is) is
  // not positive, so the BatchNorm backward kernels downstream need no mask operand
  const __nv_bfloat16* out_mask;
  int out_is_parity;
  // BatchNorm backward sums fused into a data gradient: with bn_x (the raw conv output that the
  // BatchNorm under the masked activation normalises; dense, shape of the output) `stats` receives
  // per channel sum(g) and sum(g * x) of the stored gradient g instead of sum / sum of squares
  const __nv_bfloat16* bn_x;
  // GroupNorm statistics of a forward convolution: per image and group (n_out / 8 channels) the sum
  // and the sum of squares of the stored values, [images][8][2] fp32 (accumulated).  A warp's 32
  // rows lie inside one image (or two, on the 4 x 4 grid), so the column sums of the staged tile
  // reduce over the lanes of a group and leave with one atomic pair per warp, chunk and group.
  float* gn_stats;
  int* status;
};

template <bool B_MN, int BN, int HALVES>
struct GemmCfg {
  static constexpr int A_TILE = 128 * 128;            // 128 positions x 64 channels bf16
  static constexpr int B_TILE = BN * 128;             // BN rows x 64 (K-major) == BN/64 atoms of 64 x 64
  static constexpr int STAGE = HALVES * A_TILE + B_TILE;
  // PERSISTENT kernel, one CTA per SM: the whole shared memory is pipeline.  Measured on the first
  // (one tile per CTA, two CTAs per SM, 2 stages each) version: tensor pipe 30 %, L2 21 %, issue
  // 12 % - a latency chain: ~180 KB must be in flight per SM to cover ~1 us of TMA latency at the
  // MMA's consumption rate, and prologue + epilogue were 60 % of a CTA's life.
  // one epilogue warpgroup (4 warps = the four TMEM lane quarters) per 128-row half: with a single
  // group the epilogue of a 256 x 64 tile (2 TMEM reads, stores, statistics) took longer than its
  // MMAs and the persistent kernel was epilogue-bound on the 64-channel layers
  static constexpr int EPI_WARPS = 4 * HALVES;
  static constexpr int THREADS = 128 + 32 * EPI_WARPS;
  static constexpr int STAGING = EPI_WARPS * 4096;
  // data gradients only: a second 4 KB tile per epilogue warp, TMA-loaded with the BatchNorm input
  // x of the chunk while the accumulator is read (sum(g * x) for the BatchNorm backward).  Plain
  // loads of x after the store exposed their latency once per chunk and cost more than the
  // separate reduction pass they replaced.
  static constexpr int XSTAGING = B_MN ? EPI_WARPS * 4096 : 0;
  // every byte of the 227 KB counts: the kernel is latency-bound, a fourth 48 KB stage is worth
  // more than slack (budget = 227 KB - alignment - barriers - BatchNorm sums - staging)
  static constexpr int BUDGET = 227 * 1024 - 1024 - 512 - 2 * BN * 4 - STAGING - XSTAGING;
  static constexpr int STAGES = (BUDGET / STAGE) > 8 ? 8 : (BUDGET / STAGE);
  // two accumulator sets when they fit: the epilogue of tile i overlaps the MMAs of tile i + 1
  static constexpr int ACC_STAGES = (2 * HALVES * BN <= 512) ? 2 : 1;
  static constexpr int ACC_COLS = HALVES * BN;
  static constexpr int TMEM_NEED = ACC_STAGES * ACC_COLS;
  static constexpr int TMEM_COLS = TMEM_NEED <= 32 ? 32 : TMEM_NEED <= 64 ? 64
                                   : TMEM_NEED <= 128 ? 128 : TMEM_NEED <= 256 ? 256 : 512;
  static constexpr size_t SMEM = static_cast<size_t>(STAGES) * STAGE + STAGING + XSTAGING +
                                 2 * BN * 4 + 1024 /*align*/ + 512 /*barriers*/;
  static_assert(STAGES >= 2, "pipeline needs two stages");
  static_assert(SMEM <= 227 * 1024, "shared memory budget");
};

// Tile t of a launch: t = ((cls * clients + g) * n_tiles + ntile) * m_tiles + mtile (mtile fastest, so
// the CTAs that run concurrently share a client's weight tile in L2).
struct TileCoord {
  int mtile, ntile, g, cls;
};
__device__ __forceinline__ TileCoord tile_coord(int t, int m_tiles, int n_tiles, int clients) {
  TileCoord c;
  c.mtile = t % m_tiles;
  const int r = t / m_tiles;
  c.ntile = r % n_tiles;
  const int z = r / n_tiles;
  c.g = z % clients;
  c.cls = z / clients;
  return c;
}

template <bool B_MN, int BN, int HALVES>
__global__ void __launch_bounds__(GemmCfg<B_MN, BN, HALVES>::THREADS, 1)
    rn_conv_gemm_kernel(const __grid_constant__ CUtensorMap map_a0,
                        const __grid_constant__ CUtensorMap map_a1,
                        const __grid_constant__ CUtensorMap map_b0,
                        const __grid_constant__ CUtensorMap map_b1,
                        const __grid_constant__ CUtensorMap map_o,
                        const __grid_constant__ CUtensorMap map_x, const GemmArgs args) {
  pdl_trigger();   // the next kernel of the stream may start its prologue
  using C = GemmCfg<B_MN, BN, HALVES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* staging = smem + C::STAGES * C::STAGE;
  uint8_t* xstaging = staging + C::STAGING;
  float* red = reinterpret_cast<float*>(xstaging + C::XSTAGING);        // [2][BN] BatchNorm sums
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(red + 2 * BN);
  uint64_t* empty_bar = full_bar + C::STAGES;
  uint64_t* acc_full = empty_bar + C::STAGES;          // [ACC_STAGES]
  uint64_t* acc_empty = acc_full + C::ACC_STAGES;      // [ACC_STAGES]
  uint64_t* x_bar = acc_empty + C::ACC_STAGES;         // [EPI_WARPS]: x tile of a warp has landed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(x_bar + C::EPI_WARPS);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int m_tiles = args.P / (HALVES * 128), n_tiles = args.n_out / BN;
  const int total_tiles = m_tiles * n_tiles * args.clients * args.n_cls;
  // Tiles are dealt round-robin (CTA b takes b, b + grid, ...): concurrent CTAs share a client's
  // weight tile in L2 and the parity classes of a stride-2 dgrad (1 to 5 taps per tile) interleave.
  // Contiguous ranges were tried for the sake of register-resident BatchNorm sums: faster on the
  // 64-channel layers, 25-55 % slower on the deep ones (class imbalance, weight locality).

  if (threadIdx.x == 0) {
    prefetch_tmap(&map_a0);
    prefetch_tmap(&map_b0);
    prefetch_tmap(&map_o);
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < C::ACC_STAGES; ++s) {
      mbar_init(&acc_full[s], 1);
      mbar_init(&acc_empty[s], C::EPI_WARPS);          // one arrival per epilogue warp
    }
    for (int s = 0; s < C::EPI_WARPS; ++s) mbar_init(&x_bar[s], 1);
    fence_barrier_init();
  }
  for (int i = threadIdx.x; i < 2 * BN; i += C::THREADS) red[i] = 0.f;
  if (warp == 2) tmem_alloc<C::TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();   // barriers, TMEM and tensor maps are ready: now wait for the producer of our inputs

  // k-iterations of one tile of class `cls`
  auto iters_of = [&](int cls) {
    int n = 0;
    for (int t = 0; t < args.n_taps; ++t)
      if (args.tap[t].cls < 0 || args.tap[t].cls == cls) n += args.kc[args.tap[t].src];
    return n;
  };

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      int i = 0;                                       // stage counter, runs across tiles
      bool ok = true;
      for (int tile = blockIdx.x; tile < total_tiles && ok; tile += gridDim.x) {
        const TileCoord tc_ = tile_coord(tile, m_tiles, n_tiles, args.clients);
        const int pb0 = (tc_.g * args.P + tc_.mtile * HALVES * 128) >> 7;
        int n0[HALVES], h0[HALVES];
#pragma unroll
        for (int h = 0; h < HALVES; ++h) box_origin(pb0 + h, 128, args.H, args.W, n0[h], h0[h]);
        for (int t = 0; t < args.n_taps && ok; ++t) {
          const Tap tp = args.tap[t];
          if (tp.cls >= 0 && tp.cls != tc_.cls) continue;
          const CUtensorMap* ma = tp.src ? &map_a1 : &map_a0;
          const CUtensorMap* mb = tp.src ? &map_b1 : &map_b0;
          const int kcn = args.kc[tp.src];
          for (int kc = 0; kc < kcn; ++kc, ++i) {
            const int s = i % C::STAGES;
            const uint32_t ph = (i / C::STAGES) & 1;
            if (!mbar_wait(&empty_bar[s], ph ^ 1)) { *args.status = 1; ok = false; break; }
            uint8_t* stage = smem + s * C::STAGE;
            mbar_expect_tx(&full_bar[s], C::STAGE);
#pragma unroll
            for (int h = 0; h < HALVES; ++h)
              tma_load_5d(stage + h * C::A_TILE, ma, &full_bar[s],
                          kc * 64 + tp.pw * args.cpar[tp.src], tp.dw, tp.ph, h0[h] + tp.dh, n0[h]);
            uint8_t* bdst = stage + HALVES * C::A_TILE;
            if constexpr (!B_MN) {
              tma_load_3d(bdst, mb, &full_bar[s], kc * 64, tp.wrow + tc_.ntile * BN, tc_.g);
            } else {
#pragma unroll
              for (int a = 0; a < BN / 64; ++a)
                tma_load_3d(bdst + a * 8192, mb, &full_bar[s], tc_.ntile * BN + a * 64,
                            tp.wrow + kc * 64, tc_.g);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(128, BN, 0u, B_MN ? 1u : 0u);
      int i = 0, local = 0;
      bool ok = true;
      for (int tile = blockIdx.x; tile < total_tiles && ok; tile += gridDim.x, ++local) {
        const TileCoord tc_ = tile_coord(tile, m_tiles, n_tiles, args.clients);
        const int n_iters = iters_of(tc_.cls);
        const int as = local % C::ACC_STAGES;
        const uint32_t aph = (local / C::ACC_STAGES) & 1;
        if (!mbar_wait(&acc_empty[as], aph ^ 1)) { *args.status = 1; ok = false; break; }
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * C::ACC_COLS;
        for (int it = 0; it < n_iters; ++it, ++i) {
          const int s = i % C::STAGES;
          const uint32_t ph = (i / C::STAGES) & 1;
          if (!mbar_wait(&full_bar[s], ph)) { *args.status = 1; ok = false; break; }
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + s * C::STAGE);
          const uint32_t b_addr = a_addr + HALVES * C::A_TILE;
#pragma unroll
          for (int h = 0; h < HALVES; ++h) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
              const uint64_t da = make_smem_desc(a_addr + h * C::A_TILE + kk * 32, 0, 1024, 2u);
              const uint64_t db = B_MN ? make_smem_desc(b_addr + kk * 2048, 8192, 1024, 2u)
                                       : make_smem_desc(b_addr + kk * 32, 0, 1024, 2u);
              umma_f16(d_tmem + h * BN, da, db, idesc, (it > 0 || kk > 0) ? 1u : 0u);
            }
          }
          umma_commit(&empty_bar[s]);
        }
        umma_commit(&acc_full[as]);
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int sub = warp & 3;                          // TMEM lane quarter of this warp
    const int grp = (warp - 4) >> 2;                   // = the 128-row half this warpgroup owns
    uint8_t* my_stage = staging + (warp - 4) * 4096;
    uint8_t* my_x = xstaging + (C::XSTAGING ? (warp - 4) * 4096 : 0);
    const bool bn_bwd = B_MN && args.stats && args.bn_x;
    uint32_t x_phase = 0;
    int local = 0;
    bool ok = true;
    for (int tile = blockIdx.x; tile < total_tiles && ok; tile += gridDim.x, ++local) {
      const TileCoord tc_ = tile_coord(tile, m_tiles, n_tiles, args.clients);
      const int mtile = tc_.mtile, ntile = tc_.ntile, g = tc_.g;
      const int out_ph = args.n_cls > 1 ? (tc_.cls >> 1) : args.out_ph;
      const int out_pw = args.n_cls > 1 ? (tc_.cls & 1) : args.out_pw;
      const int as = local % C::ACC_STAGES;
      const uint32_t aph = (local / C::ACC_STAGES) & 1;
      const int h = grp;
      const int pos0 = mtile * HALVES * 128 + h * 128 + sub * 32;   // within the client
      int n0, h0;
      box_origin((g * args.P + pos0) >> 5, 32, args.H, args.W, n0, h0);
      const long long row = static_cast<long long>(g) * args.P + pos0 + lane;
      // Coalesced gather of a [32 rows x 128 B] tile whose row r starts at lane r's `rp` into this
      // warp's staging buffer (same swizzle as the output tile).  Row-per-lane 16-byte loads touch
      // 32 different lines per instruction (256 tag look-ups per tile and warp: the 64-channel data
      // gradients were bound by them); here one instruction covers four whole rows.
      auto gather_rows = [&](const __nv_bfloat16* rp) {
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
        uint4 t[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int rj = j * 4 + (lane >> 3);
          const __nv_bfloat16* src = reinterpret_cast<const __nv_bfloat16*>(
              __shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(rp), rj));
          t[j] = ld_plain(src + (lane & 7) * 8);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int rj = j * 4 + (lane >> 3);
          st_vec(my_stage + rj * 128 + (((lane & 7) ^ (rj & 7)) * 16), t[j]);
        }
        __syncwarp();
      };
      // ReLU-backward mask of this warp's rows as 64 bits per chunk: fetched BEFORE the accumulator
      // is waited for (it does not depend on it), i.e. under the MMAs of this very tile
      uint32_t mbits[BN / 64][2];
      if (args.out_mask) {
        long long moff;
        if (!args.out_is_parity) {
          moff = row * args.n_out;
        } else {   // this row is pixel (2*hh + ph, 2*wh + pw) of the full-resolution output
          const int hw = args.H * args.W;
          const int n = static_cast<int>(row / hw), rem = static_cast<int>(row - static_cast<long long>(n) * hw);
          const int hh = rem / args.W, wh = rem - hh * args.W;
          moff = ((static_cast<long long>(n) * 2 * args.H + 2 * hh + out_ph) * (2 * args.W) +
                  2 * wh + out_pw) * args.out_c;
        }
        const __nv_bfloat16* mrow = args.out_mask + moff + ntile * BN;
#pragma unroll
        for (int c = 0; c < BN / 64; ++c) {
          gather_rows(mrow + c * 64);
          uint32_t lo = 0u, hi = 0u;
#pragma unroll
          for (int p = 0; p < 8; ++p) {
            const uint4 m = *reinterpret_cast<const uint4*>(my_stage + lane * 128 + ((p ^ (lane & 7)) * 16));
            const uint32_t w4[4] = {m.x, m.y, m.z, m.w};
            uint32_t byte = 0u;
#pragma unroll
            for (int q = 0; q < 4; ++q) {   // bf16 > 0  <=>  sign clear and magnitude != 0
              byte |= (((w4[q] & 0x8000u) == 0u && (w4[q] & 0x7fffu) != 0u) ? 1u : 0u) << (2 * q);
              byte |= (((w4[q] & 0x80000000u) == 0u && (w4[q] & 0x7fff0000u) != 0u) ? 1u : 0u) << (2 * q + 1);
            }
            if (p < 4) lo |= byte << (8 * p); else hi |= byte << (8 * (p - 4));
          }
          mbits[c][0] = lo; mbits[c][1] = hi;
        }
      }
      if (!mbar_wait(&acc_full[as], aph)) { if (lane == 0) *args.status = 1; ok = false; break; }
      tc_fence_after();
      const uint32_t lane_addr =
          tmem_base + as * C::ACC_COLS + (static_cast<uint32_t>(sub * 32) << 16);
      float csum[BN / 32], csq[BN / 32];   // this lane's two columns of every 64-column chunk
#pragma unroll
      for (int j = 0; j < BN / 32; ++j) csum[j] = csq[j] = 0.f;
      {
#pragma unroll
        for (int c = 0; c < BN / 64; ++c) {
          if (bn_bwd) {
            __syncwarp();          // every lane is done with the previous x tile
            if (lane == 0) {
              mbar_expect_tx(&x_bar[warp - 4], 4096);
              tma_load_5d(my_x, &map_x, &x_bar[warp - 4],
                          ntile * BN + c * 64 + out_pw * args.out_c, 0, out_ph, h0, n0);
            }
          }
          uint32_t r[64];
#pragma unroll
          for (int q = 0; q < 4; ++q)
            tmem_ld_32x32b_x16(lane_addr + h * BN + c * 64 + q * 16, r + q * 16);
          tmem_ld_wait();
          if (c == BN / 64 - 1) {
            // the accumulator set is in registers / already consumed: hand it back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[as]);
          }
          float v[64];
#pragma unroll
          for (int j = 0; j < 64; ++j) v[j] = __uint_as_float(r[j]);
          if (args.residual && !args.res_mask) {
            gather_rows(args.residual + row * args.n_out + ntile * BN + c * 64);
#pragma unroll
            for (int p = 0; p < 8; ++p) {
              float rf[8];
              unpack<__nv_bfloat16>(*reinterpret_cast<const uint4*>(
                                        my_stage + lane * 128 + ((p ^ (lane & 7)) * 16)), rf);
#pragma unroll
              for (int e = 0; e < 8; ++e) v[p * 8 + e] += rf[e];
            }
          } else if (args.residual) {
            const long long off = row * args.n_out + ntile * BN + c * 64;
#pragma unroll
            for (int p = 0; p < 8; ++p) {
              float rf[8], mf[8];
              unpack<__nv_bfloat16>(ld_plain(args.residual + off + p * 8), rf);
              unpack<__nv_bfloat16>(ld_plain(args.res_mask + off + p * 8), mf);
#pragma unroll
              for (int e = 0; e < 8; ++e) v[p * 8 + e] += mf[e] > 0.f ? rf[e] : 0.f;
            }
          }
          if (args.out_mask) {
#pragma unroll
            for (int j = 0; j < 64; ++j)
              v[j] = ((mbits[c][j >> 5] >> (j & 31)) & 1u) ? v[j] : 0.f;
          }
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
          __syncwarp();
#pragma unroll
          for (int p = 0; p < 8; ++p)
            st_vec(my_stage + lane * 128 + ((p ^ (lane & 7)) * 16), pack<__nv_bfloat16>(v + 8 * p));
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            tma_store_5d(&map_o, my_stage, ntile * BN + c * 64 + out_pw * args.out_c, 0, out_ph, h0,
                         n0);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
          if (!B_MN && args.gn_stats) {
            const int hw = args.H * args.W;                   // positions per image
            const int cg = args.n_out >> 3;                   // channels per group (8, 16, 32, 64)
            const int lpg = cg >> 1;                          // lanes per group (2 columns per lane)
            const int rpi = hw < 32 ? hw : 32;                // rows of one image in this warp's tile
            for (int r0 = 0; r0 < 32; r0 += rpi) {
              float s = 0.f, q = 0.f;
              for (int rr = r0; rr < r0 + rpi; ++rr) {
                const uint32_t w = *reinterpret_cast<const uint32_t*>(
                    my_stage + rr * 128 + (((lane >> 2) ^ (rr & 7)) * 16) + (lane & 3) * 4);
                const float a = __uint_as_float(w << 16), b = __uint_as_float(w & 0xffff0000u);
                s += a + b;
                q = fmaf(a, a, fmaf(b, b, q));
              }
              for (int o = 1; o < lpg; o <<= 1) {
                s += __shfl_xor_sync(0xffffffffu, s, o);
                q += __shfl_xor_sync(0xffffffffu, q, o);
              }
              if ((lane & (lpg - 1)) == 0) {
                const long long img = (row - lane + r0) / hw;
                const int grp = (ntile * BN + c * 64 + 2 * lane) / cg;
                atomicAdd(args.gn_stats + (img * 8 + grp) * 2, s);
                atomicAdd(args.gn_stats + (img * 8 + grp) * 2 + 1, q);
              }
            }
          }
          if (args.stats) {
            // column sums over the ROUNDED values in the staged tile: lane l owns columns 2l, 2l+1
            // of the chunk; row r's piece p sits at (p ^ (r & 7)) * 16.
            //   forward:        sum(y), sum(y * y)      (what the normalisation pass will read)
            //   with bn_x:      sum(g), sum(g * x)      (BatchNorm backward of the layer whose raw
            //                   output x produced this gradient's mask)
            if (bn_bwd) {
              if (!mbar_wait(&x_bar[warp - 4], x_phase)) {
                if (lane == 0) *args.status = 1;
                ok = false;
              }
              x_phase ^= 1;
            }
            float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll 8
            for (int rr = 0; rr < 32; ++rr) {
              const uint32_t w = *reinterpret_cast<const uint32_t*>(
                  my_stage + rr * 128 + (((lane >> 2) ^ (rr & 7)) * 16) + (lane & 3) * 4);
              const float a = __uint_as_float(w << 16), b = __uint_as_float(w & 0xffff0000u);
              float xa = a, xb = b;
              if (bn_bwd) {   // the x tile has the layout of the staged tile (same box, same swizzle)
                const uint32_t xw = *reinterpret_cast<const uint32_t*>(
                    my_x + rr * 128 + (((lane >> 2) ^ (rr & 7)) * 16) + (lane & 3) * 4);
                xa = __uint_as_float(xw << 16); xb = __uint_as_float(xw & 0xffff0000u);
              }
              s0 += a; s1 += b;
              q0 = fmaf(a, xa, q0); q1 = fmaf(b, xb, q1);
            }
            csum[2 * c] += s0; csum[2 * c + 1] += s1;
            csq[2 * c] += q0; csq[2 * c + 1] += q1;
          }
        }
      }
      if (args.stats) {
        // the tile's sums: all epilogue warps add into shared memory, then ONE fp32 atomic per
        // channel and sum leaves the CTA (a warp-level flush put 8x as many atomics onto the few
        // hundred addresses of a 64-channel layer and halved its speed)
#pragma unroll
        for (int c = 0; c < BN / 64; ++c) {
          atomicAdd(&red[c * 64 + 2 * lane], csum[2 * c]);
          atomicAdd(&red[c * 64 + 2 * lane + 1], csum[2 * c + 1]);
          atomicAdd(&red[BN + c * 64 + 2 * lane], csq[2 * c]);
          atomicAdd(&red[BN + c * 64 + 2 * lane + 1], csq[2 * c + 1]);
        }
        asm volatile("bar.sync 1, %0;" ::"n"(32 * C::EPI_WARPS) : "memory");
        float* dst = args.stats + static_cast<long long>(g) * 2 * args.n_out;
        for (int i = threadIdx.x - 128; i < 2 * BN; i += 32 * C::EPI_WARPS) {
          const int which = i / BN, col = i - which * BN;
          atomicAdd(dst + which * args.n_out + ntile * BN + col, red[i]);
          red[i] = 0.f;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(32 * C::EPI_WARPS) : "memory");
      }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<C::TMEM_COLS>(tmem_base);
}

// ---------------------------------------------------------------------------------------------
// weight gradient
// ---------------------------------------------------------------------------------------------
struct WgradArgs {
  int H, W;              // position grid of dY
  int P;                 // positions per client
  int k_split;
  int n_taps, ci_chunks; // M atoms = n_taps * ci_chunks, two per CTA
  int cpar;              // channels of X (parity views)
  Tap tap[kMaxTaps];
  int c_out, c_in;       // dWt[g][tap][c_out][c_in]
  float* out;
  long long out_gs;
  int atomic;
  float* sgd_w;            // fused SGD: tap-major fp32 master (null = write the gradient)
  __nv_bfloat16* sgd_wt;   // ... and its bf16 working copy
  float lr, wd;
  int* status;
};

template <int BN, int POSK>
struct WgradCfg {
  static constexpr int ATOM = POSK * 128;             // POSK positions x 64 channels
  static constexpr int STAGE = (2 + BN / 64) * ATOM;
  // POSK = 64: at most ~100 KB so that two CTAs share an SM (the epilogue of one - TMEM reads,
  // fp32 atomics or the fused SGD update - overlaps the k-loop of the other)
  static constexpr int BUDGET = POSK == 64 ? 100 * 1024 : 200 * 1024;
  static constexpr int STAGES = (BUDGET / STAGE) > 6 ? 6 : (BUDGET / STAGE);
  static constexpr int TMEM_COLS = BN <= 32 ? 32 : BN <= 64 ? 64 : BN <= 128 ? 128 : 256;
  static constexpr size_t SMEM = static_cast<size_t>(STAGES) * STAGE + 1024 + 256;
  static_assert(STAGES >= 2, "pipeline needs two stages");
};

template <int BN, int POSK>
__global__ void __launch_bounds__(256, 1)
    rn_conv_wgrad_kernel(const __grid_constant__ CUtensorMap map_x,
                         const __grid_constant__ CUtensorMap map_dy, const WgradArgs args) {
  pdl_trigger();   // the next kernel of the stream may start its prologue
  using C = WgradCfg<BN, POSK>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + C::STAGES * C::STAGE);
  uint64_t* empty_bar = full_bar + C::STAGES;
  uint64_t* acc_bar = empty_bar + C::STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int mtile = blockIdx.x, ntile = blockIdx.y;
  const int g = blockIdx.z / args.k_split, split = blockIdx.z % args.k_split;
  const int n_atoms = args.n_taps * args.ci_chunks;
  const int kb_total = args.P / POSK;
  const int per = (kb_total + args.k_split - 1) / args.k_split;
  const int kb_lo = split * per, kb_hi = min(kb_lo + per, kb_total);
  const int n_iters = max(kb_hi - kb_lo, 0);

  if (threadIdx.x == 0) {
    prefetch_tmap(&map_x);
    prefetch_tmap(&map_dy);
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(acc_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<C::TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();   // barriers, TMEM and tensor maps are ready: now wait for the producer of our inputs

  if (warp == 0) {
    if (elect_one()) {
      Tap tp[2];
      int cich[2];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int q = min(mtile * 2 + j, n_atoms - 1);
        tp[j] = args.tap[q / args.ci_chunks];
        cich[j] = q % args.ci_chunks;
      }
      for (int i = 0; i < n_iters; ++i) {
        const int s = i % C::STAGES;
        const uint32_t ph = (i / C::STAGES) & 1;
        if (!mbar_wait(&empty_bar[s], ph ^ 1)) { *args.status = 1; break; }
        uint8_t* stage = smem + s * C::STAGE;
        mbar_expect_tx(&full_bar[s], C::STAGE);
        int n0, h0;
        box_origin(g * kb_total + kb_lo + i, POSK, args.H, args.W, n0, h0);
#pragma unroll
        for (int j = 0; j < 2; ++j)
          tma_load_5d(stage + j * C::ATOM, &map_x, &full_bar[s],
                      cich[j] * 64 + tp[j].pw * args.cpar, tp[j].dw, tp[j].ph, h0 + tp[j].dh, n0);
#pragma unroll
        for (int a = 0; a < BN / 64; ++a)
          tma_load_5d(stage + (2 + a) * C::ATOM, &map_dy, &full_bar[s], ntile * BN + a * 64, 0, 0,
                      h0, n0);
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(128, BN, 1u, 1u);
      for (int i = 0; i < n_iters; ++i) {
        const int s = i % C::STAGES;
        const uint32_t ph = (i / C::STAGES) & 1;
        if (!mbar_wait(&full_bar[s], ph)) { *args.status = 1; break; }
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + s * C::STAGE);
        const uint32_t b_addr = a_addr + 2 * C::ATOM;
#pragma unroll
        for (int kk = 0; kk < POSK / UMMA_K; ++kk) {
          const uint64_t da = make_smem_desc(a_addr + kk * 2048, C::ATOM, 1024, 2u);
          const uint64_t db = make_smem_desc(b_addr + kk * 2048, C::ATOM, 1024, 2u);
          umma_f16(tmem_base, da, db, idesc, (i > 0 || kk > 0) ? 1u : 0u);
        }
        umma_commit(&empty_bar[s]);
      }
      umma_commit(acc_bar);
    }
  } else if (warp >= 4) {
    const int sub = warp - 4;
    bool ok = n_iters > 0 ? mbar_wait(acc_bar, 0) : false;
    if (n_iters > 0 && !ok && lane == 0) *args.status = 1;
    tc_fence_after();
    const int q = mtile * 2 + (sub >> 1);
    const bool live = ok && q < n_atoms;
    const int t = min(q, n_atoms - 1) / args.ci_chunks;
    const int ci = (min(q, n_atoms - 1) % args.ci_chunks) * 64 + (sub & 1) * 32 + lane;
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(sub * 32) << 16);
    const long long off0 = static_cast<long long>(g) * args.out_gs +
                           (static_cast<long long>(t) * args.c_out + ntile * BN) * args.c_in + ci;
    float* base = args.out + off0;
    const bool fused = args.sgd_w != nullptr;
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 16) {
      uint32_t r[16];
      tmem_ld_32x32b_x16(lane_addr + c0, r);
      tmem_ld_wait();
      if (!live) continue;
      if (fused) {
        // the gradient of these weights is complete in TMEM: apply it to the fp32 master and
        // refresh the bf16 copy; 32 lanes = 32 consecutive ci, so every access is one full line
        float w[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) w[j] = args.sgd_w[off0 + static_cast<long long>(c0 + j) * args.c_in];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const long long o = off0 + static_cast<long long>(c0 + j) * args.c_in;
          const float nw = fmaf(-args.lr, fmaf(args.wd, w[j], __uint_as_float(r[j])), w[j]);
          args.sgd_w[o] = nw;
          args.sgd_wt[o] = __float2bfloat16_rn(nw);
        }
        continue;
      }
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        float* p = base + static_cast<long long>(c0 + j) * args.c_in;
        if (args.atomic) atomicAdd(p, __uint_as_float(r[j]));
        else *p = __uint_as_float(r[j]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<C::TMEM_COLS>(tmem_base);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
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

static std::map<std::vector<long long>, CUtensorMap> g_maps;
static std::mutex g_mu;

// 5-D view of a bf16 NHWC tensor [imgs][H][W][C]: plain [C, W, 1, H, imgs] or, for stride-2
// sampling, the parity view [2C, W/2, 2, H/2, imgs]; box = `pos` positions x 64 channels.
static int act_map(const nfb_rn_act* t, int pos, CUtensorMap* out) {
  const long long H = t->H, W = t->W, Cc = t->C, N = t->imgs;
  const long long gh = t->parity ? H / 2 : H, gw = t->parity ? W / 2 : W;   // position grid
  std::vector<long long> key{reinterpret_cast<long long>(t->ptr), H, W, Cc, N, t->parity, pos};
  std::lock_guard<std::mutex> lock(g_mu);
  auto it = g_maps.find(key);
  if (it != g_maps.end()) { *out = it->second; return 0; }
  EncodeFn enc = encode_fn();
  if (!enc) return fail(NFB_ESTATE, "rn_conv: cuTensorMapEncodeTiled unavailable");
  if (pos % gw && gw % pos) return fail(NFB_EINVAL, "rn_conv: box of %d positions on a %lld-wide grid", pos, gw);
  const long long bw = gw < pos ? gw : pos;
  const long long bh = (pos / bw) < gh ? (pos / bw) : gh;
  const long long bn = pos / (bw * bh);
  if (bw * bh * bn != pos || bn > N || bn > 256)
    return fail(NFB_EINVAL, "rn_conv: cannot tile %d positions on a %lldx%lld grid", pos, gh, gw);
  cuuint64_t dims[5], strides[4];
  if (!t->parity) {
    dims[0] = Cc; dims[1] = W; dims[2] = 1; dims[3] = H; dims[4] = N;
    strides[0] = Cc * 2; strides[1] = W * Cc * 2; strides[2] = W * Cc * 2; strides[3] = H * W * Cc * 2;
  } else {
    dims[0] = 2 * Cc; dims[1] = W / 2; dims[2] = 2; dims[3] = H / 2; dims[4] = N;
    strides[0] = 2 * Cc * 2; strides[1] = W * Cc * 2; strides[2] = 2 * W * Cc * 2;
    strides[3] = H * W * Cc * 2;
  }
  cuuint32_t box[5] = {64, static_cast<cuuint32_t>(bw), 1, static_cast<cuuint32_t>(bh),
                       static_cast<cuuint32_t>(bn)};
  cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  if ((reinterpret_cast<uintptr_t>(t->ptr) & 127) || (Cc % 64))
    return fail(NFB_EINVAL, "rn_conv: activation must be 128-byte aligned with C %% 64 == 0");
  CUtensorMap m;
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(t->ptr), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(NFB_EINVAL, "rn_conv: activation tensor map rejected (%d)", static_cast<int>(r));
  if (g_maps.size() > 8192) g_maps.clear();
  g_maps[key] = m;
  *out = m;
  return 0;
}

static int weight_map(const void* w, int inner, int rows, int groups, long long gs, int box_rows,
                      CUtensorMap* out) {
  Operand op{w, inner, rows, groups, inner, gs};
  return get_map(op, 64, box_rows, 128, out);
}

template <bool B_MN, int BN, int HALVES>
static int launch_gemm(const CUtensorMap* maps, const GemmArgs& args, dim3 grid, cudaStream_t st) {
  using C = GemmCfg<B_MN, BN, HALVES>;
  auto kern = rn_conv_gemm_kernel<B_MN, BN, HALVES>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(C::SMEM));
    if (e != cudaSuccess) return fail(static_cast<int>(e), "rn_conv: smem attribute: %s",
                                      cudaGetErrorString(e));
    attr_set = true;
  }
  launch_pdl(kern, dim3(grid), dim3(C::THREADS), C::SMEM, st, maps[0], maps[1], maps[2], maps[3], maps[4], maps[5],
                                          args);
  return launch_status("rn_conv_gemm");
}

template <int BN, int POSK>
static int launch_wgrad(const CUtensorMap& mx, const CUtensorMap& mdy, const WgradArgs& args,
                        dim3 grid, cudaStream_t st) {
  using C = WgradCfg<BN, POSK>;
  auto kern = rn_conv_wgrad_kernel<BN, POSK>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(C::SMEM));
    if (e != cudaSuccess) return fail(static_cast<int>(e), "rn_conv: smem attribute: %s",
                                      cudaGetErrorString(e));
    attr_set = true;
  }
  launch_pdl(kern, dim3(grid), dim3(256), C::SMEM, st, mx, mdy, args);
  return launch_status("rn_conv_wgrad");
}

}  // namespace rn
}  // namespace nfb

extern "C" {

int nfb_rn_conv_gemm(int dgrad, const nfb_rn_act* a0, const nfb_rn_act* a1, const void* w0,
                     const void* w1, const int* w_dims /* [2][3]: inner, rows, group stride */,
                     int clients, int imgs_per_client, int H, int W, int n_out, int n_taps,
                     const int* taps, const nfb_rn_act* out, int out_ph, int out_pw, float* stats,
                     const void* residual, const void* res_mask, const void* out_mask,
                     const void* bn_x, float* gn_stats, int bn_tile, int halves, int n_cls,
                     int* status, void* stream) {
  using namespace nfb;
  using namespace nfb::rn;
  NFB_REQUIRE(a0 && w0 && w_dims && taps && out && status, NFB_EINVAL, "rn_conv_gemm: null pointer");
  NFB_REQUIRE(n_taps >= 1 && n_taps <= kMaxTaps, NFB_EINVAL, "rn_conv_gemm: 1..%d taps", kMaxTaps);
  NFB_REQUIRE(clients >= 1 && clients <= 65535, NFB_ERANGE, "rn_conv_gemm: bad client count");
  const int P = imgs_per_client * H * W;
  NFB_REQUIRE(halves == 1 || halves == 2, NFB_EINVAL, "rn_conv_gemm: halves must be 1 or 2");
  NFB_REQUIRE(P % (128 * halves) == 0, NFB_EINVAL,
              "rn_conv_gemm: %d positions per client is not a multiple of the %d-row tile", P,
              128 * halves);
  NFB_REQUIRE(n_out % bn_tile == 0 && n_out % 64 == 0, NFB_EINVAL,
              "rn_conv_gemm: %d outputs vs tile %d", n_out, bn_tile);
  NFB_REQUIRE(!residual || (!out->parity && out->C == n_out), NFB_EINVAL,
              "rn_conv_gemm: residual needs a dense output");
  GemmArgs g{};
  g.H = H; g.W = W; g.P = P; g.n_taps = n_taps; g.n_out = n_out;
  g.out_ph = out_ph; g.out_pw = out_pw; g.out_c = out->C;
  g.stats = stats;
  g.residual = static_cast<const __nv_bfloat16*>(residual);
  g.res_mask = static_cast<const __nv_bfloat16*>(res_mask);
  g.out_mask = static_cast<const __nv_bfloat16*>(out_mask);
  g.bn_x = static_cast<const __nv_bfloat16*>(bn_x);
  g.gn_stats = gn_stats;
  NFB_REQUIRE(!gn_stats || (!dgrad && !out->parity && out->C == n_out && n_out % 64 == 0 &&
                            ((H * W) % 32 == 0 || 32 % (H * W) == 0)),
              NFB_EINVAL, "rn_conv_gemm: GroupNorm statistics need a dense forward output whose "
              "images are multiples or divisors of 32 positions");
  NFB_REQUIRE(!bn_x || (dgrad && stats && out->C == n_out), NFB_EINVAL,
              "rn_conv_gemm: BatchNorm-backward sums need a data gradient, a stats buffer and a "
              "full-width output");
  g.out_is_parity = out->parity ? 1 : 0;
  NFB_REQUIRE(!out_mask || out->C == n_out, NFB_EINVAL, "rn_conv_gemm: output mask needs a full-width output");
  g.status = status;
  bool use1 = false;
  NFB_REQUIRE(n_cls == 1 || (n_cls == 4 && out->parity), NFB_EINVAL,
              "rn_conv_gemm: classes are the 4 parities of a parity-view output");
  NFB_REQUIRE(static_cast<long long>(clients) * n_cls <= 65535, NFB_ERANGE, "rn_conv_gemm: grid");
  g.clients = clients; g.n_cls = n_cls;
  for (int t = 0; t < n_taps; ++t) {
    g.tap[t] = Tap{taps[7 * t], taps[7 * t + 1], taps[7 * t + 2], taps[7 * t + 3], taps[7 * t + 4],
                   taps[7 * t + 5], n_cls > 1 ? taps[7 * t + 6] : -1};
    NFB_REQUIRE(g.tap[t].src == 0 || g.tap[t].src == 1, NFB_EINVAL, "rn_conv_gemm: tap source");
    use1 |= g.tap[t].src == 1;
  }
  NFB_REQUIRE(!use1 || (a1 && w1), NFB_EINVAL, "rn_conv_gemm: second operand pair missing");
  CUtensorMap maps[6];
  int rc = act_map(a0, 128, &maps[0]);
  if (rc) return rc;
  maps[1] = maps[0];
  if (use1 && (rc = act_map(a1, 128, &maps[1]))) return rc;
  // weights [clients][rows][inner]: forward reads [BN rows x 64 k]; dgrad reads [64 k-rows x 64 n]
  const int box_rows = dgrad ? 64 : bn_tile;
  rc = weight_map(w0, w_dims[0], w_dims[1], clients, w_dims[2], box_rows, &maps[2]);
  if (rc) return rc;
  maps[3] = maps[2];
  if (use1 && (rc = weight_map(w1, w_dims[3], w_dims[4], clients, w_dims[5], box_rows, &maps[3])))
    return rc;
  rc = act_map(out, 32, &maps[4]);
  if (rc) return rc;
  maps[5] = maps[4];
  if (bn_x) {   // same geometry as the output: the epilogue loads the tile it is about to store
    nfb_rn_act xa = *out;
    xa.ptr = bn_x;
    if ((rc = act_map(&xa, 32, &maps[5]))) return rc;
  }
  // k-chunks: forward K = input channels = weight inner; dgrad K = output channels of the conv
  g.cpar[0] = a0->C; g.cpar[1] = use1 ? a1->C : 0;
  g.kc[0] = a0->C / 64; g.kc[1] = use1 ? a1->C / 64 : 0;
  const long long tiles = static_cast<long long>(P / (128 * halves)) * (n_out / bn_tile) * clients * n_cls;
  const long long sms = sm_count();
  dim3 grid(static_cast<unsigned>(tiles < sms ? tiles : sms));
  cudaStream_t st = as_stream(stream);
#define NFB_RN_CASE(MN, BNV, HV)                                                     \
  if ((dgrad != 0) == MN && bn_tile == BNV && halves == HV)                          \
    return launch_gemm<MN, BNV, HV>(maps, g, grid, st);
  NFB_RN_CASE(false, 64, 1) NFB_RN_CASE(false, 64, 2) NFB_RN_CASE(false, 128, 1)
  NFB_RN_CASE(false, 128, 2) NFB_RN_CASE(false, 256, 1) NFB_RN_CASE(false, 256, 2)
  NFB_RN_CASE(true, 64, 1) NFB_RN_CASE(true, 64, 2) NFB_RN_CASE(true, 128, 1)
  NFB_RN_CASE(true, 128, 2) NFB_RN_CASE(true, 256, 1) NFB_RN_CASE(true, 256, 2)
#undef NFB_RN_CASE
  return fail(NFB_EINVAL, "rn_conv_gemm: no variant for tile n = %d, halves = %d", bn_tile, halves);
}

int nfb_rn_conv_wgrad(const nfb_rn_act* x, const nfb_rn_act* dy, int clients, int imgs_per_client,
                      int n_taps, const int* taps, float* out, long long out_gs, int k_split,
                      int bn_tile, int posk, float* sgd_w, void* sgd_wt, float lr, float wd,
                      int* status, void* stream) {
  using namespace nfb;
  using namespace nfb::rn;
  NFB_REQUIRE(x && dy && taps && out && status, NFB_EINVAL, "rn_conv_wgrad: null pointer");
  NFB_REQUIRE(n_taps >= 1 && n_taps <= kMaxTaps && k_split >= 1, NFB_EINVAL, "rn_conv_wgrad: taps");
  NFB_REQUIRE(!dy->parity, NFB_EINVAL, "rn_conv_wgrad: dy is a dense tensor");
  const int P = imgs_per_client * dy->H * dy->W;
  NFB_REQUIRE(P % posk == 0 && dy->C % bn_tile == 0, NFB_EINVAL, "rn_conv_wgrad: bad tiling");
  NFB_REQUIRE(static_cast<long long>(clients) * k_split <= 65535, NFB_ERANGE, "rn_conv_wgrad: grid");
  WgradArgs g{};
  g.H = dy->H; g.W = dy->W; g.P = P; g.k_split = k_split; g.n_taps = n_taps;
  g.ci_chunks = x->C / 64; g.cpar = x->C; g.c_out = dy->C; g.c_in = x->C;
  g.out = out; g.out_gs = out_gs; g.atomic = k_split > 1 ? 1 : 0; g.status = status;
  NFB_REQUIRE(!sgd_w || (k_split == 1 && sgd_wt), NFB_EINVAL,
              "rn_conv_wgrad: the fused update needs the complete gradient (k_split == 1)");
  g.sgd_w = sgd_w; g.sgd_wt = static_cast<__nv_bfloat16*>(sgd_wt); g.lr = lr; g.wd = wd;
  for (int t = 0; t < n_taps; ++t)
    g.tap[t] = Tap{taps[7 * t], taps[7 * t + 1], taps[7 * t + 2], taps[7 * t + 3], taps[7 * t + 4],
                   taps[7 * t + 5], -1};
  CUtensorMap mx, mdy;
  int rc = act_map(x, posk, &mx);
  if (rc) return rc;
  if ((rc = act_map(dy, posk, &mdy))) return rc;
  const int atoms = n_taps * g.ci_chunks;
  dim3 grid((atoms + 1) / 2, dy->C / bn_tile, clients * k_split);
  cudaStream_t st = as_stream(stream);
#define NFB_RN_CASE(BNV, PK) \
  if (bn_tile == BNV && posk == PK) return launch_wgrad<BNV, PK>(mx, mdy, g, grid, st);
  NFB_RN_CASE(64, 128) NFB_RN_CASE(128, 128) NFB_RN_CASE(256, 128)
  NFB_RN_CASE(64, 64) NFB_RN_CASE(128, 64) NFB_RN_CASE(256, 64)
#undef NFB_RN_CASE
  return fail(NFB_EINVAL, "rn_conv_wgrad: no variant for tile n = %d, k = %d", bn_tile, posk);
}

}  // extern "C"
