// K-CONV, persistent form: 3x3 convolution forward / data-gradient on the row-flattened
// NHWC grid, one CTA per (client, contiguous range of 128-position tiles).
//
// Why a second kernel next to tc_gemm.cu: there, each of the 9 taps re-loads a 128-row A
// tile (9x redundant L2->SMEM traffic, measured L2-bound).  Here
//   * the client's 9 weight tiles (36 KB bf16) are loaded ONCE per CTA and stay in SMEM;
//   * per output tile ONE slab of 128+54 activation rows is loaded by a single 4-D TMA
//     (box 8 ch x 182 rows x CH/8 chunks) in the un-swizzled "interleaved" UMMA layout
//     [channel-chunk][row][16 B], in which a row shift is just +16 B on the descriptor
//     start address — so the 9 taps are 9 descriptors into the same slab;
//   * two TMEM accumulators alternate, so the epilogue (tcgen05.ld, bias/ReLU/ReLU-mask,
//     bf16 pack, 16-byte stores) of tile i overlaps the MMAs of tile i+1.
//
// MODE 0 (forward):  out[q, co] = relu(sum_t sum_ci x[q + s_t, ci] w[co, t*32+ci] + b[co])
//                    A = x [C,R,32] (4 chunks), B_t = w[:, t*32:(t+1)*32] K-major SW64, N = 64
// MODE 2 (forward + pool): MODE 0's math on tiles of 4 grid rows (104 positions), with bias,
//                    ReLU, 2x2 max-pool and dropout in the epilogue: the conv output never
//                    reaches memory — only the pooled activations (PyTorch flatten order) and a
//                    1-byte argmax/ReLU/keep code per pooled element are written.
// MODE 1 (dgrad):    out[q, ci] = (sum_t sum_co dy[q - s_t, co] w[co, t*32+ci]) * (mask[q,ci] > 0)
//                    A = dy [C,R,64] (8 chunks), B_t = same weights read MN-major SW64, N = 32
#include <cstdlib>

#include "common.cuh"
#include "philox.cuh"
#include "tc_common.cuh"

#ifndef NFB_CONV_POOL_GROUPS
#define NFB_CONV_POOL_GROUPS 4
#endif

namespace nfb {
namespace tc {

constexpr int kTaps = 9;
constexpr int kHalo = 54;            // 2*26 + 2 : largest tap shift on the 26-wide grid
constexpr int kSlabRows = 128 + kHalo;

struct ConvArgs {
  int tiles_per_client;   // ceil(R / 128)
  int n_clients;          // clients covered by this launch
  int client0;            // first client of this launch
  int cta0;               // global index of this launch's first CTA (dropout counter)
  int rows;               // R
  int shift[kTaps];       // slab-relative row shift per tap
  void* out;              // bf16 [C, R, N]
  long long out_gs;       // elements between clients
  const float* bias;      // MODE 0: [C][64] with stride bias_gs
  long long bias_gs;
  const __nv_bfloat16* mask;  // MODE 1: [C, R, 32] ReLU mask source (same strides as out)
  // MODE 2 (fused pool) only
  __nv_bfloat16* pool_out;  // [C, B, 9216]   ch*144 + ph*12 + pw
  uint8_t* pool_idx;        // [C, B, 144, 64]
  int images;               // B
  int grid_w;               // 26
  float drop_p;
  const unsigned long long* rng_state;
  unsigned long long seed;
  // MODE 3 (dgrad + conv1 weight gradient) only: the layer-1 input images and gradient slab
  const float* x1;          // [C][B][img_w * img_w] fp32, img_w = grid_w + 2
  long long x1_gs;          // elements between clients
  float* grad1;             // gradient slab rows [C][grad1_gs]
  long long grad1_gs;
  long long w1_off, b1_off; // conv1.weight [32][9] / conv1.bias [32] inside a row
  int* status;
  int debug;  // bottleneck probes (env NFB_CONV_DEBUG): 1 = reuse first slabs, 2 = one tap, 4 = no stores
};

__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* map, uint64_t* bar,
                                            int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes "
      "[%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(smem_u32(smem_dst)),
      "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

// TMA store of a [BN x 32 rows] tile from shared memory (bulk_group completion).
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, const void* smem_src, int c0,
                                             int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(map),
      "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() {
  asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}

// Measured alternatives (B200, 32 clients x 64 images): 1 CTA/SM with 8 epilogue warps:
// fwd 77 us / dgrad 109 us; 2 CTAs/SM with 4 epilogue warps each: 84 / 120 us.  (Later probes:
// these kernels are paced by the shared-memory operand fetch of their small-N MMAs and by the
// 16-byte-piece slab loads, not by DRAM)
// Epilogue warpgroups (4 warps each = the four TMEM lane quarters); group j takes tiles
// i = j (mod GROUPS) and owns accumulator j.  MODE 2's pooling epilogue is a latency chain
// (two group barriers per tile): probes with loads / MMAs / stores disabled left its 118 us
// unchanged, so it gets four groups; the other modes measured best with two.
constexpr int kMaxEpiWarps = 16;

// MODE 3 = MODE 1 (dgrad) whose epilogue also accumulates the conv1 weight/bias gradient
template <int MODE>
constexpr bool kDgrad = (MODE == 1 || MODE == 3);

template <int MODE>
struct ConvCfg {
  static constexpr int GROUPS = MODE == 2 ? NFB_CONV_POOL_GROUPS : 2;
  static constexpr int EPI_WARPS = 4 * GROUPS;
  static constexpr int CH = !kDgrad<MODE> ? 32 : 64;   // A channels (K per tap)
  static constexpr int BN = !kDgrad<MODE> ? 64 : 32;   // output channels
  static constexpr int CHUNKS = CH / 8;
  static constexpr int SLAB = kSlabRows * CH * 2;      // bytes per stage
  static constexpr int SLAB_PAD = (SLAB + 1023) / 1024 * 1024;
  static constexpr int W_TILE = 64 * 32 * 2;           // one tap's weights (4 KB) either way
  static constexpr int W_BYTES = kTaps * W_TILE;
  // 128 or 64 (both powers of two); MODE 3 adds one [128 x 32] fp32 accumulator per epilogue
  // warp behind the two tile accumulators (64 + 8 * 32 = 320 -> 512)
  static constexpr int TMEM_COLS = MODE == 3 ? 512 : GROUPS * BN;
  // slab loads in flight per CTA
  static constexpr int STAGES = kDgrad<MODE> ? 6 : (GROUPS > 2 ? 8 : 10);
  static constexpr int OUT_ROW = BN * 2;               // bytes per output row (128 / 64)
  static constexpr int STAGING = 32 * OUT_ROW;         // per epilogue warp (4 KB / 2 KB)
  // MODE 2: per epilogue group a [128 x 64] bf16 tile (16 KB = the 4 warps' 4 KB areas) plus a
  // pooled [24 positions x 64 ch] bf16 tile (rows of 136 B)
  static constexpr int POOL_TILE = 3328;               // 24 pooled positions x 136 B, padded
  static constexpr int POOL_STAGE = MODE == 2 ? GROUPS * POOL_TILE : 0;
  static constexpr size_t SMEM = W_BYTES + STAGES * SLAB_PAD + EPI_WARPS * 4096 + POOL_STAGE + 1024 + 256;
  static_assert(SMEM <= 227 * 1024, "shared memory budget");
  static_assert(TMEM_COLS <= 512 && (TMEM_COLS & (TMEM_COLS - 1)) == 0, "TMEM columns");
};

template <int MODE>
__global__ void __launch_bounds__(128 + 32 * ConvCfg<MODE>::EPI_WARPS, 1)
    tc_conv_kernel(const __grid_constant__ CUtensorMap map_a,   // 4-D slab map
                   const __grid_constant__ CUtensorMap map_w,   // 3-D weight map (SW64)
                   const __grid_constant__ CUtensorMap map_o,   // 3-D output map, box [BN,32,1]
                   const ConvArgs args) {
  using C = ConvCfg<MODE>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* w_smem = smem;                              // 9 x 4 KB, 1024-aligned
  uint8_t* slabs = smem + C::W_BYTES;
  uint8_t* staging = slabs + C::STAGES * C::SLAB_PAD;   // EPI_WARPS x 4 KB, 1024-aligned
  uint8_t* pool_stage = staging + C::EPI_WARPS * 4096;  // MODE 2: GROUPS x 3 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(pool_stage + C::POOL_STAGE);
  uint64_t* full_bar = bars;                           // [stages]
  uint64_t* empty_bar = bars + C::STAGES;            // [stages]
  uint64_t* acc_full = bars + 2 * C::STAGES;         // [GROUPS]
  uint64_t* acc_empty = acc_full + C::GROUPS;          // [GROUPS]
  uint64_t* w_bar = acc_empty + C::GROUPS;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_bar + 1);
  uint64_t* wg_bar = w_bar + 2;                        // MODE 3: [EPI_WARPS]

  __shared__ float s_bias[64];
  __shared__ int s_wg_valid[kMaxEpiWarps];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // Work distribution: the tiles of ALL clients form one sequence (client-major) that is cut into
  // gridDim.x equal ranges, so every SM has a CTA even when the client count does not divide the
  // SM count (32 clients: 148 CTAs x 73 tiles instead of 128 x 85).  A range that crosses a client
  // boundary is processed as consecutive SEGMENTS; a segment = (client, tile range) runs the
  // whole pipeline below with that client's resident weights, then the barriers are re-initialised.
  const long long t_total = static_cast<long long>(args.n_clients) * args.tiles_per_client;
  const long long cta_lo = t_total * blockIdx.x / gridDim.x;
  const long long cta_hi = t_total * (blockIdx.x + 1) / gridDim.x;

  if (threadIdx.x == 0) {
    prefetch_tmap(&map_a);
    prefetch_tmap(&map_w);
    prefetch_tmap(&map_o);
  }
  if (warp == 2) tmem_alloc<C::TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  int tiles_done = 0;                                  // tiles of earlier segments (RNG counter)
  for (long long cur = cta_lo; cur < cta_hi;) {
  const int g_rel = static_cast<int>(cur / args.tiles_per_client);
  const int g = args.client0 + g_rel;
  const int tile_lo = static_cast<int>(cur - static_cast<long long>(g_rel) * args.tiles_per_client);
  const long long seg_end = min(cta_hi, static_cast<long long>(g_rel + 1) * args.tiles_per_client);
  const int n_tiles = static_cast<int>(seg_end - cur);
  cur = seg_end;
  if (threadIdx.x < 64)
    s_bias[threadIdx.x] = (!kDgrad<MODE> && args.bias)
                              ? args.bias[static_cast<long long>(g) * args.bias_gs + threadIdx.x] : 0.f;
  if (threadIdx.x == 0) {
    for (int s = 0; s < C::STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < C::GROUPS; ++a) { mbar_init(&acc_full[a], 1); mbar_init(&acc_empty[a], 4); }
    mbar_init(w_bar, 1);
    if (MODE == 3)
      for (int w = 0; w < C::EPI_WARPS; ++w) mbar_init(&wg_bar[w], 1);
    fence_barrier_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (warp == 0) {
    // ===================== producer =====================
    if (elect_one() && n_tiles > 0) {
      mbar_expect_tx(w_bar, C::W_BYTES);
      for (int t = 0; t < kTaps; ++t) {
        if (!kDgrad<MODE>)  // K-major box [32 k, 64 co]
          tma_load_3d(w_smem + t * C::W_TILE, &map_w, w_bar, t * 32, 0, g);
        else            // MN-major box [32 ci, 64 co-rows]: same bytes, read as B^T
          tma_load_3d(w_smem + t * C::W_TILE, &map_w, w_bar, t * 32, 0, g);
      }
      for (int i = 0; i < n_tiles; ++i) {
        const int s = i % C::STAGES;
        const uint32_t ph = (i / C::STAGES) & 1;
        if (!mbar_wait(&empty_bar[s], ph ^ 1)) { *args.status = 1; break; }
        if ((args.debug & 1) && i >= C::STAGES) { mbar_arrive(&full_bar[s]); continue; }
        mbar_expect_tx(&full_bar[s], C::SLAB);
        const int tile = tile_lo + i;
        // MODE 2 tiles are 4 grid rows of one image: 6 per image, 104 positions apart
        const int m0 = MODE == 2 ? (tile / 6) * (args.grid_w * args.grid_w) + (tile % 6) * (4 * args.grid_w)
                                 : tile * 128;
        const int row0 = !kDgrad<MODE> ? m0 : m0 - kHalo;
        tma_load_4d(slabs + s * C::SLAB_PAD, &map_a, &full_bar[s], 0, row0, 0, g);
      }
      // drain: every stage's last "slot free" arrival (tcgen05.commit) must have landed before the
      // barriers are re-initialised for the next segment
      for (int i = n_tiles; i < n_tiles + C::STAGES; ++i) {
        if (i < C::STAGES) continue;                     // this stage was never filled
        if (!mbar_wait(&empty_bar[i % C::STAGES], ((i / C::STAGES) & 1) ^ 1)) { *args.status = 1; break; }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (elect_one() && n_tiles > 0) {
      // fwd: A K-major, B K-major ; dgrad: A K-major, B MN-major
      constexpr uint32_t idesc = make_idesc_bf16(128, C::BN, 0u, !kDgrad<MODE> ? 0u : 1u);
      bool ok = mbar_wait(w_bar, 0);
      if (!ok) *args.status = 1;
      const uint32_t w_addr = smem_u32(w_smem);
      for (int i = 0; i < n_tiles && ok; ++i) {
        const int s = i % C::STAGES;
        const int acc = i % C::GROUPS;
        const uint32_t ph = (i / C::STAGES) & 1;
        const uint32_t aph = (i / C::GROUPS) & 1;
        if (!mbar_wait(&acc_empty[acc], aph ^ 1)) { *args.status = 1; ok = false; break; }
        if (!mbar_wait(&full_bar[s], ph)) { *args.status = 1; ok = false; break; }
        tc_fence_after();
        const uint32_t slab = smem_u32(slabs + s * C::SLAB_PAD);
        const uint32_t d_tmem = tmem_base + acc * C::BN;
        const int n_taps = (args.debug & 2) ? 1 : kTaps;
#pragma unroll
        for (int t = 0; t < kTaps; ++t) {
          if (t >= n_taps) break;
#pragma unroll
          for (int kk = 0; kk < C::CH / 16; ++kk) {
            // A: interleaved layout, chunk pitch = kSlabRows*16 B, 8-row group pitch = 128 B
            const uint64_t da = make_smem_desc(
                slab + args.shift[t] * 16 + kk * (2 * kSlabRows * 16), kSlabRows * 16, 128, 0);
            uint64_t db;
            if (!kDgrad<MODE>)   // K-major SW64 tile [64 co][32 k]: step K by 32 B
              db = make_smem_desc(w_addr + t * C::W_TILE + kk * 32, 0, 8 * 64, 4);
            else             // MN-major SW64 tile [64 k-rows][32 ci]: step K by 16 rows
              db = make_smem_desc(w_addr + t * C::W_TILE + kk * (16 * 64), 64 * 64, 8 * 64, 4);
            umma_f16(d_tmem, da, db, idesc, (t > 0 || kk > 0) ? 1u : 0u);
          }
        }
        umma_commit(&empty_bar[s]);
        umma_commit(&acc_full[acc]);
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    // Warps 4..7 take even tiles (accumulator 0), warps 8..11 odd tiles (accumulator 1).
    const int sub = warp & 3;                       // TMEM lane quarter this warp may read
    const int grp = (warp - 4) >> 2;
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(sub * 32) << 16);
    const __nv_bfloat16* mask = args.mask ? args.mask + static_cast<long long>(g) * args.out_gs : nullptr;
    const float* bias = args.bias ? args.bias + static_cast<long long>(g) * args.bias_gs : nullptr;
    uint8_t* my_stage = staging + (warp - 4) * 4096;
    bool ok = true;
    [[maybe_unused]] int my_tiles = 0;              // MODE 3: tiles this warp has fed to its own MMA
    // Software pipeline (dgrad): the ReLU-mask row (and MODE 3's input window) of a tile only
    // depend on the tile index, so they are requested one iteration (two tiles) ahead; fetching
    // them at the top of their own iteration left ~0.6 us of DRAM latency exposed per tile.
    [[maybe_unused]] uint4 mk_nx[C::BN / 8];
    [[maybe_unused]] float xw_nx[9];
    auto fetch_mask = [&](int row, uint4* dst) {
      if (mask && row < args.rows) {
#pragma unroll
        for (int j = 0; j < C::BN / 8; ++j)
          // L1-allocating loads: a thread reads its own 64-byte row in four 16-byte pieces, so each
          // 32-byte sector is touched by two instructions; with L1::no_allocate ncu counted twice
          // the ideal number of L2 sectors for the mask
          dst[j] = __ldg(reinterpret_cast<const uint4*>(mask + static_cast<long long>(row) * C::BN + 8 * j));
      }
    };
    auto fetch_window = [&](int row, float* dst) {
      if (row < args.rows) {
        const int gsz = args.grid_w * args.grid_w, iw = args.grid_w + 2;
        const int b = row / gsz, rem = row - b * gsz;
        const int oh = rem / args.grid_w, ow = rem - oh * args.grid_w;
        const float* px = args.x1 + static_cast<long long>(g) * args.x1_gs +
                          static_cast<long long>(b) * (iw * iw) + oh * iw + ow;
#pragma unroll
        for (int kh = 0; kh < 3; ++kh)
#pragma unroll
          for (int kw = 0; kw < 3; ++kw) dst[kh * 3 + kw] = __ldg(px + kh * iw + kw);
      }
    };
    if (kDgrad<MODE> && grp < n_tiles) {
      const int first = (tile_lo + grp) * 128 + sub * 32 + lane;
      fetch_mask(first, mk_nx);
      if (MODE == 3) fetch_window(first, xw_nx);
    }
    for (int i = grp; i < n_tiles && ok; i += C::GROUPS) {
      const int acc = grp;
      const uint32_t aph = (i / C::GROUPS) & 1;
      const int row0 = (tile_lo + i) * 128 + sub * 32;   // (MODE 0/1 tiling; MODE 2 indexes by image)
      const int m = row0 + lane;
      const bool row_ok = m < args.rows;
      uint4 mk[C::BN / 8];
      // MODE 3: this position's 3x3 window of the layer-1 input (+ a 1 for the bias column) is
      // the B row of the conv1 weight-gradient MMA
      [[maybe_unused]] float xv[16];
      if (kDgrad<MODE>) {
#pragma unroll
        for (int j = 0; j < C::BN / 8; ++j) mk[j] = mk_nx[j];
        if (MODE == 3) {
#pragma unroll
          for (int t = 0; t < 16; ++t) xv[t] = (t < 9 && row_ok) ? xw_nx[t] : 0.f;
          if (row_ok) xv[9] = 1.f;
        }
        if (i + C::GROUPS < n_tiles) {               // this group's next tile
          fetch_mask(m + 128 * C::GROUPS, mk_nx);
          if (MODE == 3) fetch_window(m + 128 * C::GROUPS, xw_nx);
        }
      }
      if (!mbar_wait(&acc_full[acc], aph)) { if (lane == 0) *args.status = 1; ok = false; break; }
      tc_fence_after();
      uint32_t r[C::BN];
#pragma unroll
      for (int c0 = 0; c0 < C::BN; c0 += 16)
        tmem_ld_32x32b_x16(lane_addr + acc * C::BN + c0, r + c0);
      tmem_ld_wait();
      // accumulator is in registers: hand it back to the MMA warp right away
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&acc_empty[acc]);
        bulk_wait_read0();          // previous TMA store has finished reading my staging tile
      }
      __syncwarp();
      if (MODE == 2) {
        // ---- fused bias + ReLU + 2x2 max-pool + dropout --------------------------------------
        // (1) the group's 4 warps stage the post-ReLU [128 x 64] tile (bf16, chunk-swizzled).
        //     No barrier is needed before overwriting it: its last readers (phase 2 of this
        //     group's previous tile) are ordered before that tile's second barrier.
        uint8_t* tile_s = staging + grp * (4 * 4096);
        // pooled tile: [24 positions][64 ch] bf16, rows of 68 elements.  (The first layout,
        // [ch][24] with one two-byte store per element, made every store an 8-way bank conflict:
        // 4.1 M of the kernel's 4.7 M excess shared-memory wavefronts in ncu.)
        __nv_bfloat16* pooled = reinterpret_cast<__nv_bfloat16*>(pool_stage + grp * C::POOL_TILE);
        const int gt = threadIdx.x - 128 - grp * 128;           // 0..127 within the group
        const int trow = sub * 32 + lane;
#pragma unroll
        for (int c0 = 0; c0 < 64; c0 += 8) {
          float v[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[c0 + j]);
          // raw accumulators are staged; bias + ReLU are applied AFTER the max (both are
          // monotonic, so max(relu(x+b)) == relu(max(x)+b)) on a quarter of the elements
          st_vec(tile_s + trow * 128 + (((c0 >> 3) ^ (trow & 7)) * 16), pack<__nv_bfloat16>(v));
        }
        asm volatile("bar.sync %0, 128;" ::"r"(1 + grp) : "memory");
        // (2) pool: 24 pooled positions x 16 channel quads = 384 items, 3 per thread
        const int tile = tile_lo + i;
        const long long img = static_cast<long long>(g) * args.images + tile / 6;
        const int prow0 = (tile % 6) * 2;
        const unsigned long long step = args.rng_state ? *args.rng_state : 0ull;
        const float keep_scale = args.drop_p > 0.f ? 1.f / (1.f - args.drop_p) : 1.f;
        // dropout: ONE Philox draw per thread and tile = 16 random bytes, one per pooled element
        // this thread produces (12 used); keep <=> byte >= round(p*256).  The keep bit is stored
        // in the code byte, so the backward pass never regenerates the stream.
        uint4 rnd = make_uint4(~0u, ~0u, ~0u, ~0u);
        if (args.drop_p > 0.f)
          rnd = Philox::draw((step << 40) + (static_cast<unsigned long long>(args.cta0 + blockIdx.x) << 20) +
                                 static_cast<unsigned long long>(tiles_done + i) * 128 + gt,
                             args.seed);
        const uint32_t thresh = static_cast<uint32_t>(args.drop_p * 256.f + 0.5f);
        const uint32_t rwords[3] = {rnd.x, rnd.y, rnd.z};
#pragma unroll
        for (int rep = 0; rep < 3; ++rep) {
          const int it = gt + rep * 128;
          const int quad = it & 15, pp = it >> 4;             // pp = prow*12 + pcol
          const int prow = pp >= 12 ? 1 : 0, pcol = pp - 12 * prow;   // pp < 24: no division
          const int r0 = (2 * prow) * args.grid_w + 2 * pcol;
          const int rows4[4] = {r0, r0 + 1, r0 + args.grid_w, r0 + args.grid_w + 1};
          // 2x2 window, four channels: max and argmax on packed bf16 pairs (the staged values
          // ARE bf16, so this is exact), only the winners are widened for bias + ReLU + dropout.
          // First maximum wins, as in a scalar `>` scan.
          uint2 raw[4];
#pragma unroll
          for (int w = 0; w < 4; ++w)
            raw[w] = *reinterpret_cast<const uint2*>(
                tile_s + rows4[w] * 128 + (((quad >> 1) ^ (rows4[w] & 7)) * 16) + (quad & 1) * 8);
          const int pos = prow0 * 12 + pp;                     // (prow0 + prow) * 12 + pcol
          uint32_t codes = 0;
          float pv[4];
#pragma unroll
          for (int h = 0; h < 2; ++h) {                        // channel pairs (0,1) and (2,3)
            __nv_bfloat162 c[4];
#pragma unroll
            for (int w = 0; w < 4; ++w) {
              const uint32_t bits = h == 0 ? raw[w].x : raw[w].y;
              c[w] = *reinterpret_cast<const __nv_bfloat162*>(&bits);
            }
            const __nv_bfloat162 mx = __hmax2(__hmax2(c[0], c[1]), __hmax2(c[2], c[3]));
            const uint32_t e0 = __heq2_mask(c[0], mx), e1 = __heq2_mask(c[1], mx),
                           e2 = __heq2_mask(c[2], mx);
            // per 16-bit half: 0 if c0 is the max, else 1 if c1, else 2 if c2, else 3
            const uint32_t arg2 = ~e0 & ((e1 & 0x00010001u) |
                                         (~e1 & ((e2 & 0x00020002u) | (~e2 & 0x00030003u))));
            const uint32_t mbits = *reinterpret_cast<const uint32_t*>(&mx);
            const float mf[2] = {__uint_as_float(mbits << 16), __uint_as_float(mbits & 0xffff0000u)};
#pragma unroll
            for (int k = 0; k < 2; ++k) {
              const int j = 2 * h + k;
              const float m = fmaxf(mf[k] + s_bias[quad * 4 + j], 0.f);
              const uint32_t a = (arg2 >> (16 * k)) & 3u;
              const bool keep = ((rwords[rep] >> (8 * j)) & 0xffu) >= thresh;
              codes |= (a | (m > 0.f ? 4u : 0u) | (keep ? 8u : 0u)) << (8 * j);
              pv[j] = keep ? m * keep_scale : 0.f;
            }
          }
          {
            const __nv_bfloat162 lo = __floats2bfloat162_rn(pv[0], pv[1]);
            const __nv_bfloat162 hi = __floats2bfloat162_rn(pv[2], pv[3]);
            uint2 pk;
            pk.x = *reinterpret_cast<const uint32_t*>(&lo);
            pk.y = *reinterpret_cast<const uint32_t*>(&hi);
            asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(smem_u32(pooled + pp * 68 + quad * 4)),
                         "r"(pk.x), "r"(pk.y) : "memory");
          }
          *reinterpret_cast<uint32_t*>(args.pool_idx + (img * 144 + pos) * 64 + quad * 4) = codes;
        }
        asm volatile("bar.sync %0, 128;" ::"r"(1 + grp) : "memory");
        // (3) write the pooled tile: per channel 24 consecutive bf16 (48 B) in flatten order
        {
          const int ch = gt >> 1, half = gt & 1;               // 64 channels x 2 halves of 12
          const uint32_t src = smem_u32(pooled + half * 12 * 68 + ch);
          uint32_t w[6];
#pragma unroll
          for (int k = 0; k < 6; ++k) {                        // two positions per 32-bit word
            uint16_t a, b;
            asm volatile("ld.shared.u16 %0, [%1];" : "=h"(a) : "r"(src + (2 * k) * 136));
            asm volatile("ld.shared.u16 %0, [%1];" : "=h"(b) : "r"(src + (2 * k + 1) * 136));
            w[k] = static_cast<uint32_t>(a) | (static_cast<uint32_t>(b) << 16);
          }
          uint2* dst = reinterpret_cast<uint2*>(args.pool_out + img * 9216 + ch * 144 + prow0 * 12 + half * 12);
          dst[0] = make_uint2(w[0], w[1]); dst[1] = make_uint2(w[2], w[3]); dst[2] = make_uint2(w[4], w[5]);
        }
        // `pooled` is rewritten only after the next tile's first barrier, i.e. after every
        // thread of the group has finished the reads above.
        continue;
      }
      if (!(args.debug & 4)) {
        if (MODE == 3 && my_tiles > 0) {
          // the tensor core must have read this warp's previous operand tiles
          if (!mbar_wait(&wg_bar[warp - 4], (my_tiles - 1) & 1)) {
            if (lane == 0) *args.status = 1;
            ok = false;
            break;
          }
        }
        // stage this warp's 32 rows in shared memory with the TMA swizzle (conflict-free
        // 16-byte stores), then ONE bulk tensor store writes them as full lines
#pragma unroll
        for (int c0 = 0; c0 < C::BN; c0 += 8) {
          float v[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[c0 + j]);
          if (MODE == 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j] + s_bias[c0 + j], 0.f);
          } else if (mask && row_ok) {
            float mf[8];
            unpack<__nv_bfloat16>(mk[c0 / 8], mf);
#pragma unroll
            for (int j = 0; j < 8; ++j)
              if (!(mf[j] > 0.f)) v[j] = 0.f;
          }
          if (MODE == 3 && !row_ok) {
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = 0.f;     // rows past R must not reach the MMA
          }
          const int chunk = c0 / 8;
          const int sw = MODE == 0 ? (chunk ^ (lane & 7)) : (chunk ^ ((lane >> 1) & 3));
          st_vec(my_stage + lane * C::OUT_ROW + sw * 16, pack<__nv_bfloat16>(v));
        }
        if (MODE == 3) {
          // ---- conv1 weight/bias gradient on the tensor core ---------------------------------
          //   dW1[ch, t] += sum_q da1[q, ch] * x[q + s_t],   t = 9 is the bias column (x = 1)
          // The staged rows ARE an MN-major 64-byte-swizzled operand (K = 32 positions, M = 32
          // channels); the windows are written next to them in the same layout (N = 32, columns
          // 0..9 used).  UMMA M is 64 (half the operand fetch of M = 128): the second 32-channel
          // atom aliases bytes inside this warp's 4 KB area (LBO 512) and only produces accumulator
          // rows 32..63, never read.  M = 64 puts row m in TMEM lane (m % 16) + 32 * (m / 16).
          uint8_t* xs = my_stage + 2048;
          st_vec(xs + lane * 64 + ((0 ^ ((lane >> 1) & 3)) * 16), pack<__nv_bfloat16>(xv));
          st_vec(xs + lane * 64 + ((1 ^ ((lane >> 1) & 3)) * 16), pack<__nv_bfloat16>(xv + 8));
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            tc_fence_after();
            constexpr uint32_t idesc_w = make_idesc_bf16(64, 32, 1u, 1u);
            const uint32_t a0 = smem_u32(my_stage), b0 = smem_u32(xs);
            const uint32_t d_w = tmem_base + 2 * C::BN + (warp - 4) * 32;
#pragma unroll
            for (int kk = 0; kk < 2; ++kk) {           // 16 positions per MMA
              const uint64_t da = make_smem_desc(a0 + kk * 1024, 512, 8 * 64, 4);
              const uint64_t db = make_smem_desc(b0 + kk * 1024, 512, 8 * 64, 4);
              umma_f16(d_w, da, db, idesc_w, (my_tiles > 0 || kk > 0) ? 1u : 0u);
            }
            umma_commit(&wg_bar[warp - 4]);
          }
          __syncwarp();
          ++my_tiles;
          continue;                                     // da1 itself is never written
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          tma_store_3d(&map_o, my_stage, 0, row0, g);   // rows >= R are clipped by the map
          bulk_commit();
        }
      }
    }
    if (lane == 0) bulk_wait_read0();
    if (MODE == 3) {
      // every warp's accumulator is complete; in each group the warps of TMEM quarters 0 and 1 add
      // up the group's four accumulators (16 channels each) and flush 32 x 10 values with atomics
      if (my_tiles > 0 && ok && !mbar_wait(&wg_bar[warp - 4], (my_tiles - 1) & 1)) {
        if (lane == 0) *args.status = 1;
        ok = false;
      }
      if (lane == 0) s_wg_valid[warp - 4] = (ok && my_tiles > 0) ? 1 : 0;
      tc_fence_before();
      asm volatile("bar.sync 1, %0;" ::"n"(32 * C::EPI_WARPS) : "memory");
      tc_fence_after();
      if (sub < 2) {                                   // channels 16*sub + lane, lanes 0..15
        float sum[10];
#pragma unroll
        for (int t = 0; t < 10; ++t) sum[t] = 0.f;
        for (int w = grp * 4; w < grp * 4 + 4; ++w) {
          if (!s_wg_valid[w]) continue;                 // warp-uniform
          uint32_t rr[16];
          tmem_ld_32x32b_x16(lane_addr + 2 * C::BN + w * 32, rr);
          tmem_ld_wait();
#pragma unroll
          for (int t = 0; t < 10; ++t) sum[t] += __uint_as_float(rr[t]);
        }
        float* gd = args.grad1 + static_cast<long long>(g) * args.grad1_gs;
        if (lane < 16) {
          const int ch = sub * 16 + lane;
#pragma unroll
          for (int t = 0; t < 9; ++t) atomicAdd(gd + args.w1_off + ch * 9 + t, sum[t]);
          atomicAdd(gd + args.b1_off + ch, sum[9]);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  tiles_done += n_tiles;
  }  // segments

  if (warp == 2) tmem_dealloc<C::TMEM_COLS>(tmem_base);
}

// -------------------------------------------------------------------------------------
// Weight gradient, persistent form: one CTA per (client, contiguous range of 64-row
// k-blocks).  For kernel row kh:   D_kh[co, (kw,ci)] += sum_q dy[q, co] * x[q + kh*W + kw, ci]
//   A (M = 64 output channels): dy, MN-major, 128-byte swizzle, one 64-wide atom.
//   B (N = 96 = kw x 32 input channels): x, MN-major, 64-byte swizzle, ONE tile of 120 rows per
//     k-block.  The three kw taps are three 32-channel N atoms of the same tile whose start
//     addresses differ by one row (descriptor LBO = 64 B); the kernel row is a start-address offset
//     of kh*W rows.  (Reading a swizzled tile at whole-row offsets is legal because the hardware
//     swizzle is a function of the absolute shared-memory address — verified on B200: a
//     128-B-swizzled slab read at start + s*128 B gives exact results with base_offset = 0.)
//   UMMA M = 64 puts accumulator row m in TMEM lane (m % 16) + 32 * (m / 16) (verified).
// 3 accumulators x 96 TMEM columns live for the CTA's whole k-range; 12 MMAs of 64x96x16 per
// k-block.  History (B200, 32 clients x 64 images): generic kernel 256 us -> x as 16-byte-piece
// interleaved slabs 166 us (TMA piece rate ~1.4/ns was the limit) -> x loaded three times
// (kw-shifted) as the M operand, M = 96 -> 128 with a zeroed pad atom: 91 us -> one x tile, atoms
// one row apart: 91 us (loads were not the limit) -> roles swapped (M = 64): 87 us.
// Probes (NFB_CONV_DEBUG): without loads 87 us, with 1 of 12 MMAs 57 us: the kernel is paced by the
// MMAs at ~85 cycles each, for both operand arrangements — about (A bytes + B bytes) / 64 B per
// cycle, i.e. the shared-memory operand fetch of small-N MMAs, not the tensor pipe (48 cycles).
// Epilogue: fp32 atomics into the OIHW gradient inside the flat slab.
// -------------------------------------------------------------------------------------
constexpr int kWgRows = 120;                         // 64 k-block rows + 52 (two kernel rows) + 3, to 8
constexpr int kWgStages = 8;
constexpr int kWgB = 64 * 128;                       // dy tile [64 rows x 128 B], first in a stage
constexpr int kWgX = kWgRows * 64;                   // ONE x tile [120 rows x 64 B]
constexpr int kWgStage = kWgB + 8192;                // 16 KB (x tile padded to keep 1 KB alignment)
constexpr size_t kWgSmem = static_cast<size_t>(kWgStages) * kWgStage + 1024 + 256;

struct WgradArgs {
  int kblocks_per_client;  // ceil(R / 64)
  int n_clients;           // clients covered by this launch
  int client0;             // first client of this launch
  int shift[kTaps];        // [kh] = kh * width (entries 0..2 used)
  float* grad;             // first client's [64][32][3][3] block inside the gradient slab
  long long grad_gs;
  int* status;
  int debug;               // probes (env NFB_CONV_DEBUG): 1 = reuse first stages, 2 = 1/12 MMAs
  unsigned int x_off[12];  // descriptor address field of x for (kh, kk), 16-byte units (host)
};

__global__ void __launch_bounds__(256, 1)
    tc_conv_wgrad_kernel(const __grid_constant__ CUtensorMap map_dy,  // 3-D, box [64, 64, 1], SW128
                         const __grid_constant__ CUtensorMap map_x,   // 3-D, box [32, 116, 1], SW64
                         const WgradArgs args) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + kWgStages * kWgStage);
  uint64_t* empty_bar = full_bar + kWgStages;
  uint64_t* acc_bar = empty_bar + kWgStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // one sequence of k-blocks over all clients, cut into gridDim.x equal ranges; a range crossing a
  // client boundary runs as consecutive segments (see tc_conv_kernel)
  const long long t_total = static_cast<long long>(args.n_clients) * args.kblocks_per_client;
  const long long cta_lo = t_total * blockIdx.x / gridDim.x;
  const long long cta_hi = t_total * (blockIdx.x + 1) / gridDim.x;

  if (threadIdx.x == 0) {
    prefetch_tmap(&map_dy);
    prefetch_tmap(&map_x);
  }
  if (warp == 2) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  for (long long cur = cta_lo; cur < cta_hi;) {
  const int g_rel = static_cast<int>(cur / args.kblocks_per_client);
  const int g = args.client0 + g_rel;
  const int kb_lo = static_cast<int>(cur - static_cast<long long>(g_rel) * args.kblocks_per_client);
  const long long seg_end = min(cta_hi, static_cast<long long>(g_rel + 1) * args.kblocks_per_client);
  const int n_kb = static_cast<int>(seg_end - cur);
  cur = seg_end;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kWgStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(acc_bar, 1);
    fence_barrier_init();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (warp == 0) {
    if (elect_one()) {
      for (int i = 0; i < n_kb; ++i) {
        const int s = i % kWgStages;
        const uint32_t ph = (i / kWgStages) & 1;
        if (!mbar_wait(&empty_bar[s], ph ^ 1)) { *args.status = 1; break; }
        uint8_t* stage = smem + s * kWgStage;
        if ((args.debug & 1) && i >= kWgStages) { mbar_arrive(&full_bar[s]); continue; }
        const int q0 = (kb_lo + i) * 64;
        mbar_expect_tx(&full_bar[s], kWgB + kWgX);
        tma_load_3d(stage, &map_dy, &full_bar[s], 0, q0, g);                 // dy [64 co x 64 rows]
        tma_load_3d(stage + kWgB, &map_x, &full_bar[s], 0, q0, g);           // x  [32 ci x 120 rows]
      }
      for (int i = n_kb; i < n_kb + kWgStages; ++i) {   // drain the "slot free" arrivals
        if (i < kWgStages) continue;
        if (!mbar_wait(&empty_bar[i % kWgStages], ((i / kWgStages) & 1) ^ 1)) { *args.status = 1; break; }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc_bf16(64, 96, 1u, 1u);
      // Descriptor prototypes with a zero address field; per MMA only the address is added (the
      // issuing thread is the critical path: 12 MMAs per 64 positions).
      //   A (M = 64 output channels): dy, MN-major SW128, one 64-wide atom, 16 rows per K step.
      //   B (N = 96 = kw x 32 input channels): x, MN-major SW64, 8-k-row groups 512 B apart.  The
      //     three 32-channel N atoms are the SAME tile one row (64 B) apart (LBO = 64): atom kw
      //     reads x shifted by kw positions, so one copy of x serves kw = 0..2.  Kernel row kh and
      //     the K step are row offsets as well (x_off, computed on the host).
      const uint64_t a_proto = make_smem_desc(0, 64 * 128, 8 * 128, 2);
      const uint64_t b_proto = make_smem_desc(0, 64, 8 * 64, 4);
      bool ok = true;
      for (int i = 0; i < n_kb && ok; ++i) {
        const int s = i % kWgStages;
        const uint32_t ph = (i / kWgStages) & 1;
        if (!mbar_wait(&full_bar[s], ph)) { *args.status = 1; ok = false; break; }
        tc_fence_after();
        const uint32_t st_lo = (smem_u32(smem + s * kWgStage) & 0x3FFFFu) >> 4;
        if (args.debug & 2) {                          // probe: 1 of the 12 MMAs
          umma_f16(tmem_base, a_proto | st_lo, b_proto | (st_lo + (kWgB >> 4) + args.x_off[0]),
                   idesc, i > 0 ? 1u : 0u);
        } else {
#pragma unroll
          for (int kh = 0; kh < 3; ++kh) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              umma_f16(tmem_base + kh * 96, a_proto | (st_lo + kk * (16 * 128 / 16)),
                       b_proto | (st_lo + (kWgB >> 4) + args.x_off[kh * 4 + kk]), idesc,
                       (i > 0 || kk > 0) ? 1u : 0u);
          }
        }
        umma_commit(&empty_bar[s]);
      }
      umma_commit(acc_bar);
    }
  } else if (warp >= 4) {
    const int sub = warp - 4;
    bool ok = true;
    if (n_kb > 0) ok = mbar_wait(acc_bar, 0);
    if (!ok && lane == 0) *args.status = 1;
    tc_fence_after();
    // UMMA M = 64 puts accumulator row m in TMEM lane (m % 16) + 32 * (m / 16): every warp owns
    // 16 output channels in its lanes 0..15 (lanes 16..31 of each quarter are unused)
    const int co = sub * 16 + (lane & 15);
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(sub * 32) << 16);
    float* gout = args.grad + static_cast<long long>(g) * args.grad_gs + co * 288;
#pragma unroll 1
    for (int kh = 0; kh < 3; ++kh) {
#pragma unroll 1
      for (int c0 = 0; c0 < 96; c0 += 16) {          // column n = kw * 32 + ci
        uint32_t r[16];
        tmem_ld_32x32b_x16(lane_addr + kh * 96 + c0, r);
        tmem_ld_wait();
        if (ok && n_kb > 0 && lane < 16) {
          const int kw = c0 >> 5, ci0 = c0 & 31;
#pragma unroll
          for (int j = 0; j < 16; ++j)
            atomicAdd(gout + (ci0 + j) * 9 + kh * 3 + kw, __uint_as_float(r[j]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  }  // segments
  if (warp == 2) tmem_dealloc<512>(tmem_base);
}

}  // namespace tc
}  // namespace nfb

// =====================================================================================
// Host side
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

static EncodeFn conv_encode_fn() {
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

using ConvKey = std::tuple<const void*, long long, long long, long long, int, int>;
static std::map<ConvKey, CUtensorMap> g_conv_maps;
static std::mutex g_conv_mu;

// 4-D slab map over a bf16 [C, rows, ch] tensor: (8 elems, rows, ch/8 chunks, clients)
static int slab_map(const void* ptr, long long rows, long long groups, long long group_stride,
                    int ch, int box_rows, CUtensorMap* out) {
  ConvKey key{ptr, rows, groups, group_stride, ch, box_rows};
  std::lock_guard<std::mutex> lock(g_conv_mu);
  auto it = g_conv_maps.find(key);
  if (it != g_conv_maps.end()) { *out = it->second; return 0; }
  EncodeFn enc = conv_encode_fn();
  if (!enc) return fail(NFB_ESTATE, "tc_conv: cuTensorMapEncodeTiled unavailable");
  cuuint64_t dims[4] = {8, static_cast<cuuint64_t>(rows), static_cast<cuuint64_t>(ch / 8),
                        static_cast<cuuint64_t>(groups)};
  cuuint64_t strides[3] = {static_cast<cuuint64_t>(ch) * 2, 16,
                           static_cast<cuuint64_t>(group_stride) * 2};
  cuuint32_t box[4] = {8, static_cast<cuuint32_t>(box_rows), static_cast<cuuint32_t>(ch / 8), 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUtensorMap m;
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(NFB_EINVAL, "tc_conv: slab tensor map rejected (CUresult %d)", static_cast<int>(r));
  if (g_conv_maps.size() > 1024) g_conv_maps.clear();
  g_conv_maps[key] = m;
  *out = m;
  return 0;
}

static int weight_map(const void* ptr, long long groups, long long group_stride, CUtensorMap* out) {
  ConvKey key{ptr, 64, groups, group_stride, -1, 0};
  std::lock_guard<std::mutex> lock(g_conv_mu);
  auto it = g_conv_maps.find(key);
  if (it != g_conv_maps.end()) { *out = it->second; return 0; }
  EncodeFn enc = conv_encode_fn();
  if (!enc) return fail(NFB_ESTATE, "tc_conv: cuTensorMapEncodeTiled unavailable");
  cuuint64_t dims[3] = {288, 64, static_cast<cuuint64_t>(groups)};
  cuuint64_t strides[2] = {288 * 2, static_cast<cuuint64_t>(group_stride) * 2};
  cuuint32_t box[3] = {32, 64, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUtensorMap m;
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(NFB_EINVAL, "tc_conv: weight tensor map rejected (CUresult %d)", static_cast<int>(r));
  g_conv_maps[key] = m;
  *out = m;
  return 0;
}

// 3-D output map over bf16 [C, rows, bn]: box [bn, 32, 1], swizzle = row bytes (128 / 64)
static int out_map(const void* ptr, long long rows, long long groups, long long group_stride,
                   int bn, CUtensorMap* out) {
  ConvKey key{ptr, rows, groups, group_stride, -3, bn};
  std::lock_guard<std::mutex> lock(g_conv_mu);
  auto it = g_conv_maps.find(key);
  if (it != g_conv_maps.end()) { *out = it->second; return 0; }
  EncodeFn enc = conv_encode_fn();
  if (!enc) return fail(NFB_ESTATE, "tc_conv: cuTensorMapEncodeTiled unavailable");
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(bn), static_cast<cuuint64_t>(rows),
                        static_cast<cuuint64_t>(groups)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(bn) * 2,
                           static_cast<cuuint64_t>(group_stride) * 2};
  cuuint32_t box[3] = {static_cast<cuuint32_t>(bn), 32, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUtensorMap m;
  CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides,
                   box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   bn == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                   CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return fail(NFB_EINVAL, "tc_conv: output tensor map rejected (CUresult %d)", static_cast<int>(r));
  g_conv_maps[key] = m;
  *out = m;
  return 0;
}

// Grid of the persistent conv kernels: one CTA per SM (or per work unit if there are fewer).
// History: one CTA range per client left 20 of 148 SMs idle with 32 clients; covering the cohort
// with 2 / 4 back-to-back launches of 16 / 8 clients was SLOWER (0.547 / 0.641 vs 0.527 ms per
// local step: a persistent kernel's drain + the next prologue cost more than the idle SMs);
// ranges that cross client boundaries inside ONE launch (segments) are what is used now.
static int conv_grid(int clients, int units_per_client) {
  const long long total = static_cast<long long>(clients) * units_per_client;
  const long long sms = sm_count();
  return static_cast<int>(total < sms ? total : sms);
}

template <int MODE>
static int launch_conv(const CUtensorMap& ma, const CUtensorMap& mw, const CUtensorMap& mo,
                       int ctas, const ConvArgs& a_in, cudaStream_t st) {
  using C = ConvCfg<MODE>;
  auto kern = tc_conv_kernel<MODE>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(C::SMEM));
    if (e != cudaSuccess)
      return fail(static_cast<int>(e), "tc_conv: smem attribute: %s", cudaGetErrorString(e));
    attr_set = true;
  }
  (void)ctas;
  ConvArgs a = a_in;
  a.client0 = 0;
  a.cta0 = 0;
  kern<<<conv_grid(a.n_clients, a.tiles_per_client), 128 + 32 * C::EPI_WARPS, C::SMEM, st>>>(ma, mw, mo, a);
  return launch_status("tc_conv");
}

}  // namespace tc
}  // namespace nfb

// mode 0: forward (x [C,R,32] -> out [C,R,64], bias+ReLU); mode 1: dgrad (dy [C,R,64] -> out
// [C,R,32], optional ReLU mask).  w = bf16 [C][64][288] (tap-major, channel-minor).
extern "C" int nfb_tc_conv3x3(int mode, const void* a, long long rows, long long a_group_stride,
                              const void* w, long long w_group_stride, int clients, int width,
                              void* out, long long out_group_stride, const float* bias,
                              long long bias_group_stride, const void* mask, int* status,
                              void* stream) {
  using namespace nfb;
  using namespace nfb::tc;
  NFB_REQUIRE(a && w && out && status, NFB_EINVAL, "tc_conv3x3: null pointer");
  NFB_REQUIRE(mode == 0 || mode == 1, NFB_EINVAL, "tc_conv3x3: mode must be 0 or 1");
  NFB_REQUIRE(clients >= 1 && rows >= 1, NFB_EINVAL, "tc_conv3x3: bad sizes");
  NFB_REQUIRE(2 * width + 2 <= kHalo, NFB_ERANGE,
              "tc_conv3x3: grid width %d needs a halo of %d rows, kernel has %d", width,
              2 * width + 2, kHalo);
  const int ch = mode == 0 ? 32 : 64;
  CUtensorMap ma, mw;
  int rc = slab_map(a, rows, clients, a_group_stride, ch, kSlabRows, &ma);
  if (rc) return rc;
  rc = weight_map(w, clients, w_group_stride, &mw);
  if (rc) return rc;
  CUtensorMap mo;
  rc = out_map(out, rows, clients, out_group_stride, mode == 0 ? 64 : 32, &mo);
  if (rc) return rc;
  ConvArgs args{};
  args.tiles_per_client = static_cast<int>((rows + 127) / 128);
  args.n_clients = clients;
  const int per_client = 1;   // (grid is chosen in launch_conv)
  args.rows = static_cast<int>(rows);
  for (int kh = 0; kh < 3; ++kh)
    for (int kw = 0; kw < 3; ++kw) {
      const int s = kh * width + kw;
      args.shift[kh * 3 + kw] = mode == 0 ? s : kHalo - s;
    }
  args.out = out;
  args.out_gs = out_group_stride;
  args.bias = bias;
  args.bias_gs = bias_group_stride;
  args.mask = static_cast<const __nv_bfloat16*>(mask);
  args.status = status;
  {
    const char* dbg = getenv("NFB_CONV_DEBUG");
    args.debug = dbg ? atoi(dbg) : 0;
  }
  const int ctas = clients * per_client;
  cudaStream_t st = as_stream(stream);
  return mode == 0 ? launch_conv<0>(ma, mw, mo, ctas, args, st)
                   : launch_conv<1>(ma, mw, mo, ctas, args, st);
}

// Data gradient of the 32->64 3x3 convolution fused with the weight/bias gradient of the
// 1->32 3x3 convolution in front of it (mode 3 of the persistent kernel):
//   da1[q, ci]  = (sum_t sum_co dy[q - s_t, co] w[co, t*32+ci]) * (mask[q, ci] > 0)      (not stored)
//   grad1[c][w1_off + ci*9 + t] += sum_q da1[q, ci] * x1[c, image(q), pixel(q) + tap t]
//   grad1[c][b1_off + ci]       += sum_q da1[q, ci]
// x1 = fp32 [C][B][(width+2)^2] layer-1 inputs, B = rows / width^2.  The caller zeroes grad1's
// target entries; accumulation is fp32 (tensor core, then 320 atomics per CTA).
extern "C" int nfb_tc_conv3x3_dgrad_wgrad1(const void* dy, long long rows,
                                           long long dy_group_stride, const void* w,
                                           long long w_group_stride, int clients, int width,
                                           const void* mask, long long mask_group_stride,
                                           const float* x1, long long x1_group_stride,
                                           float* grad1, long long grad1_group_stride,
                                           long long w1_off, long long b1_off, int* status,
                                           void* stream) {
  using namespace nfb;
  using namespace nfb::tc;
  NFB_REQUIRE(dy && w && mask && x1 && grad1 && status, NFB_EINVAL,
              "tc_conv3x3_dgrad_wgrad1: null pointer");
  NFB_REQUIRE(clients >= 1 && rows >= 1 && width >= 1 && rows % (width * width) == 0, NFB_EINVAL,
              "tc_conv3x3_dgrad_wgrad1: rows must be whole width x width images");
  NFB_REQUIRE(2 * width + 2 <= kHalo, NFB_ERANGE,
              "tc_conv3x3_dgrad_wgrad1: grid width %d needs a halo of %d rows, kernel has %d",
              width, 2 * width + 2, kHalo);
  CUtensorMap ma, mw;
  int rc = slab_map(dy, rows, clients, dy_group_stride, 64, kSlabRows, &ma);
  if (rc) return rc;
  rc = weight_map(w, clients, w_group_stride, &mw);
  if (rc) return rc;
  ConvArgs args{};
  args.tiles_per_client = static_cast<int>((rows + 127) / 128);
  args.n_clients = clients;
  const int per_client = 1;   // (grid is chosen in launch_conv)
  args.rows = static_cast<int>(rows);
  for (int kh = 0; kh < 3; ++kh)
    for (int kw = 0; kw < 3; ++kw) args.shift[kh * 3 + kw] = kHalo - (kh * width + kw);
  args.out = nullptr;
  args.out_gs = mask_group_stride;       // the kernel addresses the mask with the output strides
  args.mask = static_cast<const __nv_bfloat16*>(mask);
  args.grid_w = width;
  args.x1 = x1;
  args.x1_gs = x1_group_stride;
  args.grad1 = grad1;
  args.grad1_gs = grad1_group_stride;
  args.w1_off = w1_off;
  args.b1_off = b1_off;
  args.status = status;
  {
    const char* dbg = getenv("NFB_CONV_DEBUG");
    args.debug = dbg ? atoi(dbg) : 0;
  }
  return launch_conv<3>(ma, mw, ma /*no output map*/, clients * per_client, args,
                        as_stream(stream));
}

// Weight gradient: grad[c][co][ci][t] += sum_q dy[c,q,co] * x[c,q+s_t,ci]  (fp32 atomics; the
// caller zeroes the target).  dy = bf16 [C,R,64] (zero outside the valid window), x = bf16 [C,R,32].
extern "C" int nfb_tc_conv3x3_wgrad(const void* dy, const void* x, long long rows,
                                    long long dy_group_stride, long long x_group_stride,
                                    int clients, int width, float* grad, long long grad_group_stride,
                                    int* status, void* stream) {
  using namespace nfb;
  using namespace nfb::tc;
  NFB_REQUIRE(dy && x && grad && status, NFB_EINVAL, "tc_conv3x3_wgrad: null pointer");
  NFB_REQUIRE(clients >= 1 && rows >= 1, NFB_EINVAL, "tc_conv3x3_wgrad: bad sizes");
  NFB_REQUIRE(2 * width <= 52, NFB_ERANGE, "tc_conv3x3_wgrad: grid width %d too large", width);
  CUtensorMap mdy, mx;
  {
    // dy as MN-major operand: 3-D map (64 co, rows, clients), box [64, 64, 1], 128-B swizzle
    ConvKey key{dy, rows, clients, dy_group_stride, -2, 0};
    std::lock_guard<std::mutex> lock(g_conv_mu);
    auto it = g_conv_maps.find(key);
    if (it != g_conv_maps.end()) {
      mdy = it->second;
    } else {
      EncodeFn enc = conv_encode_fn();
      if (!enc) return fail(NFB_ESTATE, "tc_conv3x3_wgrad: cuTensorMapEncodeTiled unavailable");
      cuuint64_t dims[3] = {64, static_cast<cuuint64_t>(rows), static_cast<cuuint64_t>(clients)};
      cuuint64_t strides[2] = {128, static_cast<cuuint64_t>(dy_group_stride) * 2};
      cuuint32_t box[3] = {64, 64, 1};
      cuuint32_t estr[3] = {1, 1, 1};
      CUresult r = enc(&mdy, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(dy), dims,
                       strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS)
        return fail(NFB_EINVAL, "tc_conv3x3_wgrad: dy tensor map rejected (%d)", static_cast<int>(r));
      g_conv_maps[key] = mdy;
    }
  }
  {
    ConvKey key{x, rows, clients, x_group_stride, -5, kWgRows};
    std::lock_guard<std::mutex> lock(g_conv_mu);
    auto it = g_conv_maps.find(key);
    if (it != g_conv_maps.end()) {
      mx = it->second;
    } else {
      EncodeFn enc = conv_encode_fn();
      if (!enc) return fail(NFB_ESTATE, "tc_conv3x3_wgrad: cuTensorMapEncodeTiled unavailable");
      cuuint64_t dims[3] = {32, static_cast<cuuint64_t>(rows), static_cast<cuuint64_t>(clients)};
      cuuint64_t strides[2] = {64, static_cast<cuuint64_t>(x_group_stride) * 2};
      cuuint32_t box[3] = {32, static_cast<cuuint32_t>(kWgRows), 1};
      cuuint32_t estr[3] = {1, 1, 1};
      CUresult r = enc(&mx, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(x), dims,
                       strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                       CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
      if (r != CUDA_SUCCESS)
        return fail(NFB_EINVAL, "tc_conv3x3_wgrad: x tensor map rejected (%d)", static_cast<int>(r));
      g_conv_maps[key] = mx;
    }
  }
  WgradArgs a{};
  a.kblocks_per_client = static_cast<int>((rows + 63) / 64);
  a.n_clients = clients;
  a.client0 = 0;
  for (int kh = 0; kh < 3; ++kh) a.shift[kh] = kh * width;
  for (int kh = 0; kh < 3; ++kh)
    for (int kk = 0; kk < 4; ++kk)       // rows of 64 B = 4 descriptor units each
      a.x_off[kh * 4 + kk] = static_cast<unsigned>((kh * width + kk * 16) * 4);
  a.grad = grad;
  a.grad_gs = grad_group_stride;
  a.status = status;
  {
    const char* dbg = getenv("NFB_CONV_DEBUG");
    a.debug = dbg ? atoi(dbg) : 0;
  }
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(tc_conv_wgrad_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(kWgSmem));
    if (e != cudaSuccess)
      return fail(static_cast<int>(e), "tc_conv3x3_wgrad: smem attribute: %s", cudaGetErrorString(e));
    attr_set = true;
  }
  tc_conv_wgrad_kernel<<<conv_grid(clients, a.kblocks_per_client), 256, kWgSmem, as_stream(stream)>>>(
      mdy, mx, a);
  return launch_status("tc_conv3x3_wgrad");
}

// Forward 3x3 conv (32 -> 64 ch) + bias + ReLU + 2x2 max-pool + dropout in ONE kernel; the conv
// output is never written.  x = bf16 [C, B*26*26, 32] on the 26x26 grid; pool_out = bf16
// [C, B, 9216] in PyTorch flatten order (ch*144 + ph*12 + pw); pool_idx = u8 [C, B, 144, 64]
// (bits 0-1 argmax, bit 2 max > 0, bit 3 dropout keep) for the backward pass.
extern "C" int nfb_tc_conv3x3_pool_fwd(const void* x, long long rows, long long x_group_stride,
                                       const void* w, long long w_group_stride, int clients,
                                       int images, int width, const float* bias,
                                       long long bias_group_stride, void* pool_out,
                                       uint8_t* pool_idx, float drop_p, const void* rng_state,
                                       unsigned long long seed, int* status, void* stream) {
  using namespace nfb;
  using namespace nfb::tc;
  NFB_REQUIRE(x && w && pool_out && pool_idx && status, NFB_EINVAL,
              "tc_conv3x3_pool_fwd: null pointer");
  NFB_REQUIRE(width == 26 && rows == static_cast<long long>(images) * 26 * 26, NFB_ERANGE,
              "tc_conv3x3_pool_fwd: only the 26x26 grid (24x24 conv output, 12x12 pooled) is "
              "implemented (got width %d, rows %lld for %d images)", width, rows, images);
  CUtensorMap ma, mw;
  int rc = slab_map(x, rows, clients, x_group_stride, 32, kSlabRows, &ma);
  if (rc) return rc;
  rc = weight_map(w, clients, w_group_stride, &mw);
  if (rc) return rc;
  ConvArgs args{};
  args.tiles_per_client = images * 6;
  args.n_clients = clients;
  const int per_client = 1;   // (grid is chosen in launch_conv)
  args.rows = static_cast<int>(rows);
  for (int kh = 0; kh < 3; ++kh)
    for (int kw = 0; kw < 3; ++kw) args.shift[kh * 3 + kw] = kh * width + kw;
  args.bias = bias;
  args.bias_gs = bias_group_stride;
  args.pool_out = static_cast<__nv_bfloat16*>(pool_out);
  args.pool_idx = pool_idx;
  args.images = images;
  args.grid_w = width;
  args.drop_p = drop_p;
  args.rng_state = static_cast<const unsigned long long*>(rng_state);
  args.seed = seed;
  args.status = status;
  return launch_conv<2>(ma, mw, ma /*unused output map*/, clients * per_client, args,
                        as_stream(stream));
}
