// K-GEMM, persistent form for the wide fully-connected layer's data gradient:
//
//     out[g][b, i] = sum_j dy[g][b, j] * w[g][j, i]        (b <= 128, j = 128, i = thousands)
//
// tc_gemm.cu runs this as one CTA per 128-column tile: 2304 CTAs for 32 clients, each a latency
// chain (barrier/TMEM setup -> TMA -> 8 MMAs -> epilogue -> store) that ncu shows at 7 % tensor
// and 37 % issue utilisation, 65 us for 112 MB of traffic.  Here one CTA owns a contiguous range
// of column tiles of ONE client:
//   * dy (A, K-major, 128-byte swizzle, 32 KB) is loaded once and stays in shared memory;
//   * the weight tiles (B, MN-major: rows = j, 128 columns of i = two 64-element swizzle atoms)
//     stream through a 4-deep TMA ring;
//   * two TMEM accumulators alternate, so the epilogue of tile t (tcgen05.ld -> bf16 -> swizzled
//     staging -> TMA store) overlaps the MMAs of tile t+1 and the loads of tiles t+2..t+4.
// Warp roles (384 threads): 0 = TMA producer, 1 = MMA issuer, 2 = TMEM allocator,
// 4..7 = epilogue of even tiles, 8..11 = epilogue of odd tiles (warp % 4 = TMEM lane quarter).
#include "common.cuh"
#include "tc_common.cuh"

namespace nfb {
namespace tc {

constexpr int kFcJ = 128;                         // reduction length (hidden width)
constexpr int kFcKBlocks = kFcJ / 64;             // 64-element k-blocks
constexpr int kFcBN = 128;                        // output columns per tile
constexpr int kFcStages = 4;
constexpr int kFcATile = 128 * 64 * 2;            // one k-block of A: 128 rows x 128 B
constexpr int kFcABytes = kFcKBlocks * kFcATile;  // 32 KB
constexpr int kFcBAtom = 64 * 128;                // 64 j-rows x 128 B (64 columns of i)
constexpr int kFcBStage = kFcKBlocks * 2 * kFcBAtom;  // 32 KB per column tile
constexpr int kFcEpiWarps = 8;
constexpr size_t kFcSmem = kFcABytes + kFcStages * kFcBStage + kFcEpiWarps * 4096 + 1024 + 256;

struct FcDgradArgs {
  int n_tiles;          // column tiles per client
  int ctas_per_client;
  int m_valid;          // batch rows (<= 128)
  int n_valid;          // columns
  float scale;          // applied to the fp32 accumulators before the bf16 rounding
  int* status;
};

__global__ void __launch_bounds__(128 + 32 * kFcEpiWarps, 1)
    tc_fc_dgrad_kernel(const __grid_constant__ CUtensorMap map_a,   // dy  [g][b][j], box [64, 128]
                       const __grid_constant__ CUtensorMap map_b,   // w   [g][j][i], box [64, 64]
                       const __grid_constant__ CUtensorMap map_o,   // out [g][b][i], box [64, 32]
                       const FcDgradArgs args) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>(
      (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* a_smem = smem;
  uint8_t* ring = smem + kFcABytes;
  uint8_t* staging = ring + kFcStages * kFcBStage;
  uint64_t* bars = reinterpret_cast<uint64_t*>(staging + kFcEpiWarps * 4096);
  uint64_t* full_bar = bars;                      // [stages]
  uint64_t* empty_bar = bars + kFcStages;         // [stages]
  uint64_t* acc_full = bars + 2 * kFcStages;      // [2]
  uint64_t* acc_empty = acc_full + 2;             // [2]
  uint64_t* a_bar = acc_empty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(a_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = blockIdx.x / args.ctas_per_client;
  const int part = blockIdx.x % args.ctas_per_client;
  const int per = (args.n_tiles + args.ctas_per_client - 1) / args.ctas_per_client;
  const int tile_lo = part * per;
  const int tile_hi = min(tile_lo + per, args.n_tiles);
  const int n_tiles = max(tile_hi - tile_lo, 0);

  if (threadIdx.x == 0) {
    prefetch_tmap(&map_a);
    prefetch_tmap(&map_b);
    prefetch_tmap(&map_o);
    for (int s = 0; s < kFcStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(&acc_full[a], 1); mbar_init(&acc_empty[a], 4); }
    mbar_init(a_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<2 * kFcBN>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== producer =====================
    if (elect_one() && n_tiles > 0) {
      mbar_expect_tx(a_bar, kFcABytes);
      for (int kb = 0; kb < kFcKBlocks; ++kb)     // rows >= m_valid are zero-filled by TMA
        tma_load_3d(a_smem + kb * kFcATile, &map_a, a_bar, kb * 64, 0, g);
      for (int i = 0; i < n_tiles; ++i) {
        const int s = i % kFcStages;
        const uint32_t ph = (i / kFcStages) & 1;
        if (!mbar_wait(&empty_bar[s], ph ^ 1)) { *args.status = 1; break; }
        mbar_expect_tx(&full_bar[s], kFcBStage);
        const int n0 = (tile_lo + i) * kFcBN;
        uint8_t* st = ring + s * kFcBStage;
        for (int kb = 0; kb < kFcKBlocks; ++kb)
          for (int a = 0; a < 2; ++a)             // one box per 64-column swizzle atom
            tma_load_3d(st + (kb * 2 + a) * kFcBAtom, &map_b, &full_bar[s], n0 + a * 64, kb * 64, g);
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (elect_one() && n_tiles > 0) {
      constexpr uint32_t idesc = make_idesc_bf16(128, kFcBN, 0u, 1u);
      bool ok = mbar_wait(a_bar, 0);
      if (!ok) *args.status = 1;
      const uint32_t a_addr = smem_u32(a_smem);
      for (int i = 0; i < n_tiles && ok; ++i) {
        const int s = i % kFcStages;
        const int acc = i & 1;
        const uint32_t ph = (i / kFcStages) & 1;
        const uint32_t aph = (i >> 1) & 1;
        if (!mbar_wait(&acc_empty[acc], aph ^ 1)) { *args.status = 1; ok = false; break; }
        if (!mbar_wait(&full_bar[s], ph)) { *args.status = 1; ok = false; break; }
        tc_fence_after();
        const uint32_t b_addr = smem_u32(ring + s * kFcBStage);
        const uint32_t d_tmem = tmem_base + acc * kFcBN;
#pragma unroll
        for (int kb = 0; kb < kFcKBlocks; ++kb) {
#pragma unroll
          for (int kk = 0; kk < 4; ++kk) {
            // A: K-major SW128 (8-row groups 1024 B apart, 32 B per K step)
            const uint64_t da = make_smem_desc(a_addr + kb * kFcATile + kk * 32, 0, 1024, 2);
            // B: MN-major SW128 (16 j-rows per K step, atoms 8 KB apart, 8-row groups 1024 B)
            const uint64_t db = make_smem_desc(b_addr + kb * 2 * kFcBAtom + kk * (16 * 128),
                                               kFcBAtom, 1024, 2);
            umma_f16(d_tmem, da, db, idesc, (kb > 0 || kk > 0) ? 1u : 0u);
          }
        }
        umma_commit(&empty_bar[s]);
        umma_commit(&acc_full[acc]);
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue =====================
    const int sub = warp & 3;
    const int grp = (warp - 4) >> 2;
    const uint32_t lane_addr = tmem_base + (static_cast<uint32_t>(sub * 32) << 16);
    uint8_t* my_stage = staging + (warp - 4) * 4096;
    const bool rows_live = sub * 32 < args.m_valid;      // warp-uniform: any valid batch row here?
    bool ok = true;
    for (int i = grp; i < n_tiles && ok; i += 2) {
      const int acc = grp;
      const uint32_t aph = (i >> 1) & 1;
      const int n0 = (tile_lo + i) * kFcBN;
      if (!mbar_wait(&acc_full[acc], aph)) { if (lane == 0) *args.status = 1; ok = false; break; }
      tc_fence_after();
      if (rows_live) {
#pragma unroll 1
        for (int c0 = 0; c0 < kFcBN; c0 += 64) {
          if (n0 + c0 >= args.n_valid) break;
          uint32_t r[64];
#pragma unroll
          for (int q = 0; q < 4; ++q) tmem_ld_32x32b_x16(lane_addr + acc * kFcBN + c0 + q * 16, r + q * 16);
          tmem_ld_wait();
          if (c0 + 64 >= kFcBN || n0 + c0 + 64 >= args.n_valid) {
            // last read of this accumulator: hand it back before the stores
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[acc]);
          }
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
          __syncwarp();
#pragma unroll
          for (int p = 0; p < 8; ++p) {
            float v[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(r[8 * p + e]) * args.scale;
            st_vec(my_stage + lane * 128 + ((p ^ (lane & 7)) * 16), pack<__nv_bfloat16>(v));
          }
          fence_proxy_async();
          __syncwarp();
          if (lane == 0) {
            asm volatile(
                "cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
                    &map_o),
                "r"(smem_u32(my_stage)), "r"(n0 + c0), "r"(sub * 32), "r"(g)
                : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
        }
      } else {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&acc_empty[acc]);
      }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<2 * kFcBN>(tmem_base);
}

}  // namespace tc
}  // namespace nfb

extern "C" {

// out[g][b, i] = scale * sum_j dy[g][b, j] * w[g][j, i]; dy bf16 [G][B <= 128][128], w bf16 [G][128][I]
// (row stride I, free group stride), out bf16 [G][B][I] with row stride `out_rs` and group stride
// `out_gs` (elements).  I must be a multiple of 64.
int nfb_tc_fc_dgrad(const nfb_tc_operand* dy, const nfb_tc_operand* w, void* out,
                    long long out_rs, long long out_gs, float scale, int* status, void* stream) {
  using namespace nfb;
  using namespace nfb::tc;
  NFB_REQUIRE(dy && w && out && status, NFB_EINVAL, "tc_fc_dgrad: null pointer");
  NFB_REQUIRE(dy->inner == kFcJ && w->rows == kFcJ, NFB_EINVAL,
              "tc_fc_dgrad: hidden width must be %d", kFcJ);
  NFB_REQUIRE(dy->rows >= 1 && dy->rows <= 128, NFB_ERANGE, "tc_fc_dgrad: batch must be 1..128");
  NFB_REQUIRE(w->inner >= 64 && w->inner % 64 == 0, NFB_EINVAL,
              "tc_fc_dgrad: output width must be a multiple of 64");
  NFB_REQUIRE(dy->groups == w->groups && dy->groups >= 1, NFB_EINVAL, "tc_fc_dgrad: group mismatch");
  const int groups = static_cast<int>(dy->groups);
  Operand a{dy->ptr, dy->inner, dy->rows, dy->groups, dy->row_stride, dy->group_stride};
  Operand b{w->ptr, w->inner, w->rows, w->groups, w->row_stride, w->group_stride};
  CUtensorMap ma, mb, mo;
  int rc = get_map(a, 64, 128, 128, &ma);          // K-major: box [64 k, 128 rows]
  if (rc) return rc;
  rc = get_map(b, 64, 64, 128, &mb);               // MN-major: box [64 columns, 64 j-rows]
  if (rc) return rc;
  rc = get_out_map(out, w->inner, dy->rows, groups, out_rs, out_gs, true, &mo);
  if (rc) return rc;
  FcDgradArgs args{};
  args.n_tiles = static_cast<int>((w->inner + kFcBN - 1) / kFcBN);
  int per_client = sm_count() / groups;
  if (per_client < 1) per_client = 1;
  if (per_client > args.n_tiles) per_client = args.n_tiles;
  args.ctas_per_client = per_client;
  args.m_valid = static_cast<int>(dy->rows);
  args.n_valid = static_cast<int>(w->inner);
  args.scale = scale;
  args.status = status;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(tc_fc_dgrad_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(kFcSmem));
    if (e != cudaSuccess)
      return fail(static_cast<int>(e), "tc_fc_dgrad: smem attribute: %s", cudaGetErrorString(e));
    attr_set = true;
  }
  tc_fc_dgrad_kernel<<<groups * per_client, 128 + 32 * kFcEpiWarps, kFcSmem, as_stream(stream)>>>(
      ma, mb, mo, args);
  return launch_status("tc_fc_dgrad");
}

}  // extern "C"
