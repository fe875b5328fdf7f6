// conv1 (1 -> 32 channels, 3x3) + bias + ReLU on the tensor core.
//
// K = 9 looks too thin for UMMA, and the first version of this layer ran on CUDA cores (45 us per
// step for 32 clients x 64 images, LSU-bound: 18 LDS.128 per 144 FMA).  As a GEMM it is
//     a1[q, ch] = relu( sum_t X[q, t] * W[ch, t] ),   X[q, t] = x[pixel(q) + tap t],  t = 0..8
// and the fp32 bias added in the epilogue.  bf16 operands alone would cost this layer its fp32
// accuracy (the ReLU pattern of a1 is the mask of the backward pass; conv1.bias' gradient moved
// from 8 % to 11 % off the fp32 reference), so both operands are split into bf16 high and low
// parts and K carries three products:  x_hi*w_hi + x_lo*w_hi + x_hi*w_lo  (27 of K = 32, two
// tcgen05.mma of 128 x 32 x 16 per 128 output positions; the dropped x_lo*w_lo term is 2^-16
// relative).  The im2col rows are built in shared memory by the threads that later read the
// result (9 cached loads + four 16-byte stores each); the kernel is bound by its 88 MB of output.
//
// 256 threads = two independent groups of 4 warps (warp % 4 = TMEM lane quarter).  Per tile a
// group: writes its 128 window rows -> group barrier -> one thread issues the MMA and commits to
// the group's mbarrier -> all wait, tcgen05.ld 32 columns, ReLU, bf16 -> swizzled staging ->
// TMA store of 32 rows per warp.  Operands use the un-swizzled K-major layout
// [8-element chunk][row][16 B] (LBO = chunk pitch, SBO = 128 B).
#include "common.cuh"
#include "tc_common.cuh"

namespace nfb {
namespace tc {

struct Conv1Args {
  const float* x;          // [C][B][img_w * img_w] fp32
  long long x_gs;          // elements between clients
  const float* slots;      // fp32 parameter rows
  long long slot_gs;
  long long w_off, b_off;  // conv1.weight [32][9], conv1.bias [32] inside a row
  int rows;                // B * grid_w * grid_w output positions per client
  int grid_w;              // 26
  int tiles_per_client;
  int ctas_per_client;
  int* status;
};

constexpr int kC1Chunks = 4;                     // K = 32 = four 8-element chunks
constexpr int kC1ATile = kC1Chunks * 128 * 16;   // chunks x 128 rows x 16 B
constexpr int kC1BTile = kC1Chunks * 32 * 16;    // chunks x 32 channels x 16 B

// K slot k of an A row / B row:  k in [0,9): x_hi[t] / w_hi[t];  [9,18): x_lo[t] / w_hi[t];
// [18,27): x_hi[t] / w_lo[t];  [27,32): zero.
__device__ __forceinline__ void split_bf16(float v, float& hi, float& lo) {
  hi = __bfloat162float(__float2bfloat16_rn(v));
  lo = v - hi;
}

__global__ void __launch_bounds__(256, 3)
    tc_conv1_fwd_kernel(const __grid_constant__ CUtensorMap map_o,   // a1 [C][rows][32] bf16, box [32, 32]
                        const Conv1Args args) {
  __shared__ __align__(1024) uint8_t a_tiles[2][kC1ATile];
  __shared__ __align__(1024) uint8_t staging[8][2048];
  __shared__ __align__(128) uint8_t b_tile[kC1BTile];
  __shared__ __align__(16) float s_bias[32];
  __shared__ __align__(8) uint64_t mma_bar[2];
  __shared__ uint32_t tmem_slot;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int grp = warp >> 2, sub = warp & 3;
  const int g = blockIdx.x / args.ctas_per_client;
  const int part = blockIdx.x % args.ctas_per_client;
  const int per = (args.tiles_per_client + args.ctas_per_client - 1) / args.ctas_per_client;
  const int tile_lo = part * per;
  const int tile_hi = min(tile_lo + per, args.tiles_per_client);

  if (threadIdx.x == 0) {
    prefetch_tmap(&map_o);
    mbar_init(&mma_bar[0], 1);
    mbar_init(&mma_bar[1], 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<64>(&tmem_slot);
  if (threadIdx.x >= 128 && threadIdx.x < 160)
    s_bias[threadIdx.x - 128] =
        args.slots[static_cast<long long>(g) * args.slot_gs + args.b_off + threadIdx.x - 128];
  // weights: one (channel, chunk) per thread
  if (threadIdx.x < 32 * kC1Chunks) {
    const int ch = threadIdx.x & 31, chunk = threadIdx.x >> 5;
    const float* w = args.slots + static_cast<long long>(g) * args.slot_gs + args.w_off + ch * 9;
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int k = chunk * 8 + e;
      float hi = 0.f, lo = 0.f;
      if (k < 27) split_bf16(w[k % 9], hi, lo);
      v[e] = k < 18 ? hi : lo;                      // (k >= 27: both zero)
    }
    st_vec(b_tile + chunk * 512 + ch * 16, pack<__nv_bfloat16>(v));
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const uint32_t d_tmem = tmem_base + grp * 32;
  const uint32_t lane_addr = d_tmem + (static_cast<uint32_t>(sub * 32) << 16);
  uint8_t* a_tile = a_tiles[grp];
  uint8_t* my_stage = staging[warp];
  const int gsz = args.grid_w * args.grid_w, iw = args.grid_w + 2;
  const float* xg = args.x + static_cast<long long>(g) * args.x_gs;
  constexpr uint32_t idesc = make_idesc_bf16(128, 32, 0u, 0u);
  const int trow = sub * 32 + lane;                // row of the group's tile this thread owns

  // this thread's 3x3 input window for a tile row (zeros past the last position)
  auto fetch_window = [&](int m, float* dst) {
#pragma unroll
    for (int t = 0; t < 9; ++t) dst[t] = 0.f;
    if (m < args.rows) {
      const int b = m / gsz, rem = m - b * gsz;
      const int oh = rem / args.grid_w, ow = rem - oh * args.grid_w;
      const float* px = xg + static_cast<long long>(b) * (iw * iw) + oh * iw + ow;
#pragma unroll
      for (int kh = 0; kh < 3; ++kh)
#pragma unroll
        for (int kw = 0; kw < 3; ++kw) dst[kh * 3 + kw] = __ldg(px + kh * iw + kw);
    }
  };
  float win[9];
  if (tile_lo + grp < tile_hi) fetch_window((tile_lo + grp) * 128 + trow, win);
  int n_done = 0;
  for (int tile = tile_lo + grp; tile < tile_hi; tile += 2, ++n_done) {
    float row[32];
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      float hi, lo;
      split_bf16(win[t], hi, lo);
      row[t] = hi; row[9 + t] = lo; row[18 + t] = hi;
    }
#pragma unroll
    for (int k = 27; k < 32; ++k) row[k] = 0.f;
    // (the previous MMA of this group has completed: everyone waited on its barrier below)
#pragma unroll
    for (int c = 0; c < kC1Chunks; ++c)
      st_vec(a_tile + c * 2048 + trow * 16, pack<__nv_bfloat16>(row + 8 * c));
    fence_proxy_async();
    tc_fence_before();                              // orders last tile's tcgen05.ld before the MMA
    asm volatile("bar.sync %0, 128;" ::"r"(1 + grp) : "memory");
    if (sub == 0 && lane == 0) {
      tc_fence_after();
#pragma unroll
      for (int kk = 0; kk < 2; ++kk) {              // 16 K slots = two chunks per MMA
        const uint64_t da = make_smem_desc(smem_u32(a_tile) + kk * 4096, 2048, 128, 0);
        const uint64_t db = make_smem_desc(smem_u32(b_tile) + kk * 1024, 512, 128, 0);
        umma_f16(d_tmem, da, db, idesc, kk);
      }
      umma_commit(&mma_bar[grp]);
    }
    // the next tile's window loads fly while the tensor core works
    if (tile + 2 < tile_hi) fetch_window((tile + 2) * 128 + trow, win);
    if (!mbar_wait(&mma_bar[grp], n_done & 1)) {
      if (lane == 0) *args.status = 1;
      break;
    }
    tc_fence_after();
    uint32_t r[32];
    tmem_ld_32x32b_x16(lane_addr, r);
    tmem_ld_32x32b_x16(lane_addr + 16, r + 16);
    tmem_ld_wait();
    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    __syncwarp();
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float v[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) v[e] = fmaxf(__uint_as_float(r[8 * c + e]) + s_bias[8 * c + e], 0.f);
      st_vec(my_stage + lane * 64 + ((c ^ ((lane >> 1) & 3)) * 16), pack<__nv_bfloat16>(v));
    }
    fence_proxy_async();
    __syncwarp();
    if (lane == 0) {
      asm volatile(
          "cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(
              &map_o),
          "r"(smem_u32(my_stage)), "r"(0), "r"(tile * 128 + sub * 32), "r"(g)
          : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
  }
  if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<64>(tmem_base);
}

}  // namespace tc
}  // namespace nfb

// a1[c][q][ch] = relu(conv1(x)[q][ch] + bias[ch]) for every position q of the width x width output
// grid of each of the B images (q = b * width^2 + oh * width + ow); x fp32 [C][B][(width+2)^2].
extern "C" int nfb_tc_conv1_fwd(const float* x, long long x_group_stride, const float* slots,
                                long long slot_group_stride, long long w_off, long long b_off,
                                int clients, int images, int width, void* a1,
                                long long a1_group_stride, int* status, void* stream) {
  using namespace nfb;
  using namespace nfb::tc;
  NFB_REQUIRE(x && slots && a1 && status, NFB_EINVAL, "tc_conv1_fwd: null pointer");
  NFB_REQUIRE(clients >= 1 && images >= 1 && width >= 1, NFB_EINVAL, "tc_conv1_fwd: bad sizes");
  const long long rows = static_cast<long long>(images) * width * width;
  CUtensorMap mo;
  int rc = get_lowp_map(a1, 32, rows, clients, 32, a1_group_stride, &mo);
  if (rc) return rc;
  Conv1Args args{};
  args.x = x;
  args.x_gs = x_group_stride;
  args.slots = slots;
  args.slot_gs = slot_group_stride;
  args.w_off = w_off;
  args.b_off = b_off;
  args.rows = static_cast<int>(rows);
  args.grid_w = width;
  args.tiles_per_client = static_cast<int>((rows + 127) / 128);
  int per_client = (3 * sm_count()) / clients;          // three resident CTAs per SM
  if (per_client < 1) per_client = 1;
  if (per_client > (args.tiles_per_client + 1) / 2) per_client = (args.tiles_per_client + 1) / 2;
  args.ctas_per_client = per_client;
  args.status = status;
  tc_conv1_fwd_kernel<<<clients * per_client, 256, 0, as_stream(stream)>>>(mo, args);
  return launch_status("tc_conv1_fwd");
}
