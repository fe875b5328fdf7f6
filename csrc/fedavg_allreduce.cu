// K-AGG + K-BCAST, multi-GPU form: ONE kernel per rank that
//   A. reduces this GPU's clients into an unnormalised partial  P_r = sum_k n_k m_k c_k x_k
//      (same column_reduce as the single-GPU kernel; HBM-bound, K_g*N*b bytes),
//   B. crosses a flag barrier with the other ranks (release/acquire on words in peer
//      memory — no NCCL, no host),
//   C. reduce-scatters over NVLink from inside the kernel: rank r owns columns
//      [r*N/G, (r+1)*N/G) and pulls that shard of every peer's partial — either with
//      plain P2P loads (ld.relaxed.sys on mapped peer pointers) or, through NVSwitch,
//      with ONE multimem.ld_reduce per 16 bytes that returns the sum already reduced in
//      the switch (NVLS),
//   D. finishes the FedAvg math on the shard (1/sum n, folded DP noise, casts) and
//      all-gathers by STORING the result into every rank's global-model buffers (P2P
//      st.global to each peer, or one multimem.st that the switch replicates) — this is
//      the broadcast of the new global model, so no separate K-BCAST transfer exists,
//   E. crosses a second flag barrier so every rank owns the complete new model on exit.
//
// Per-GPU NVLink traffic: 4N(G-1)/G bytes in (C) and 4N(G-1)/G (+2N(G-1)/G bf16) out (D).
//
// Emulation: `ranks_in_grid` > 1 runs G emulated ranks as blockIdx.y of ONE cooperative
// launch on one GPU (all pointers then refer to the same device).  This is how the kernel
// is tested on a single B200 without launching kernels that wait on each other.
//
// Safety: every spin is bounded by `spin_limit_ns` of %globaltimer; on expiry the kernel
// sets status[0] = 1 and returns instead of hanging the GPU.
#include <cooperative_groups.h>

#include "common.cuh"
#include "fedavg_device.cuh"
#include "philox.cuh"

namespace nfb {

struct RankView {
  // local inputs (only dereferenced by the owning rank)
  const void* slots;
  const float* counts;
  const float* mask;
  const float* sumsq;
  // peer-visible buffers (dereferenced by every rank)
  float* partial;        // [N]
  float* scal;           // [4 + 2G]: own sum n, sum n^2; then every rank's pair (pushed with flag 1)
  unsigned* flags;       // [(1 + kMaxChunks) * G]: epoch words written by peers (slot 0 = done)
  float* master;         // [N] new global model
  __nv_bfloat16* lowp;   // [N] bf16 copy (may be null)
  unsigned* counters;    // [1 + kMaxChunks] local arrival counters (done, per-chunk), zero between calls
  float* stats_all;      // [G][stats_len] every rank's per-client round metrics (may be null)
};

struct AllreduceArgs {
  int G, K, x_dtype;
  int64_t N, row_stride;
  float clip_norm, noise_std;
  const uint8_t* noise_mask;  // local copy on each rank (same content)
  uint64_t seed, group0;
  unsigned epoch;             // generation, strictly increasing per call
  int producers;              // CTAs [0, producers) reduce locally, the rest reduce across ranks
  int n_chunks;               // column chunks (<= kMaxChunks)
  int64_t chunk_elems;        // elements per chunk (multiple of 8)
  unsigned long long spin_limit_ns;
  const float* mc_partial;    // multicast address of partial[] (null = P2P loads)
  float* mc_master;           // multicast address of master[]  (null = P2P stores)
  __nv_bfloat16* mc_lowp;
  int* status;                // [1] local: set to 1 on barrier timeout
  // Streaming waves (K * N * b larger than HBM): `accumulate` adds this call's clients to the
  // partial (and to sum n, sum n^2) left by earlier calls instead of overwriting them;
  // `local_only` stops after the local phase - no flag, no peer traffic - so only the LAST wave
  // pays the cross-GPU reduce-scatter / all-gather.
  int accumulate, local_only;
  const float* my_stats;      // [stats_len] local per-client metrics, copied to every rank's stats_all
  int stats_len;
  // K-BCAST as the tail of the same kernel: once this rank owns the complete new model (second
  // flag barrier), EVERY CTA replicates it into the local client rows (fp32 + bf16) and zeroes the
  // momentum rows.  Emulated ranks (one launch) use consecutive targets of `bc_rank_stride` rows.
  nfb_bcast_target bc;
  // optional phase timestamps (%globaltimer, ns) for latency studies: [0] kernel start (CTA 0),
  // [1] last producer CTA done (flag published), [2] consumer CTA 0 saw every peer's chunk-0 flag,
  // [3] consumer CTA 0 has read the peers' scalars, [4] last consumer CTA done (before its signal),
  // [5] completion flags of all peers seen, [6] broadcast tail finished (CTA 0).  Null = off.
  unsigned long long* dbg;
};

__device__ __forceinline__ unsigned long long now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_acquire_gpu_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_relaxed_sys_f4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ float4 multimem_ld_reduce_f4(const float* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(mc)
               : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f4(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc),
               "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void multimem_st_b64(void* mc, uint2 v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1,%2};" ::"l"(mc),
               "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y))
               : "memory");
}

constexpr int kMaxChunks = 64;
constexpr int kGoSlot = 1 + kMaxChunks;   // counters[kGoSlot]: epoch of the last completed call

// Tail phase: replicate master[] into the client rows.  Called by every CTA of the rank after the
// completion barrier; `part` / `parts` = this CTA's index / the rank's CTA count.
__device__ __forceinline__ void bcast_tail(const float* master, int64_t N, const nfb_bcast_target& bc,
                                           int part, int parts) {
  const int64_t groups = N >> 2;
  for (int64_t g = part * static_cast<int64_t>(blockDim.x) + threadIdx.x; g < groups;
       g += static_cast<int64_t>(parts) * blockDim.x) {
    const float4 v = __ldcg(reinterpret_cast<const float4*>(master + 4 * g));   // L2: peers wrote it
    const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
    uint2 pk;
    pk.x = *reinterpret_cast<const uint32_t*>(&lo);
    pk.y = *reinterpret_cast<const uint32_t*>(&hi);
    for (int64_t k = 0; k < bc.rows; ++k) {
      if (bc.slots) *reinterpret_cast<float4*>(bc.slots + k * bc.slots_stride + 4 * g) = v;
      if (bc.lowp)
        *reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(bc.lowp) + k * bc.lowp_stride + 4 * g) = pk;
      if (bc.momentum)
        *reinterpret_cast<float4*>(bc.momentum + k * bc.momentum_stride + 4 * g) =
            make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
}

// Every CTA of the rank waits (bounded) until the last consumer CTA has seen all peers' completion
// flags and published this call's epoch in counters[kGoSlot].
__device__ __forceinline__ bool wait_go(const unsigned* go, unsigned epoch, unsigned long long limit,
                                        int* status) {
  __shared__ int s_go;
  if (threadIdx.x == 0) {
    s_go = 1;
    const unsigned long long t0 = now_ns();
    unsigned v;
    do {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(go) : "memory");
      if (v >= epoch) break;
      if (now_ns() - t0 > limit) { *status = 1; s_go = 0; break; }
    } while (true);
  }
  __syncthreads();
  return s_go != 0;
}

// One thread per peer: publish `value` into slot `slot` of every peer's flag array.
__device__ __forceinline__ void signal_peers(const RankView* views, int rank, int G, int slot,
                                             unsigned value) {
  if (threadIdx.x < G) {
    __threadfence_system();
    st_release_sys(views[threadIdx.x].flags + slot * G + rank, value);
  }
}
// One thread per peer: wait until every peer has published `value` into my slot `slot`.
__device__ __forceinline__ bool wait_peers(const RankView* views, int rank, int G, int slot,
                                           unsigned value, unsigned long long limit, int* status) {
  __shared__ int s_ok;
  if (threadIdx.x == 0) s_ok = 1;
  __syncthreads();
  if (threadIdx.x < G) {
    const unsigned* mine = views[rank].flags + slot * G + threadIdx.x;
    const unsigned long long t0 = now_ns();
    while (ld_acquire_sys(mine) < value) {
      if (now_ns() - t0 > limit) { *status = 1; s_ok = 0; break; }
    }
  }
  __syncthreads();
  return s_ok != 0;
}

// The kernel.  gridDim.x = P producer CTAs followed by Q consumer CTAs (all co-resident):
//   producers walk the column chunks in order: weighted reduce of this GPU's clients into the
//     (peer-visible) partial buffer; the LAST producer CTA to finish chunk c publishes flag
//     1+c to every peer — no grid-wide barrier, just one arrival counter per chunk;
//   consumers wait for flag 1+c from every peer, then reduce THIS rank's shard of chunk c over
//     all ranks (P2P loads or multimem.ld_reduce), apply 1/sum(n), the folded DP noise and the
//     casts, and store the result into every rank's model buffers (P2P stores or multimem.st);
//   so chunk c crosses NVLink while chunk c+1 is still being read from HBM;
//   the last consumer CTA to finish publishes flag 0 ("my stores are done") and waits for flag
//   0 from every peer, which keeps the kernel alive until this rank owns the complete model.
template <typename T>
__global__ void __launch_bounds__(256)
    fedavg_allreduce_kernel(const RankView* __restrict__ views, int my_rank, AllreduceArgs a) {
  extern __shared__ float coef[];
  __shared__ float red[33];
  __shared__ unsigned s_last;
  constexpr int V = Vec16<T>::N;
  const int rank = (gridDim.y > 1) ? static_cast<int>(blockIdx.y) : my_rank;
  const RankView me = views[rank];
  const int G = a.G;
  const int64_t chunk = a.chunk_elems;              // multiple of 8*G... see host wrapper
  const int n_chunks = a.n_chunks;
  // Latency regime (one chunk): there is nothing to pipeline, so EVERY CTA first reduces its share
  // of the local clients and then turns consumer - the local phase runs on the whole grid instead
  // of the producer half (2 GPUs x 32 clients x 1.2 M: 42 -> 25 us of a 79 us kernel).
  const bool single = n_chunks == 1 && !a.local_only;
  const int P = (a.local_only || single) ? static_cast<int>(gridDim.x) : a.producers;
  const int Q = single ? static_cast<int>(gridDim.x) : static_cast<int>(gridDim.x) - P;
  const bool producer = static_cast<int>(blockIdx.x) < P;

  if (a.dbg && blockIdx.x == 0 && threadIdx.x == 0) a.dbg[0] = now_ns();
  if (producer) {
    // ---- A. local unnormalised partial, chunk by chunk -----------------------------------
    const float q_local = agg_prologue(coef, a.K, me.counts, me.sumsq, me.mask, a.clip_norm,
                                       /*normalise=*/false, red);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      float s = 0.f;
      for (int k = 0; k < a.K; ++k) s += me.counts[k] * (me.mask ? me.mask[k] : 1.f);
      me.scal[0] = (a.accumulate ? me.scal[0] : 0.f) + s;
      me.scal[1] = (a.accumulate ? me.scal[1] : 0.f) + q_local;
    }
    const T* x = static_cast<const T*>(me.slots);
    for (int c = 0; c < n_chunks; ++c) {
      const int64_t lo = c * chunk, hi = min(lo + chunk, a.N);
      const int64_t g_lo = lo / V, groups = (hi - lo) / V;
      for (int64_t g = blockIdx.x * static_cast<int64_t>(blockDim.x) + threadIdx.x; g < groups;
           g += static_cast<int64_t>(P) * blockDim.x) {
        float acc[V];
        column_reduce<T>(x, a.row_stride, a.K, coef, (g_lo + g) * V, acc);
#pragma unroll
        for (int q = 0; q < V / 4; ++q) {
          float4* dst = reinterpret_cast<float4*>(me.partial + (g_lo + g) * V + 4 * q);
          float4 v = make_float4(acc[4 * q], acc[4 * q + 1], acc[4 * q + 2], acc[4 * q + 3]);
          if (a.accumulate) {
            const float4 prev = *dst;
            v.x += prev.x; v.y += prev.y; v.z += prev.z; v.w += prev.w;
          }
          *dst = v;
        }
      }
      if (a.local_only) continue;          // nobody waits for this wave
      __threadfence_system();
      __syncthreads();
      if (threadIdx.x == 0) {
        const unsigned prev = atomicAdd(me.counters + 1 + c, 1u);
        s_last = (prev == static_cast<unsigned>(P) - 1u) ? 1u : 0u;
        if (s_last) me.counters[1 + c] = 0u;       // self-clean for the next call
      }
      __syncthreads();
      if (s_last) {
        __threadfence();
        // the weight sums travel WITH the first flag: written into every peer's scal[4 + 2 * rank]
        // before the release store, so consumers read them from local memory (one NVLink round
        // trip less on the critical path)
        if (c == 0 && threadIdx.x < G) {
          float* dst = views[threadIdx.x].scal + 4 + 2 * rank;
          dst[0] = me.scal[0];
          dst[1] = me.scal[1];
        }
        signal_peers(views, rank, G, 1 + c, a.epoch);
        if (a.dbg && threadIdx.x == 0 && c == n_chunks - 1) a.dbg[1] = now_ns();
      }
      __syncthreads();
    }
    if (!single && a.bc.rows > 0 && !a.local_only &&
        wait_go(me.counters + kGoSlot, a.epoch, a.spin_limit_ns, a.status)) {
      nfb_bcast_target bc = a.bc;
      if (gridDim.y > 1) {   // emulation: rank r's rows follow rank r-1's
        if (bc.slots) bc.slots += static_cast<int64_t>(rank) * bc.rows * bc.slots_stride;
        if (bc.lowp) bc.lowp = static_cast<__nv_bfloat16*>(bc.lowp) + static_cast<int64_t>(rank) * bc.rows * bc.lowp_stride;
        if (bc.momentum) bc.momentum += static_cast<int64_t>(rank) * bc.rows * bc.momentum_stride;
      }
      bcast_tail(me.master, a.N, bc, static_cast<int>(blockIdx.x), static_cast<int>(gridDim.x));
    }
    if (!single) return;
  }

  // ---- consumers -----------------------------------------------------------------------------
  const int cb = single ? static_cast<int>(blockIdx.x) : static_cast<int>(blockIdx.x) - P;
  float s_tot = 0.f, q_tot = 0.f;
  bool have_scal = false;
  bool ok = true;
  for (int c = 0; c < n_chunks && ok; ++c) {
    ok = wait_peers(views, rank, G, 1 + c, a.epoch, a.spin_limit_ns, a.status);
    if (!ok) break;
    if (a.dbg && cb == 0 && threadIdx.x == 0 && c == 0) a.dbg[2] = now_ns();
    if (!have_scal) {
      // every rank pushed its (sum n, sum n^2) into my scal[4 + 2r] before its flag: local reads
      for (int r = 0; r < G; ++r) {
        float s_r, q_r;
        asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(s_r) : "l"(me.scal + 4 + 2 * r) : "memory");
        asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(q_r) : "l"(me.scal + 5 + 2 * r) : "memory");
        s_tot += s_r;
        q_tot += q_r;
      }
      have_scal = true;
      if (a.dbg && cb == 0 && threadIdx.x == 0) a.dbg[3] = now_ns();
    }
    // every client of every rank masked out: keep the previous global model (same decision on
    // all ranks, s_tot is the global sum)
    if (!(s_tot > 0.f)) continue;
    const float inv = 1.f / s_tot;
    const float sigma = a.noise_std > 0.f ? a.noise_std * sqrtf(q_tot) * inv : 0.f;

    const int64_t lo = c * chunk, hi = min(lo + chunk, a.N);
    const int64_t quads = (hi - lo) >> 2;
    const int64_t per = (quads + G - 1) / G;
    const int64_t q_lo = (lo >> 2) + per * rank;
    const int64_t q_hi = min(q_lo + per, (lo >> 2) + quads);
    for (int64_t q = q_lo + cb * static_cast<int64_t>(blockDim.x) + threadIdx.x; q < q_hi;
         q += static_cast<int64_t>(Q) * blockDim.x) {
      const int64_t col = q << 2;
      float4 sum;
      if (a.mc_partial) {
        sum = multimem_ld_reduce_f4(a.mc_partial + col);
      } else {
        float4 v[8];
        sum = make_float4(0.f, 0.f, 0.f, 0.f);
        // fixed rank order => identical bits on every rank; up to 8 peer loads in flight
        for (int i0 = 0; i0 < G; i0 += 8) {
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (i0 + i < G) v[i] = ld_relaxed_sys_f4(views[i0 + i].partial + col);
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (i0 + i < G) { sum.x += v[i].x; sum.y += v[i].y; sum.z += v[i].z; sum.w += v[i].w; }
        }
      }
      float o[4] = {sum.x * inv, sum.y * inv, sum.z * inv, sum.w * inv};
      if (sigma > 0.f) {
        const float4 z = Philox::normal4(a.group0 + q, a.seed);
        float zz[4] = {z.x, z.y, z.z, z.w};
        if (a.noise_mask) {
          const uchar4 m = *reinterpret_cast<const uchar4*>(a.noise_mask + col);
          zz[0] *= m.x ? 1.f : 0.f; zz[1] *= m.y ? 1.f : 0.f;
          zz[2] *= m.z ? 1.f : 0.f; zz[3] *= m.w ? 1.f : 0.f;
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) o[e] = fmaf(sigma, zz[e], o[e]);
      }
      const float4 ov = make_float4(o[0], o[1], o[2], o[3]);
      const __nv_bfloat162 lo2 = __floats2bfloat162_rn(o[0], o[1]);
      const __nv_bfloat162 hi2 = __floats2bfloat162_rn(o[2], o[3]);
      uint2 packed;
      packed.x = *reinterpret_cast<const uint32_t*>(&lo2);
      packed.y = *reinterpret_cast<const uint32_t*>(&hi2);
      if (a.mc_master) {
        multimem_st_f4(a.mc_master + col, ov);
        if (a.mc_lowp) multimem_st_b64(a.mc_lowp + col, packed);
      } else {
        for (int i = 0; i < G; ++i) {
          const int r = (i + rank) % G;
          *reinterpret_cast<float4*>(views[r].master + col) = ov;
          if (views[r].lowp) *reinterpret_cast<uint2*>(views[r].lowp + col) = packed;
        }
      }
    }
  }

  // ---- E. completion: last consumer CTA tells the peers and waits for theirs ---------------
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned prev = atomicAdd(me.counters, 1u);
    s_last = (prev == static_cast<unsigned>(Q) - 1u) ? 1u : 0u;
    if (s_last) me.counters[0] = 0u;
  }
  __syncthreads();
  if (s_last) {
    if (a.dbg && threadIdx.x == 0) a.dbg[4] = now_ns();
    // job-wide round metrics ride along: my per-client (loss, correct, samples) rows go into every
    // rank's stats_all[rank] before the completion flag, so each rank leaves with all of them
    if (a.my_stats && me.stats_all) {
      // (emulated ranks of one launch read consecutive rows of one array)
      const float* mine = a.my_stats + (gridDim.y > 1 ? static_cast<int64_t>(rank) * a.stats_len : 0);
      for (int i = threadIdx.x; i < a.stats_len * G; i += blockDim.x) {
        const int r = i / a.stats_len, j = i - r * a.stats_len;
        views[r].stats_all[static_cast<int64_t>(rank) * a.stats_len + j] = mine[j];
      }
      __threadfence_system();
      __syncthreads();
    }
    __threadfence();
    signal_peers(views, rank, G, 0, a.epoch);
    const bool done = wait_peers(views, rank, G, 0, a.epoch, a.spin_limit_ns, a.status);
    if (a.dbg && threadIdx.x == 0) a.dbg[5] = now_ns();
    if (a.bc.rows > 0 && done && threadIdx.x == 0) {
      __threadfence();
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(me.counters + kGoSlot), "r"(a.epoch)
                   : "memory");
    }
  }
  if (a.bc.rows > 0 && ok && wait_go(me.counters + kGoSlot, a.epoch, a.spin_limit_ns, a.status)) {
    nfb_bcast_target bc = a.bc;
    if (gridDim.y > 1) {
      if (bc.slots) bc.slots += static_cast<int64_t>(rank) * bc.rows * bc.slots_stride;
      if (bc.lowp) bc.lowp = static_cast<__nv_bfloat16*>(bc.lowp) + static_cast<int64_t>(rank) * bc.rows * bc.lowp_stride;
      if (bc.momentum) bc.momentum += static_cast<int64_t>(rank) * bc.rows * bc.momentum_stride;
    }
    bcast_tail(me.master, a.N, bc, static_cast<int>(blockIdx.x), static_cast<int>(gridDim.x));
    if (a.dbg && cb == 0 && threadIdx.x == 0) a.dbg[6] = now_ns();
  }
}

}  // namespace nfb

namespace nfb {
// Host-side mirror of RankView for the C ABI (11 pointers, same order).
static_assert(sizeof(RankView) == 11 * sizeof(void*), "RankView must stay 11 pointers");

template <typename T>
static int launch_allreduce(const RankView* views, int my_rank, int ranks_in_grid,
                            int grid_blocks, const AllreduceArgs& a, cudaStream_t st) {
  const size_t smem = static_cast<size_t>(a.K) * sizeof(float);
  int per_sm = 0;
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(
      &per_sm, fedavg_allreduce_kernel<T>, 256, smem);
  if (e != cudaSuccess) return fail(static_cast<int>(e), "allreduce occupancy query: %s",
                                    cudaGetErrorString(e));
  const int64_t resident = static_cast<int64_t>(per_sm) * sm_count();
  if (static_cast<int64_t>(grid_blocks) * ranks_in_grid > resident)
    return fail(NFB_ERANGE, "fedavg_allreduce: %d x %d CTAs exceed the %lld that can be "
                "co-resident", grid_blocks, ranks_in_grid, static_cast<long long>(resident));
  dim3 grid(static_cast<unsigned>(grid_blocks), static_cast<unsigned>(ranks_in_grid));
  AllreduceArgs args = a;
  void* params[] = {const_cast<RankView**>(&views), &my_rank, &args};
  e = cudaLaunchCooperativeKernel(reinterpret_cast<void*>(fedavg_allreduce_kernel<T>), grid,
                                  dim3(256), params, smem, st);
  if (e != cudaSuccess) {
    cudaGetLastError();
    return fail(static_cast<int>(e), "fedavg_allreduce launch: %s", cudaGetErrorString(e));
  }
  return 0;
}

}  // namespace nfb

extern "C" int nfb_allreduce_max_blocks(int x_dtype, int64_t K) {
  using namespace nfb;
  const size_t smem = static_cast<size_t>(K) * sizeof(float);
  int per_sm = 0;
  cudaError_t e;
  if (x_dtype == NFB_DT_F32)
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fedavg_allreduce_kernel<float>,
                                                      256, smem);
  else if (x_dtype == NFB_DT_BF16)
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(
        &per_sm, fedavg_allreduce_kernel<__nv_bfloat16>, 256, smem);
  else
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fedavg_allreduce_kernel<__half>,
                                                      256, smem);
  if (e != cudaSuccess) return fail(-static_cast<int>(e), "occupancy query: %s",
                                    cudaGetErrorString(e));
  return per_sm * sm_count();
}

extern "C" int nfb_fedavg_allreduce(const void* views_dev, int my_rank, int G, int ranks_in_grid,
                                    int x_dtype, int64_t K, int64_t N, int64_t row_stride,
                                    float clip_norm, float noise_std, const uint8_t* noise_mask,
                                    uint64_t seed, uint64_t offset, uint32_t epoch,
                                    int producers, int64_t chunk_elems, int grid_blocks,
                                    const float* mc_partial,
                                    float* mc_master, void* mc_lowp, int* status,
                                    uint64_t spin_limit_ns, int flags, const float* my_stats,
                                    int stats_len, const void* bcast, unsigned long long* dbg,
                                    void* stream) {
  using namespace nfb;
  NFB_REQUIRE(views_dev && status, NFB_EINVAL, "fedavg_allreduce: null pointer");
  NFB_REQUIRE(G >= 1 && G <= 32 && my_rank >= 0 && my_rank < G, NFB_EINVAL,
              "fedavg_allreduce: bad rank %d of %d", my_rank, G);
  NFB_REQUIRE(ranks_in_grid == 1 || ranks_in_grid == G, NFB_EINVAL,
              "fedavg_allreduce: ranks_in_grid must be 1 or G");
  NFB_REQUIRE(K >= 1 && K <= 12000, NFB_ERANGE, "fedavg_allreduce: K outside [1, 12000]");
  NFB_REQUIRE((N & 7) == 0 && (row_stride & 7) == 0 && N > 0, NFB_EINVAL,
              "fedavg_allreduce: N and row_stride must be positive multiples of 8");
  NFB_REQUIRE((offset & 3u) == 0, NFB_EINVAL, "fedavg_allreduce: offset must be multiple of 4");
  NFB_REQUIRE(epoch >= 1 && grid_blocks >= 2, NFB_EINVAL, "fedavg_allreduce: bad epoch/grid");
  NFB_REQUIRE(producers >= 1 && producers < grid_blocks, NFB_EINVAL,
              "fedavg_allreduce: need at least one producer and one consumer CTA");
  NFB_REQUIRE(chunk_elems >= 8 && (chunk_elems & 7) == 0 &&
                  (N + chunk_elems - 1) / chunk_elems <= kMaxChunks,
              NFB_ERANGE, "fedavg_allreduce: chunk size must be a multiple of 8 giving <= %d chunks",
              kMaxChunks);
  AllreduceArgs a;
  a.G = G; a.K = static_cast<int>(K); a.x_dtype = x_dtype;
  a.N = N; a.row_stride = row_stride;
  a.clip_norm = clip_norm; a.noise_std = noise_std; a.noise_mask = noise_mask;
  a.seed = seed; a.group0 = offset >> 2;
  a.epoch = epoch;
  a.producers = producers;
  a.chunk_elems = chunk_elems;
  a.n_chunks = static_cast<int>((N + chunk_elems - 1) / chunk_elems);
  a.spin_limit_ns = spin_limit_ns ? spin_limit_ns : 5000000000ull;
  a.mc_partial = mc_partial; a.mc_master = mc_master;
  a.mc_lowp = static_cast<__nv_bfloat16*>(mc_lowp);
  a.status = status;
  a.accumulate = (flags & 1) ? 1 : 0;
  a.local_only = (flags & 2) ? 1 : 0;
  a.my_stats = my_stats;
  a.stats_len = my_stats ? stats_len : 0;
  a.dbg = dbg;
  a.bc = nfb_bcast_target{};
  if (bcast && !a.local_only) a.bc = *static_cast<const nfb_bcast_target*>(bcast);
  const RankView* views = static_cast<const RankView*>(views_dev);
  cudaStream_t st = as_stream(stream);
  return dispatch_dtype(x_dtype, [&](auto tag) {
    using T = decltype(tag);
    return launch_allreduce<T>(views, my_rank, ranks_in_grid, grid_blocks, a, st);
  });
}
