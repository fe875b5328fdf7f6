// K-STATS: deterministic sum-of-squares / non-finite reductions.
//
// Shape of every kernel here: `rows` independent reductions, each split over
// SLICES thread blocks.  A block writes its partial into scratch, bumps the row's
// arrival counter, and the LAST block to arrive adds the SLICES partials in index
// order — so the result does not depend on block scheduling (no float atomics) —
// then zeroes the counter again so the scratch can be reused by the next launch.
//
// Scratch layout depends only on the BUFFER SIZE, never on the call's row count, so a
// buffer stays valid ("arrived[] all zero") across calls with different shapes:
//   cap = bytes / (4 + SLICES*8) rows ;  unsigned arrived[cap] | float partial[cap][SLICES]
//   | int bad[cap][SLICES].   Always pass the same (pointer, size) pair for one buffer.
#include <mutex>

#include "common.cuh"

namespace nfb {

constexpr int kSlices = 32;      // blocks per reduction row
constexpr int kThreads = 256;

struct Scratch {
  float* partial;
  int* bad;
  unsigned* arrived;
};

static size_t scratch_bytes(int64_t rows) {
  return static_cast<size_t>(rows) * kSlices * (sizeof(float) + sizeof(int)) +
         static_cast<size_t>(rows) * sizeof(unsigned);
}

static int64_t scratch_capacity(size_t bytes) {
  return static_cast<int64_t>(bytes / (sizeof(unsigned) + kSlices * (sizeof(float) + sizeof(int))));
}

static Scratch carve(void* base, size_t bytes) {
  const int64_t cap = scratch_capacity(bytes);
  Scratch s;
  s.arrived = static_cast<unsigned*>(base);
  s.partial = reinterpret_cast<float*>(s.arrived + cap);
  s.bad = reinterpret_cast<int*>(s.partial + cap * kSlices);
  return s;
}

// Sum of squares of `len` elements at `p` handled by slice `slice` of kSlices.
// Vector body over the 16-byte-aligned middle, scalar head/tail.
template <typename T>
__device__ __forceinline__ void slice_sumsq(const T* __restrict__ p, int64_t len, int slice,
                                            float& acc, int& bad) {
  constexpr int V = Vec16<T>::N;
  const uintptr_t addr = reinterpret_cast<uintptr_t>(p);
  int64_t head = ((16 - (addr & 15)) & 15) / sizeof(T);
  if (head > len) head = len;
  const int64_t nvec = (len - head) / V;
  const int64_t tail0 = head + nvec * V;

  // scalar head + tail belong to slice 0
  if (slice == 0) {
    for (int64_t i = threadIdx.x; i < head; i += blockDim.x) {
      const float v = to_f32(p[i]);
      acc += v * v;
      bad |= !finite_f(v);
    }
    for (int64_t i = tail0 + threadIdx.x; i < len; i += blockDim.x) {
      const float v = to_f32(p[i]);
      acc += v * v;
      bad |= !finite_f(v);
    }
  }
  const uint4* vp = reinterpret_cast<const uint4*>(p + head);
  const int64_t stride = static_cast<int64_t>(kSlices) * blockDim.x;
  int64_t i = static_cast<int64_t>(slice) * blockDim.x + threadIdx.x;
  // 8 independent 16-byte loads in flight per thread (few rows => few CTAs: the bytes in
  // flight per SM must come from the unroll depth)
  for (; i + 7 * stride < nvec; i += 8 * stride) {
    uint4 r[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) r[u] = ld_stream(vp + i + u * stride);
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      float f[V];
      unpack<T>(r[u], f);
#pragma unroll
      for (int e = 0; e < V; ++e) {
        acc += f[e] * f[e];
        bad |= !finite_f(f[e]);
      }
    }
  }
  for (; i < nvec; i += stride) {
    float f[V];
    unpack<T>(ld_stream(vp + i), f);
#pragma unroll
    for (int e = 0; e < V; ++e) {
      acc += f[e] * f[e];
      bad |= !finite_f(f[e]);
    }
  }
}

// Publish this block's partial; the last block of the row finalises in fixed order.
__device__ __forceinline__ void finish_row(float acc, int bad, int64_t row, int slice,
                                           Scratch s, float* __restrict__ out_sum,
                                           int* __restrict__ out_bad, float* smem) {
  const float total = block_sum(acc, smem);
  const int any_bad = __syncthreads_or(bad);
  __shared__ bool is_last;
  if (threadIdx.x == 0) {
    s.partial[row * kSlices + slice] = total;
    s.bad[row * kSlices + slice] = any_bad;
    __threadfence();
    const unsigned ticket = atomicAdd(&s.arrived[row], 1u);
    is_last = (ticket == kSlices - 1);
  }
  __syncthreads();
  if (is_last && threadIdx.x < 32) {
    __threadfence();
    // kSlices == 32: one partial per lane, summed by a fixed shuffle tree.
    const volatile float* part = s.partial + row * kSlices;
    const volatile int* pb = s.bad + row * kSlices;
    float v = part[threadIdx.x];
    int b = pb[threadIdx.x];
    v = warp_sum(v);
    b = __any_sync(0xffffffffu, b != 0);
    if (threadIdx.x == 0) {
      out_sum[row] = v;
      if (out_bad) out_bad[row] = b ? 1 : 0;
      s.arrived[row] = 0u;  // self-clean for the next launch
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads) rows_sumsq_kernel(const T* __restrict__ x,
                                                              int64_t n, int64_t row_stride,
                                                              float* __restrict__ sumsq,
                                                              int* __restrict__ bad, Scratch s) {
  __shared__ float smem[33];
  const int64_t row = blockIdx.y;
  float acc = 0.f;
  int b = 0;
  slice_sumsq<T>(x + row * row_stride, n, blockIdx.x, acc, b);
  finish_row(acc, b, row, blockIdx.x, s, sumsq, bad, smem);
}

template <typename T>
__global__ void __launch_bounds__(kThreads) table_sumsq_kernel(const void* const* __restrict__ ptrs,
                                                               const int64_t* __restrict__ lens,
                                                               float* __restrict__ sumsq,
                                                               int* __restrict__ bad, Scratch s) {
  __shared__ float smem[33];
  const int64_t t = blockIdx.y;
  float acc = 0.f;
  int b = 0;
  slice_sumsq<T>(static_cast<const T*>(ptrs[t]), lens[t], blockIdx.x, acc, b);
  finish_row(acc, b, t, blockIdx.x, s, sumsq, bad, smem);
}

template <typename T>
__global__ void __launch_bounds__(kThreads) segment_sumsq_kernel(const T* __restrict__ x,
                                                                 int64_t row_stride,
                                                                 const int64_t* __restrict__ segs,
                                                                 int S, float* __restrict__ out,
                                                                 Scratch s) {
  __shared__ float smem[33];
  const int64_t k = blockIdx.z, seg = blockIdx.y;
  float acc = 0.f;
  int b = 0;
  slice_sumsq<T>(x + k * row_stride + segs[2 * seg], segs[2 * seg + 1], blockIdx.x, acc, b);
  finish_row(acc, b, k * S + seg, blockIdx.x, s, out, nullptr, smem);
}

}  // namespace nfb

extern "C" size_t nfb_stats_scratch_bytes(int64_t rows) {
  return nfb::scratch_bytes(rows < 1 ? 1 : rows);
}

extern "C" int nfb_update_stats(const void* x, int x_dtype, int64_t K, int64_t N,
                                int64_t row_stride, float* sumsq, int* nonfinite, void* scratch,
                                size_t scratch_bytes, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(x && sumsq && scratch, NFB_EINVAL, "update_stats: null pointer");
  NFB_REQUIRE(K >= 1 && K <= 65535, NFB_ERANGE, "update_stats: K=%lld outside [1, 65535]",
              static_cast<long long>(K));
  NFB_REQUIRE(N >= 0 && row_stride >= N, NFB_EINVAL, "update_stats: bad N/row_stride");
  NFB_REQUIRE(scratch_bytes >= nfb::scratch_bytes(K), NFB_ERANGE,
              "update_stats: scratch has %zu bytes, needs %zu", scratch_bytes,
              nfb::scratch_bytes(K));
  const Scratch s = carve(scratch, scratch_bytes);
  const dim3 grid(kSlices, static_cast<unsigned>(K));
  cudaStream_t st = as_stream(stream);
  return dispatch_dtype(x_dtype, [&](auto tag) {
    using T = decltype(tag);
    rows_sumsq_kernel<T><<<grid, kThreads, 0, st>>>(static_cast<const T*>(x), N, row_stride,
                                                    sumsq, nonfinite, s);
    return launch_status("update_stats");
  });
}

extern "C" int nfb_multi_sumsq(const void* const* ptrs, const int64_t* lens, int T_, int dtype,
                               float* sumsq, int* nonfinite, void* scratch, size_t scratch_bytes,
                               void* stream) {
  using namespace nfb;
  NFB_REQUIRE(ptrs && lens && sumsq && scratch, NFB_EINVAL, "multi_sumsq: null pointer");
  NFB_REQUIRE(T_ >= 1 && T_ <= 65535, NFB_ERANGE, "multi_sumsq: T=%d outside [1, 65535]", T_);
  NFB_REQUIRE(scratch_bytes >= nfb::scratch_bytes(T_), NFB_ERANGE,
              "multi_sumsq: scratch has %zu bytes, needs %zu", scratch_bytes,
              nfb::scratch_bytes(T_));
  const Scratch s = carve(scratch, scratch_bytes);
  const dim3 grid(kSlices, static_cast<unsigned>(T_));
  cudaStream_t st = as_stream(stream);
  return dispatch_dtype(dtype, [&](auto tag) {
    using T = decltype(tag);
    table_sumsq_kernel<T><<<grid, kThreads, 0, st>>>(ptrs, lens, sumsq, nonfinite, s);
    return launch_status("multi_sumsq");
  });
}

extern "C" int nfb_segment_sumsq(const void* x, int x_dtype, int64_t K, int64_t row_stride,
                                 const int64_t* segs, int S, float* out, void* scratch,
                                 size_t scratch_bytes, void* stream) {
  using namespace nfb;
  NFB_REQUIRE(x && segs && out && scratch, NFB_EINVAL, "segment_sumsq: null pointer");
  NFB_REQUIRE(K >= 1 && K <= 65535 && S >= 1 && S <= 65535, NFB_ERANGE,
              "segment_sumsq: K or S outside [1, 65535]");
  NFB_REQUIRE(scratch_bytes >= nfb::scratch_bytes(K * S), NFB_ERANGE,
              "segment_sumsq: scratch has %zu bytes, needs %zu", scratch_bytes,
              nfb::scratch_bytes(K * S));
  const Scratch s = carve(scratch, scratch_bytes);
  const dim3 grid(kSlices, static_cast<unsigned>(S), static_cast<unsigned>(K));
  cudaStream_t st = as_stream(stream);
  return dispatch_dtype(x_dtype, [&](auto tag) {
    using T = decltype(tag);
    segment_sumsq_kernel<T><<<grid, kThreads, 0, st>>>(static_cast<const T*>(x), row_stride,
                                                       segs, S, out, s);
    return launch_status("segment_sumsq");
  });
}

// -------------------------------------------------------------------------------------
// nfb_update_stats_static: nfb_update_stats for callers that cannot allocate.
//
// Some hosts of this library have no way to obtain device memory of their own: a plugin
// that is handed input/output buffers by its framework and is forbidden to call the
// allocator, a real-time or safety-reviewed service whose memory is fixed at start-up, a
// process running after its pool has been sealed, or an error/out-of-memory path that must
// still be able to screen an update.  For them the reduction scratch lives here, inside the
// library image, as a fixed array in device memory.
//
//  * Limit: K <= NFB_STATS_STATIC_MAX_ROWS (1024) rows per call; larger K is rejected with
//    NFB_ERANGE.  The array costs 260 KiB of device memory on EVERY GPU that loads the
//    library, from load until the process exits.
//  * In host code the name of a __device__ array is only a stand-in, not a device pointer;
//    handing it to a kernel faults, or on machines where the GPU can address host memory
//    silently uses the wrong storage.  The real address is fetched with
//    cudaGetSymbolAddress once per device, with that device current, and cached.
//  * The FIRST call on a device performs that look-up.  Make that first call outside any
//    CUDA stream capture (e.g. once at start-up); later calls only enqueue work.
//  * One array per GPU is shared by every call.  It keeps whatever the previous call left
//    there, so each call first clears the arrival counters it is about to use (an async
//    memset on the caller's stream).  Two calls that may overlap on the GPU — different
//    streams, or different host threads without ordering — race on it; serialise them.
//  * Calls no cudaMalloc / cudaFree / any allocating routine.
// -------------------------------------------------------------------------------------
namespace nfb {

constexpr int64_t kStaticMaxRows = 1024;
constexpr size_t kStaticScratchBytes =
    static_cast<size_t>(kStaticMaxRows) * (sizeof(unsigned) + kSlices * (sizeof(float) + sizeof(int)));

static __device__ __align__(256) unsigned char g_stats_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
static void* g_static_scratch_ptr[kMaxDevices] = {nullptr};
static std::mutex g_static_scratch_mu;

// Device address of the static scratch for the CURRENT device (resolved once per device).
static int static_scratch_for_current_device(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess)
    return fail(static_cast<int>(e), "update_stats_static: cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices)
    return fail(NFB_ERANGE, "update_stats_static: device ordinal %d not supported (max %d)", dev,
                kMaxDevices - 1);
  std::lock_guard<std::mutex> lock(g_static_scratch_mu);
  if (g_static_scratch_ptr[dev] == nullptr) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_stats_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      cudaGetLastError();
      return fail(e != cudaSuccess ? static_cast<int>(e) : NFB_ESTATE,
                  "update_stats_static: cudaGetSymbolAddress failed on device %d: %s", dev,
                  cudaGetErrorString(e));
    }
    g_static_scratch_ptr[dev] = p;
  }
  *out = g_static_scratch_ptr[dev];
  return 0;
}

}  // namespace nfb

extern "C" int nfb_stats_static_max_rows(void) { return static_cast<int>(nfb::kStaticMaxRows); }

extern "C" int nfb_update_stats_static(const void* x, int x_dtype, int64_t K, int64_t N,
                                       int64_t row_stride, float* sumsq, int* nonfinite,
                                       void* stream) {
  using namespace nfb;
  NFB_REQUIRE(x && sumsq, NFB_EINVAL, "update_stats_static: null pointer");
  NFB_REQUIRE(K >= 1, NFB_EINVAL, "update_stats_static: K must be positive");
  NFB_REQUIRE(K <= kStaticMaxRows, NFB_ERANGE,
              "update_stats_static: K=%lld exceeds the built-in scratch (max %lld rows); use "
              "nfb_update_stats with caller-provided scratch",
              static_cast<long long>(K), static_cast<long long>(kStaticMaxRows));
  NFB_REQUIRE(N >= 0 && row_stride >= N, NFB_EINVAL, "update_stats_static: bad N/row_stride");
  void* scratch = nullptr;
  const int rc = static_scratch_for_current_device(&scratch);
  if (rc != 0) return rc;
  cudaStream_t st = as_stream(stream);
  const Scratch s = carve(scratch, kStaticScratchBytes);
  // the array is shared and keeps the last call's contents: the counters must start at zero
  cudaError_t e = cudaMemsetAsync(s.arrived, 0, static_cast<size_t>(K) * sizeof(unsigned), st);
  if (e != cudaSuccess)
    return fail(static_cast<int>(e), "update_stats_static: clearing counters: %s",
                cudaGetErrorString(e));
  const dim3 grid(kSlices, static_cast<unsigned>(K));
  return dispatch_dtype(x_dtype, [&](auto tag) {
    using T = decltype(tag);
    rows_sumsq_kernel<T><<<grid, kThreads, 0, st>>>(static_cast<const T*>(x), N, row_stride, sumsq,
                                                    nonfinite, s);
    return launch_status("update_stats_static");
  });
}
