// Shared device/host helpers for the nanofed_b200 kernels (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>

#include "../include/nanofed_b200.h"

namespace nfb {

// ---- error reporting ---------------------------------------------------------
char* error_buffer();
int fail(int code, const char* fmt, ...);
int launch_status(const char* what);

#define NFB_REQUIRE(cond, code, ...)            \
  do {                                          \
    if (!(cond)) return ::nfb::fail((code), __VA_ARGS__); \
  } while (0)

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

int sm_count();

// ---- programmatic dependent launch ------------------------------------------------------------
// The kernels of a local step form one dependency chain inside a CUDA graph; between two of them
// the GPU idles for the launch latency plus the next kernel's prologue (barrier init, TMEM
// allocation, tensor-map fetch: ~3 us, 98 times per ResNet step).  With the attribute below a
// kernel's CTAs may become resident while the previous kernel's last wave is still running; every
// such kernel calls pdl_trigger() first thing (lets ITS successor do the same) and pdl_wait()
// before its first access to global memory (returns when the predecessor grid has completed and
// its writes are visible).  Both instructions are no-ops in a launch without the attribute.
// MEASURED on the ResNet-18 step (98 kernels in one graph): 138.3 k client-samples/s with the
// attribute, 141.9 k without - the graph's own kernel-to-kernel gaps are already shorter than what
// the early-resident CTAs cost the predecessor's last wave.  Off unless NFB_PDL=1.
__device__ __forceinline__ void pdl_trigger() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

bool pdl_enabled();   // NFB_PDL=1 switches it on

template <typename... Params, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(Params...), dim3 grid, dim3 block, size_t smem,
                              cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<Params>(args)...);
}

// ---- dtype tags ---------------------------------------------------------------
template <int DT> struct DType;
template <> struct DType<NFB_DT_F32> { using type = float; };
template <> struct DType<NFB_DT_BF16> { using type = __nv_bfloat16; };
template <> struct DType<NFB_DT_F16> { using type = __half; };

__device__ __forceinline__ float to_f32(float v) { return v; }
__device__ __forceinline__ float to_f32(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float to_f32(__half v) { return __half2float(v); }

template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }

// ---- 16-byte vectors -----------------------------------------------------------
// One 16-byte global access = 4 fp32 or 8 bf16/fp16 elements.
template <typename T> struct Vec16 { static constexpr int N = 16 / sizeof(T); };

// Streaming load: read-once data should not displace reusable lines in L1.
__device__ __forceinline__ uint4 ld_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ld_plain(const void* p) { return *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ void st_vec(void* p, uint4 v) { *reinterpret_cast<uint4*>(p) = v; }

// Unpack a 16-byte register quad into Vec16<T>::N floats.
template <typename T> __device__ __forceinline__ void unpack(const uint4& raw, float* f);
template <> __device__ __forceinline__ void unpack<float>(const uint4& raw, float* f) {
  f[0] = __uint_as_float(raw.x); f[1] = __uint_as_float(raw.y);
  f[2] = __uint_as_float(raw.z); f[3] = __uint_as_float(raw.w);
}
template <> __device__ __forceinline__ void unpack<__nv_bfloat16>(const uint4& raw, float* f) {
  // bf16 -> fp32 is a 16-bit left shift.
  f[0] = __uint_as_float(raw.x << 16); f[1] = __uint_as_float(raw.x & 0xffff0000u);
  f[2] = __uint_as_float(raw.y << 16); f[3] = __uint_as_float(raw.y & 0xffff0000u);
  f[4] = __uint_as_float(raw.z << 16); f[5] = __uint_as_float(raw.z & 0xffff0000u);
  f[6] = __uint_as_float(raw.w << 16); f[7] = __uint_as_float(raw.w & 0xffff0000u);
}
template <> __device__ __forceinline__ void unpack<__half>(const uint4& raw, float* f) {
  const __half2* h = reinterpret_cast<const __half2*>(&raw);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 v = __half22float2(h[i]);
    f[2 * i] = v.x; f[2 * i + 1] = v.y;
  }
}

template <typename T> __device__ __forceinline__ uint4 pack(const float* f);
template <> __device__ __forceinline__ uint4 pack<float>(const float* f) {
  return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]),
                    __float_as_uint(f[3]));
}
template <> __device__ __forceinline__ uint4 pack<__nv_bfloat16>(const float* f) {
  uint4 r;
  __nv_bfloat162 a = __floats2bfloat162_rn(f[0], f[1]);
  __nv_bfloat162 b = __floats2bfloat162_rn(f[2], f[3]);
  __nv_bfloat162 c = __floats2bfloat162_rn(f[4], f[5]);
  __nv_bfloat162 d = __floats2bfloat162_rn(f[6], f[7]);
  r.x = *reinterpret_cast<uint32_t*>(&a); r.y = *reinterpret_cast<uint32_t*>(&b);
  r.z = *reinterpret_cast<uint32_t*>(&c); r.w = *reinterpret_cast<uint32_t*>(&d);
  return r;
}
template <> __device__ __forceinline__ uint4 pack<__half>(const float* f) {
  uint4 r;
  __half2 a = __floats2half2_rn(f[0], f[1]);
  __half2 b = __floats2half2_rn(f[2], f[3]);
  __half2 c = __floats2half2_rn(f[4], f[5]);
  __half2 d = __floats2half2_rn(f[6], f[7]);
  r.x = *reinterpret_cast<uint32_t*>(&a); r.y = *reinterpret_cast<uint32_t*>(&b);
  r.z = *reinterpret_cast<uint32_t*>(&c); r.w = *reinterpret_cast<uint32_t*>(&d);
  return r;
}

// ---- reductions ------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum; result valid in every thread.  `smem` needs 33 floats.
__device__ __forceinline__ float block_sum(float v, float* smem) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  const int nwarps = (blockDim.x + 31) >> 5;
  float t = (threadIdx.x < nwarps) ? smem[threadIdx.x] : 0.f;
  if (warp == 0) {
    t = warp_sum(t);
    if (lane == 0) smem[32] = t;
  }
  __syncthreads();
  t = smem[32];
  __syncthreads();
  return t;
}

__device__ __forceinline__ bool finite_f(float v) { return fabsf(v) <= 3.402823466e38f; }

template <typename F> int dispatch_dtype(int dt, F&& f) {
  switch (dt) {
    case NFB_DT_F32: return f(float{});
    case NFB_DT_BF16: return f(__nv_bfloat16{});
    case NFB_DT_F16: return f(__half{});
    default: return fail(NFB_EINVAL, "unknown dtype code %d", dt);
  }
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

}  // namespace nfb
