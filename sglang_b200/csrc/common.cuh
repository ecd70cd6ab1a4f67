// Shared device/host helpers for libb200attn (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <utility>
#include <cstdarg>
#include <cstdio>

#include "../../include/b200attn.h"

namespace b200 {

// ---------------------------------------------------------------- host side
void set_error(const char* fmt, ...);

#define B200_CHECK_ARG(cond, ...)        \
  do {                                   \
    if (!(cond)) {                       \
      ::b200::set_error(__VA_ARGS__);    \
      return B200_EINVAL;                \
    }                                    \
  } while (0)

#define B200_CHECK_CUDA(expr)                                                          \
  do {                                                                                 \
    cudaError_t _e = (expr);                                                           \
    if (_e != cudaSuccess) {                                                           \
      ::b200::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                        __LINE__);                                                     \
      return B200_ECUDA;                                                               \
    }                                                                                  \
  } while (0)

int sm_count();

// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-DEVICE attribute: remember it per device (bit d of `done`),
// thread-safe, so a process that drives several GPUs configures every one of them.
template <typename Kern>
inline cudaError_t set_max_dynamic_smem(Kern kern, int bytes, std::atomic<uint64_t>& done) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const uint64_t bit = 1ull << (dev & 63);
  if (done.load(std::memory_order_acquire) & bit) return cudaSuccess;
  e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess) done.fetch_or(bit, std::memory_order_release);
  return e;
}

// Programmatic dependent launch (PDL): the kernels of the per-layer decode chain (store -> attention -> merge) are
// launched with cudaLaunchAttributeProgrammaticStreamSerialization, call pdl_launch_dependents() first thing and
// pdl_wait() before they touch global memory.  The next kernel's CTAs are then resident (launch latency paid) while
// the previous grid still runs; griddepcontrol.wait returns once every prerequisite grid has completed and its
// writes are visible, so the data flow is the stream order's.  Launched without the attribute both are no-ops.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}

// -------------------------------------------------------------- device side
constexpr int kWarp = 32;

__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

struct e4m3_t { uint8_t x; };
struct e5m2_t { uint8_t x; };

template <typename T> struct dtype_of;
template <> struct dtype_of<__nv_bfloat16> { static constexpr int value = B200_BF16; };
template <> struct dtype_of<__half> { static constexpr int value = B200_F16; };
template <> struct dtype_of<e4m3_t> { static constexpr int value = B200_FP8_E4M3; };
template <> struct dtype_of<e5m2_t> { static constexpr int value = B200_FP8_E5M2; };

__device__ __forceinline__ float to_float(__nv_bfloat16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float to_float(__half v) { return __half2float(v); }
__device__ __forceinline__ float to_float(e4m3_t v) {
  // e4m3 -> f16 is exact (hardware cvt), f16 -> f32 is exact.
  uint32_t h2;
  uint16_t in = v.x;
  asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"(in));
  return __half2float(__ushort_as_half((unsigned short)(h2 & 0xffffu)));
}
__device__ __forceinline__ float to_float(e5m2_t v) {
  // e5m2 is the top byte of an f16.
  return __half2float(__ushort_as_half((unsigned short)((uint16_t)v.x << 8)));
}

template <typename T> __device__ __forceinline__ T from_float(float f);
template <> __device__ __forceinline__ __nv_bfloat16 from_float<__nv_bfloat16>(float f) {
  return __float2bfloat16_rn(f);
}
template <> __device__ __forceinline__ __half from_float<__half>(float f) { return __float2half_rn(f); }

// torch's float -> float8_e4m3fn: round-to-nearest-even, NOT saturating.
// Everything that would round to the 0x7f encoding (|f| > 464) or is NaN/Inf
// becomes NaN (0x7f | sign); 464 itself ties to even = 448.
__device__ __forceinline__ uint8_t float_to_e4m3_bits(float f) {
  uint32_t b = __float_as_uint(f);
  const uint32_t sign = b & 0x80000000u;
  b ^= sign;
  uint8_t r;
  if (b >= 0x43F00000u) {  // >= 480, inf, nan
    r = 0x7f;
  } else if (b < 0x3C800000u) {  // < 2^-6: subnormal target, add the magic 2^14 .. bias trick
    const float denorm = __uint_as_float(141u << 23);
    uint32_t t = __float_as_uint(__uint_as_float(b) + denorm);
    r = (uint8_t)(t - (141u << 23));
  } else {
    const uint32_t odd = (b >> 20) & 1u;
    b += ((uint32_t)(7 - 127) << 23) + 0x7FFFFu;
    b += odd;
    r = (uint8_t)(b >> 20);
  }
  return r | (uint8_t)(sign >> 24);
}

// torch's float -> float8_e5m2 (RNE; overflow -> inf, NaN stays NaN).
__device__ __forceinline__ uint8_t float_to_e5m2_bits(float f) {
  uint32_t b = __float_as_uint(f);
  const uint32_t sign = b & 0x80000000u;
  b ^= sign;
  uint8_t r;
  if (b >= (143u << 23)) {  // >= 65536
    r = b > (255u << 23) ? 0x7f : 0x7c;
  } else if (b < (113u << 23)) {  // < 2^-14
    const float denorm = __uint_as_float(134u << 23);
    uint32_t t = __float_as_uint(__uint_as_float(b) + denorm);
    r = (uint8_t)(t - (134u << 23));
  } else {
    const uint32_t odd = (b >> 21) & 1u;
    b += ((uint32_t)(15 - 127) << 23) + 0xFFFFFu;
    b += odd;
    r = (uint8_t)(b >> 21);
  }
  return r | (uint8_t)(sign >> 24);
}

template <> __device__ __forceinline__ e4m3_t from_float<e4m3_t>(float f) {
  return e4m3_t{float_to_e4m3_bits(f)};
}
template <> __device__ __forceinline__ e5m2_t from_float<e5m2_t>(float f) {
  return e5m2_t{float_to_e5m2_bits(f)};
}

__device__ __forceinline__ int64_t load_index(const void* p, int is_i64, int64_t i) {
  return is_i64 ? reinterpret_cast<const int64_t*>(p)[i]
                : (int64_t) reinterpret_cast<const int32_t*>(p)[i];
}

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

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fast_tanh(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// tanh from one ex2 and one rcp (both MUFU): absolute error ~1e-6, unlike tanh.approx (2^-11 relative), which
// at logit_cap = 30..50 would shift the capped logits by up to 0.02
__device__ __forceinline__ float tanh_ex2(float x) {
  const float e = fast_exp2(x * (2.f * 1.4426950408889634f));
  return 1.f - __fdividef(2.f, e + 1.f);
}

// two floats -> packed 16-bit pair (lo in the low half)
template <typename CT> __device__ __forceinline__ uint32_t pack2(float lo, float hi);
template <> __device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
template <> __device__ __forceinline__ uint32_t pack2<__half>(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

// plan buffer layout shared by the planner kernels and their consumers
constexpr int kPlanHeader = 8;
constexpr int kPlanPageRows = 4;   // extend plan: rows per verified page box (0 = page table not verified)

}  // namespace b200
