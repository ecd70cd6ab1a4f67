// KV-store scatter for the token-granular KV pool (sm_100a).
//
// Replaces: MHATokenToKVPool.set_kv_buffer -> _set_kv_buffer_impl -> store_cache
// (reference python/sglang/srt/mem_cache/memory_pool.py:141-192, 2305-2381;
//  kernels/jit/csrc/elementwise/kvcache.cuh:188-219) and the fused FP8 variant
// (kernels/jit/csrc/attention/fused_fp8_qkv_kv_cache.cuh:58-90).
//
// Layout: one warp moves one 256-element segment of one K row and the matching
// V row segment; every lane owns 8 consecutive elements of each = two 16-byte loads in
// flight and two 16-byte (same dtype) or 8-byte (FP8) stores.  The grid is tokens x segments.
// FP8-e4m3 uses the hardware RNE pack (cvt.rn.satfinite.e4m3x2.f32) and falls back to the scalar
// bit-exact restatement of torch's non-saturating cast only for vectors holding |x| > 464 / Inf / NaN.
// HBM-bound byte work: algorithmic bytes = T * (Ek + Ev) * (src + dst itemsize) + T * 8.
#include "common.cuh"

namespace b200 {
namespace {

constexpr int kStoreWarps = 4;
constexpr int kVec = 8;                       // elements per lane per segment
constexpr int kSegElems = kWarp * kVec;       // 256 elements per warp

struct StoreParams {
  const void* k;
  const void* v;
  void* k_cache;
  void* v_cache;
  const void* loc;
  int64_t k_src_stride, v_src_stride, k_cache_stride, v_cache_stride;
  int64_t num_tokens;
  int64_t size_limit, skip_index;
  int32_t k_row_elems, v_row_elems;
  int32_t k_segs, segs;                        // segments in a K row, warps (segments) per token
  float k_scale, v_scale;
  int32_t loc_is_i64;
};

__device__ __forceinline__ uint4 ld_stream_16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_16(void* p, uint4 v) {
  asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}
__device__ __forceinline__ void st_8(void* p, uint2 v) {
  asm volatile("st.global.v2.u32 [%0], {%1,%2};" ::"l"(p), "r"(v.x), "r"(v.y) : "memory");
}

template <typename SRC> __device__ __forceinline__ float src_bits_to_float(uint16_t b);
template <> __device__ __forceinline__ float src_bits_to_float<__nv_bfloat16>(uint16_t b) {
  return __uint_as_float((uint32_t)b << 16);
}
template <> __device__ __forceinline__ float src_bits_to_float<__half>(uint16_t b) {
  // cvt.f32.f16 canonicalises NaNs (drops the sign); torch keeps it, and the FP8 cast copies it
  if ((b & 0x7fffu) > 0x7c00u) return __uint_as_float(((uint32_t)(b & 0x8000u) << 16) | 0x7fc00000u);
  return __half2float(__ushort_as_half(b));
}
template <typename SRC> __device__ __forceinline__ float round_to_src(float f);
template <> __device__ __forceinline__ float round_to_src<__nv_bfloat16>(float f) {
  return __bfloat162float(__float2bfloat16_rn(f));
}
template <> __device__ __forceinline__ float round_to_src<__half>(float f) {
  return __half2float(__float2half_rn(f));
}

// Convert one source element (as raw 16 bits) to one destination element.
template <typename SRC, typename DST>
__device__ __forceinline__ auto convert_elem(uint16_t bits, float scale, bool scaled) {
  float f = src_bits_to_float<SRC>(bits);
  if (scaled) f = round_to_src<SRC>(__fdiv_rn(f, scale));  // the pool's in-place div_ in src dtype
  return from_float<DST>(f);
}

// two floats -> two e4m3 bytes (hi -> bits 15:8, lo -> bits 7:0), RNE, saturating (the caller has already
// ruled out everything torch's non-saturating cast would turn into NaN, i.e. |x| > 464)
__device__ __forceinline__ uint32_t pack_e4m3x2_sat(float hi, float lo) {
  uint16_t r;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(r) : "f"(hi), "f"(lo));
  return r;
}

// Bit-exact scalar pack of 8 floats to e4m3 for vectors holding |x| > 464, Inf or NaN (never seen with sane K/V).
// Deliberately NOT inlined: eight copies of the scalar cast at every call site made the fused RoPE+store kernel
// 6 500 instructions long and 36 % of its warp-stall samples were instruction fetches (ncu, r1_store_kernels).
// Arguments are passed BY VALUE (ABI registers): an array reference put the caller's f[8] in local memory and cost the
// plain FP8 store 30 % (5.7 -> 3.9 TB/s on cfg5).
__device__ __noinline__ uint32_t e4m3_pack4_exact(float f0, float f1, float f2, float f3) {
  return (uint32_t)float_to_e4m3_bits(f0) | ((uint32_t)float_to_e4m3_bits(f1) << 8) |
         ((uint32_t)float_to_e4m3_bits(f2) << 16) | ((uint32_t)float_to_e4m3_bits(f3) << 24);
}

// Convert + store the 8 elements of one 16-byte source vector (already loaded).
template <typename SRC, typename DST>
__device__ __forceinline__ void finish_vec8(const uint4 raw, DST* dst, float scale, bool scaled) {
  if constexpr (sizeof(DST) == 2 && dtype_of<SRC>::value == dtype_of<DST>::value) {
    st_16(dst, raw);  // byte copy
  } else if constexpr (sizeof(DST) == 1) {
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
    uint32_t out[2] = {0u, 0u};
    if constexpr (dtype_of<DST>::value == B200_FP8_E4M3) {
      // Hardware RNE pack is bit-identical to torch's cast for |x| <= 464 (464 ties to 448 in both); larger
      // magnitudes, Inf and NaN must become NaN | sign, which the saturating instruction cannot produce:
      // those vectors (never seen with sane K/V) take the bit-exact scalar restatement below.
      float f[8];
      uint32_t amax = 0u;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        f[i] = src_bits_to_float<SRC>((uint16_t)(w[i >> 1] >> ((i & 1) * 16)));
        if (scaled) f[i] = round_to_src<SRC>(__fdiv_rn(f[i], scale));
        amax = max(amax, __float_as_uint(f[i]) & 0x7fffffffu);
      }
      if (amax <= 0x43E80000u) {  // 464.0f
        out[0] = pack_e4m3x2_sat(f[1], f[0]) | (pack_e4m3x2_sat(f[3], f[2]) << 16);
        out[1] = pack_e4m3x2_sat(f[5], f[4]) | (pack_e4m3x2_sat(f[7], f[6]) << 16);
      } else {
        out[0] = e4m3_pack4_exact(f[0], f[1], f[2], f[3]);   // out of line: keeps the hot loop small
        out[1] = e4m3_pack4_exact(f[4], f[5], f[6], f[7]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint16_t bits = (uint16_t)(w[i >> 1] >> ((i & 1) * 16));
        const uint32_t byte = convert_elem<SRC, DST>(bits, scale, scaled).x;
        out[i >> 2] |= byte << ((i & 3) * 8);
      }
    }
    st_8(dst, make_uint2(out[0], out[1]));
  } else {
    // 16-bit to a different 16-bit type (bf16 <-> f16)
    const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
    uint32_t out[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      DST lo = convert_elem<SRC, DST>((uint16_t)(w[i] & 0xffffu), scale, scaled);
      DST hi = convert_elem<SRC, DST>((uint16_t)(w[i] >> 16), scale, scaled);
      out[i] = (uint32_t)(*reinterpret_cast<uint16_t*>(&lo)) |
               ((uint32_t)(*reinterpret_cast<uint16_t*>(&hi)) << 16);
    }
    st_16(dst, make_uint4(out[0], out[1], out[2], out[3]));
  }
}

template <typename SRC, typename DST>
__device__ __forceinline__ void move_scalar(const SRC* src, DST* dst, int n, int lane, float scale,
                                            bool scaled) {
  for (int i = lane; i < n; i += kWarp) {
    const uint16_t bits = *reinterpret_cast<const uint16_t*>(src + i);
    if constexpr (sizeof(DST) == 2 && dtype_of<SRC>::value == dtype_of<DST>::value) {
      *reinterpret_cast<uint16_t*>(dst + i) = bits;
    } else {
      dst[i] = convert_elem<SRC, DST>(bits, scale, scaled);
    }
  }
}

// kVecOK: rows, strides and base pointers are all multiples of 8 elements.
// One warp = segment `s` of BOTH the K row and the V row of one token: the two 16-byte loads are issued
// back to back (the slot lookup, needed only for the stores, overlaps them).
template <typename SRC, typename DST, bool kVecOK>
__global__ void __launch_bounds__(kStoreWarps* kWarp) store_kv_kernel(const StoreParams p) {
  pdl_launch_dependents();          // the next kernel of the layer (attention) may take its seats
  pdl_wait();                       // ... while this one waits for whatever ran before it
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * kStoreWarps + (threadIdx.x >> 5);
  const int64_t token = warp / p.segs;
  if (token >= p.num_tokens) return;
  const int s = (int)(warp - token * p.segs);
  const int k_rem = p.k_row_elems - s * kSegElems, v_rem = p.v_row_elems - s * kSegElems;
  const SRC* k_src = reinterpret_cast<const SRC*>(p.k) + token * p.k_src_stride + (int64_t)s * kSegElems;
  const SRC* v_src = reinterpret_cast<const SRC*>(p.v) + token * p.v_src_stride + (int64_t)s * kSegElems;
  uint4 k_raw = make_uint4(0, 0, 0, 0), v_raw = make_uint4(0, 0, 0, 0);
  const bool k_on = lane * kVec < k_rem, v_on = lane * kVec < v_rem;
  if constexpr (kVecOK) {
    if (k_on) k_raw = ld_stream_16(k_src + lane * kVec);
    if (v_on) v_raw = ld_stream_16(v_src + lane * kVec);
  }
  const int64_t slot = load_index(p.loc, p.loc_is_i64, token);
  if (slot == p.skip_index) return;  // reserved sink slot: padded CUDA-graph rows land here
  if (slot < 0 || slot >= p.size_limit) {
    if (lane == 0) printf("b200attn store_kv: slot %lld out of [0, %lld)\n", (long long)slot,
                          (long long)p.size_limit);
    __trap();
  }
  DST* k_dst = reinterpret_cast<DST*>(p.k_cache) + slot * p.k_cache_stride + (int64_t)s * kSegElems;
  DST* v_dst = reinterpret_cast<DST*>(p.v_cache) + slot * p.v_cache_stride + (int64_t)s * kSegElems;
  if constexpr (kVecOK) {
    if (k_on) finish_vec8<SRC, DST>(k_raw, k_dst + lane * kVec, p.k_scale, p.k_scale != 1.0f);
    if (v_on) finish_vec8<SRC, DST>(v_raw, v_dst + lane * kVec, p.v_scale, p.v_scale != 1.0f);
  } else {
    if (k_rem > 0)
      move_scalar<SRC, DST>(k_src, k_dst, k_rem < kSegElems ? k_rem : kSegElems, lane, p.k_scale, p.k_scale != 1.0f);
    if (v_rem > 0)
      move_scalar<SRC, DST>(v_src, v_dst, v_rem < kSegElems ? v_rem : kSegElems, lane, p.v_scale, p.v_scale != 1.0f);
  }
}

template <typename SRC, typename DST>
int launch_store(const b200_store_kv_args& a, cudaStream_t stream) {
  StoreParams p;
  p.k = a.k; p.v = a.v; p.k_cache = a.k_cache; p.v_cache = a.v_cache; p.loc = a.loc;
  p.k_src_stride = a.k_src_stride; p.v_src_stride = a.v_src_stride;
  p.k_cache_stride = a.k_cache_stride; p.v_cache_stride = a.v_cache_stride;
  p.num_tokens = a.num_tokens; p.size_limit = a.size_limit; p.skip_index = a.skip_index;
  p.k_row_elems = a.k_row_elems; p.v_row_elems = a.v_row_elems;
  p.k_segs = (a.k_row_elems + kSegElems - 1) / kSegElems;
  const int v_segs = (a.v_row_elems + kSegElems - 1) / kSegElems;
  p.segs = p.k_segs > v_segs ? p.k_segs : v_segs;          // warps per token: one per 256-element segment
  p.k_scale = a.k_scale; p.v_scale = a.v_scale; p.loc_is_i64 = a.loc_is_i64;
  const int64_t warps = a.num_tokens * p.segs;
  const int64_t blocks = (warps + kStoreWarps - 1) / kStoreWarps;
  B200_CHECK_ARG(blocks < (int64_t)0x7fffffff, "store_kv: too many rows (%lld)", (long long)a.num_tokens);
  auto aligned = [](const void* ptr, int64_t bytes) { return ((uintptr_t)ptr % bytes) == 0; };
  const bool vec_ok = a.k_row_elems % kVec == 0 && a.v_row_elems % kVec == 0 &&
                      a.k_src_stride % kVec == 0 && a.v_src_stride % kVec == 0 &&
                      a.k_cache_stride % kVec == 0 && a.v_cache_stride % kVec == 0 &&
                      aligned(a.k, 16) && aligned(a.v, 16) &&
                      aligned(a.k_cache, sizeof(DST) * kVec) && aligned(a.v_cache, sizeof(DST) * kVec);
  if (vec_ok) {
    B200_CHECK_CUDA(launch_pdl(store_kv_kernel<SRC, DST, true>, dim3((unsigned)blocks), dim3(kStoreWarps * kWarp), 0, stream, p));
  } else {
    B200_CHECK_CUDA(launch_pdl(store_kv_kernel<SRC, DST, false>, dim3((unsigned)blocks), dim3(kStoreWarps * kWarp), 0, stream, p));
  }
  B200_CHECK_CUDA(cudaGetLastError());
  return B200_OK;
}

#include "rope_store.cuh"

}  // namespace
}  // namespace b200

extern "C" int b200attn_store_kv(const b200_store_kv_args* a, b200_stream_t stream_) {
  using namespace b200;
  B200_CHECK_ARG(a != nullptr, "store_kv: null args");
  if (a->num_tokens == 0) return B200_OK;
  B200_CHECK_ARG(a->k && a->v && a->k_cache && a->v_cache && a->loc, "store_kv: null pointer");
  B200_CHECK_ARG(a->num_tokens > 0 && a->k_row_elems > 0 && a->v_row_elems > 0,
                 "store_kv: bad sizes T=%lld Ek=%d Ev=%d", (long long)a->num_tokens, a->k_row_elems,
                 a->v_row_elems);
  B200_CHECK_ARG(a->k_scale > 0.f && a->v_scale > 0.f, "store_kv: scales must be positive");
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  const int s = a->src_dtype, d = a->dst_dtype;
  if (s == B200_BF16 && d == B200_BF16) return launch_store<__nv_bfloat16, __nv_bfloat16>(*a, stream);
  if (s == B200_F16 && d == B200_F16) return launch_store<__half, __half>(*a, stream);
  if (s == B200_BF16 && d == B200_F16) return launch_store<__nv_bfloat16, __half>(*a, stream);
  if (s == B200_F16 && d == B200_BF16) return launch_store<__half, __nv_bfloat16>(*a, stream);
  if (s == B200_BF16 && d == B200_FP8_E4M3) return launch_store<__nv_bfloat16, e4m3_t>(*a, stream);
  if (s == B200_F16 && d == B200_FP8_E4M3) return launch_store<__half, e4m3_t>(*a, stream);
  if (s == B200_BF16 && d == B200_FP8_E5M2) return launch_store<__nv_bfloat16, e5m2_t>(*a, stream);
  if (s == B200_F16 && d == B200_FP8_E5M2) return launch_store<__half, e5m2_t>(*a, stream);
  set_error("store_kv: unsupported dtype pair src=%d dst=%d", s, d);
  return B200_EINVAL;
}

extern "C" int b200attn_rope_store_kv(const b200_rope_store_args* a, b200_stream_t stream_) {
  using namespace b200;
  B200_CHECK_ARG(a != nullptr, "rope_store_kv: null args");
  if (a->num_tokens == 0) return B200_OK;
  B200_CHECK_ARG(a->q && a->k && a->positions && a->cos_sin_cache, "rope_store_kv: null pointer");
  B200_CHECK_ARG(a->k_cache == nullptr || (a->v && a->v_cache && a->loc), "rope_store_kv: null v / v_cache / loc");
  B200_CHECK_ARG(a->num_tokens > 0 && a->num_q_heads > 0 && a->num_kv_heads > 0, "rope_store_kv: bad sizes");
  B200_CHECK_ARG(a->src_dtype == B200_BF16 || a->src_dtype == B200_F16, "rope_store_kv: src dtype %d", a->src_dtype);
  B200_CHECK_ARG(a->k_scale > 0.f && a->v_scale > 0.f, "rope_store_kv: scales must be positive");
  const int D = a->head_dim, Dv = a->v_head_dim, rot = a->rotary_dim;
  B200_CHECK_ARG(D > 0 && D % kVec == 0 && Dv > 0 && Dv % kVec == 0, "rope_store_kv: head dims must be multiples of 8");
  B200_CHECK_ARG(rot > 0 && rot <= D && rot % (a->is_neox ? 2 * kVec : kVec) == 0,
                 "rope_store_kv: rotary_dim %d must be a multiple of %d and <= head_dim", rot,
                 a->is_neox ? 2 * kVec : kVec);
  auto a16 = [](const void* ptr) { return ((uintptr_t)ptr & 15) == 0; };
  B200_CHECK_ARG(a16(a->q) && a16(a->k) && a16(a->v) && a->q_stride_t % kVec == 0 && a->k_stride_t % kVec == 0 &&
                     a->v_stride_t % kVec == 0 && a->k_cache_stride % 16 == 0 && a->v_cache_stride % 16 == 0 &&
                     a16(a->k_cache) && a16(a->v_cache),
                 "rope_store_kv: pointers must be 16-byte aligned and strides multiples of 8 (16 for the pool)");
  RopeParams p;
  p.q = a->q; p.k = a->k; p.v = a->v;
  p.q_stride_t = a->q_stride_t; p.k_stride_t = a->k_stride_t; p.v_stride_t = a->v_stride_t;
  p.positions = a->positions; p.cache = a->cos_sin_cache; p.max_position = a->max_position;
  p.k_cache = a->k_cache; p.v_cache = a->v_cache;
  p.k_cache_stride = a->k_cache_stride; p.v_cache_stride = a->v_cache_stride;
  p.loc = a->loc; p.num_tokens = a->num_tokens; p.size_limit = a->size_limit; p.skip_index = a->skip_index;
  p.positions_is_i64 = a->positions_is_i64; p.loc_is_i64 = a->loc_is_i64;
  p.cache_is_f32 = a->cache_dtype == B200_CACHE_F32;
  p.rot = rot; p.D = D; p.Dv = Dv; p.Hq = a->num_q_heads; p.Hkv = a->num_kv_heads;
  p.is_neox = a->is_neox; p.native = a->rounding == B200_ROPE_NATIVE; p.write_k = a->write_k;
  const int n_rot_units = a->is_neox ? rot / (2 * kVec) : rot / kVec;
  p.uq = p.uk = n_rot_units + (D - rot) / kVec;
  p.uv = a->k_cache ? Dv / kVec : 0;
  const int hpt = rope_heads_per_thread(p.Hq, p.Hkv);
  p.units_per_token = (p.Hq / hpt) * p.uq + (p.Hkv / hpt) * p.uk + p.Hkv * p.uv;   // threads per token
  p.k_scale = a->k_scale; p.v_scale = a->v_scale;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  const int s = a->src_dtype, d = a->k_cache ? a->dst_dtype : a->src_dtype;
  if (s == B200_BF16 && d == B200_BF16) return launch_rope_any<__nv_bfloat16, __nv_bfloat16>(p, stream);
  if (s == B200_F16 && d == B200_F16) return launch_rope_any<__half, __half>(p, stream);
  if (s == B200_BF16 && d == B200_FP8_E4M3) return launch_rope_any<__nv_bfloat16, e4m3_t>(p, stream);
  if (s == B200_F16 && d == B200_FP8_E4M3) return launch_rope_any<__half, e4m3_t>(p, stream);
  if (s == B200_BF16 && d == B200_FP8_E5M2) return launch_rope_any<__nv_bfloat16, e5m2_t>(p, stream);
  if (s == B200_F16 && d == B200_FP8_E5M2) return launch_rope_any<__half, e5m2_t>(p, stream);
  set_error("rope_store_kv: unsupported dtype pair src=%d dst=%d", s, d);
  return B200_EINVAL;
}
