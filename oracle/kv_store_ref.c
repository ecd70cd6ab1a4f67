/*
 * ORACLE (test infrastructure, never shipped or timed as the product):
 * plain-C restatement of the KV-store byte work of the reference's CPU path.
 *
 *   ref_store_kv      <-  MHATokenToKVPool.set_kv_buffer cast/scale/view rules
 *                         (python/sglang/srt/mem_cache/memory_pool.py:2339-2349) and the
 *                         `k_cache[indices] = k` scatter (memory_pool.py:187-192), with the
 *                         optional reserved-slot skip of store_cache
 *                         (kernels/ops/kvcache/kvcache.py:81-84, kvcache.cuh:215).
 *   ref_f32_to_e4m3 / ref_f32_to_e5m2
 *                     <-  the dtype cast `cache_k.to(torch.float8_e4m3fn)` executes.  That cast
 *                         lives in PyTorch (third-party, torch 2.11.0, c10/util/Float8_e4m3fn.h
 *                         and Float8_e5m2.h): round-to-nearest-even on the fp32 bit pattern,
 *                         no saturation -- e4m3: |x| >= 480 or NaN -> 0x7f, results that round
 *                         to the 0x7f code are NaN as well.  Restated from the published
 *                         algorithm and pinned exhaustively against torch itself for all 65536
 *                         bf16 and f16 inputs (tests/golden/fp8_cast_tables.npz, made by
 *                         oracle/gen_golden.py).
 *
 * Build: gcc -O2 -fPIC -shared -o oracle/_build/liboracle_kv.so oracle/kv_store_ref.c
 */
#include <stdint.h>
#include <string.h>

enum { REF_BF16 = 0, REF_F16 = 1, REF_E4M3 = 2, REF_E5M2 = 3 };

static uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

float ref_bf16_to_f32(uint16_t b) { return u2f((uint32_t)b << 16); }

uint16_t ref_f32_to_bf16(float f) {
  uint32_t u = f2u(f);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc0; /* NaN */
  uint32_t lsb = (u >> 16) & 1u;
  u += 0x7fffu + lsb;
  return (uint16_t)(u >> 16);
}

float ref_f16_to_f32(uint16_t h) {
  uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
  uint32_t exp = (h >> 10) & 0x1fu;
  uint32_t man = h & 0x3ffu;
  if (exp == 0) {
    if (man == 0) return u2f(sign);
    /* subnormal: value = man * 2^-24 */
    float v = (float)man * (1.0f / 16777216.0f);
    return u2f(f2u(v) | sign);
  }
  if (exp == 31) return u2f(sign | 0x7f800000u | (man << 13));
  return u2f(sign | ((exp + 112u) << 23) | (man << 13));
}

uint16_t ref_f32_to_f16(float f) {
  uint32_t u = f2u(f);
  uint16_t sign = (uint16_t)((u >> 16) & 0x8000u);
  u &= 0x7fffffffu;
  if (u > 0x7f800000u) return sign | 0x7e00; /* NaN */
  if (u >= 0x477ff000u) return sign | 0x7c00; /* >= 65520 rounds to inf */
  if (u < 0x38800000u) {                      /* < 2^-14: subnormal half */
    /* add 0.5f so the fp32 adder performs the RNE rounding at the 2^-24 grid */
    float v = u2f(u) + 0.5f;
    return sign | (uint16_t)(f2u(v) - f2u(0.5f));
  }
  uint32_t lsb = (u >> 13) & 1u;
  u += ((uint32_t)(15 - 127) << 23) + 0xfffu + lsb;
  return sign | (uint16_t)(u >> 13);
}

uint8_t ref_f32_to_e4m3(float f) {
  uint32_t b = f2u(f);
  const uint32_t sign = b & 0x80000000u;
  uint8_t r;
  b ^= sign;
  if (b >= 0x43f00000u) {               /* |f| >= 480, inf, NaN */
    r = 0x7f;
  } else if (b < (121u << 23)) {        /* below the smallest normal 2^-6 */
    const uint32_t magic = 141u << 23;
    r = (uint8_t)(f2u(u2f(b) + u2f(magic)) - magic);
  } else {
    const uint32_t odd = (b >> 20) & 1u;
    b += ((uint32_t)(7 - 127) << 23) + 0x7ffffu + odd;
    r = (uint8_t)(b >> 20);
  }
  return (uint8_t)(r | (sign >> 24));
}

uint8_t ref_f32_to_e5m2(float f) {
  uint32_t b = f2u(f);
  const uint32_t sign = b & 0x80000000u;
  uint8_t r;
  b ^= sign;
  if (b >= (143u << 23)) {              /* |f| >= 65536 */
    r = b > (255u << 23) ? 0x7f : 0x7c;
  } else if (b < (113u << 23)) {        /* below 2^-14 */
    const uint32_t magic = 134u << 23;
    r = (uint8_t)(f2u(u2f(b) + u2f(magic)) - magic);
  } else {
    const uint32_t odd = (b >> 21) & 1u;
    b += ((uint32_t)(15 - 127) << 23) + 0xfffffu + odd;
    r = (uint8_t)(b >> 21);
  }
  return (uint8_t)(r | (sign >> 24));
}

static float load_src(const void* p, int64_t i, int dtype) {
  const uint16_t bits = ((const uint16_t*)p)[i];
  return dtype == REF_BF16 ? ref_bf16_to_f32(bits) : ref_f16_to_f32(bits);
}
static float round_src(float f, int dtype) {
  return dtype == REF_BF16 ? ref_bf16_to_f32(ref_f32_to_bf16(f)) : ref_f16_to_f32(ref_f32_to_f16(f));
}

static void store_row(const void* src, void* dst, int64_t n, int src_dtype, int dst_dtype, float scale) {
  for (int64_t i = 0; i < n; ++i) {
    if (src_dtype == dst_dtype) { /* same dtype: byte copy, the pool applies no scale */
      ((uint16_t*)dst)[i] = ((const uint16_t*)src)[i];
      continue;
    }
    float f = load_src(src, i, src_dtype);
    if (scale != 1.0f) f = round_src(f / scale, src_dtype); /* cache_k.div_(k_scale) in the source dtype */
    switch (dst_dtype) {
      case REF_BF16: ((uint16_t*)dst)[i] = ref_f32_to_bf16(f); break;
      case REF_F16: ((uint16_t*)dst)[i] = ref_f32_to_f16(f); break;
      case REF_E4M3: ((uint8_t*)dst)[i] = ref_f32_to_e4m3(f); break;
      default: ((uint8_t*)dst)[i] = ref_f32_to_e5m2(f); break;
    }
  }
}

/* returns the number of rows written, or -(i+1) when loc[i] is out of range */
int64_t ref_store_kv(const void* k, const void* v, int64_t k_src_stride, int64_t v_src_stride, void* k_cache,
                     void* v_cache, int64_t k_cache_stride, int64_t v_cache_stride, const int64_t* loc,
                     int64_t num_tokens, int64_t k_row_elems, int64_t v_row_elems, int src_dtype, int dst_dtype,
                     float k_scale, float v_scale, int64_t size_limit, int64_t skip_index) {
  const int64_t ssz = 2, dsz = (dst_dtype == REF_E4M3 || dst_dtype == REF_E5M2) ? 1 : 2;
  int64_t written = 0;
  for (int64_t i = 0; i < num_tokens; ++i) {
    const int64_t slot = loc[i];
    if (slot == skip_index) continue;
    if (slot < 0 || slot >= size_limit) return -(i + 1);
    store_row((const char*)k + i * k_src_stride * ssz, (char*)k_cache + slot * k_cache_stride * dsz, k_row_elems,
              src_dtype, dst_dtype, k_scale);
    store_row((const char*)v + i * v_src_stride * ssz, (char*)v_cache + slot * v_cache_stride * dsz, v_row_elems,
              src_dtype, dst_dtype, v_scale);
    ++written;
  }
  return written;
}

/* table helpers for the exhaustive pin: out[i] = cast(src16 bit pattern i) */
void ref_cast_table(int src_dtype, int dst_dtype, uint8_t* out) {
  for (uint32_t i = 0; i < 65536u; ++i) {
    const uint16_t bits = (uint16_t)i;
    const float f = load_src(&bits, 0, src_dtype);
    out[i] = dst_dtype == REF_E4M3 ? ref_f32_to_e4m3(f) : ref_f32_to_e5m2(f);
  }
}
