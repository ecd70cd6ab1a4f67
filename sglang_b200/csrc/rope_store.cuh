// Fused rotary embedding + KV-store kernel (included by kv_store.cu inside namespace b200 { namespace {).
//
// Replaces  q, k = rotary_emb(positions, q, k); set_kv_buffer(layer, loc, k, v)  (reference
// srt/layers/rotary_embedding/base.py:234-275,363-385, utils.py:36-60, kernels/ops/kvcache/rope_cache.py)
// with one pass: q, k, v are read once, q (and optionally k) are rewritten in place, the pool rows are
// written with the cast / scale / skip rules of store_kv_kernel.  HBM-bound byte work:
//   bytes = T * [Hq*D*2*2 + Hkv*D*(2 + 2*write_k + s_kv) + Hkv*Dv*(2 + s_kv)] + T * 16.
//
// A thread owns one "unit" of one head of one token: 8 consecutive elements (16 bytes) plus, for the neox
// pairing, the partner 8 elements half a rotary_dim away.  Units of a token are laid out so that the lanes
// of a warp touch consecutive 16-byte pieces.

struct RopeParams {
  void* q;
  void* k;
  const void* v;
  int64_t q_stride_t, k_stride_t, v_stride_t;
  const void* positions;
  const void* cache;
  int64_t max_position;
  void* k_cache;
  void* v_cache;
  int64_t k_cache_stride, v_cache_stride;
  const void* loc;
  int64_t num_tokens, size_limit, skip_index;
  int32_t positions_is_i64, loc_is_i64, cache_is_f32;
  int32_t rot, D, Dv, Hq, Hkv;
  int32_t is_neox, native, write_k;
  int32_t uq, uk, uv;          // units per q head / k head / v head
  int32_t units_per_token;
  float k_scale, v_scale;
};

// Heads one thread rotates per cos/sin load.  Measured on B200 (131072 tokens, Llama-8B heads): 1 -> 4.44 TB/s,
// 4 -> 4.21 TB/s (bf16 pool) and 3.92 vs 2.33 TB/s (FP8 pool): the extra registers cost more than the saved
// L1 hits, so the dispatcher keeps 1; the wider instances stay for shapes where that changes.
inline int rope_heads_per_thread(int hq, int hkv) { (void)hq; (void)hkv; return 1; }

template <typename SRC> __device__ __forceinline__ uint16_t float_to_src_bits(float f);
template <> __device__ __forceinline__ uint16_t float_to_src_bits<__nv_bfloat16>(float f) {
  return __bfloat16_as_ushort(__float2bfloat16_rn(f));
}
template <> __device__ __forceinline__ uint16_t float_to_src_bits<__half>(float f) {
  return __half_as_ushort(__float2half_rn(f));
}

template <typename SRC> __device__ __forceinline__ void unpack8(const uint4 raw, float (&f)[8]);

// cos / sin of pair index i for this token (already offset to the token's cache row)
template <typename SRC>
__device__ __forceinline__ void load_cs8(const RopeParams& p, const void* row, int i0, float (&c)[8], float (&s)[8]) {
  const int half = p.rot >> 1;
  if (p.cache_is_f32) {   // i0 is a multiple of 8 and rows are rot * 4 bytes: 16-byte vector loads
    const float4* rc = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(row) + i0);
    const float4* rs = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(row) + half + i0);
    const float4 c0 = __ldg(rc), c1 = __ldg(rc + 1), s0 = __ldg(rs), s1 = __ldg(rs + 1);
    c[0] = c0.x; c[1] = c0.y; c[2] = c0.z; c[3] = c0.w; c[4] = c1.x; c[5] = c1.y; c[6] = c1.z; c[7] = c1.w;
    s[0] = s0.x; s[1] = s0.y; s[2] = s0.z; s[3] = s0.w; s[4] = s1.x; s[5] = s1.y; s[6] = s1.z; s[7] = s1.w;
  } else {
    float cc[8], ss[8];
    unpack8<SRC>(__ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const SRC*>(row) + i0)), cc);
    unpack8<SRC>(__ldg(reinterpret_cast<const uint4*>(reinterpret_cast<const SRC*>(row) + half + i0)), ss);
#pragma unroll
    for (int j = 0; j < 8; ++j) { c[j] = cc[j]; s[j] = ss[j]; }
  }
  if (p.native) {   // forward_native casts the cache to the activation dtype before multiplying
#pragma unroll
    for (int j = 0; j < 8; ++j) { c[j] = round_to_src<SRC>(c[j]); s[j] = round_to_src<SRC>(s[j]); }
  }
}

// (x1, x2) -> (x1 c - x2 s, x2 c + x1 s)
template <typename SRC>
__device__ __forceinline__ void rotate_pair(bool native, float x1, float x2, float c, float s, float& o1, float& o2) {
  if (native) {   // op-by-op rounding in the model dtype, no contraction
    const float a = round_to_src<SRC>(__fmul_rn(x1, c)), b = round_to_src<SRC>(__fmul_rn(x2, s));
    const float d = round_to_src<SRC>(__fmul_rn(x2, c)), e = round_to_src<SRC>(__fmul_rn(x1, s));
    o1 = __fsub_rn(a, b);
    o2 = __fadd_rn(d, e);
  } else {
    o1 = __fmaf_rn(-x2, s, __fmul_rn(x1, c));
    o2 = __fmaf_rn(x1, s, __fmul_rn(x2, c));
  }
}

__device__ __forceinline__ uint4 ld_16(const void* p) { return *reinterpret_cast<const uint4*>(p); }

template <typename SRC>
__device__ __forceinline__ void unpack8(const uint4 raw, float (&f)[8]) {
  const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
  for (int i = 0; i < 8; ++i) f[i] = src_bits_to_float<SRC>((uint16_t)(w[i >> 1] >> ((i & 1) * 16)));
}
template <typename SRC>
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
    w[i] = (uint32_t)float_to_src_bits<SRC>(f[2 * i]) | ((uint32_t)float_to_src_bits<SRC>(f[2 * i + 1]) << 16);
  return make_uint4(w[0], w[1], w[2], w[3]);
}

// HPT: consecutive heads (of the same tensor) one thread rotates with one load of the cos / sin values;
// all 2 * HPT data loads are issued before the first dependent use.
template <typename SRC, typename DST, int HPT>
__global__ void __launch_bounds__(256) rope_store_kernel(const RopeParams p) {
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t token = gid / p.units_per_token;
  if (token >= p.num_tokens) return;
  int u = (int)(gid - token * p.units_per_token);
  const int half = p.rot >> 1;
  const int n_rot_units = p.is_neox ? half / kVec : p.rot / kVec;

  // ---- which tensor / head / unit ----
  int kind = 0;                                  // 0 = q, 1 = k, 2 = v
  int head, unit;
  const int q_units = (p.Hq / HPT) * p.uq, k_units = (p.Hkv / HPT) * p.uk;
  if (u < q_units) {
    head = u / p.uq; unit = u - head * p.uq; head *= HPT;
  } else {
    u -= q_units;
    if (u < k_units) { kind = 1; head = u / p.uk; unit = u - head * p.uk; head *= HPT; }
    else { u -= k_units; kind = 2; head = u / p.uv; unit = u - head * p.uv; }
  }
  if (kind == 0 && unit >= n_rot_units) return;  // pass-through part of q: nothing to do

  // ---- pool slot (only k / v units store) ----
  int64_t slot = -1;
  bool store = false;
  if (kind != 0 && p.k_cache != nullptr) {
    slot = load_index(p.loc, p.loc_is_i64, token);
    store = slot != p.skip_index;
    if (store && (slot < 0 || slot >= p.size_limit)) {
      printf("b200attn rope_store_kv: slot %lld out of [0, %lld)\n", (long long)slot, (long long)p.size_limit);
      __trap();
    }
  }

  if (kind == 2) {                               // V: plain store of one 16-byte chunk
    if (!store) return;
    const SRC* src = reinterpret_cast<const SRC*>(p.v) + token * p.v_stride_t + (int64_t)head * p.Dv + unit * kVec;
    DST* dst = reinterpret_cast<DST*>(p.v_cache) + slot * p.v_cache_stride + (int64_t)head * p.Dv + unit * kVec;
    finish_vec8<SRC, DST>(ld_stream_16(src), dst, p.v_scale, p.v_scale != 1.0f);
    return;
  }

  SRC* x = (kind == 0 ? reinterpret_cast<SRC*>(p.q) + token * p.q_stride_t
                      : reinterpret_cast<SRC*>(p.k) + token * p.k_stride_t) + (int64_t)head * p.D;
  DST* kdst = store ? reinterpret_cast<DST*>(p.k_cache) + slot * p.k_cache_stride + (int64_t)head * p.D : nullptr;
  const bool scaled = p.k_scale != 1.0f;

  if (unit >= n_rot_units) {                     // pass-through chunk of k: pool store only
    const int e0 = p.rot + (unit - n_rot_units) * kVec;
    if (store) {
#pragma unroll
      for (int h = 0; h < HPT; ++h)
        finish_vec8<SRC, DST>(ld_16(x + h * p.D + e0), kdst + h * p.D + e0, p.k_scale, scaled);
    }
    return;
  }

  const int64_t pos = load_index(p.positions, p.positions_is_i64, token);
  if (pos < 0 || pos >= p.max_position) {
    printf("b200attn rope_store_kv: position %lld out of [0, %lld)\n", (long long)pos, (long long)p.max_position);
    __trap();
  }
  const uint8_t* row = reinterpret_cast<const uint8_t*>(p.cache) + pos * p.rot * (p.cache_is_f32 ? 4 : 2);

  if (p.is_neox) {
    const int e1 = unit * kVec, e2 = half + unit * kVec;   // pair (i, i + rot/2), i = e1 .. e1+7
    uint4 a[HPT], b[HPT];
#pragma unroll
    for (int h = 0; h < HPT; ++h) { a[h] = ld_16(x + h * p.D + e1); b[h] = ld_16(x + h * p.D + e2); }
    float c[8], s[8];
    load_cs8<SRC>(p, row, e1, c, s);
#pragma unroll
    for (int h = 0; h < HPT; ++h) {
      float x1[8], x2[8], o1[8], o2[8];
      unpack8<SRC>(a[h], x1);
      unpack8<SRC>(b[h], x2);
#pragma unroll
      for (int j = 0; j < 8; ++j) rotate_pair<SRC>(p.native != 0, x1[j], x2[j], c[j], s[j], o1[j], o2[j]);
      const uint4 r1 = pack8<SRC>(o1), r2 = pack8<SRC>(o2);
      if (kind == 0 || p.write_k) {
        *reinterpret_cast<uint4*>(x + h * p.D + e1) = r1;
        *reinterpret_cast<uint4*>(x + h * p.D + e2) = r2;
      }
      if (store) {
        finish_vec8<SRC, DST>(r1, kdst + h * p.D + e1, p.k_scale, scaled);
        finish_vec8<SRC, DST>(r2, kdst + h * p.D + e2, p.k_scale, scaled);
      }
    }
  } else {
    const int e0 = unit * kVec;                            // pairs (2i, 2i+1), i = e0/2 .. e0/2+3
    uint4 a[HPT];
#pragma unroll
    for (int h = 0; h < HPT; ++h) a[h] = ld_16(x + h * p.D + e0);
    float c[4], s[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int i = e0 / 2 + j;
      if (p.cache_is_f32) {
        c[j] = reinterpret_cast<const float*>(row)[i]; s[j] = reinterpret_cast<const float*>(row)[half + i];
      } else {
        c[j] = to_float(reinterpret_cast<const SRC*>(row)[i]); s[j] = to_float(reinterpret_cast<const SRC*>(row)[half + i]);
      }
      if (p.native) { c[j] = round_to_src<SRC>(c[j]); s[j] = round_to_src<SRC>(s[j]); }
    }
#pragma unroll
    for (int h = 0; h < HPT; ++h) {
      float xv[8], o[8];
      unpack8<SRC>(a[h], xv);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        rotate_pair<SRC>(p.native != 0, xv[2 * j], xv[2 * j + 1], c[j], s[j], o[2 * j], o[2 * j + 1]);
      const uint4 r = pack8<SRC>(o);
      if (kind == 0 || p.write_k) *reinterpret_cast<uint4*>(x + h * p.D + e0) = r;
      if (store) finish_vec8<SRC, DST>(r, kdst + h * p.D + e0, p.k_scale, scaled);
    }
  }
}

template <typename SRC, typename DST>
int launch_rope_store(const RopeParams& p, cudaStream_t stream) {
  const int64_t threads = p.num_tokens * p.units_per_token;
  const int64_t blocks = (threads + 255) / 256;
  B200_CHECK_ARG(blocks > 0 && blocks < (int64_t)0x7fffffff, "rope_store_kv: too many rows (%lld)",
                 (long long)p.num_tokens);
  const int hpt = rope_heads_per_thread(p.Hq, p.Hkv);
  if (hpt == 4) rope_store_kernel<SRC, DST, 4><<<(unsigned)blocks, 256, 0, stream>>>(p);
  else if (hpt == 2) rope_store_kernel<SRC, DST, 2><<<(unsigned)blocks, 256, 0, stream>>>(p);
  else rope_store_kernel<SRC, DST, 1><<<(unsigned)blocks, 256, 0, stream>>>(p);
  B200_CHECK_CUDA(cudaGetLastError());
  return B200_OK;
}

// ------------------------------------------------------------------------------------------------------------
// Fast path: full rotary (rotary_dim == head_dim == v_head_dim, D in {64, 128, 256}).
//
// The first kernel above spends its time on per-thread index arithmetic (a 64-bit division and several 32-bit
// ones per 32 bytes) and re-reads the cos/sin row once per head.  Here a WARP owns a slice of one token:
//   * lane -> (head within the pass, 16-byte chunk j of the first half of the head): TPH = D/16 lanes per head,
//     HPP = 32/TPH heads per warp pass; the lane also owns chunk j + TPH (the neox partner, or simply the second
//     half of the head for the interleaved layout) -- both loads of a pass hit whole 128-byte lines;
//   * the cos/sin values of the lane's chunk are loaded ONCE per token and reused for every q and k head;
//   * passes of a token: ceil(Hq/HPP) q passes, ceil(Hkv/HPP) k passes, then v passes (plain 2 x 512-byte copies);
//     kBatch passes are loaded before the first dependent use (the in-place q/k stores would otherwise order
//     every load behind the previous store);
//   * `slices` warps share a token when there are few tokens (decode), one warp takes the whole token at prefill.
// ------------------------------------------------------------------------------------------------------------
struct RopeFastParams {
  RopeParams b;
  int32_t nq_pass, nk_pass, nv_pass;   // passes per token
  int32_t slices, pass_per_slice;
};

constexpr int kRopeBatch = 4;
constexpr int kRopeWarps = 4;

// SCALED = false: k_scale == v_scale == 1 (no IEEE divisions compiled in: 128 inlined div.rn sequences made the FP8
// instance 5 000+ instructions long and instruction-fetch bound).
template <typename SRC, typename DST, int D, bool NEOX, bool SCALED>
__global__ void __launch_bounds__(kRopeWarps * 32) rope_store_fast_kernel(const RopeFastParams fp) {
  constexpr int TPH = D / 16, HPP = 32 / TPH;
  const RopeParams& p = fp.b;
  const int lane = threadIdx.x & 31;
  const int64_t item = (int64_t)blockIdx.x * kRopeWarps + (threadIdx.x >> 5);
  const int64_t token = item / fp.slices;
  if (token >= p.num_tokens) return;
  const int slice = (int)(item - token * fp.slices);
  const int n_pass = fp.nq_pass + fp.nk_pass + fp.nv_pass;
  const int p_begin = slice * fp.pass_per_slice;
  const int p_end = min(n_pass, p_begin + fp.pass_per_slice);
  if (p_begin >= p_end) return;
  const int j = lane % TPH, hsub = lane / TPH;

  // ---- pool slot (k / v passes only) ----
  int64_t slot = -1;
  bool store = false;
  if (p.k_cache != nullptr && p_end > fp.nq_pass) {
    slot = load_index(p.loc, p.loc_is_i64, token);
    store = slot != p.skip_index;
    if (store && (slot < 0 || slot >= p.size_limit)) {
      if (lane == 0) printf("b200attn rope_store_kv: slot %lld out of [0, %lld)\n", (long long)slot, (long long)p.size_limit);
      __trap();
    }
  }

  // ---- cos / sin of this lane's chunks, once per token ----
  float c[8], s[8];    // neox: pair index 8j + e.   interleaved: [0,4) pairs of chunk j, [4,8) pairs of chunk j + TPH
  if (p_begin < fp.nq_pass + fp.nk_pass) {
    const int64_t pos = load_index(p.positions, p.positions_is_i64, token);
    if (pos < 0 || pos >= p.max_position) {
      if (lane == 0) printf("b200attn rope_store_kv: position %lld out of [0, %lld)\n", (long long)pos, (long long)p.max_position);
      __trap();
    }
    constexpr int half = D / 2;
    if (p.cache_is_f32) {
      const float* row = reinterpret_cast<const float*>(p.cache) + pos * D;
      if constexpr (NEOX) {
        const float4 c0 = __ldg(reinterpret_cast<const float4*>(row + 8 * j)), c1 = __ldg(reinterpret_cast<const float4*>(row + 8 * j + 4));
        const float4 s0 = __ldg(reinterpret_cast<const float4*>(row + half + 8 * j)), s1 = __ldg(reinterpret_cast<const float4*>(row + half + 8 * j + 4));
        c[0] = c0.x; c[1] = c0.y; c[2] = c0.z; c[3] = c0.w; c[4] = c1.x; c[5] = c1.y; c[6] = c1.z; c[7] = c1.w;
        s[0] = s0.x; s[1] = s0.y; s[2] = s0.z; s[3] = s0.w; s[4] = s1.x; s[5] = s1.y; s[6] = s1.z; s[7] = s1.w;
      } else {
        const float4 c0 = __ldg(reinterpret_cast<const float4*>(row + 4 * j)), c1 = __ldg(reinterpret_cast<const float4*>(row + 4 * (j + TPH)));
        const float4 s0 = __ldg(reinterpret_cast<const float4*>(row + half + 4 * j)), s1 = __ldg(reinterpret_cast<const float4*>(row + half + 4 * (j + TPH)));
        c[0] = c0.x; c[1] = c0.y; c[2] = c0.z; c[3] = c0.w; c[4] = c1.x; c[5] = c1.y; c[6] = c1.z; c[7] = c1.w;
        s[0] = s0.x; s[1] = s0.y; s[2] = s0.z; s[3] = s0.w; s[4] = s1.x; s[5] = s1.y; s[6] = s1.z; s[7] = s1.w;
      }
    } else {
      const SRC* row = reinterpret_cast<const SRC*>(p.cache) + pos * D;
      if constexpr (NEOX) {
        unpack8<SRC>(__ldg(reinterpret_cast<const uint4*>(row + 8 * j)), c);
        unpack8<SRC>(__ldg(reinterpret_cast<const uint4*>(row + half + 8 * j)), s);
      } else {
        const uint2 c0 = __ldg(reinterpret_cast<const uint2*>(row + 4 * j)), c1 = __ldg(reinterpret_cast<const uint2*>(row + 4 * (j + TPH)));
        const uint2 s0 = __ldg(reinterpret_cast<const uint2*>(row + half + 4 * j)), s1 = __ldg(reinterpret_cast<const uint2*>(row + half + 4 * (j + TPH)));
        unpack8<SRC>(make_uint4(c0.x, c0.y, c1.x, c1.y), c);
        unpack8<SRC>(make_uint4(s0.x, s0.y, s1.x, s1.y), s);
      }
    }
    if (p.native) {   // forward_native casts the cache to the activation dtype before multiplying
#pragma unroll
      for (int e = 0; e < 8; ++e) { c[e] = round_to_src<SRC>(c[e]); s[e] = round_to_src<SRC>(s[e]); }
    }
  }

  SRC* qrow = reinterpret_cast<SRC*>(p.q) + token * p.q_stride_t;
  SRC* krow = reinterpret_cast<SRC*>(p.k) + token * p.k_stride_t;
  const SRC* vrow = reinterpret_cast<const SRC*>(p.v) + token * p.v_stride_t;
  DST* kdst = store ? reinterpret_cast<DST*>(p.k_cache) + slot * p.k_cache_stride : nullptr;
  DST* vdst = store ? reinterpret_cast<DST*>(p.v_cache) + slot * p.v_cache_stride : nullptr;
  const bool k_scaled = SCALED && p.k_scale != 1.0f, v_scaled = SCALED && p.v_scale != 1.0f;
  const int v_chunks = p.Hkv * (D / 8);                       // 16-byte chunks of the token's V row

  for (int p0 = p_begin; p0 < p_end; p0 += kRopeBatch) {
    uint4 a[kRopeBatch], b[kRopeBatch];
    int off[kRopeBatch];          // element offset of the lane's first chunk inside the token row; -1 = idle
    // ---- issue every load of the batch ----
#pragma unroll
    for (int u = 0; u < kRopeBatch; ++u) {
      const int ps = p0 + u;
      off[u] = -1;
      a[u] = b[u] = make_uint4(0u, 0u, 0u, 0u);
      if (ps >= p_end) continue;
      if (ps < fp.nq_pass + fp.nk_pass) {
        const bool is_q = ps < fp.nq_pass;
        const int head = (is_q ? ps : ps - fp.nq_pass) * HPP + hsub;
        if (head < (is_q ? p.Hq : p.Hkv)) {
          off[u] = head * D + 8 * j;
          const SRC* src = (is_q ? qrow : krow) + off[u];
          a[u] = ld_16(src);
          b[u] = ld_16(src + D / 2);
        }
      } else if (store) {
        const int ch = (ps - fp.nq_pass - fp.nk_pass) * 64 + lane;      // chunks ch and ch + 32
        if (ch < v_chunks) {
          off[u] = ch * 8;
          a[u] = ld_stream_16(vrow + off[u]);
          if (ch + 32 < v_chunks) b[u] = ld_stream_16(vrow + off[u] + 256);
        }
      }
    }
    // ---- rotate / convert / store ----
#pragma unroll
    for (int u = 0; u < kRopeBatch; ++u) {
      const int ps = p0 + u;
      if (off[u] < 0) continue;
      if (ps >= fp.nq_pass + fp.nk_pass) {                    // V: pool store only
        finish_vec8<SRC, DST>(a[u], vdst + off[u], p.v_scale, v_scaled);
        if (off[u] / 8 + 32 < v_chunks) finish_vec8<SRC, DST>(b[u], vdst + off[u] + 256, p.v_scale, v_scaled);
        continue;
      }
      const bool is_q = ps < fp.nq_pass;
      float x1[8], x2[8], o1[8], o2[8];
      unpack8<SRC>(a[u], x1);
      unpack8<SRC>(b[u], x2);
      if constexpr (NEOX) {
#pragma unroll
        for (int e = 0; e < 8; ++e) rotate_pair<SRC>(p.native != 0, x1[e], x2[e], c[e], s[e], o1[e], o2[e]);
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          rotate_pair<SRC>(p.native != 0, x1[2 * e], x1[2 * e + 1], c[e], s[e], o1[2 * e], o1[2 * e + 1]);
          rotate_pair<SRC>(p.native != 0, x2[2 * e], x2[2 * e + 1], c[4 + e], s[4 + e], o2[2 * e], o2[2 * e + 1]);
        }
      }
      const uint4 r1 = pack8<SRC>(o1), r2 = pack8<SRC>(o2);
      if (is_q || p.write_k) {
        SRC* dst = (is_q ? qrow : krow) + off[u];
        *reinterpret_cast<uint4*>(dst) = r1;
        *reinterpret_cast<uint4*>(dst + D / 2) = r2;
      }
      if (!is_q && store) {
        finish_vec8<SRC, DST>(r1, kdst + off[u], p.k_scale, k_scaled);
        finish_vec8<SRC, DST>(r2, kdst + off[u] + D / 2, p.k_scale, k_scaled);
      }
    }
  }
}

// -DB200_DEV builds only: B200ATTN_ROPE_FAST=0 forces the generic kernel (A/B measurements)
inline bool rope_fast_enabled() {
#ifdef B200_DEV
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200ATTN_ROPE_FAST");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v != 0;
#else
  return true;
#endif
}

template <typename SRC, typename DST, int D, bool NEOX>
int launch_rope_fast_inst(const RopeParams& p, cudaStream_t stream) {
  constexpr int HPP = 32 / (D / 16);
  RopeFastParams fp;
  fp.b = p;
  fp.nq_pass = (p.Hq + HPP - 1) / HPP;
  fp.nk_pass = (p.Hkv + HPP - 1) / HPP;
  fp.nv_pass = p.k_cache != nullptr ? (p.Hkv * (D / 8) + 63) / 64 : 0;
  const int n_pass = fp.nq_pass + fp.nk_pass + fp.nv_pass;
  // enough warps to fill the machine: one per token at prefill, several per token for small batches
  const int64_t want_warps = (int64_t)sm_count() * 32;
  int slices = 1;
  if (p.num_tokens < want_warps) {
    slices = (int)((want_warps + p.num_tokens - 1) / p.num_tokens);
    if (slices > n_pass) slices = n_pass;
  }
  fp.pass_per_slice = (n_pass + slices - 1) / slices;
  fp.slices = (n_pass + fp.pass_per_slice - 1) / fp.pass_per_slice;
  const int64_t warps = p.num_tokens * fp.slices;
  const int64_t blocks = (warps + kRopeWarps - 1) / kRopeWarps;
  B200_CHECK_ARG(blocks > 0 && blocks < (int64_t)0x7fffffff, "rope_store_kv: too many rows (%lld)",
                 (long long)p.num_tokens);
  if (p.k_cache != nullptr && (p.k_scale != 1.0f || p.v_scale != 1.0f))
    rope_store_fast_kernel<SRC, DST, D, NEOX, true><<<(unsigned)blocks, kRopeWarps * 32, 0, stream>>>(fp);
  else
    rope_store_fast_kernel<SRC, DST, D, NEOX, false><<<(unsigned)blocks, kRopeWarps * 32, 0, stream>>>(fp);
  B200_CHECK_CUDA(cudaGetLastError());
  return B200_OK;
}

// full rotary with D = Dv in {64, 128, 256}: the warp-per-token kernel; everything else: the per-unit kernel
template <typename SRC, typename DST>
int launch_rope_any(const RopeParams& p, cudaStream_t stream) {
  const bool fast = rope_fast_enabled() && p.rot == p.D && (p.k_cache == nullptr || p.Dv == p.D) &&
                    (p.D == 64 || p.D == 128 || p.D == 256);
  if (!fast) return launch_rope_store<SRC, DST>(p, stream);
  if (p.is_neox) {
    if (p.D == 64) return launch_rope_fast_inst<SRC, DST, 64, true>(p, stream);
    if (p.D == 128) return launch_rope_fast_inst<SRC, DST, 128, true>(p, stream);
    return launch_rope_fast_inst<SRC, DST, 256, true>(p, stream);
  }
  if (p.D == 64) return launch_rope_fast_inst<SRC, DST, 64, false>(p, stream);
  if (p.D == 128) return launch_rope_fast_inst<SRC, DST, 128, false>(p, stream);
  return launch_rope_fast_inst<SRC, DST, 256, false>(p, stream);
}
