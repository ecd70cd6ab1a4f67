// Fast split-KV decode kernel (included by decode.cu inside namespace b200).
//
// One CTA = 4 independent warps, persistent over work items (chunk, kv head).
// Every warp streams its own contiguous slice of the chunk through a private
// cp.async ring (no block barrier in the steady state):
//   page table  --coalesced int32 loads, one 32-token block ahead-->  registers
//   K/V rows    --16-byte cp.async gathers, XOR-swizzled-->           shared memory
//   shared      --ldmatrix(.trans)-->  mma.sync m16n8k16 (bf16/f16, fp32 accumulate)
// The 16 MMA rows hold the `group` query heads that share this kv head, so K and V
// are read from HBM exactly once per group.  FP8-e4m3 pools are dequantised in
// registers (cvt.f16x2.e4m3x2) after ldmatrix and multiplied in f16.
// The kernel is HBM-bound: the MMA work is ~4 mma per KiB of KV.

namespace {

constexpr int kFastWarps = 4;
constexpr int kTileTok = 16;            // tokens per warp pipeline stage
constexpr int kWarpSmemTarget = 24576;  // ring bytes per warp

__device__ __forceinline__ void cp_async_16(uint32_t dst, const void* src, bool valid) {
  const int sz = valid ? 16 : 0;  // src-size 0 -> zero fill (masked tokens must not be NaN for P*V)
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
template <typename CT>
__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1) {
  if constexpr (sizeof(CT) == 2 && dtype_of<CT>::value == B200_BF16) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  } else {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
}
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}
// two FP8 bytes (low half of the argument) -> f16x2, exact for both formats: e4m3 through the hardware converter,
// e5m2 by byte placement (an e5m2 byte is the high byte of its f16)
template <int KV> __device__ __forceinline__ uint32_t fp8x2_to_f16x2(uint32_t two_bytes) {
  if constexpr (KV == B200_FP8_E5M2) {
    return prmt(two_bytes, 0u, 0x1404);
  } else {
    uint32_t r;
    asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(r) : "h"((uint16_t)two_bytes));
    return r;
  }
}

template <int D, int KV>
struct FastCfg {
  static constexpr bool kFp8 = KV == B200_FP8_E4M3 || KV == B200_FP8_E5M2;
  static constexpr int kEsz = kFp8 ? 1 : 2;
  static constexpr int kRowB = D * kEsz;                 // bytes of one K (or V) row of one head
  static constexpr int kCpr = kRowB / 16;                // 16-byte chunks per row
  static constexpr int kTileB = kTileTok * kRowB;        // K (or V) bytes per stage
  static constexpr int kStageB = 2 * kTileB;
  static constexpr int kStagesRaw = (kFp8 ? 16384 : kWarpSmemTarget) / kStageB;   // FP8: 3 CTAs per SM
  static constexpr int kStages = kStagesRaw < 3 ? 3 : (kStagesRaw > 8 ? 8 : kStagesRaw);
  static constexpr int kWarpSmem = kStages * kStageB;
  static constexpr int kCtaSmem = kFastWarps * kWarpSmem;
  static constexpr int kLdInstr = kTileTok * kCpr / 32;  // cp.async instructions per tile per tensor
  static constexpr int kNT = D / 8;                      // output n-tiles
  static_assert(kCpr >= 8, "rows shorter than 128 bytes use the generic kernel");
  static_assert(16 * D * 4 + 256 <= kWarpSmem, "epilogue staging must fit the ring");
};

// Split merge: one CTA per (request, query head).  out = sum_c 2^(lse_c - M) o_c / sum_c 2^(lse_c - M).
// Lanes own D/32 consecutive output dims (vector loads); the 4 warps take every 4th chunk and are
// combined through shared memory, so even 100+ chunks cost only a few dependent L2 round trips.
template <typename QT, int D>
__global__ void __launch_bounds__(kFastWarps* kWarp) decode_merge_fast_kernel(const DecodeParams p) {
  pdl_launch_dependents();          // next layer's KV store
  pdl_wait();                       // every partial of this layer has been written
  constexpr int kVec = D / 32;
  __shared__ float s_acc[kFastWarps][D];
  __shared__ float s_ml[kFastWarps][2];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int Hq = p.num_q_heads;
  const int req = blockIdx.x / Hq, h = blockIdx.x - req * Hq;
  const int c0 = p.plan[kPlanHeader + req], n = p.plan[kPlanHeader + req + 1] - c0;
  if (n == 1) return;   // single-chunk requests were written by the attention kernel
  if (n <= 0) {         // a request with no attended token: zeros and lse = -inf, like decode_merge_kernel
    QT* orow = reinterpret_cast<QT*>(p.o) + (int64_t)req * p.o_stride_t + (int64_t)h * p.o_stride_h;
    for (int d = threadIdx.x; d < D; d += kFastWarps * kWarp) orow[d] = from_float<QT>(0.f);
    if (p.lse != nullptr && threadIdx.x == 0) p.lse[(int64_t)req * Hq + h] = -INFINITY;
    return;
  }
  const float* lse = p.ws_lse + (int64_t)c0 * Hq + h;
  const float* po = p.ws_o + ((int64_t)c0 * Hq + h) * D + lane * kVec;
  float m = -INFINITY;
  for (int i = lane; i < n; i += kWarp) m = fmaxf(m, lse[(int64_t)i * Hq]);
  m = warp_max(m);                                   // every warp computes the global max (n is small)
  float acc[kVec];
#pragma unroll
  for (int e = 0; e < kVec; ++e) acc[e] = 0.f;
  float den = 0.f;
#pragma unroll 2
  for (int i = warp; i < n; i += kFastWarps) {
    const float w = fast_exp2(lse[(int64_t)i * Hq] - m);
    const float* src = po + (int64_t)i * Hq * D;
    float v[kVec];
    if constexpr (kVec == 2) {
      const float2 t = *reinterpret_cast<const float2*>(src);
      v[0] = t.x; v[1] = t.y;
    } else {
#pragma unroll
      for (int e = 0; e < kVec; e += 4) {
        const float4 t = *reinterpret_cast<const float4*>(src + e);
        v[e] = t.x; v[e + 1] = t.y; v[e + 2] = t.z; v[e + 3] = t.w;
      }
    }
    den += w;
#pragma unroll
    for (int e = 0; e < kVec; ++e) acc[e] = fmaf(w, v[e], acc[e]);
  }
#pragma unroll
  for (int e = 0; e < kVec; ++e) s_acc[warp][lane * kVec + e] = acc[e];
  if (lane == 0) s_ml[warp][0] = den;
  __syncthreads();
  float dsum = 0.f;
#pragma unroll
  for (int w = 0; w < kFastWarps; ++w) dsum += s_ml[w][0];
  dsum += exp2f(sink_log2(p, h) - m);                  // attention sink: only the denominator grows
  if (p.lse != nullptr && threadIdx.x == 0) p.lse[(int64_t)req * Hq + h] = (m + log2f(dsum)) * kLn2;
  const float inv = p.v_scale / dsum;
  QT* orow = reinterpret_cast<QT*>(p.o) + (int64_t)req * p.o_stride_t + (int64_t)h * p.o_stride_h;
  for (int d = threadIdx.x; d < D; d += kFastWarps * kWarp) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < kFastWarps; ++w) t += s_acc[w][d];
    orow[d] = from_float<QT>(t * inv);
  }
}

template <int D, int KV, typename QT, bool UPPER>
__global__ void __launch_bounds__(kFastWarps* kWarp, (FastCfg<D, KV>::kCtaSmem > 110 * 1024) ? 1 : ((FastCfg<D, KV>::kCtaSmem <= 72 * 1024) ? 3 : 2))
    decode_mma_kernel(const DecodeParams p) {
  pdl_launch_dependents();          // the merge kernel's CTAs may become resident now
  pdl_wait();                       // the KV store of this layer (and the planner) have completed
  using C = FastCfg<D, KV>;
  using CT = typename std::conditional<C::kFp8, __half, QT>::type;  // MMA operand type
  constexpr int kRows = UPPER ? 16 : 8;
  extern __shared__ __align__(128) uint8_t smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int quad = lane & 3, qrow = lane >> 2;
  uint8_t* wsm = smem + warp * C::kWarpSmem;
  const uint32_t wsm_u = (uint32_t)__cvta_generic_to_shared(wsm);

  const int total_chunks = p.plan[1];
  const int Hkv = p.num_kv_heads, g = p.group, Hq = p.num_q_heads;
  const int n_items = total_chunks * Hkv;
  const float scale_log2 = p.sm_scale * p.k_scale * kLog2e;
  const bool capped = p.logit_cap > 0.f;               // s = cap * tanh(s * sm_scale / cap)
  const float cap_in = capped ? p.sm_scale * p.k_scale / p.logit_cap : 0.f, cap_out = p.logit_cap * kLog2e;
  const float score_mul = capped ? 1.f : scale_log2;
  const uint8_t* kc = reinterpret_cast<const uint8_t*>(p.k_cache);
  const uint8_t* vc = reinterpret_cast<const uint8_t*>(p.v_cache);
  const int64_t k_slot_b = p.k_stride_s * C::kEsz, v_slot_b = p.v_stride_s * C::kEsz;

  for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
    const int c = item / Hkv, kvh = item - c * Hkv;
    const ChunkInfo ci = load_chunk(p, c);
    const int n_tiles = (ci.n_tok + kTileTok - 1) / kTileTok;
    const int tpw = (n_tiles + kFastWarps - 1) / kFastWarps;
    const int tb = warp * tpw;
    const int my_tiles = max(0, min(n_tiles, tb + tpw) - tb);
    const int w_ntok = max(0, min(ci.n_tok - tb * kTileTok, my_tiles * kTileTok));  // valid tokens of this warp

    // ---- Q fragments (A operand, rows = query heads of this kv head) ----
    uint32_t qa[D / 16][UPPER ? 4 : 2];
    {
      const QT* qb = reinterpret_cast<const QT*>(p.q) + (int64_t)ci.req * p.q_stride_t +
                     (int64_t)(kvh * g) * p.q_stride_h;
#pragma unroll
      for (int s = 0; s < D / 16; ++s) {
#pragma unroll
        for (int half = 0; half < (UPPER ? 2 : 1); ++half) {
          const int row = qrow + 8 * half;
          uint32_t lo = 0u, hi = 0u;
          if (row < g) {
            const QT* qr = qb + (int64_t)row * p.q_stride_h;
            if constexpr (!C::kFp8) {
              lo = *reinterpret_cast<const uint32_t*>(qr + 16 * s + 2 * quad);
              hi = *reinterpret_cast<const uint32_t*>(qr + 16 * s + 8 + 2 * quad);
            } else {
              // k index permutation matching the fp8 K fragments: 4 consecutive d per lane
              const uint2 raw = *reinterpret_cast<const uint2*>(qr + 16 * s + 4 * quad);
              if constexpr (dtype_of<QT>::value == B200_F16) {
                lo = raw.x; hi = raw.y;
              } else {
                lo = pack2<__half>(__uint_as_float(raw.x << 16), __uint_as_float(raw.x & 0xffff0000u));
                hi = pack2<__half>(__uint_as_float(raw.y << 16), __uint_as_float(raw.y & 0xffff0000u));
              }
            }
          }
          qa[s][half] = lo;                       // a0 (rows 0-7) / a1 (rows 8-15)
          qa[s][(UPPER ? 2 : 1) + half] = hi;     // a2 / a3
        }
      }
    }
    auto QA = [&](int s, int idx) -> uint32_t {  // idx = a0..a3
      if constexpr (UPPER) return qa[s][(idx & 1) + 2 * (idx >> 1)];
      else return (idx & 1) ? 0u : qa[s][idx >> 1];
    };

    // ---- per-warp online-softmax state ----
    float o[C::kNT][4];
#pragma unroll
    for (int i = 0; i < C::kNT; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};

    // ---- page table registers: lane holds the slot of token (32 * blk + lane) of this warp ----
    const int32_t* pt = p.req_to_token + ci.pool_row * p.req_to_token_stride + ci.start + tb * kTileTok;
    auto load_pt = [&](int blk) -> int {
      const int t = blk * 32 + lane;
      return t < w_ntok ? pt[t] : 0;
    };
    int pt_cur = 0, pt_nxt = 0, blk_cur = 0;
    if (my_tiles > 0) { pt_cur = load_pt(0); pt_nxt = load_pt(1); }

    auto issue = [&](int T) {
      if ((T >> 1) != blk_cur) { pt_cur = pt_nxt; ++blk_cur; pt_nxt = load_pt(blk_cur + 1); }
      const int stage = T % C::kStages;
      const uint32_t kst = wsm_u + stage * C::kStageB, vst = kst + C::kTileB;
      const int sub = (T & 1) * 16;
#pragma unroll
      for (int i = 0; i < C::kLdInstr; ++i) {
        const int idx = i * 32 + lane;
        const int tok = idx / C::kCpr, ch = idx % C::kCpr;
        const int64_t slot = __shfl_sync(0xffffffffu, pt_cur, sub + tok);
        const bool valid = T * kTileTok + tok < w_ntok;
        const uint32_t off = tok * C::kRowB + ((ch ^ (tok & 7)) << 4);
        cp_async_16(kst + off, kc + slot * k_slot_b + (int64_t)kvh * p.k_stride_h * C::kEsz + ch * 16, valid);
        cp_async_16(vst + off, vc + slot * v_slot_b + (int64_t)kvh * p.v_stride_h * C::kEsz + ch * 16, valid);
      }
    };

    // prologue
#pragma unroll
    for (int s = 0; s < C::kStages - 1; ++s) {
      if (s < my_tiles) issue(s);
      cp_async_commit();
    }

    for (int T = 0; T < my_tiles; ++T) {
      cp_async_wait<C::kStages - 2>();
      __syncwarp();
      if (T + C::kStages - 1 < my_tiles) issue(T + C::kStages - 1);
      cp_async_commit();

      const int stage = T % C::kStages;
      const uint32_t kst = wsm_u + stage * C::kStageB, vst = kst + C::kTileB;

      // ---- S = Q K^T (16 heads x 16 tokens) ----
      float s_acc[2][4];
#pragma unroll
      for (int i = 0; i < 2; ++i) s_acc[i][0] = s_acc[i][1] = s_acc[i][2] = s_acc[i][3] = 0.f;
      if constexpr (!C::kFp8) {
#pragma unroll
        for (int ks = 0; ks < D / 16; ++ks) {
          const int r = lane & 7, mi = lane >> 3;
          const int tok = r + 8 * (mi >> 1), ch = 2 * ks + (mi & 1);
          uint32_t b[4];
          ldmatrix_x4(b, kst + tok * C::kRowB + ((ch ^ r) << 4));
          mma16816<CT>(s_acc[0], QA(ks, 0), QA(ks, 1), QA(ks, 2), QA(ks, 3), b[0], b[1]);
          mma16816<CT>(s_acc[1], QA(ks, 0), QA(ks, 1), QA(ks, 2), QA(ks, 3), b[2], b[3]);
        }
      } else {
#pragma unroll
        for (int c2 = 0; c2 < C::kCpr; c2 += 2) {
          const int r = lane & 7, mi = lane >> 3;
          const int tok = r + 8 * (mi & 1), ch = c2 + (mi >> 1);
          uint32_t b[4];
          ldmatrix_x4(b, kst + tok * C::kRowB + ((ch ^ r) << 4));
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            const int ks = c2 + u;
            mma16816<CT>(s_acc[0], QA(ks, 0), QA(ks, 1), QA(ks, 2), QA(ks, 3),
                         fp8x2_to_f16x2<KV>(b[2 * u] & 0xffffu), fp8x2_to_f16x2<KV>(b[2 * u] >> 16));
            mma16816<CT>(s_acc[1], QA(ks, 0), QA(ks, 1), QA(ks, 2), QA(ks, 3),
                         fp8x2_to_f16x2<KV>(b[2 * u + 1] & 0xffffu), fp8x2_to_f16x2<KV>(b[2 * u + 1] >> 16));
          }
        }
      }

      // ---- online softmax (base 2) ----
      const int tok0 = T * kTileTok + 2 * quad;  // token (within this warp's slice) of s_acc[0][0]
      float pmax[2] = {-INFINITY, -INFINITY};
      if (capped) {                                        // warp-uniform; the common path below is unchanged
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
          for (int e = 0; e < 4; ++e) s_acc[nt][e] = cap_out * tanh_ex2(s_acc[nt][e] * cap_in);
      }
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int tok = tok0 + 8 * nt + (e & 1);
          float v = s_acc[nt][e] * score_mul;
          if (tok >= w_ntok) v = -INFINITY;
          s_acc[nt][e] = v;
          if (UPPER || e < 2) pmax[e >> 1] = fmaxf(pmax[e >> 1], v);
        }
      }
      float corr[2] = {1.f, 1.f};
#pragma unroll
      for (int h = 0; h < (UPPER ? 2 : 1); ++h) {
        pmax[h] = fmaxf(pmax[h], __shfl_xor_sync(0xffffffffu, pmax[h], 1));
        pmax[h] = fmaxf(pmax[h], __shfl_xor_sync(0xffffffffu, pmax[h], 2));
        const float m_new = fmaxf(m_run[h], pmax[h]);
        corr[h] = fast_exp2(m_run[h] - m_new);
        m_run[h] = m_new;
      }
      uint32_t pa[4] = {0u, 0u, 0u, 0u};
      {
        float psum[2] = {0.f, 0.f};
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          const float p0 = fast_exp2(s_acc[nt][0] - m_run[0]);
          const float p1 = fast_exp2(s_acc[nt][1] - m_run[0]);
          psum[0] += p0 + p1;
          pa[2 * nt] = pack2<CT>(p0, p1);
          if constexpr (UPPER) {
            const float p2 = fast_exp2(s_acc[nt][2] - m_run[1]);
            const float p3 = fast_exp2(s_acc[nt][3] - m_run[1]);
            psum[1] += p2 + p3;
            pa[2 * nt + 1] = pack2<CT>(p2, p3);
          }
        }
        l_run[0] = l_run[0] * corr[0] + psum[0];
        l_run[1] = l_run[1] * corr[1] + psum[1];
      }
      const bool need_rescale = corr[0] != 1.f || (UPPER && corr[1] != 1.f);
      if (__any_sync(0xffffffffu, need_rescale)) {
#pragma unroll
        for (int i = 0; i < C::kNT; ++i) {
          o[i][0] *= corr[0]; o[i][1] *= corr[0];
          if constexpr (UPPER) { o[i][2] *= corr[1]; o[i][3] *= corr[1]; }
        }
      }

      // ---- O += P V ----
      if constexpr (!C::kFp8) {
#pragma unroll
        for (int j = 0; j < C::kNT; j += 2) {
          const int r = lane & 7;
          const int tok = r + 8 * ((lane >> 3) & 1), ch = j + (lane >> 4);
          uint32_t b[4];
          ldmatrix_x4_trans(b, vst + tok * C::kRowB + ((ch ^ r) << 4));
          mma16816<CT>(o[j], pa[0], pa[1], pa[2], pa[3], b[0], b[1]);
          mma16816<CT>(o[j + 1], pa[0], pa[1], pa[2], pa[3], b[2], b[3]);
        }
      } else {
#pragma unroll
        for (int j = 0; j < C::kCpr; j += 2) {
          const int r = lane & 7;
          const int tok = r + 8 * ((lane >> 3) & 1), ch = j + (lane >> 4);
          uint32_t b[4];
          ldmatrix_x4_trans(b, vst + tok * C::kRowB + ((ch ^ r) << 4));
#pragma unroll
          for (int u = 0; u < 2; ++u) {
            // bytes [t0 d, t0 d+1, t1 d, t1 d+1] -> [t0 d, t1 d | t0 d+1, t1 d+1]
            const uint32_t x0 = prmt(b[2 * u], 0u, 0x3120u), x1 = prmt(b[2 * u + 1], 0u, 0x3120u);
            const int nt = 2 * (j + u);
            mma16816<CT>(o[nt], pa[0], pa[1], pa[2], pa[3], fp8x2_to_f16x2<KV>(x0 & 0xffffu),
                         fp8x2_to_f16x2<KV>(x1 & 0xffffu));
            mma16816<CT>(o[nt + 1], pa[0], pa[1], pa[2], pa[3], fp8x2_to_f16x2<KV>(x0 >> 16),
                         fp8x2_to_f16x2<KV>(x1 >> 16));
          }
        }
      }
    }

    // ---- epilogue: stage per-warp partials in shared memory, combine the 4 warps ----
    cp_async_wait<0>();
    __syncwarp();
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 1);
      l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 2);
    }
    float* o_s = reinterpret_cast<float*>(wsm);          // [kRows][D]
    float* ml_s = o_s + 16 * D;                          // [2][16]
#pragma unroll
    for (int nt = 0; nt < C::kNT; ++nt) {
#pragma unroll
      for (int e = 0; e < (UPPER ? 4 : 2); ++e) {
        const int row = qrow + 8 * (e >> 1);
        int d;
        if constexpr (!C::kFp8) d = 8 * nt + 2 * quad + (e & 1);
        else d = 16 * (nt >> 1) + 4 * quad + 2 * (e & 1) + (nt & 1);
        o_s[row * D + d] = o[nt][e];
      }
    }
    if (quad == 0) {
      ml_s[qrow] = m_run[0]; ml_s[16 + qrow] = l_run[0];
      if constexpr (UPPER) { ml_s[8 + qrow] = m_run[1]; ml_s[16 + 8 + qrow] = l_run[1]; }
    }
    __syncthreads();
    const bool final_out = ci.n_chunks == 1;
    for (int idx = threadIdx.x; idx < g * D; idx += kFastWarps * kWarp) {
      const int row = idx / D, d = idx - row * D;
      float mw[kFastWarps], M = -INFINITY;
#pragma unroll
      for (int w = 0; w < kFastWarps; ++w) {
        mw[w] = reinterpret_cast<const float*>(smem + w * C::kWarpSmem)[16 * D + row];
        M = fmaxf(M, mw[w]);
      }
      float num = 0.f, den = 0.f;
#pragma unroll
      for (int w = 0; w < kFastWarps; ++w) {
        const float* ow = reinterpret_cast<const float*>(smem + w * C::kWarpSmem);
        const float wgt = fast_exp2(mw[w] - M);
        num = fmaf(wgt, ow[row * D + d], num);
        den = fmaf(wgt, ow[16 * D + 16 + row], den);
      }
      const int h = kvh * g + row;
      if (final_out) {
        const float den_fin = den + exp2f(sink_log2(p, h) - M);
        reinterpret_cast<QT*>(p.o)[(int64_t)ci.req * p.o_stride_t + (int64_t)h * p.o_stride_h + d] =
            from_float<QT>(num / den_fin * p.v_scale);
        if (p.lse != nullptr && d == 0) p.lse[(int64_t)ci.req * Hq + h] = (M + log2f(den_fin)) * kLn2;
      } else {
        const float val = num / den;
        p.ws_o[((int64_t)c * Hq + h) * D + d] = val;
        if (d == 0) p.ws_lse[(int64_t)c * Hq + h] = M + log2f(den);
      }
    }
    __syncthreads();
    (void)kRows;
  }
}

bool decode_fast_supported(int dqk, int dv, int group, int q_dtype, int kv_dtype) {
  if (dqk != dv) return false;
  if (dqk != 64 && dqk != 128 && dqk != 256) return false;
  if (group < 1 || group > 16) return false;
  if (q_dtype != B200_BF16 && q_dtype != B200_F16) return false;
  if (kv_dtype == q_dtype) return true;
  if (kv_dtype == B200_FP8_E4M3) return dqk >= 128;
  if (kv_dtype == B200_FP8_E5M2) return dqk == 128;
  return false;
}

int decode_fast_ctas_per_sm(int head_dim, int kv_dtype) {
  if (head_dim >= 256) return 1;
  return ((kv_dtype == B200_FP8_E4M3 || kv_dtype == B200_FP8_E5M2) && head_dim == 128) ? 3 : 2;
}

bool decode_fast_layout_ok(const DecodeParams& p, int kv_dtype) {
  const int esz = (kv_dtype == B200_FP8_E4M3 || kv_dtype == B200_FP8_E5M2) ? 1 : 2;
  auto a16 = [](const void* ptr) { return ((uintptr_t)ptr & 15) == 0; };
  return a16(p.k_cache) && a16(p.v_cache) && a16(p.q) && (p.k_stride_s * esz) % 16 == 0 &&
         (p.k_stride_h * esz) % 16 == 0 && (p.v_stride_s * esz) % 16 == 0 && (p.v_stride_h * esz) % 16 == 0 &&
         p.q_stride_t % 8 == 0 && p.q_stride_h % 8 == 0;
}

template <int D, int KV, typename QT, bool UPPER>
int launch_decode_fast_inst(const DecodeParams& p, int num_ctas, cudaStream_t stream) {
  using C = FastCfg<D, KV>;
  static std::atomic<uint64_t> configured{0};
  auto kern = decode_mma_kernel<D, KV, QT, UPPER>;
  B200_CHECK_CUDA(set_max_dynamic_smem(kern, C::kCtaSmem, configured));
  B200_CHECK_CUDA(launch_pdl(kern, dim3(num_ctas), dim3(kFastWarps * kWarp), C::kCtaSmem, stream, p));
  B200_CHECK_CUDA(cudaGetLastError());
  B200_CHECK_CUDA(launch_pdl(decode_merge_fast_kernel<QT, D>, dim3(p.batch_size * p.num_q_heads), dim3(kFastWarps * kWarp), 0,
                             stream, p));
  B200_CHECK_CUDA(cudaGetLastError());
  return B200_OK;
}

template <int D, int KV, typename QT>
int launch_decode_fast_d(const DecodeParams& p, int num_ctas, cudaStream_t stream) {
  if (p.group > 8) return launch_decode_fast_inst<D, KV, QT, true>(p, num_ctas, stream);
  return launch_decode_fast_inst<D, KV, QT, false>(p, num_ctas, stream);
}

template <typename QT>
int launch_decode_fast_q(const DecodeParams& p, int kv_dtype, int num_ctas, cudaStream_t stream) {
  constexpr int kSame = dtype_of<QT>::value;
  const int D = p.head_dim_qk;
  if (kv_dtype == B200_FP8_E4M3) {
    if (D == 128) return launch_decode_fast_d<128, B200_FP8_E4M3, QT>(p, num_ctas, stream);
    if (D == 256) return launch_decode_fast_d<256, B200_FP8_E4M3, QT>(p, num_ctas, stream);
  } else if (kv_dtype == B200_FP8_E5M2) {
    if (D == 128) return launch_decode_fast_d<128, B200_FP8_E5M2, QT>(p, num_ctas, stream);
  } else {
    if (D == 64) return launch_decode_fast_d<64, kSame, QT>(p, num_ctas, stream);
    if (D == 128) return launch_decode_fast_d<128, kSame, QT>(p, num_ctas, stream);
    if (D == 256) return launch_decode_fast_d<256, kSame, QT>(p, num_ctas, stream);
  }
  set_error("decode fast: no instance for D=%d kv=%d", D, kv_dtype);
  return B200_EINVAL;
}

int launch_decode_fast(const DecodeParams& p, int q_dtype, int kv_dtype, int num_ctas, cudaStream_t stream) {
  if (q_dtype == B200_BF16) return launch_decode_fast_q<__nv_bfloat16>(p, kv_dtype, num_ctas, stream);
  return launch_decode_fast_q<__half>(p, kv_dtype, num_ctas, stream);
}

}  // namespace
