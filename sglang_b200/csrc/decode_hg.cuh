// Head-group split-KV decode kernel (included by decode.cu after decode_mma.cuh, inside namespace b200).
//
// Why it exists: at page_size 1 every (token, kv head) row of K or V sits at a random pool slot, and a
// gather of random 128-byte segments tops out at ~4.7 TB/s on B200 (256-byte segments: 6.4 TB/s,
// 512+ bytes: 6.8-7.0 TB/s; benchmarks/gather_rate.cu).  decode_mma_kernel reads one head per work item,
// i.e. 128-byte (FP8, D=128) or 256-byte (bf16) segments.  Here a work item is (chunk, group of 4 adjacent
// kv heads): the CTA's 4 warps cooperatively gather the 4-head slice of each token row -- 512 contiguous
// bytes for FP8, 1 KiB for bf16, one or two full-warp cp.async per token -- into a shared ring, and warp w
// then runs the same ldmatrix / mma.sync / online-softmax core as decode_mma_kernel on head w's slice.
//   full[s]  mbarrier, 128 arrivals via cp.async.mbarrier.arrive (data of stage s has landed)
//   empty[s] mbarrier, 4 arrivals (every warp is done reading stage s)
// No block barrier anywhere in the steady state or between work items; each warp owns its head's
// result and writes it straight from registers.
namespace {

constexpr int kHG = 4;   // kv heads per work item == warps per CTA

template <int D, int KV>
struct HgCfg {
  static constexpr bool kFp8 = KV == B200_FP8_E4M3 || KV == B200_FP8_E5M2;
  static constexpr int kEsz = kFp8 ? 1 : 2;
  static constexpr int kRowB = D * kEsz;                 // one head's K (or V) row
  static constexpr int kCpr = kRowB / 16;                // 16-byte chunks per head row
  static constexpr int kTokB = kHG * kRowB;              // the 4-head slice of a token row (contiguous in HBM)
  static constexpr int kTileB = kTileTok * kTokB;        // K (or V) bytes per stage
  static constexpr int kStageB = 2 * kTileB;
  static constexpr int kStages = kStageB <= 16 * 1024 ? 4 : 3;
  static constexpr int kRingB = kStages * kStageB;
  static constexpr int kCtaSmem = kRingB + 16 * kStages;
  static constexpr int kLdPerTok = kTokB / 512;          // full-warp cp.async per token per tensor (0: half a warp)
  static constexpr int kTokPerWarp = kTileTok / kHG;     // tokens each warp gathers per stage
  static constexpr int kNT = D / 8;                      // output n-tiles
#ifndef B200_HG_SMALL_BLOCKS
#define B200_HG_SMALL_BLOCKS 4
#endif
  // 8 KiB stages (D = 64 FP8): half the bytes per tile behind the same per-tile dependency chain, so more resident
  // warps are what hides it
  static constexpr int kMinBlocks =
      kCtaSmem > 110 * 1024 ? 1 : (kCtaSmem <= 36 * 1024 ? B200_HG_SMALL_BLOCKS : (kCtaSmem <= 72 * 1024 ? 3 : 2));
  // D = 64 FP8 pools: a head row is 64 bytes (4 chunks), the 4-head slice 256 bytes = HALF a warp's cp.async, so one
  // copy instruction covers two tokens and the swizzle runs over chunk indices that span two heads
  static constexpr bool kHalfWarpTok = kTokB == 256;
  // D = 80 / 96 / 112 with 16-bit pools (Phi-class heads): rows of 160 / 192 / 224 bytes -- the 4-head slice is still
  // contiguous and a multiple of 128 bytes, so the gather walks it as a flat list of 16-byte chunks (kFlatGather) and
  // a head's first chunk goes through the swizzle XOR like the 64-byte FP8 rows
  static constexpr bool kFlatGather = !kFp8 && kCpr % 8 != 0;
  static constexpr int kTokChunks = kTokB / 16;
  static_assert(kCpr % 8 == 0 || (kCpr == 4 && kFp8) || (kFlatGather && kCpr % 2 == 0 && kTokChunks % 8 == 0),
                "head rows: whole 128-byte lines, 64-byte FP8 rows, or 16-bit rows of a multiple of 32 bytes");
  static_assert(kTokB % 512 == 0 || kHalfWarpTok || (kFlatGather && (kTokPerWarp * kTokChunks) % 32 == 0),
                "a token slice must be whole (or exactly half) full-warp copies, or a flat chunk list");
  static_assert(!kHalfWarpTok || kTokPerWarp % 2 == 0, "half-warp copies take tokens in pairs");
};

template <int D, int KV, typename QT, bool UPPER>
__global__ void __launch_bounds__(kHG* kWarp, HgCfg<D, KV>::kMinBlocks) decode_hg_kernel(const DecodeParams p) {
  pdl_launch_dependents();          // the merge kernel's CTAs may become resident now
  pdl_wait();                       // the KV store of this layer (and the planner) have completed
  using C = HgCfg<D, KV>;
  using CT = typename std::conditional<C::kFp8, __half, QT>::type;  // MMA operand type
  extern __shared__ __align__(128) uint8_t smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int quad = lane & 3, qrow = lane >> 2;
  const uint32_t ring = (uint32_t)__cvta_generic_to_shared(smem);
  const uint32_t bar_full = ring + C::kRingB, bar_empty = bar_full + 8 * C::kStages;
  if (threadIdx.x == 0) {
    for (int s = 0; s < C::kStages; ++s) {
      ptx::mbar_init(bar_full + 8 * s, kHG * kWarp);
      ptx::mbar_init(bar_empty + 8 * s, kHG);
    }
    ptx::mbar_fence_init();
  }
  __syncthreads();

  const int total_chunks = p.plan[1];
  const int Hkv = p.num_kv_heads, g = p.group, Hq = p.num_q_heads;
  const int n_groups = Hkv / kHG;
  const int n_items = total_chunks * n_groups;
  const float scale_log2 = p.sm_scale * p.k_scale * kLog2e;
  const bool capped = p.logit_cap > 0.f;               // s = cap * tanh(s * sm_scale / cap)
  const float cap_in = capped ? p.sm_scale * p.k_scale / p.logit_cap : 0.f, cap_out = p.logit_cap * kLog2e;
  const float score_mul = capped ? 1.f : scale_log2;
  const int64_t k_slot_b = p.k_stride_s * C::kEsz, v_slot_b = p.v_stride_s * C::kEsz;
  uint32_t n_issue = 0, n_cons = 0;   // tiles issued / consumed by this warp since kernel start (ring phases)

  for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
    const int c = item / n_groups, hg = item - c * n_groups;
    const int kvh = hg * kHG + warp;
    const ChunkInfo ci = load_chunk(p, c);
    const int n_tok = ci.n_tok;
    const int n_tiles = (n_tok + kTileTok - 1) / kTileTok;
    const uint8_t* kc = reinterpret_cast<const uint8_t*>(p.k_cache) + (int64_t)hg * kHG * p.k_stride_h * C::kEsz;
    const uint8_t* vc = reinterpret_cast<const uint8_t*>(p.v_cache) + (int64_t)hg * kHG * p.v_stride_h * C::kEsz;

    // ---- Q fragments (A operand, rows = query heads of this warp's kv head) ----
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

    // ---- online-softmax state of this warp's head ----
    float o[C::kNT][4];
#pragma unroll
    for (int i = 0; i < C::kNT; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};

    // ---- page table registers: lane holds the slot of token (32 * blk + lane) of the chunk ----
    const int32_t* pt = p.req_to_token + ci.pool_row * p.req_to_token_stride + ci.start;
    auto load_pt = [&](int blk) -> int {
      const int t = blk * 32 + lane;
      return t < n_tok ? __ldg(pt + t) : 0;
    };
    int pt_cur = load_pt(0), pt_nxt = load_pt(1), blk_cur = 0;

    // gather this warp's 4 tokens of tile T (all 4 heads of each) into the ring
    auto issue = [&](int T) {
      if ((T >> 1) != blk_cur) { pt_cur = pt_nxt; ++blk_cur; pt_nxt = load_pt(blk_cur + 1); }
      const uint32_t stage = n_issue % C::kStages;
      if (n_issue >= (uint32_t)C::kStages)
        ptx::mbar_wait(bar_empty + 8 * stage, ((n_issue / C::kStages) - 1) & 1, 100 + stage);
      const uint32_t kst = ring + stage * C::kStageB, vst = kst + C::kTileB;
      if constexpr (C::kHalfWarpTok) {
#pragma unroll
        for (int k = 0; k < C::kTokPerWarp / 2; ++k) {                     // lanes 0-15: one token, 16-31: the next
          const int tok = warp * C::kTokPerWarp + 2 * k + (lane >> 4);
          const int ch = lane & 15;                                        // 16-byte chunk of the 4-head slice
          const int64_t slot = __shfl_sync(0xffffffffu, pt_cur, (T & 1) * 16 + tok);
          const bool valid = T * kTileTok + tok < n_tok;
          const uint32_t off = tok * C::kTokB + ((ch ^ (tok & 7)) << 4);
          cp_async_16(kst + off, kc + slot * k_slot_b + ch * 16, valid);
          cp_async_16(vst + off, vc + slot * v_slot_b + ch * 16, valid);
        }
      } else if constexpr (C::kFlatGather) {
#pragma unroll
        for (int i = 0; i < C::kTokPerWarp * C::kTokChunks / 32; ++i) {    // this warp's tokens as one list of chunks
          const int f = i * 32 + lane;
          const int t = f / C::kTokChunks, ch = f - t * C::kTokChunks;
          const int tok = warp * C::kTokPerWarp + t;
          const int64_t slot = __shfl_sync(0xffffffffu, pt_cur, (T & 1) * 16 + tok);
          const bool valid = T * kTileTok + tok < n_tok;
          const uint32_t off = tok * C::kTokB + ((ch ^ (tok & 7)) << 4);
          cp_async_16(kst + off, kc + slot * k_slot_b + ch * 16, valid);
          cp_async_16(vst + off, vc + slot * v_slot_b + ch * 16, valid);
        }
      } else {
#pragma unroll
      for (int k = 0; k < C::kTokPerWarp; ++k) {
        const int tok = warp * C::kTokPerWarp + k;
        const int64_t slot = __shfl_sync(0xffffffffu, pt_cur, (T & 1) * 16 + tok);
        const bool valid = T * kTileTok + tok < n_tok;
#pragma unroll
        for (int i = 0; i < C::kLdPerTok; ++i) {
          const int ch = i * 32 + lane;                                  // 16-byte chunk of the 4-head slice
          const uint32_t off = tok * C::kTokB + ((ch ^ (tok & 7)) << 4);
          cp_async_16(kst + off, kc + slot * k_slot_b + ch * 16, valid);
          cp_async_16(vst + off, vc + slot * v_slot_b + ch * 16, valid);
        }
      }
      }
      ptx::cp_async_mbar_arrive_noinc(bar_full + 8 * stage);
      ++n_issue;
    };

#pragma unroll
    for (int s = 0; s < C::kStages - 1; ++s)
      if (s < n_tiles) issue(s);

    for (int T = 0; T < n_tiles; ++T) {
      if (T + C::kStages - 1 < n_tiles) issue(T + C::kStages - 1);
      const uint32_t stage = n_cons % C::kStages;
      ptx::mbar_wait(bar_full + 8 * stage, (n_cons / C::kStages) & 1, 110 + stage);
      // this warp's head inside the stage: whole 128-byte lines sit beside each other, so the head offset is added
      // outside the swizzle; 64-byte rows (D = 64 FP8) share a line with the neighbouring head and the head's first
      // chunk goes through the XOR with the token's row instead (soff)
      const uint32_t kst = ring + stage * C::kStageB + (C::kCpr % 8 == 0 ? warp * C::kRowB : 0), vst = kst + C::kTileB;
      auto soff = [&](int tok, int ch, int r) -> uint32_t {     // r == tok & 7
        if constexpr (C::kCpr % 8 == 0) return tok * C::kTokB + ((ch ^ r) << 4);
        else return tok * C::kTokB + (((warp * C::kCpr + ch) ^ r) << 4);
      };

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
          ldmatrix_x4(b, kst + soff(tok, ch, r));
          mma16816<CT>(s_acc[0], QA(ks, 0), QA(ks, 1), QA(ks, 2), QA(ks, 3), b[0], b[1]);
          mma16816<CT>(s_acc[1], QA(ks, 0), QA(ks, 1), QA(ks, 2), QA(ks, 3), b[2], b[3]);
        }
      } else {
#pragma unroll
        for (int c2 = 0; c2 < C::kCpr; c2 += 2) {
          const int r = lane & 7, mi = lane >> 3;
          const int tok = r + 8 * (mi & 1), ch = c2 + (mi >> 1);
          uint32_t b[4];
          ldmatrix_x4(b, kst + soff(tok, ch, r));
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
      const int tok0 = T * kTileTok + 2 * quad;  // token (within the chunk) of s_acc[0][0]
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
          if (tok >= n_tok) v = -INFINITY;
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
          ldmatrix_x4_trans(b, vst + soff(tok, ch, r));
          mma16816<CT>(o[j], pa[0], pa[1], pa[2], pa[3], b[0], b[1]);
          mma16816<CT>(o[j + 1], pa[0], pa[1], pa[2], pa[3], b[2], b[3]);
        }
      } else {
#pragma unroll
        for (int j = 0; j < C::kCpr; j += 2) {
          const int r = lane & 7;
          const int tok = r + 8 * ((lane >> 3) & 1), ch = j + (lane >> 4);
          uint32_t b[4];
          ldmatrix_x4_trans(b, vst + soff(tok, ch, r));
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
      // done reading this stage
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(bar_empty + 8 * stage);
      ++n_cons;
    }

    // ---- epilogue: this warp's head, straight from registers ----
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 1);
      l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 2);
    }
    const bool final_out = ci.n_chunks == 1;
#pragma unroll
    for (int half = 0; half < (UPPER ? 2 : 1); ++half) {
      const int row = qrow + 8 * half;
      if (row < g) {
        const int h = kvh * g + row;
        if (final_out) {                                   // attention sink + LSE of single-chunk requests
          l_run[half] += exp2f(sink_log2(p, h) - m_run[half]);
          if (p.lse != nullptr && quad == 0)
            p.lse[(int64_t)ci.req * Hq + h] = (m_run[half] + log2f(l_run[half])) * kLn2;
        }
        const float inv = (final_out ? p.v_scale : 1.f) / l_run[half];
        QT* orow = reinterpret_cast<QT*>(p.o) + (int64_t)ci.req * p.o_stride_t + (int64_t)h * p.o_stride_h;
        float* wrow = p.ws_o + ((int64_t)c * Hq + h) * D;
        if constexpr (!C::kFp8) {
#pragma unroll
          for (int nt = 0; nt < C::kNT; ++nt) {   // d = 8 nt + 2 quad, +1
            const float v0 = o[nt][2 * half] * inv, v1 = o[nt][2 * half + 1] * inv;
            const int d = 8 * nt + 2 * quad;
            if (final_out) *reinterpret_cast<uint32_t*>(orow + d) = pack2<QT>(v0, v1);
            else *reinterpret_cast<float2*>(wrow + d) = make_float2(v0, v1);
          }
        } else {
#pragma unroll
          for (int m = 0; m < C::kNT / 2; ++m) {  // d = 16 m + 4 quad + 2 (e & 1) + (nt & 1): 4 consecutive
            const float v0 = o[2 * m][2 * half] * inv, v1 = o[2 * m + 1][2 * half] * inv;
            const float v2 = o[2 * m][2 * half + 1] * inv, v3 = o[2 * m + 1][2 * half + 1] * inv;
            const int d = 16 * m + 4 * quad;
            if (final_out) {
              *reinterpret_cast<uint2*>(orow + d) = make_uint2(pack2<QT>(v0, v1), pack2<QT>(v2, v3));
            } else {
              *reinterpret_cast<float4*>(wrow + d) = make_float4(v0, v1, v2, v3);
            }
          }
        }
        if (!final_out && quad == 0) p.ws_lse[(int64_t)c * Hq + h] = m_run[half] + log2f(l_run[half]);
      }
    }
  }
}

// head dims / dtypes the head-group kernel has an instance for: everything the per-head kernel covers plus D = 64
// with FP8 pools (64-byte rows, below the 128-byte ldmatrix line the per-head kernel is laid out for; here the
// 4-head slice of a token is 256 contiguous bytes, which is all the layout needs)
bool decode_hg_dims_supported(int dqk, int dv, int group, int q_dtype, int kv_dtype) {
  if (decode_fast_supported(dqk, dv, group, q_dtype, kv_dtype)) return true;
  const bool fp8 = kv_dtype == B200_FP8_E4M3 || kv_dtype == B200_FP8_E5M2;
  if (dqk != dv || group < 1 || group > 16 || (q_dtype != B200_BF16 && q_dtype != B200_F16)) return false;
  if (fp8) return dqk == 64;
  return kv_dtype == q_dtype && (dqk == 80 || dqk == 96 || dqk == 112);     // Phi-class heads, 16-bit pools
}

bool decode_hg_supported(const DecodeParams& p, int q_dtype, int kv_dtype) {
  const int D = p.head_dim_qk;
  if (p.num_kv_heads % kHG != 0) return false;
  if (!decode_hg_dims_supported(D, p.head_dim_v, p.group, q_dtype, kv_dtype)) return false;
  if (D == 256 && kv_dtype != B200_FP8_E4M3) return false;     // 192 KiB ring: keep the per-head kernel
  const int esz = (kv_dtype == B200_FP8_E4M3 || kv_dtype == B200_FP8_E5M2) ? 1 : 2;
  // the 4 heads of a group must be contiguous in a pool row; outputs must take 8-byte vector stores
  return p.k_stride_h == D && p.v_stride_h == D && (p.k_stride_s * esz) % 16 == 0 && (p.v_stride_s * esz) % 16 == 0 &&
         p.o_stride_t % 4 == 0 && p.o_stride_h % 4 == 0 && ((uintptr_t)p.o & 7) == 0;
}

template <int D, int KV, typename QT, bool UPPER>
int launch_decode_hg_inst(const DecodeParams& p, int num_ctas, cudaStream_t stream) {
  using C = HgCfg<D, KV>;
  static std::atomic<uint64_t> configured{0};
  auto kern = decode_hg_kernel<D, KV, QT, UPPER>;
  B200_CHECK_CUDA(set_max_dynamic_smem(kern, C::kCtaSmem, configured));
  B200_CHECK_CUDA(launch_pdl(kern, dim3(num_ctas), dim3(kHG * kWarp), C::kCtaSmem, stream, p));
  B200_CHECK_CUDA(cudaGetLastError());
  if constexpr (D % 64 == 0) {
    B200_CHECK_CUDA(launch_pdl(decode_merge_fast_kernel<QT, D>, dim3(p.batch_size * p.num_q_heads), dim3(kFastWarps * kWarp), 0,
                               stream, p));
  } else {          // D = 80 / 96 / 112: the generic split merge (same partial format)
    const int64_t rows = (int64_t)p.batch_size * p.num_q_heads;
    decode_merge_kernel<QT><<<(unsigned)((rows + kMergeWarps - 1) / kMergeWarps), kMergeWarps * kWarp, 0, stream>>>(p);
  }
  B200_CHECK_CUDA(cudaGetLastError());
  return B200_OK;
}

template <int D, int KV, typename QT>
int launch_decode_hg_d(const DecodeParams& p, int num_ctas, cudaStream_t stream) {
  if (p.group > 8) return launch_decode_hg_inst<D, KV, QT, true>(p, num_ctas, stream);
  return launch_decode_hg_inst<D, KV, QT, false>(p, num_ctas, stream);
}

template <typename QT>
int launch_decode_hg_q(const DecodeParams& p, int kv_dtype, int num_ctas, cudaStream_t stream) {
  constexpr int kSame = dtype_of<QT>::value;
  const int D = p.head_dim_qk;
  if (kv_dtype == B200_FP8_E4M3) {
    if (D == 64) return launch_decode_hg_d<64, B200_FP8_E4M3, QT>(p, num_ctas, stream);
    if (D == 128) return launch_decode_hg_d<128, B200_FP8_E4M3, QT>(p, num_ctas, stream);
    if (D == 256) return launch_decode_hg_d<256, B200_FP8_E4M3, QT>(p, num_ctas, stream);
  } else if (kv_dtype == B200_FP8_E5M2) {
    if (D == 64) return launch_decode_hg_d<64, B200_FP8_E5M2, QT>(p, num_ctas, stream);
    if (D == 128) return launch_decode_hg_d<128, B200_FP8_E5M2, QT>(p, num_ctas, stream);
  } else {
    if (D == 64) return launch_decode_hg_d<64, kSame, QT>(p, num_ctas, stream);
    if (D == 128) return launch_decode_hg_d<128, kSame, QT>(p, num_ctas, stream);
    if (D == 80) return launch_decode_hg_d<80, kSame, QT>(p, num_ctas, stream);
    if (D == 96) return launch_decode_hg_d<96, kSame, QT>(p, num_ctas, stream);
    if (D == 112) return launch_decode_hg_d<112, kSame, QT>(p, num_ctas, stream);
  }
  set_error("decode head-group: no instance for D=%d kv=%d", D, kv_dtype);
  return B200_EINVAL;
}

int launch_decode_hg(const DecodeParams& p, int q_dtype, int kv_dtype, int num_ctas, cudaStream_t stream) {
  if (q_dtype == B200_BF16) return launch_decode_hg_q<__nv_bfloat16>(p, kv_dtype, num_ctas, stream);
  return launch_decode_hg_q<__half>(p, kv_dtype, num_ctas, stream);
}

// resident CTAs per SM of the instance the dispatcher would pick (grid = SMs x this)
int decode_hg_ctas_per_sm(int head_dim, int kv_dtype) {
  const bool fp8 = kv_dtype == B200_FP8_E4M3 || kv_dtype == B200_FP8_E5M2;
  const int stage = 2 * kTileTok * kHG * head_dim * (fp8 ? 1 : 2);
  const int stages = stage <= 16 * 1024 ? 4 : 3;
  const int smem = stages * stage + 16 * stages;          // HgCfg::kCtaSmem
  return smem > 110 * 1024 ? 1 : (smem <= 36 * 1024 ? B200_HG_SMALL_BLOCKS : (smem <= 72 * 1024 ? 3 : 2));
}

}  // namespace
