// tcgen05 / TMEM extend (prefill over a cached prefix) kernel, included by extend.cu inside namespace b200.
//
// One CTA per work item = (request, kv head, block of 2*BQ query tokens).  The 128 MMA rows of a
// query tile hold BQ = 128/group tokens x `group` query heads, so K/V are read once per GQA group.
// Two query tiles per CTA ping-pong on the tensor core while the other one is in softmax:
//
//   warps 0-3   softmax / correction / epilogue for query tile 0 (one thread per TMEM lane = row)
//   warps 4-7   the same for query tile 1
//   warp  8     TMEM allocation + the single thread that issues tcgen05.mma
//   warps 9-12  producers: page-table gather of K/V rows (and Q once) with 16-byte cp.async into
//               128B-swizzled shared-memory tiles that UMMA descriptors read directly
//
//   S_t = Q_t K_j^T   (M128 x N64 x K128, A/B K-major)        accumulator: TMEM, 64 fp32 columns
//   O_t += P_t V_j    (M128 x N128 x K64, A K-major from smem, B = V rows, MN-major)   TMEM, 128 columns
//
// TMEM map (512 columns allocated): S0 [0,64)  S1 [64,128)  O0 [128,256)  O1 [256,384).
// Shared memory: Q0,Q1 32 KiB each | 3 stages x (K 16 KiB + V 16 KiB) | P0,P1 16 KiB each | barriers.


namespace {
namespace tc {

constexpr int kD = 128;            // head dim of this instance
constexpr int kBlockKV = 64;       // kv tokens per pipeline stage
constexpr int kStages = 3;
constexpr int kQTiles = 2;
constexpr int kSoftmaxThreads = 128;                 // per query tile
constexpr int kMmaWarp = 8;
constexpr int kProducerWarp0 = 9;
constexpr int kProducerWarps = 4;
constexpr int kProducerThreads = kProducerWarps * 32;
constexpr int kThreads = (kProducerWarp0 + kProducerWarps) * 32;   // 416

constexpr int kRowBytes = kD * 2;                    // one K/V/Q row of one head (16-bit elements)
constexpr int kQTileBytes = 128 * kRowBytes;         // 32 KiB: two 128-row x 128-byte swizzle blocks
constexpr int kKTileBytes = kBlockKV * kRowBytes;    // 16 KiB: two 64-row x 128-byte blocks
constexpr int kStageBytes = 2 * kKTileBytes;         // K then V
constexpr int kPTileBytes = 128 * kBlockKV * 2;      // 16 KiB: one 128-row x 128-byte block

constexpr int kSmemQ = 0;
constexpr int kSmemKV = kSmemQ + kQTiles * kQTileBytes;
constexpr int kSmemP = kSmemKV + kStages * kStageBytes;
constexpr int kSmemBar = kSmemP + kQTiles * kPTileBytes;
constexpr int kSmemBytes = kSmemBar + 256 + 1024;    // barriers + slack for the 1024-byte alignment

constexpr int kTmemCols = 512;
constexpr int kTmemS = 0, kTmemO = 128;

// barrier slots (8 bytes each)
enum {
  kBarQFull = 0,
  kBarKVFull = 1,                    // kStages
  kBarKVEmpty = kBarKVFull + kStages,
  kBarSFull = kBarKVEmpty + kStages, // kQTiles each below
  kBarSFree = kBarSFull + kQTiles,
  kBarPFull = kBarSFree + kQTiles,
  kBarPVDone = kBarPFull + kQTiles,
  kNumBars = kBarPVDone + kQTiles,
};

// byte offset of (row, 16-byte chunk) inside a tile made of 128-byte-wide swizzle blocks of `rows` rows
__device__ __forceinline__ uint32_t swz_off(int row, int chunk, int rows) {
  return (uint32_t)((chunk >> 3) * rows * 128 + row * 128 + (((chunk & 7) ^ (row & 7)) << 4));
}

}  // namespace tc

template <typename QT>
__global__ void __launch_bounds__(tc::kThreads, 1) extend_tc_kernel(const ExtendParams p, const int max_tiles) {
  using namespace tc;
  using namespace ptx;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  auto bar = [&](int i) -> uint32_t { return smem_base + kSmemBar + 8 * i; };
  const uint32_t tmem_slot = smem_base + kSmemBar + 8 * kNumBars;

  // ---- work item -----------------------------------------------------------------------------
  const int Hkv = p.num_kv_heads, g = p.group;
  const int item_tile = blockIdx.x / Hkv, kvh = blockIdx.x - item_tile * Hkv;
  const int num_tiles = p.plan[0];
  if (item_tile >= num_tiles || item_tile >= max_tiles) return;
  const int32_t* qo_indptr = p.plan + kPlanHeader;
  const int32_t* tiles = p.plan + kPlanHeader + p.max_bs + 1;
  const int req = tiles[2 * item_tile], q_start = tiles[2 * item_tile + 1];
  const int BQ = 128 / g;                                 // query tokens per 128-row tile
  const int P = (int)load_index(p.extend_prefix_lens, p.lens_is_i64, req);
  const int E = (int)load_index(p.extend_seq_lens, p.lens_is_i64, req);
  const int64_t pool_row = load_index(p.req_pool_indices, p.req_pool_indices_is_i64, req);
  const int q_tok0 = qo_indptr[req] + q_start;            // first global q row of this item
  const int n_q = min(kQTiles * BQ, E - q_start);          // valid query tokens in this item
  const int kv_end = p.causal ? P + q_start + n_q : P + E; // kv positions [0, kv_end) are needed
  const int n_kv_tiles = (kv_end + kBlockKV - 1) / kBlockKV;

  // ---- one-time setup --------------------------------------------------------------------------
  if (threadIdx.x == 0) {
    mbar_init(bar(kBarQFull), kProducerThreads);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(bar(kBarKVFull + s), kProducerThreads);
      mbar_init(bar(kBarKVEmpty + s), 1);
    }
    for (int t = 0; t < kQTiles; ++t) {
      mbar_init(bar(kBarSFull + t), 1);
      mbar_init(bar(kBarSFree + t), kSoftmaxThreads);
      mbar_init(bar(kBarPFull + t), kSoftmaxThreads);
      mbar_init(bar(kBarPVDone + t), 1);
    }
    mbar_fence_init();
  }
  if (warp == kMmaWarp) tmem_alloc<kTmemCols>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  const uint8_t* kc = reinterpret_cast<const uint8_t*>(p.k_cache);
  const uint8_t* vc = reinterpret_cast<const uint8_t*>(p.v_cache);

  if (warp >= kProducerWarp0) {
    // =========================== producers: Q once, then the K/V ring ===========================
    const int pw = warp - kProducerWarp0;
    const int sub = lane >> 4, chunk = lane & 15;
    {
      const uint8_t* qb = reinterpret_cast<const uint8_t*>(p.q);
#pragma unroll 1
      for (int t = 0; t < kQTiles; ++t) {
#pragma unroll 4
        for (int i = 0; i < 16; ++i) {
          const int row = 32 * pw + 2 * i + sub;
          const int tok = row / g, head = row - tok * g;
          const bool valid = row < BQ * g && t * BQ + tok < n_q;
          const int64_t src = ((int64_t)(q_tok0 + t * BQ + tok) * p.q_stride_t + (int64_t)(kvh * g + head) * p.q_stride_h) * 2;
          cp_async16(smem_base + kSmemQ + t * kQTileBytes + swz_off(row, chunk, 128),
                     qb + (valid ? src : 0) + chunk * 16, valid);
        }
      }
      cp_async_commit_group();
    }
    const int32_t* pt = p.req_to_token + pool_row * p.req_to_token_stride;
    auto load_slot = [&](int j) -> int {
      const int kv = j * kBlockKV + 16 * pw + (lane & 15);
      return (j < n_kv_tiles && kv < kv_end) ? pt[kv] : 0;
    };
    int slot_next = load_slot(0);
    const int64_t k_slot_b = p.k_stride_s * 2, v_slot_b = p.v_stride_s * 2;
    const int64_t k_head_b = (int64_t)kvh * p.k_stride_h * 2, v_head_b = (int64_t)kvh * p.v_stride_h * 2;
#pragma unroll 1
    for (int j = 0; j <= n_kv_tiles; ++j) {
      if (j < n_kv_tiles) {
        const int s = j % kStages;
        if (j >= kStages) mbar_wait(bar(kBarKVEmpty + s), ((j / kStages) - 1) & 1, 10 + s);
        const int slot_cur = slot_next;
        slot_next = load_slot(j + 1);
        const uint32_t kst = smem_base + kSmemKV + s * kStageBytes, vst = kst + kKTileBytes;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int rl = 2 * i + sub, row = 16 * pw + rl;
          const int64_t slot = __shfl_sync(0xffffffffu, slot_cur, rl);
          const bool valid = j * kBlockKV + row < kv_end;
          const uint32_t off = swz_off(row, chunk, kBlockKV);
          cp_async16(kst + off, kc + slot * k_slot_b + k_head_b + chunk * 16, valid);
          cp_async16(vst + off, vc + slot * v_slot_b + v_head_b + chunk * 16, valid);
        }
      }
      cp_async_commit_group();
      cp_async_wait_group<1>();           // everything but the newest group has landed
      fence_proxy_async_smem();           // generic-proxy writes -> visible to the tensor core (async proxy)
      if (j == 0) mbar_arrive(bar(kBarQFull));
      else mbar_arrive(bar(kBarKVFull + (j - 1) % kStages));
    }
    cp_async_wait_group<0>();
  } else if (warp == kMmaWarp) {
    // =========================== MMA issuer =====================================================
    // All 32 lanes run the (warp-uniform) control flow and descriptor arithmetic so it stays on the
    // uniform datapath; one elected lane issues tcgen05.mma / tcgen05.commit.
    {
      constexpr int kFmt = dtype_of<QT>::value == B200_BF16 ? 1 : 0;
      constexpr uint32_t idesc_s = umma_idesc_f16(128, kBlockKV, kFmt, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc_f16(128, kD, kFmt, 0, 1);
      // descriptor = {hi: SBO 1024 B | version 1 | SWIZZLE_128B, lo: LBO | start address >> 4}
      constexpr uint32_t kHi = (1024u >> 4) | (1u << 14) | (2u << 29);
      constexpr uint32_t kLboK = (16u >> 4) << 16;                    // unused by swizzled K-major operands
      constexpr uint32_t kLboV = ((kBlockKV * 128u) >> 4) << 16;      // V (MN-major): 64-wide d halves 8 KiB apart
      auto mk = [](uint32_t lo, uint32_t hi) -> uint64_t { return ((uint64_t)hi << 32) | lo; };
      const bool leader = elect_one();
      const uint32_t q_lo = ((smem_base + kSmemQ) >> 4) | kLboK;
      const uint32_t p_lo = ((smem_base + kSmemP) >> 4) | kLboK;
      const uint32_t kv_lo = (smem_base + kSmemKV) >> 4;
      auto issue_pv = [&](int jj) {
        const int s = jj % kStages;
        const uint32_t v_lo = (kv_lo + s * (kStageBytes >> 4) + (kKTileBytes >> 4)) | kLboV;
#pragma unroll
        for (int t = 0; t < kQTiles; ++t) {
          mbar_wait(bar(kBarPFull + t), jj & 1, 20 + t);
          tc_fence_after_sync();
#pragma unroll
          for (int ks = 0; ks < kBlockKV / 16; ++ks) {
            // A = P_t (K-major, 32 B per k-step); B = 16 V rows per k-step (2 KiB)
            const uint64_t a = mk(p_lo + t * (kPTileBytes >> 4) + ks * 2, kHi);
            const uint64_t b = mk(v_lo + ks * (2048 >> 4), kHi);
            if (leader) umma_f16_ss(tmem_base + kTmemO + t * kD, a, b, idesc_pv, (jj > 0 || ks > 0) ? 1u : 0u);
          }
          if (leader) umma_commit(bar(kBarPVDone + t));
        }
        if (leader) umma_commit(bar(kBarKVEmpty + s));
      };
      mbar_wait(bar(kBarQFull), 0, 30);
      tc_fence_after_sync();
#pragma unroll 1
      for (int j = 0; j < n_kv_tiles; ++j) {
        const int s = j % kStages;
        mbar_wait(bar(kBarKVFull + s), (j / kStages) & 1, 31 + s);
        tc_fence_after_sync();
        const uint32_t k_lo = (kv_lo + s * (kStageBytes >> 4)) | kLboK;
#pragma unroll
        for (int t = 0; t < kQTiles; ++t) {
          if (j > 0) {
            mbar_wait(bar(kBarSFree + t), (j - 1) & 1, 40 + t);
            tc_fence_after_sync();
          }
#pragma unroll
          for (int ks = 0; ks < kD / 16; ++ks) {
            const uint64_t a = mk(q_lo + t * (kQTileBytes >> 4) + (ks >> 2) * ((128 * 128) >> 4) + (ks & 3) * 2, kHi);
            const uint64_t b = mk(k_lo + (ks >> 2) * ((kBlockKV * 128) >> 4) + (ks & 3) * 2, kHi);
            if (leader) umma_f16_ss(tmem_base + kTmemS + t * kBlockKV, a, b, idesc_s, ks > 0 ? 1u : 0u);
          }
          if (leader) umma_commit(bar(kBarSFull + t));
        }
        if (j > 0) issue_pv(j - 1);
      }
      issue_pv(n_kv_tiles - 1);
    }
    __syncwarp();
  } else {
    // =========================== softmax / correction / epilogue ================================
    const int t = warp >> 2;                                  // query tile of this warpgroup
    const int row = (warp & 3) * 32 + lane;                   // MMA row == TMEM lane
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    const int tok = row / g, head = row - tok * g;
    const bool row_valid = row < BQ * g && t * BQ + tok < n_q;
    const int limit = (row_valid && p.causal) ? P + q_start + t * BQ + tok : kv_end - 1;  // last visible kv index
    const float scale_log2 = p.sm_scale * p.k_scale * kLog2e;
    const uint32_t s_addr = tmem_base + lane_base + kTmemS + t * kBlockKV;
    const uint32_t o_addr = tmem_base + lane_base + kTmemO + t * kD;
    const uint32_t p_row = smem_base + kSmemP + t * kPTileBytes + row * 128;
    float m_run = -INFINITY, l_run = 0.f;   // m_run in raw (unscaled) score units

#pragma unroll 1
    for (int j = 0; j < n_kv_tiles; ++j) {
      mbar_wait(bar(kBarSFull + t), j & 1, 50 + t);
      tc_fence_after_sync();
      uint32_t sv[2][32];
      tmem_ld32(s_addr, sv[0]);
      tmem_ld32(s_addr + 32, sv[1]);
      tmem_ld_wait();
      tc_fence_before_sync();
      mbar_arrive(bar(kBarSFree + t));                        // S_t may be overwritten by the next QK^T

      const int kv0 = j * kBlockKV;
      if (kv0 + kBlockKV - 1 > limit) {                       // only the diagonal / ragged tiles need the mask
#pragma unroll
        for (int i = 0; i < 64; ++i)
          if (kv0 + i > limit) sv[i >> 5][i & 31] = 0xff800000u;   // -inf
      }
      // running max in raw score units (scale > 0); 3-input max halves the instruction count
      float mloc = __uint_as_float(sv[0][0]);
#pragma unroll
      for (int i = 1; i < 63; i += 2)
        mloc = fmaxf(mloc, fmaxf(__uint_as_float(sv[i >> 5][i & 31]), __uint_as_float(sv[(i + 1) >> 5][(i + 1) & 31])));
      mloc = fmaxf(mloc, __uint_as_float(sv[1][31]));
      float m_new = fmaxf(m_run, mloc);
      // lazy rescale: keep the old reference max while the new one is < 2^8 larger (P stays <= 256)
      if ((m_new - m_run) * scale_log2 <= 8.f) m_new = m_run;
      const float corr = fast_exp2((m_run - m_new) * scale_log2);
      const float neg = -m_new * scale_log2;
      float2 sum2 = make_float2(0.f, 0.f);
      uint32_t pk[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        const int a = 2 * i, b = 2 * i + 1;
        float2 e = ffma2(make_float2(__uint_as_float(sv[a >> 5][a & 31]), __uint_as_float(sv[b >> 5][b & 31])),
                         make_float2(scale_log2, scale_log2), make_float2(neg, neg));
        e.x = fast_exp2(e.x);
        e.y = fast_exp2(e.y);
        sum2 = fadd2(sum2, e);
        pk[i] = pack2<QT>(e.x, e.y);
      }
      l_run = l_run * corr + (sum2.x + sum2.y);
      m_run = m_new;

      if (j > 0) {
        mbar_wait(bar(kBarPVDone + t), (j - 1) & 1, 60 + t);  // O_t stable, P_t buffer free
        tc_fence_after_sync();
        if (__any_sync(0xffffffffu, corr != 1.f)) {           // tcgen05.ld/st are warp-collective
#pragma unroll 1
          for (int c = 0; c < kD / 32; ++c) {
            uint32_t ov[32];
            tmem_ld32(o_addr + 32 * c, ov);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) ov[i] = __float_as_uint(__uint_as_float(ov[i]) * corr);
            tmem_st32(o_addr + 32 * c, ov);
          }
          tmem_st_wait();
        }
      }
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(p_row + ((c ^ (row & 7)) << 4)),
                     "r"(pk[4 * c]), "r"(pk[4 * c + 1]), "r"(pk[4 * c + 2]), "r"(pk[4 * c + 3])
                     : "memory");
      }
      fence_proxy_async_smem();
      tc_fence_before_sync();
      mbar_arrive(bar(kBarPFull + t));
    }

    // ---- epilogue: O_t / l -> global ----
    mbar_wait(bar(kBarPVDone + t), (n_kv_tiles - 1) & 1, 70 + t);
    tc_fence_after_sync();
    const float inv = p.v_scale / l_run;
    QT* orow = reinterpret_cast<QT*>(p.o) + (int64_t)(q_tok0 + t * BQ + tok) * p.o_stride_t +
               (int64_t)(kvh * g + head) * p.o_stride_h;
#pragma unroll 1
    for (int c = 0; c < kD / 32; ++c) {
      uint32_t ov[32];
      tmem_ld32(o_addr + 32 * c, ov);
      tmem_ld_wait();
      if (row_valid) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          uint4 w;
          w.x = pack2<QT>(__uint_as_float(ov[8 * i]) * inv, __uint_as_float(ov[8 * i + 1]) * inv);
          w.y = pack2<QT>(__uint_as_float(ov[8 * i + 2]) * inv, __uint_as_float(ov[8 * i + 3]) * inv);
          w.z = pack2<QT>(__uint_as_float(ov[8 * i + 4]) * inv, __uint_as_float(ov[8 * i + 5]) * inv);
          w.w = pack2<QT>(__uint_as_float(ov[8 * i + 6]) * inv, __uint_as_float(ov[8 * i + 7]) * inv);
          *reinterpret_cast<uint4*>(orow + 32 * c + 8 * i) = w;
        }
      }
    }
    if (p.lse != nullptr && row_valid)
      p.lse[(int64_t)(q_tok0 + t * BQ + tok) * p.num_q_heads + kvh * g + head] = (m_run * scale_log2 + log2f(l_run)) * kLn2;
    tc_fence_before_sync();
  }

  // ---- teardown ----
  __syncthreads();
  if (warp == kMmaWarp) {
    tc_fence_after_sync();
    tmem_dealloc<kTmemCols>(tmem_base);
  }
}

bool extend_fast_supported(int dqk, int dv, int group, int q_dtype, int kv_dtype) {
  if (dqk != tc::kD || dv != tc::kD) return false;
  if (group < 1 || group > 128) return false;
  if (q_dtype != B200_BF16 && q_dtype != B200_F16) return false;
  return kv_dtype == q_dtype;   // FP8 pools take the generic kernel until the dequantising producer lands
}

// query tokens per work item (two 128-row tiles of 128/group tokens each)
int extend_fast_block_q(int group) { return tc::kQTiles * (128 / group); }

bool extend_fast_layout_ok(const ExtendParams& p, int kv_dtype) {
  (void)kv_dtype;
  auto a16 = [](const void* ptr) { return ((uintptr_t)ptr & 15) == 0; };
  return a16(p.k_cache) && a16(p.v_cache) && a16(p.q) && a16(p.o) && p.k_stride_s % 8 == 0 &&
         p.k_stride_h % 8 == 0 && p.v_stride_s % 8 == 0 && p.v_stride_h % 8 == 0 && p.q_stride_t % 8 == 0 &&
         p.q_stride_h % 8 == 0 && p.o_stride_t % 8 == 0 && p.o_stride_h % 8 == 0;
}

template <typename QT>
int launch_extend_fast_t(const ExtendParams& p, const b200_extend_args& a, cudaStream_t stream) {
  static bool configured = false;
  auto kern = extend_tc_kernel<QT>;
  if (!configured) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::kSmemBytes));
    configured = true;
  }
  const int block_q = extend_fast_block_q(p.group);
  // host-side bound on the number of work items: sum_r ceil(E_r / block_q) <= T / block_q + bs
  int64_t tiles = a.num_q_tokens / block_q + a.batch_size;
  if (tiles > a.max_tiles) tiles = a.max_tiles;
  const int64_t grid = tiles * p.num_kv_heads;
  B200_CHECK_ARG(grid > 0 && grid < 0x7fffffff, "extend: bad grid %lld", (long long)grid);
  kern<<<(unsigned)grid, tc::kThreads, tc::kSmemBytes, stream>>>(p, (int)tiles);
  B200_CHECK_CUDA(cudaGetLastError());
  return B200_OK;
}

int launch_extend_fast(const ExtendParams& p, const b200_extend_args& a, cudaStream_t stream) {
  if (a.q_dtype == B200_BF16) return launch_extend_fast_t<__nv_bfloat16>(p, a, stream);
  return launch_extend_fast_t<__half>(p, a, stream);
}

}  // namespace
