// CTA-pair flavour of the tcgen05 extend kernel (textually included by extend_tc.cuh; 16-bit pools, D = 128).
//
// A cluster of two CTAs (the two SMs of one TPC) works on ONE item = (request, kv head, 4*BQ query tokens): CTA c
// owns query tokens [2 c BQ, 2 c BQ + 2 BQ) of the item -- its own two ping-pong query tiles, softmax warpgroups,
// TMEM accumulators and epilogue, exactly as in the one-CTA kernel -- but the K/V stream is SHARED: every MMA is a
// `tcgen05.mma.cta_group::2` of M = 256 (rows [0,128) from CTA 0, [128,256) from CTA 1) whose B operand comes half
// from each CTA's shared memory:
//
//   S_t = Q_t K_j^T   B = K_j, N = 128 kv rows:  CTA c holds rows [64 c, 64 c + 64)           (K-major, SW128)
//   O_t += P_t V_j    B = V_j, N = 128 d:        CTA c holds d in [64 c, 64 c + 64), 128 rows  (MN-major, SW128)
//
// so each CTA gathers HALF of every K and V block (16 KiB ring slots instead of 32 KiB, 10 slots deep): half the
// LDGSTS instructions and L2->SM bytes per SM, and the SS-mode QK reads 6 KiB instead of 8 KiB of shared memory per
// k-step (it sat exactly on the 128 B/clk limit).  One thread of CTA 0 issues all MMAs; `tcgen05.commit` is multicast
// to the same barrier offsets in both CTAs.  What CTA 0 must learn from CTA 1 arrives on CTA 0's barriers through
// the cluster: the peer's "slot filled" (a relay lane in CTA 1's otherwise idle MMA warp: wait local Full -> proxy
// fence -> remote arrive) and the peer's "P_t written" (its softmax warps arrive remotely).
//
// Causal rule: both CTAs walk the kv range of the pair's LAST query; for CTA 0 the (at most one) extra block is fully
// masked, which the online softmax handles without a special case (block 0 always holds a visible key).

namespace tc {

struct PairLayout {
  static constexpr int kSlots = 10;                       // ring slots, half a K or V block each
  static constexpr int kSlotBytes = kBlockKV * kRowBytes / 2;   // 16 KiB
  static constexpr int kKHalfBytes = 64 * 128;            // K half block: [d half][64 rows][128 B]
  static constexpr int kSmemQ = 0;
  static constexpr int kSmemKV = kSmemQ + kQTiles * kQTileBytes;
  static constexpr int kSmemBar = kSmemKV + kSlots * kSlotBytes;
  static constexpr int kSmemBytes = kSmemBar + 256 + 1024;
  static constexpr int kBarQFull = 0;
  static constexpr int kBarFull = 1;
  static constexpr int kBarEmpty = kBarFull + kSlots;
  static constexpr int kBarSFull = kBarEmpty + kSlots;
  static constexpr int kBarPFull = kBarSFull + kQTiles;
  static constexpr int kBarODone = kBarPFull + kQTiles;
  static constexpr int kNumBars = kBarODone + kQTiles;
  static_assert(kNumBars * 8 + 8 <= 256, "barrier area");
  static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
};

}  // namespace tc

template <typename QT, bool kOpts>
__global__ void __launch_bounds__(tc::kThreads, 1) extend_tc_pair_kernel(const ExtendParams p, const int max_tiles) {
  using namespace tc;
  using namespace ptx;
  using L = PairLayout;
  using CT = QT;
  constexpr int kHD = kD, kEmu = kEmuEighths;
  constexpr int kSlots = L::kSlots, kSlotBytes = L::kSlotBytes, kSmemQ = L::kSmemQ, kSmemKV = L::kSmemKV,
                kSmemBar = L::kSmemBar, kNumBars = L::kNumBars;
  constexpr int kBarQFull = L::kBarQFull, kBarFull = L::kBarFull, kBarEmpty = L::kBarEmpty,
                kBarSFull = L::kBarSFull, kBarPFull = L::kBarPFull, kBarODone = L::kBarODone;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;   // same offset in both CTAs of the pair
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  auto bar = [&](int i) -> uint32_t { return smem_base + kSmemBar + 8 * i; };
  const uint32_t tmem_slot = smem_base + kSmemBar + 8 * kNumBars;
  const uint32_t rank = cluster_ctarank();                 // 0 = leader (issues the MMAs)

  // ---- work item (identical in both CTAs) ----------------------------------------------------------------
  const int Hkv = p.num_kv_heads, g = p.group;
  const int cid = blockIdx.x >> 1;
  const int item_tile = cid / Hkv, kvh = cid - item_tile * Hkv;
  const int num_tiles = p.plan[0];
  if (item_tile >= num_tiles || item_tile >= max_tiles) return;
  const int32_t* qo_indptr = p.plan + kPlanHeader;
  const int32_t* tiles = p.plan + kPlanHeader + p.max_bs + 1;
  const int req = tiles[2 * item_tile], pair_q_start = tiles[2 * item_tile + 1];
  const int BQ = 128 / g;
  if (p.plan[1] != 2 * kQTiles * BQ) {                     // the work list was planned for another kernel's tile size
    if (threadIdx.x == 0) printf("b200attn extend (pair): plan block_q %d != %d\n", p.plan[1], 2 * kQTiles * BQ);
    __trap();
  }
  const int P = (int)load_index(p.extend_prefix_lens, p.lens_is_i64, req);
  const int E = (int)load_index(p.extend_seq_lens, p.lens_is_i64, req);
  const int64_t pool_row = load_index(p.req_pool_indices, p.req_pool_indices_is_i64, req);
  const int pair_n_q = min(2 * kQTiles * BQ, E - pair_q_start);
  // this CTA's own query tokens
  const int q_start = pair_q_start + (int)rank * kQTiles * BQ;
  const int n_q = max(0, min(kQTiles * BQ, pair_n_q - (int)rank * kQTiles * BQ));
  const int q_tok0 = qo_indptr[req] + q_start;
  const bool masked = kOpts && p.custom_mask != nullptr;
  const int kv_end = (p.causal && !masked) ? P + pair_q_start + pair_n_q : P + E;
  const int kv_first = (kOpts && p.window_keys > 0 && P + pair_q_start + 1 > p.window_keys)
                           ? P + pair_q_start + 1 - p.window_keys : 0;
  const int kv_off = kv_first / kBlockKV * kBlockKV;
  const int n_blocks = (kv_end - kv_off + kBlockKV - 1) / kBlockKV;

  // ---- one-time setup ---------------------------------------------------------------------------------------
  if (threadIdx.x == 0) {
    const int peer = rank == 0 ? 1 : 0;                    // the leader's fill barriers also count the peer's relay
    mbar_init(bar(kBarQFull), kProducerThreads + peer);
    for (int s = 0; s < kSlots; ++s) {
      mbar_init(bar(kBarFull + s), kProducerThreads + peer);
      mbar_init(bar(kBarEmpty + s), 1);
    }
    for (int t = 0; t < kQTiles; ++t) {
      mbar_init(bar(kBarSFull + t), 1);
      mbar_init(bar(kBarPFull + t), 2 * kSoftmaxWarps);    // both CTAs' softmax warps (used in the leader only)
      mbar_init(bar(kBarODone + t), 1);
    }
    mbar_fence_init();
  }
  if (warp == kMmaWarp) tmem_alloc_pair<kTmemCols>(tmem_slot);
  tc_fence_before_sync();
  cluster_sync_all();                                      // barriers of both CTAs exist before any remote arrive
  tc_fence_after_sync();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp >= kProducerWarp0) {
    setmaxnreg_dec<kRegsOther>();
    // ---------------------------------------------------------------- producers: Q, then HALF of every K / V block
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
          const int64_t src =
              ((int64_t)(q_tok0 + t * BQ + tok) * p.q_stride_t + (int64_t)(kvh * g + head) * p.q_stride_h) * 2;
          cp_async16(smem_base + kSmemQ + t * kQTileBytes + swz_off(row, chunk, 128),
                     qb + (valid ? src : 0) + chunk * 16, valid);
        }
      }
      cp_async_mbar_arrive_noinc(bar(kBarQFull));
    }
    const int32_t* pt = p.req_to_token + pool_row * p.req_to_token_stride;
    // K: this CTA's 64 rows of the block; warp pw gathers rows [16 pw, 16 pw + 16), a half-warp per 256-byte row.
    // V: all 128 rows, this CTA's 128-byte d half; warp pw gathers rows [32 pw, 32 pw + 32), a quarter-warp per row.
    auto load_k_slot = [&](int j) -> int {
      const int kv = kv_off + j * kBlockKV + 64 * (int)rank + 16 * pw + (lane & 15);
      return (j < n_blocks && kv < kv_end) ? __ldg(pt + kv) : 0;
    };
    auto load_v_slot = [&](int j) -> int {
      const int kv = kv_off + j * kBlockKV + 32 * pw + lane;
      return (j < n_blocks && kv < kv_end) ? __ldg(pt + kv) : 0;
    };
    const uint64_t kc64 = reinterpret_cast<uint64_t>(p.k_cache) + (uint64_t)((int64_t)kvh * p.k_stride_h * 2) + chunk * 16;
    const int vq = lane >> 3, vch = lane & 7;              // V: row within a group of 4, 16-byte chunk of the d half
    const uint64_t vc64 = reinterpret_cast<uint64_t>(p.v_cache) + (uint64_t)((int64_t)kvh * p.v_stride_h * 2) +
                          (8 * rank + vch) * 16;
    const uint64_t k_slot_b = (uint64_t)p.k_stride_s * 2, v_slot_b = (uint64_t)p.v_stride_s * 2;
    // destination offsets inside a ring slot (lane-dependent part; the row loop adds immediates)
    //   K row r = 16 pw + 2 i + sub:  (chunk >> 3) * 8 KiB + r * 128 + (((chunk & 7) ^ (r & 7)) << 4)
    //   V row r = 32 pw + 4 i + vq:   r * 128 + ((vch ^ (r & 7)) << 4)
    uint32_t k_dst[4], v_dst[2];
#pragma unroll
    for (int k = 0; k < 4; ++k)
      k_dst[k] = smem_base + kSmemKV + (chunk >> 3) * L::kKHalfBytes + (16 * pw + sub) * 128 +
                 ((((chunk & 7) ^ sub) ^ (2 * k)) << 4);
#pragma unroll
    for (int k = 0; k < 2; ++k)
      v_dst[k] = smem_base + kSmemKV + (32 * pw + vq) * 128 + (((vch ^ vq) ^ (4 * k)) << 4);
    int ks_next = load_k_slot(0), vs_next = load_v_slot(0);
    int fill = 0;                                          // ring sequence number: 2j = K_j, 2j+1 = V_j
#pragma unroll 1
    for (int j = 0; j < n_blocks; ++j) {
      const int ks_cur = ks_next, vs_cur = vs_next;
      ks_next = load_k_slot(j + 1);
      vs_next = load_v_slot(j + 1);
      {   // ---- K half block
        const int s = fill % kSlots;
        if (fill >= kSlots) mbar_wait(bar(kBarEmpty + s), ((fill / kSlots) - 1) & 1, 10 + s);
        const uint32_t slot_off = s * kSlotBytes;
        const int rows_left = kv_end - kv_off - j * kBlockKV - 64 * (int)rank - 16 * pw;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t slot = __shfl_sync(0xffffffffu, ks_cur, 2 * i + sub);
          cp_async16(k_dst[i & 3] + slot_off + i * 256, reinterpret_cast<const void*>(kc64 + (uint64_t)slot * k_slot_b),
                     2 * i + sub < rows_left);
        }
        cp_async_mbar_arrive_noinc(bar(kBarFull + s));
        ++fill;
      }
      {   // ---- V half block
        const int s = fill % kSlots;
        if (fill >= kSlots) mbar_wait(bar(kBarEmpty + s), ((fill / kSlots) - 1) & 1, 10 + s);
        const uint32_t slot_off = s * kSlotBytes;
        const int rows_left = kv_end - kv_off - j * kBlockKV - 32 * pw;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint32_t slot = __shfl_sync(0xffffffffu, vs_cur, 4 * i + vq);
          cp_async16(v_dst[i & 1] + slot_off + i * 512, reinterpret_cast<const void*>(vc64 + (uint64_t)slot * v_slot_b),
                     4 * i + vq < rows_left);
        }
        cp_async_mbar_arrive_noinc(bar(kBarFull + s));
        ++fill;
      }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
  } else if (warp >= kMmaWarp) {
    setmaxnreg_dec<kRegsOther>();
    if (warp == kMmaWarp && rank != 0) {
      // ---------------------------------------------------------------- peer relay: "my half is in" -> leader
      const bool one = elect_one();
      auto relay = [&](int b, uint32_t parity, int tag) {
        mbar_wait(bar(b), parity, tag);
        if (one) {
          fence_proxy_async_smem();      // my producers' cp.async writes -> the leader's tensor-core reads of my smem
          mbar_arrive_remote(mapa_shared(bar(b), 0));
        }
        __syncwarp();
      };
      relay(kBarQFull, 0, 80);
#pragma unroll 1
      for (int fill = 0; fill < 2 * n_blocks; ++fill) relay(kBarFull + fill % kSlots, (fill / kSlots) & 1, 81);
    } else if (warp == kMmaWarp) {
      // ---------------------------------------------------------------- leader: issues every MMA of the pair
      constexpr int kFmt = dtype_of<CT>::value == B200_BF16 ? 1 : 0;
      constexpr uint32_t idesc_s = umma_idesc_f16(256, kBlockKV, kFmt, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc_f16(256, kHD, kFmt, 0, 1);
      constexpr uint32_t kHi = (1024u >> 4) | (1u << 14) | (2u << 29);      // SBO 1024 B | version 1 | SWIZZLE_128B
      constexpr uint32_t kLbo = (16u >> 4) << 16;                           // unused: one swizzle atom per N / K extent
      auto mk = [](uint32_t lo, uint32_t hi) -> uint64_t { return ((uint64_t)hi << 32) | lo; };
      const bool leader = elect_one();
      const uint32_t q_lo = ((smem_base + kSmemQ) >> 4) | kLbo;
      const uint32_t kv_lo = (smem_base + kSmemKV) >> 4;
      auto issue_qk = [&](int t, int s) {
        const uint32_t k_lo = (kv_lo + s * (kSlotBytes >> 4)) | kLbo;
#pragma unroll
        for (int ks = 0; ks < kHD / 16; ++ks) {
          const uint32_t qoff = (ks >> 2) * (kHalfBytes >> 4) + (ks & 3) * 2;          // Q tile: 128-row d halves
          const uint32_t koff = (ks >> 2) * (L::kKHalfBytes >> 4) + (ks & 3) * 2;      // K half block: 64-row d halves
          if (leader)
            umma_f16_ss_pair(tmem_base + kTmemS + t * kBlockKV, mk(q_lo + t * (kQTileBytes >> 4) + qoff, kHi),
                             mk(k_lo + koff, kHi), idesc_s, ks > 0 ? 1u : 0u);
        }
        if (leader) umma_commit_pair(bar(kBarSFull + t));
      };
      auto issue_pv = [&](int t, int s, uint32_t acc) {
        const uint32_t v_lo = (kv_lo + s * (kSlotBytes >> 4)) | kLbo;
#pragma unroll
        for (int ks = 0; ks < kBlockKV / 16; ++ks) {
          if (leader)
            umma_f16_ts_pair(tmem_base + kTmemO + t * kD, tmem_base + kTmemS + t * kBlockKV + ks * 8,
                             mk(v_lo + ks * (2048 >> 4), kHi), idesc_pv, (acc | (uint32_t)(ks > 0)));
        }
      };
      auto wait_full = [&](int fill) {
        mbar_wait_cluster(bar(kBarFull + fill % kSlots), (fill / kSlots) & 1, 30 + fill % kSlots);
        fence_proxy_async_smem();
        tc_fence_after_sync();
      };
      mbar_wait_cluster(bar(kBarQFull), 0, 29);
      fence_proxy_async_smem();
      wait_full(0);
      issue_qk(0, 0);
      issue_qk(1, 0);
      if (leader) umma_commit_pair(bar(kBarEmpty + 0));
#pragma unroll 1
      for (int j = 0; j < n_blocks; ++j) {
        const int fv = 2 * j + 1, fk = 2 * j + 2;
        const bool more = j + 1 < n_blocks;
        wait_full(fv);
        if (more) wait_full(fk);
#pragma unroll
        for (int t = 0; t < kQTiles; ++t) {
          mbar_wait_cluster(bar(kBarPFull + t), j & 1, 40 + t);
          tc_fence_after_sync();
          issue_pv(t, fv % kSlots, j > 0 ? 1u : 0u);
          if (t == kQTiles - 1 && leader) umma_commit_pair(bar(kBarEmpty + fv % kSlots));
          if (more) issue_qk(t, fk % kSlots);
          else if (leader) umma_commit_pair(bar(kBarODone + t));
        }
        if (more && leader) umma_commit_pair(bar(kBarEmpty + fk % kSlots));
      }
      __syncwarp();
    }
  } else {
    setmaxnreg_inc<kRegsSoftmax>();
    const uint32_t p_full_leader = mapa_shared(bar(kBarPFull), 0);
    auto signal_p_ready = [&](int t_) {
      if (rank == 0) mbar_arrive(bar(kBarPFull + t_));
      else mbar_arrive_remote(p_full_leader + 8 * t_);
    };
#include "extend_tc_softmax.inc"
  }

  // ---- teardown: the leader's MMAs read the peer's shared memory and write its TMEM until the very end ----
  tc_fence_before_sync();
  cluster_sync_all();
  if (warp == kMmaWarp) {
    tc_fence_after_sync();
    tmem_dealloc_pair<kTmemCols>(tmem_base);
  }
}

// query tokens per PAIR item
int extend_pair_block_q(int group) { return 2 * tc::kQTiles * (128 / group); }

template <typename QT, bool kOpts>
int launch_extend_pair_t(const ExtendParams& p, const b200_extend_args& a, cudaStream_t stream) {
  static std::atomic<uint64_t> configured{0};
  constexpr int kSmem = tc::PairLayout::kSmemBytes;
  auto kern = extend_tc_pair_kernel<QT, kOpts>;
  B200_CHECK_CUDA(set_max_dynamic_smem(kern, kSmem, configured));
  const int block_q = extend_pair_block_q(p.group);
  int64_t tiles = a.num_q_tokens / block_q + a.batch_size;
  if (tiles > a.max_tiles) tiles = a.max_tiles;
  const int64_t grid = tiles * p.num_kv_heads * 2;
  B200_CHECK_ARG(grid > 0 && grid < 0x7fffffff, "extend: bad grid %lld", (long long)grid);
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)grid, 1, 1);
  cfg.blockDim = dim3(tc::kThreads, 1, 1);
  cfg.dynamicSmemBytes = kSmem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B200_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, p, (int)tiles));
  return B200_OK;
}
