// tcgen05 / TMEM extend (prefill over a cached prefix) kernel, included by extend.cu inside namespace b200.
//
// One CTA per work item = (request, kv head, block of 2*BQ query tokens).  The 128 MMA rows of a
// query tile hold BQ = 128/group tokens x `group` query heads, so K/V are read once per GQA group.
// Both GEMMs of a kv block are M128 x N128 tcgen05.mma (measured on B200: an N=64 instruction costs
// 48 cycles instead of 32, N=128 runs at the 64-cycle floor from shared memory or TMEM alike):
//
//   S_t = Q_t K_j^T    A = Q_t (smem, K-major)   B = K_j (smem, K-major)   D = S_t  (TMEM, 128 fp32 columns)
//   O_t += P_t V_j     A = P_t (TMEM, 16-bit, aliased onto the first 64 columns of S_t)
//                      B = V_j rows (smem, MN-major: no transpose, K and V share one fill pattern)
//
// Two query tiles ping-pong: the tensor pipe runs  PV_0(j) QK_0(j+1) PV_1(j) QK_1(j+1)  while the
// softmax warpgroup of the other tile turns S into P.  512 threads, register budget re-balanced with
// setmaxnreg (the softmax threads hold a whole 128-column row of S):
//
//   warps 0-3    softmax / O correction / epilogue of query tile 0 (one thread per TMEM lane = MMA row)
//   warps 4-7    the same for query tile 1
//   warp  8      TMEM allocation + MMA issue (warp-uniform control flow, one elected lane issues)
//   warps 9-11   idle (they only give their registers away)
//   warps 12-15  producers.  Q once, with 16-byte cp.async into 128B-swizzled tiles.  K/V ring:
//                * 16-bit pools with page_size >= 16 (kTma): ONE warp issues TMA `cp.async.bulk.tensor.2d` -- a
//                  page's one-head slice is a strided box [page rows x 64 words] of the pool, landed with the
//                  hardware 128B swizzle, completion counted in bytes on the slot's mbarrier (no generic-proxy
//                  fence, no LDGSTS issue slots)
//                * page_size 1 (SGLang's default), FP8 pools, strided pools: page-table gather with 16-byte cp.async
//                  by all four warps (FP8 bytes are widened to f16 on the way), completion by
//                  cp.async.mbarrier.arrive.  (tile::gather4 was measured 2.35x slower than LDGSTS at 128-byte
//                  rows, profiles/r2_extend_experiments.md.)
//
// TMEM (512 columns): S0 [0,128)  S1 [128,256)  O0 [256,384)  O1 [384,512);  P_t = S_t columns [0,64).
// Shared memory: Q0,Q1 32 KiB each | ring of 5 x 32 KiB slots filled K0 V0 K1 V1 ... | barriers.


// Timing-aid ablations exist only in -DB200_DEV builds (benchmarks/_ab/build_variant.sh): skip the ring refills /
// the exponentials.  The shipped library compiles neither.
#if !defined(B200_DEV) || !defined(B200_ABLATE_FILL)
#undef B200_ABLATE_FILL
#define B200_ABLATE_FILL 0
#endif
#if !defined(B200_DEV) || !defined(B200_ABLATE_EXP)
#undef B200_ABLATE_EXP
#define B200_ABLATE_EXP 0
#endif

// Timeline trace of one CTA (dev builds with -DB200_TRACE=<block index>): lane 0 of selected warps stamps clock64 at
// each hand-off of every kv block; read back with b200attn_dev_trace_read (benchmarks/_ab/trace_extend.py).
// -DB200_TRACE_WARPS instead stamps every softmax warp's P hand-off (benchmarks/_ab/trace_warps.py).
#if defined(B200_DEV) && defined(B200_TRACE)
__device__ long long g_ext_trace[16 * 512];
#define B200_TR_CORE(slot, j)                                                                \
  do {                                                                                       \
    if (blockIdx.x == (B200_TRACE) && (threadIdx.x & 31) == 0 && (j) < 512) {                \
      long long c_;                                                                          \
      asm volatile("mov.u64 %0, %%clock64;" : "=l"(c_)::"memory");                           \
      g_ext_trace[(slot) * 512 + (j)] = c_;                                                  \
    }                                                                                        \
  } while (0)
#if defined(B200_TRACE_WARPS)
#define B200_TR2(slot, j) B200_TR_CORE(slot, j)
#define B200_TR(slot, j)
#else
#define B200_TR(slot, j) B200_TR_CORE(slot, j)
#define B200_TR2(slot, j)
#endif
#else
#define B200_TR(slot, j)
#define B200_TR2(slot, j)
#endif

// tuning switches of the softmax pass (defaults = the shipped choice; dev builds override them for A/B runs)
#ifndef B200_SUM_LATE
#define B200_SUM_LATE 1
#endif
#ifndef B200_EMU
#define B200_EMU 4      // sixteenths of the column pairs that take the polynomial exp2
#endif
#ifndef B200_RING_FENCE
#define B200_RING_FENCE 1   // fence.proxy.async in the MMA warp after every ring-slot wait (A/B switch of dev builds)
#endif
#ifndef B200_EMU_ALT
#define B200_EMU_ALT 1      // spread them over the row instead of taking the first ones of every 16
#endif

namespace {
namespace tc {

constexpr int kD = 128;            // head dim the tiles are laid out for (kHD = 64 uses the first 64-wide half)
constexpr int kBlockKV = 128;      // kv tokens per MMA block
constexpr int kQTiles = 2;
constexpr int kSoftmaxWarps = 4;                     // per query tile
constexpr int kMmaWarp = 8;
constexpr int kProducerWarp0 = 12;
constexpr int kProducerWarps = 4;
constexpr int kProducerThreads = kProducerWarps * 32;
#ifndef B200_ROWS_WARP0
#define B200_ROWS_WARP0 8
#endif
constexpr int kRowsWarp0 = B200_ROWS_WARP0;          // K/V rows per block gathered by the producer warp next to the MMA warp
constexpr int kRowsOther = (kBlockKV - kRowsWarp0) / 3;   // (0: the others take 48 / 40 / 40)
static_assert(kRowsWarp0 == 0 || (kRowsWarp0 % 8 == 0 && kRowsOther % 8 == 0 && kRowsWarp0 + 3 * kRowsOther == kBlockKV),
              "row split");
constexpr int kThreads = 512;

constexpr int kRowBytes = kD * 2;                    // one K/V/Q row of one head (16-bit elements)
constexpr int kQTileBytes = 128 * kRowBytes;         // 32 KiB: two 128-row x 128-byte swizzle blocks
constexpr int kSlotBytes = kBlockKV * kRowBytes;     // 32 KiB: two 128-row x 128-byte blocks
constexpr int kHalfBytes = kBlockKV * 128;           // distance between the two 64-wide d halves of a tile

constexpr int kRawBytes = kBlockKV * kD;             // one FP8 K or V block as it sits in the pool (16 KiB)

constexpr int kTmemCols = 512;
constexpr int kTmemS = 0, kTmemO = 256;

constexpr int kRegsSoftmax = 200, kRegsOther = 56;   // 256 x 200 + 256 x 56 = 64 Ki registers
// Of every 8 column pairs, this many take the polynomial exp2 on the FMA pipe instead of MUFU.EX2 (16/clk/SM would
// otherwise cost as many cycles as the MMAs); tuned on B200, both paths agree to 1e-4 relative.
constexpr int kEmuEighths = B200_EMU;

// Shared-memory map and barrier slots.  16-bit pools: 5 ring slots.  FP8 pools: 4 ring slots (the MMA
// operands are f16 tiles the producers write) + 2 raw 16 KiB landing buffers for the cp.async gathers.
template <bool kFp8>
struct Layout {
  static constexpr int kSlots = kFp8 ? 4 : 5;          // ring slots, one K or V block each
  static constexpr int kRawBufs = kFp8 ? 2 : 0;
  static constexpr int kSmemQ = 0;
  static constexpr int kSmemKV = kSmemQ + kQTiles * kQTileBytes;
  static constexpr int kSmemRaw = kSmemKV + kSlots * kSlotBytes;
  static constexpr int kSmemBar = kSmemRaw + kRawBufs * kRawBytes;
  static constexpr int kSmemBytes = kSmemBar + 256 + 1024;    // barriers + slack for the 1024-byte alignment
  // barrier slots (8 bytes each)
  static constexpr int kBarQFull = 0;
  static constexpr int kBarFull = 1;                     // kSlots
  static constexpr int kBarEmpty = kBarFull + kSlots;    // kSlots
  static constexpr int kBarSFull = kBarEmpty + kSlots;   // kQTiles each below
  static constexpr int kBarPFull = kBarSFull + kQTiles;
  static constexpr int kBarODone = kBarPFull + kQTiles;
  static constexpr int kNumBars = kBarODone + kQTiles;
  static_assert(kNumBars * 8 + 8 <= 256, "barrier area");
  static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
};

// byte offset of (row, 16-byte chunk) inside a tile made of 128-byte-wide swizzle blocks of `rows` rows
__device__ __forceinline__ uint32_t swz_off(int row, int chunk, int rows) {
  return (uint32_t)((chunk >> 3) * rows * 128 + row * 128 + (((chunk & 7) ^ (row & 7)) << 4));
}

}  // namespace tc

// kEmu: how many of every 8 column pairs take the polynomial exp2 (FMA pipe) instead of MUFU.EX2.
// kFp8: the pool holds FP8 bytes (e4m3, or e5m2 when p.kv_e5m2); the MMAs then run in f16 (both widenings are exact) on tiles the producers
// dequantise, Q is converted to f16 on its way into shared memory and P is packed as f16.
// kOpts: sliding window / logit cap / attention sinks / custom masks compiled in (the plain causal instance stays
// untouched).
// kHD: head dim, 128 or 64.  The 64 instance keeps the 128 instance's tile geometry and barrier protocol and simply
// never touches the second 64-wide d half: the producers fill 128-byte rows, QK runs 4 k-steps instead of 8, PV is
// an M128 x N64 MMA into 64 O columns.  (Its softmax cost per kv block is the same as for D = 128, so it is
// exp-bound rather than MMA-bound.)
// kTma: the K/V ring may be filled by TMA page boxes through the two tensor maps (16-bit pools, page_size >= 16);
// the instance also carries the row-gather producer and picks per launch from the plan's page check.
// kOneTile: ONE query tile of 128 rows per work item (the plan was built with block_q = 128 / group) -- the instance
// for short extends (TARGET_VERIFY / DRAFT_EXTEND, a few new tokens after a long cached prefix), where the second
// tile of the two-tile kernel would be empty.  The two S regions of TMEM then double-buffer the SAME tile across kv
// blocks: QK(j+2) is issued right behind PV(j), so S(j+1) is already waiting when the softmax warps finish block j
// and the tile's chain is the softmax pass alone instead of softmax + MMAs.  Valid for any extend length.
template <typename QT, int kEmu, bool kFp8, bool kOpts, int kHD, bool kTma, bool kOneTile>
__global__ void __launch_bounds__(tc::kThreads, 1)
extend_tc_kernel(const ExtendParams p, const int max_tiles, const __grid_constant__ CUtensorMap tm_k,
                 const __grid_constant__ CUtensorMap tm_v) {
  static_assert(!(kTma && kFp8), "the FP8 producers widen bytes themselves");
  static_assert(!(kTma && kOneTile), "the one-tile instance keeps the row gather");
  static_assert(kHD != 256 || kOneTile, "D = 256 exists as the one-tile instance only");
  constexpr int kNT = kOneTile ? 1 : tc::kQTiles;           // query tiles of this instance
  using namespace tc;
  using namespace ptx;
  // Tile geometry of this instance (these shadow tc::).  D = 256: rows are 512 bytes = four 64-wide d groups, a kv
  // block is 64 rows so that K_j (or V_j) still fills one 32 KiB ring slot, the Q tile takes both Q areas (64 KiB),
  // S is 128 x 64 and O takes both O regions of TMEM (256 columns): S0 S1 O = 512 columns.
  constexpr int kBlockKV = kHD == 256 ? 64 : tc::kBlockKV;  // kv rows per block
  constexpr int kHalfBytes = kBlockKV * 128;                // distance between the 64-wide d groups of a K/V tile
  constexpr int kQHalfBytes = 128 * 128;                    // ... of a Q tile (always 128 rows)
  using L = Layout<kFp8>;
  using CT = typename std::conditional<kFp8, __half, QT>::type;   // MMA operand type
  constexpr int kSlots = L::kSlots, kSmemQ = L::kSmemQ, kSmemKV = L::kSmemKV, kSmemRaw = L::kSmemRaw,
                kSmemBar = L::kSmemBar, kNumBars = L::kNumBars;
  constexpr int kBarQFull = L::kBarQFull, kBarFull = L::kBarFull, kBarEmpty = L::kBarEmpty,
                kBarSFull = L::kBarSFull, kBarPFull = L::kBarPFull, kBarODone = L::kBarODone;
  (void)kSmemRaw;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  auto bar = [&](int i) -> uint32_t { return smem_base + kSmemBar + 8 * i; };
  const uint32_t tmem_slot = smem_base + kSmemBar + 8 * kNumBars;

  // ---- work item -----------------------------------------------------------------------------
  const int Hkv = p.num_kv_heads, g = p.group;
  // split-KV (one-tile instances only): blockIdx.x = (item_tile * num_splits + split) * Hkv + kvh
  const int n_splits = kOneTile ? p.num_splits : 1;
  const int item_part = blockIdx.x / Hkv, kvh = blockIdx.x - item_part * Hkv;
  const int tile_local = kOneTile ? item_part / n_splits : item_part;       // tile inside this launch's tile range
  const int split = kOneTile ? item_part - tile_local * n_splits : 0;
  const int item_tile = kOneTile ? tile_local + p.tile_offset : item_part;
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
  const int n_q = min(kNT * BQ, E - q_start);              // valid query tokens in this item
  // a custom (tree) mask replaces the causal rule on the new tokens: every row may see all P + E positions
  const bool masked = kOpts && p.custom_mask != nullptr;
  const int kv_end = (p.causal && !masked) ? P + q_start + n_q : P + E;   // kv positions [kv_off, kv_end) are needed
  // sliding window: the first query of the item sees nothing before P + q_start + 1 - window_keys, so whole
  // 128-token blocks below that are skipped; all three roles count blocks from kv_off
  const int kv_first =
      (kOpts && p.window_keys > 0 && P + q_start + 1 > p.window_keys) ? P + q_start + 1 - p.window_keys : 0;
  const int kv_off_item = kv_first / kBlockKV * kBlockKV;
  const int n_blocks_item = (kv_end - kv_off_item + kBlockKV - 1) / kBlockKV;
  // This CTA's part of the item's kv blocks.  The split count is trimmed per item so that a part keeps at least
  // kSplitMinBlocks blocks; parts beyond it only report "saw nothing" (LSE -inf), and an item left with ONE part is
  // finished by split 0 exactly like an unsplit item (it flags its rows with LSE +inf so that the merge skips them).
  int kv_off_part = kv_off_item, n_blocks_part = n_blocks_item;
  bool partial = false;                                   // this CTA writes a partial to the split workspace
  if constexpr (kOneTile) {
    if (n_splits > 1) {
      const int s_eff = max(1, min(n_splits, n_blocks_item / kSplitMinBlocks));
      partial = s_eff > 1;
      const int b0 = (int)((int64_t)n_blocks_item * split / s_eff), b1 = (int)((int64_t)n_blocks_item * (split + 1) / s_eff);
      if (split >= s_eff || !partial) {
        // rows of this item in the LSE block of this split: -inf (no data), or +inf from split 0 of an unsplit item
        const int r = threadIdx.x;
        const int tok = r / g;
        if (r < 128 && r < BQ * g && tok < n_q)
          p.split_lse[((int64_t)split * p.num_q_tokens + q_tok0 + tok) * p.num_q_heads + kvh * g + (r - tok * g)] =
              (split == 0) ? INFINITY : -INFINITY;
        if (split > 0) return;
      } else {
        kv_off_part = kv_off_item + b0 * kBlockKV;
        n_blocks_part = b1 - b0;
      }
    }
  }
  const int kv_off = kv_off_part, n_blocks = n_blocks_part;
  // TMA page boxes only when the plan has verified that every box-sized run of positions of every request sits on
  // consecutive pool rows (b200attn_extend_plan); any other page table is read row by row
  const bool use_tma = kTma && p.plan[kPlanPageRows] == p.tma_rows;

  // ---- one-time setup --------------------------------------------------------------------------
  // D = 256 with FP8 pools is producer-bound (two widened fills per 1 280 MMA cycles): the one-tile layout's idle
  // warps 4-7 (the second tile's softmax warpgroup) fill and widen as well -- 8 producer warps, 8 kv rows each
  constexpr bool kWideProd = kFp8 && kHD == 256;
  constexpr int kProdThreadsAll = kWideProd ? 2 * kProducerThreads : kProducerThreads;
  if (threadIdx.x == 0) {
    mbar_init(bar(kBarQFull), kProdThreadsAll);
    for (int s = 0; s < kSlots; ++s) {
      mbar_init(bar(kBarFull + s), use_tma ? 1 : kProdThreadsAll);
      mbar_init(bar(kBarEmpty + s), 1);
    }
    for (int t = 0; t < kQTiles; ++t) {                   // (one-tile instance: S buffer t; ODone + 1 = "PV done")
      mbar_init(bar(kBarSFull + t), 1);
      mbar_init(bar(kBarPFull + t), kSoftmaxWarps);
      mbar_init(bar(kBarODone + t), 1);
    }
    mbar_fence_init();
  }
  if (warp == kMmaWarp) tmem_alloc<kTmemCols>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp >= kProducerWarp0 || (kWideProd && warp >= kSoftmaxWarps && warp < 2 * kSoftmaxWarps)) {
    setmaxnreg_dec<kRegsOther>();
    if constexpr (kFp8) {
#include "extend_tc_producer_fp8.inc"
    } else if (use_tma) {
      if constexpr (kTma) {
#include "extend_tc_producer_tma.inc"
      }
    } else {
#include "extend_tc_producer.inc"
    }
  } else if (warp >= kMmaWarp) {
    setmaxnreg_dec<kRegsOther>();
    if (warp == kMmaWarp) {
#include "extend_tc_mma.inc"
    }
  } else {
    setmaxnreg_inc<kRegsSoftmax>();
#include "extend_tc_softmax.inc"
  }

  // ---- teardown ----
  __syncthreads();
  if (warp == kMmaWarp) {
    tc_fence_after_sync();
    tmem_dealloc<kTmemCols>(tmem_base);
  }
}


// head dims that run zero-padded on the D = 128 one-tile instance (16-bit pools)
bool extend_padded_dim(int d) { return d == 80 || d == 96 || d == 112; }

bool extend_fast_supported(int dqk, int dv, int group, int q_dtype, int kv_dtype) {
  if (dqk == dv && extend_padded_dim(dqk))
    return group >= 1 && group <= 128 && (q_dtype == B200_BF16 || q_dtype == B200_F16) && kv_dtype == q_dtype;
  if (dqk != dv || (dqk != 128 && dqk != 64 && dqk != 256)) return false;
  if (group < 1 || group > 128) return false;
  if (q_dtype != B200_BF16 && q_dtype != B200_F16) return false;
  return kv_dtype == q_dtype || kv_dtype == B200_FP8_E4M3 || kv_dtype == B200_FP8_E5M2;
}

// query tokens per work item (two 128-row tiles of 128/group tokens each; D = 256 has the one-tile instance only)
int extend_fast_block_q(int group, int head_dim) {
  return ((head_dim == 256 || extend_padded_dim(head_dim)) ? 1 : tc::kQTiles) * (128 / group);
}
// ... or one 128-row tile (the kOneTile instance): the host picks it when no request extends by more tokens
int extend_fast_block_q_one_tile(int group) { return 128 / group; }

bool extend_fast_layout_ok(const ExtendParams& p, int kv_dtype) {
  const int kv_align = (kv_dtype == B200_FP8_E4M3 || kv_dtype == B200_FP8_E5M2) ? 16 : 8;   // 16-byte pieces
  auto a16 = [](const void* ptr) { return ((uintptr_t)ptr & 15) == 0; };
  return a16(p.k_cache) && a16(p.v_cache) && a16(p.q) && a16(p.o) && p.k_stride_s % kv_align == 0 &&
         p.k_stride_h % kv_align == 0 && p.v_stride_s % kv_align == 0 && p.v_stride_h % kv_align == 0 &&
         p.q_stride_t % 8 == 0 &&
         p.q_stride_h % 8 == 0 && p.o_stride_t % 8 == 0 && p.o_stride_h % 8 == 0 &&
         p.k_stride_s > 0 && p.v_stride_s > 0 && p.k_stride_s < (1ll << 30) && p.v_stride_s < (1ll << 30);
}

// ---- TMA descriptors (host) -------------------------------------------------------------------------------------
// The pool buffer [slots, Hkv * D] as a 2-D tensor of 16-bit words; box = 64 words x `box_rows` rows, SWIZZLE_128B:
// one copy lands the 128-byte rows of a page's one-head d half exactly where the UMMA descriptors expect them.  cuTensorMapEncodeTiled is
// resolved through the runtime (no link-time dependency on libcuda).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn tensor_map_encoder() {
  static std::atomic<void*> cached{nullptr};
  void* fn = cached.load(std::memory_order_acquire);
  if (fn == nullptr) {
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      return nullptr;
    cached.store(fn, std::memory_order_release);
  }
  return reinterpret_cast<EncodeTiledFn>(fn);
}

// rows x width 16-bit words, row pitch `stride_elems` words
bool make_pool_tensor_map(CUtensorMap* tm, const void* base, int64_t rows, int64_t width, int64_t stride_elems,
                          int box_rows) {
  EncodeTiledFn enc = tensor_map_encoder();
  if (enc == nullptr) return false;
  const cuuint64_t gdim[2] = {(cuuint64_t)width, (cuuint64_t)rows};
  const cuuint64_t gstride[1] = {(cuuint64_t)stride_elems * 2};
  const cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  return enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// rows per TMA box: the page (power of two >= 16), at most one 128-row kv block
int extend_tma_rows(int page_size) {
  if (page_size < 16 || (page_size & (page_size - 1)) != 0) return 0;
  return page_size < tc::kBlockKV ? page_size : tc::kBlockKV;
}

// TMA needs: whole pages, heads contiguous inside a pool row, 16-byte aligned base and row pitch, int32 row ids
bool extend_tma_ok(const ExtendParams& p, const b200_extend_args& a) {
  return extend_tma_rows(a.page_size) > 0 && p.k_stride_h == p.head_dim_qk && p.v_stride_h == p.head_dim_v && a.pool_slots > 0 &&
         a.pool_slots < 0x7fffffff && p.k_stride_s % 8 == 0 && p.v_stride_s % 8 == 0 &&
         p.k_stride_s >= (int64_t)p.num_kv_heads * p.head_dim_qk && p.v_stride_s >= (int64_t)p.num_kv_heads * p.head_dim_v;
}

template <typename QT, bool kFp8, bool kOpts, int kHD, bool kTma, bool kOneTile = false>
int launch_extend_fast_t(const ExtendParams& p_in, const b200_extend_args& a, cudaStream_t stream) {
  ExtendParams p = p_in;
  p.tma_rows = kTma ? extend_tma_rows(a.page_size) : 0;
  static std::atomic<uint64_t> configured{0};
  constexpr int kSmem = tc::Layout<kFp8>::kSmemBytes;
  auto kern = extend_tc_kernel<QT, tc::kEmuEighths, kFp8, kOpts, kHD, kTma, kOneTile>;
  B200_CHECK_CUDA(set_max_dynamic_smem(kern, kSmem, configured));
  CUtensorMap tm_k, tm_v;
  memset(&tm_k, 0, sizeof(tm_k));
  memset(&tm_v, 0, sizeof(tm_v));
  if (kTma) {
    const int rows = extend_tma_rows(a.page_size);
    const bool ok =
        make_pool_tensor_map(&tm_k, p.k_cache, a.pool_slots, (int64_t)p.num_kv_heads * kHD, p.k_stride_s, rows) &&
        make_pool_tensor_map(&tm_v, p.v_cache, a.pool_slots, (int64_t)p.num_kv_heads * kHD, p.v_stride_s, rows);
    if (!ok) {
      set_error("extend: cuTensorMapEncodeTiled failed for the KV pool (%lld slots)", (long long)a.pool_slots);
      return B200_ECUDA;
    }
  }
  const int block_q = kOneTile ? extend_fast_block_q_one_tile(p.group) : extend_fast_block_q(p.group, kHD);
  // host-side bound on the number of work items: sum_r ceil(E_r / block_q) <= T / block_q + bs
  int64_t tiles = a.num_q_tokens / block_q + a.batch_size;
  if (tiles > a.max_tiles) tiles = a.max_tiles;
  const int64_t tiles_launch = (kOneTile && p.tile_count > 0) ? p.tile_count : tiles;
  const int64_t grid = tiles_launch * p.num_kv_heads * (kOneTile ? p.num_splits : 1);
  B200_CHECK_ARG(grid > 0 && grid < 0x7fffffff, "extend: bad grid %lld", (long long)grid);
  B200_CHECK_ARG(kOneTile || p.num_splits == 1, "extend: split-KV needs the one-tile instance");
  kern<<<(unsigned)grid, tc::kThreads, kSmem, stream>>>(p, (int)tiles, tm_k, tm_v);
  B200_CHECK_CUDA(cudaGetLastError());
  return B200_OK;
}

template <typename QT, bool kFp8, int kHD>
int launch_extend_fast_d(const ExtendParams& p, const b200_extend_args& a, cudaStream_t stream) {
  const bool opts = p.window_keys > 0 || p.logit_cap > 0.f || p.sinks != nullptr || p.custom_mask != nullptr;
  if (a.block_q == extend_fast_block_q_one_tile(p.group) || extend_padded_dim(p.head_dim_qk))   // one-tile items
    return opts ? launch_extend_fast_t<QT, kFp8, true, kHD, false, true>(p, a, stream)
                : launch_extend_fast_t<QT, kFp8, false, kHD, false, true>(p, a, stream);
  if constexpr (!kFp8) {
    if (extend_tma_ok(p, a))      // page_size >= 16: TMA page boxes; page_size 1 keeps the LDGSTS gather
      return opts ? launch_extend_fast_t<QT, false, true, kHD, true>(p, a, stream)
                  : launch_extend_fast_t<QT, false, false, kHD, true>(p, a, stream);
  }
  return opts ? launch_extend_fast_t<QT, kFp8, true, kHD, false>(p, a, stream)
              : launch_extend_fast_t<QT, kFp8, false, kHD, false>(p, a, stream);
}

template <typename QT, bool kFp8>
int launch_extend_fast_q(const ExtendParams& p, const b200_extend_args& a, cudaStream_t stream) {
  if (p.head_dim_qk == 256) {
    const bool opts = p.window_keys > 0 || p.logit_cap > 0.f || p.sinks != nullptr || p.custom_mask != nullptr;
    return opts ? launch_extend_fast_t<QT, kFp8, true, 256, false, true>(p, a, stream)
                : launch_extend_fast_t<QT, kFp8, false, 256, false, true>(p, a, stream);
  }
  return p.head_dim_qk == 64 ? launch_extend_fast_d<QT, kFp8, 64>(p, a, stream)
                             : launch_extend_fast_d<QT, kFp8, 128>(p, a, stream);     // 128, and 80 / 96 / 112 padded
}

int launch_extend_fast(const ExtendParams& p, const b200_extend_args& a, cudaStream_t stream) {
  const bool fp8 = a.kv_dtype == B200_FP8_E4M3 || a.kv_dtype == B200_FP8_E5M2;
  const int bq1 = extend_fast_block_q(p.group, p.head_dim_qk);
  const int bq = a.block_q > 0 ? a.block_q : bq1;
  B200_CHECK_ARG(bq == bq1 || bq == extend_fast_block_q_one_tile(p.group),
                 "extend: the plan was built with block_q %d, the tcgen05 kernel needs %d (or %d: one-tile items)", bq, bq1,
                 extend_fast_block_q_one_tile(p.group));
  if (a.q_dtype == B200_BF16)
    return fp8 ? launch_extend_fast_q<__nv_bfloat16, true>(p, a, stream)
               : launch_extend_fast_q<__nv_bfloat16, false>(p, a, stream);
  return fp8 ? launch_extend_fast_q<__half, true>(p, a, stream) : launch_extend_fast_q<__half, false>(p, a, stream);
}

}  // namespace
