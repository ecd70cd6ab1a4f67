// Extend (prefill over a cached prefix) attention for the RadixAttention hot path (sm_100a).
//
// Replaces (reference python/sglang/...):
//   srt/layers/attention/torch_native_backend.py:61-174    per-request gather + causal SDPA (oracle semantics)
//   kernels/ops/attention/extend_attention.py:241-661,664-812  Triton 2-stage extend kernel
//   srt/layers/attention/triton_backend.py:860-900          qo_indptr / prefix kv_indices metadata
//
// Query token i of request r attends pool rows tok(r, 0 .. P_r + i): the new tokens were
// stored into the pool before this launch, so prefix and new K/V are both read through
// the page table, exactly like the torch_native oracle.
// Algorithmic FLOPs per launch: 4 * Hq * D * sum_r (E_r * P_r + E_r * (E_r + 1) / 2).
#include <cuda.h>   // CUtensorMap types only; the encoder is resolved through cudaGetDriverEntryPoint

#include <algorithm>
#include <cstring>
#include <type_traits>

#include "common.cuh"
#include "generic_attn.cuh"
#include "sm100_ptx.cuh"

namespace b200 {
namespace {

// ----------------------------------------------------------------------------
// planner: qo_indptr (exclusive scan of E_r) + (request, first q row) per q tile.
// Tiles of one request are listed last-first: under the causal mask the last q tile
// has the longest KV range, so the persistent kernel starts the heaviest work first.
// ----------------------------------------------------------------------------
struct ExtPlanParams {
  const void* extend_seq_lens;
  int32_t* plan;
  int32_t lens_is_i64, batch_size, max_bs, max_tiles, block_q;
  int32_t page_rows;            // rows per page box to verify (0: no page check, plan[kPlanPageRows] = 0)
  // page check only
  const void* extend_prefix_lens;
  const int32_t* req_to_token;
  int64_t req_to_token_stride;
  const void* req_pool_indices;
  int32_t req_pool_indices_is_i64;
};

constexpr int kPlanThreads = 1024;

__global__ void __launch_bounds__(kPlanThreads) extend_plan_kernel(const ExtPlanParams p) {
  __shared__ int32_t s_wq[32], s_wt[32];
  __shared__ int32_t s_carry_q, s_carry_t, s_pass_q, s_pass_t;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int bs = p.batch_size;
  int32_t* qo_indptr = p.plan + kPlanHeader;
  int32_t* tiles = p.plan + kPlanHeader + p.max_bs + 1;
  if (tid == 0) { s_carry_q = 0; s_carry_t = 0; }
  __syncthreads();
  for (int base = 0; base < bs; base += kPlanThreads) {
    const int r = base + tid;
    int32_t e = 0;
    if (r < bs) {
      const int64_t E = load_index(p.extend_seq_lens, p.lens_is_i64, r);
      e = E > 0 ? (int32_t)E : 0;
    }
    const int32_t nt = (e + p.block_q - 1) / p.block_q;
    int32_t iq = e, it = nt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int32_t a = __shfl_up_sync(0xffffffffu, iq, o);
      const int32_t b = __shfl_up_sync(0xffffffffu, it, o);
      if (lane >= o) { iq += a; it += b; }
    }
    if (lane == 31) { s_wq[warp] = iq; s_wt[warp] = it; }
    __syncthreads();
    if (warp == 0) {
      const int32_t wq = s_wq[lane], wt = s_wt[lane];
      int32_t aq = wq, at = wt;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int32_t a = __shfl_up_sync(0xffffffffu, aq, o);
        const int32_t b = __shfl_up_sync(0xffffffffu, at, o);
        if (lane >= o) { aq += a; at += b; }
      }
      s_wq[lane] = aq - wq; s_wt[lane] = at - wt;
      if (lane == 31) { s_pass_q = aq; s_pass_t = at; }
    }
    __syncthreads();
    const int32_t cq = s_carry_q, ct = s_carry_t;
    if (r < bs) {
      const int32_t q0 = cq + s_wq[warp] + iq - e;
      const int32_t t0 = ct + s_wt[warp] + it - nt;
      qo_indptr[r] = q0;
      for (int i = 0; i < nt; ++i) {
        if (t0 + i < p.max_tiles) {
          tiles[2 * (t0 + i)] = r;
          tiles[2 * (t0 + i) + 1] = (nt - 1 - i) * p.block_q;
        }
      }
    }
    __syncthreads();
    if (tid == 0) { s_carry_q = cq + s_pass_q; s_carry_t = ct + s_pass_t; }
    __syncthreads();
  }
  if (tid == 0) {
    qo_indptr[bs] = s_carry_q;
    p.plan[0] = s_carry_t < p.max_tiles ? s_carry_t : p.max_tiles;
    p.plan[1] = p.block_q;
    p.plan[2] = bs;
    p.plan[3] = s_carry_q;
    p.plan[kPlanPageRows] = p.page_rows;   // cleared by extend_page_check_kernel when a box would not be contiguous
  }
}

// Page check: the TMA fill of the tcgen05 kernel reads positions [b * R, (b + 1) * R) of a request as ONE box of R
// consecutive pool rows starting at the slot of position b * R.  SGLang's paged allocator guarantees that for
// R | page_size, but req_to_token itself is token-granular and the reference's kernels never assume more (its own
// unit-test kit scatters the new tokens of a request inside their page, dense_attention.py:685-703), so the
// plan verifies it per batch: grid (chunks, bs), one comparison per position, any mismatch clears the flag.
constexpr int kPageCheckThreads = 256, kPageCheckChunks = 8;

__global__ void __launch_bounds__(kPageCheckThreads) extend_page_check_kernel(const ExtPlanParams p) {
  bool bad = false;
  for (int r = blockIdx.y; r < p.batch_size; r += gridDim.y) {
  const int64_t E = load_index(p.extend_seq_lens, p.lens_is_i64, r);
  const int64_t P = load_index(p.extend_prefix_lens, p.lens_is_i64, r);
  const int L = (int)((P > 0 ? P : 0) + (E > 0 ? E : 0));
  const int32_t* pt = p.req_to_token + load_index(p.req_pool_indices, p.req_pool_indices_is_i64, r) * p.req_to_token_stride;
  const int R = p.page_rows;
  for (int pos = blockIdx.x * kPageCheckThreads + threadIdx.x; pos < L; pos += kPageCheckChunks * kPageCheckThreads) {
    const int in_box = pos & (R - 1);
    bad |= __ldg(pt + pos) != __ldg(pt + pos - in_box) + in_box;
  }
  }
  if (__syncthreads_or(bad) && threadIdx.x == 0) p.plan[kPlanPageRows] = 0;
}

struct ExtendParams {
  const void* q;
  void* o;
  const void* k_cache;
  const void* v_cache;
  const int32_t* req_to_token;
  const void* req_pool_indices;
  const void* extend_prefix_lens;
  const void* extend_seq_lens;
  const int32_t* plan;
  float* lse;
  int64_t q_stride_t, q_stride_h, o_stride_t, o_stride_h;
  int64_t k_stride_s, k_stride_h, v_stride_s, v_stride_h;
  int64_t req_to_token_stride;
  int64_t num_q_tokens;
  int32_t req_pool_indices_is_i64, lens_is_i64;
  int32_t batch_size, num_q_heads, num_kv_heads, group;
  int32_t head_dim_qk, head_dim_v;
  int32_t max_bs, causal;
  int32_t window_keys;          // query at position x attends (x - window_keys, x]; 0 = off
  int32_t mask_skip_prefix;
  int32_t kv_e5m2;              // FP8 pool bytes are e5m2 (else e4m3); read by the tcgen05 kernel's FP8 producers only
  int32_t pool_slots;           // rows of the pool buffers (TMA: a box at row >= pool_slots is zero-filled)
  int32_t tma_rows;             // rows per TMA box (min(page_size, 128)); 0 when the fill is LDGSTS
  float sm_scale, k_scale, v_scale, logit_cap;
  const float* sinks;           // [Hq] natural-log sink logits or nullptr
  const uint8_t* custom_mask;   // see b200_extend_args
  const int64_t* mask_indptr;
  // split-KV of one-tile items (num_splits > 1): partial s of q row i, head h lives at
  //   split_o  [((s * num_q_tokens + i) * Hq + h) * Dv ...]   normalised fp32
  //   split_lse[ (s * num_q_tokens + i) * Hq + h ]            base-2 LSE of the part; -inf: the part saw nothing;
  //                                                           +inf (split 0 only): o / lse already hold the final row
  int32_t num_splits;
  float* split_o;
  float* split_lse;
  // one-tile launches over a sub-range of the plan's tiles (tail-only split): tiles [tile_offset, tile_offset +
  // tile_count), tile_count 0 = all
  int32_t tile_offset, tile_count;
  // 16-byte chunks of a real Q/K/V row (head_dim / 8).  Head dims 80 / 96 / 112 run on the D = 128 one-tile instance
  // with the remaining chunks of every tile row zero-filled (exact: the padding multiplies by zero).
  int32_t d_chunks;
};

constexpr int kSplitMinBlocks = 4;   // a split part keeps at least this many kv blocks (the device trims num_splits per item)

// ----------------------------------------------------------------------------
// generic kernel: one warp per (q token, q head); any head_dim <= 512 / dtype pair.
// ----------------------------------------------------------------------------
constexpr int kGenWarps = 4;

template <typename QT, typename KVT>
__global__ void __launch_bounds__(kGenWarps* kWarp) extend_generic_kernel(const ExtendParams p) {
  extern __shared__ float smem_f[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* q_s = smem_f + warp * p.head_dim_qk;
  const int Dk = p.head_dim_qk, Dv = p.head_dim_v, Hq = p.num_q_heads, bs = p.batch_size;
  const int32_t* qo_indptr = p.plan + kPlanHeader;
  const int64_t n_rows = (int64_t)qo_indptr[bs] * Hq;
  const QT* q = reinterpret_cast<const QT*>(p.q);
  const KVT* kc = reinterpret_cast<const KVT*>(p.k_cache);
  const KVT* vc = reinterpret_cast<const KVT*>(p.v_cache);
  const float qk_scale = p.sm_scale * p.k_scale;

  for (int64_t row = (int64_t)blockIdx.x * kGenWarps + warp; row < n_rows;
       row += (int64_t)gridDim.x * kGenWarps) {
    const int t = (int)(row / Hq), h = (int)(row % Hq);
    int lo = 0, hi = bs;  // request owning q token t
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (qo_indptr[mid] <= t) lo = mid; else hi = mid;
    }
    const int r = lo, i = t - qo_indptr[r];
    const int64_t P = load_index(p.extend_prefix_lens, p.lens_is_i64, r);
    const int64_t E = load_index(p.extend_seq_lens, p.lens_is_i64, r);
    const bool masked = p.custom_mask != nullptr;      // the custom mask replaces the causal rule on the new tokens
    const int kv_len = (int)((p.causal && !masked) ? P + i + 1 : P + E);
    int kv_lo = 0;                                     // sliding window: positions (x - window_keys, x], x = P + i
    if (p.window_keys > 0 && P + i + 1 > p.window_keys) kv_lo = (int)(P + i + 1 - p.window_keys);
    const uint8_t* mrow = masked ? p.custom_mask + p.mask_indptr[r] + (int64_t)i * (P + E) + kv_lo : nullptr;
    const int mask_from = p.mask_skip_prefix ? (int)P - kv_lo : 0;
    const int64_t pool_row = load_index(p.req_pool_indices, p.req_pool_indices_is_i64, r);
    const int kvh = h / p.group;
    const QT* qrow = q + (int64_t)t * p.q_stride_t + (int64_t)h * p.q_stride_h;
    __syncwarp();
    for (int d = lane; d < Dk; d += kWarp) q_s[d] = to_float(qrow[d]);
    __syncwarp();
    RowState st;
#pragma unroll
    for (int j = 0; j < kGenAcc; ++j) st.acc[j] = 0.f;
    attend_row_generic<KVT>(st, q_s, p.req_to_token + pool_row * p.req_to_token_stride + kv_lo, kv_len - kv_lo,
                            kc, vc, p.k_stride_s, p.v_stride_s, (int64_t)kvh * p.k_stride_h,
                            (int64_t)kvh * p.v_stride_h, Dk, Dv, qk_scale, p.logit_cap, mrow, mask_from);
    if (p.sinks != nullptr) st.l += exp2f(p.sinks[h] * kLog2e - st.m);   // sink: denominator only
    const float inv_l = st.l > 0.f ? p.v_scale / st.l : 0.f;
    QT* orow = reinterpret_cast<QT*>(p.o) + (int64_t)t * p.o_stride_t + (int64_t)h * p.o_stride_h;
#pragma unroll
    for (int j = 0; j < kGenAcc; ++j) {
      const int d = lane + j * kWarp;
      if (d < Dv) orow[d] = from_float<QT>(st.acc[j] * inv_l);
    }
    if (p.lse != nullptr && lane == 0) p.lse[(int64_t)t * Hq + h] = (st.m + log2f(st.l)) * kLn2;
  }
}

template <typename QT>
int launch_generic(const ExtendParams& p, int kv_dtype, int num_ctas, cudaStream_t stream) {
  const size_t smem = (size_t)kGenWarps * p.head_dim_qk * sizeof(float);
  const int grid = num_ctas * 8;
  switch (kv_dtype) {
    case B200_BF16: extend_generic_kernel<QT, __nv_bfloat16><<<grid, kGenWarps * kWarp, smem, stream>>>(p); break;
    case B200_F16: extend_generic_kernel<QT, __half><<<grid, kGenWarps * kWarp, smem, stream>>>(p); break;
    case B200_FP8_E4M3: extend_generic_kernel<QT, e4m3_t><<<grid, kGenWarps * kWarp, smem, stream>>>(p); break;
    case B200_FP8_E5M2: extend_generic_kernel<QT, e5m2_t><<<grid, kGenWarps * kWarp, smem, stream>>>(p); break;
    default: set_error("extend: unsupported kv dtype %d", kv_dtype); return B200_EINVAL;
  }
  B200_CHECK_CUDA(cudaGetLastError());
  return B200_OK;
}

}  // namespace

#include "extend_tc.cuh"

}  // namespace b200

extern "C" int64_t b200attn_extend_plan_ints(int32_t max_bs, int32_t max_tiles) {
  return (int64_t)b200::kPlanHeader + (int64_t)max_bs + 1 + 2 * (int64_t)max_tiles;
}

extern "C" int b200attn_extend_plan(const b200_extend_plan_args* a, b200_stream_t stream_) {
  using namespace b200;
  B200_CHECK_ARG(a && a->extend_seq_lens && a->plan, "extend_plan: null pointer");
  B200_CHECK_ARG(a->batch_size >= 0 && a->batch_size <= a->max_bs, "extend_plan: bs %d > max_bs %d",
                 a->batch_size, a->max_bs);
  B200_CHECK_ARG(a->block_q > 0 && a->max_tiles >= 0, "extend_plan: bad block_q/max_tiles");
  ExtPlanParams p;
  p.extend_seq_lens = a->extend_seq_lens; p.plan = a->plan; p.lens_is_i64 = a->lens_is_i64;
  p.batch_size = a->batch_size; p.max_bs = a->max_bs; p.max_tiles = a->max_tiles; p.block_q = a->block_q;
  p.page_rows = a->req_to_token != nullptr ? extend_tma_rows(a->page_size) : 0;
  p.extend_prefix_lens = a->extend_prefix_lens; p.req_to_token = a->req_to_token;
  p.req_to_token_stride = a->req_to_token_stride; p.req_pool_indices = a->req_pool_indices;
  p.req_pool_indices_is_i64 = a->req_pool_indices_is_i64;
  B200_CHECK_ARG(p.page_rows == 0 || (a->extend_prefix_lens && a->req_pool_indices),
                 "extend_plan: the page check needs extend_prefix_lens and req_pool_indices");
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  extend_plan_kernel<<<1, kPlanThreads, 0, stream>>>(p);
  B200_CHECK_CUDA(cudaGetLastError());
  if (p.page_rows > 0 && a->batch_size > 0) {
    extend_page_check_kernel<<<dim3(kPageCheckChunks, a->batch_size < 32768 ? a->batch_size : 32768), kPageCheckThreads, 0,
                               stream>>>(p);
    B200_CHECK_CUDA(cudaGetLastError());
  }
  return B200_OK;
}

namespace b200 {
namespace {
// Merge of the split-KV partials of one-tile items: one warp per (q row, q head); lane s holds split s's LSE.
//   out = v_scale * sum_s 2^(lse_s - m) o_s / (sum_s 2^(lse_s - m) + 2^(sink - m))
constexpr int kSplitMergeWarps = 4;

template <typename QT>
__global__ void __launch_bounds__(kSplitMergeWarps* kWarp) extend_split_merge_kernel(const ExtendParams p) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int Hq = p.num_q_heads, D = p.head_dim_v, S = p.num_splits;
  const int64_t rows = (int64_t)p.plan[3] * Hq;                       // sum E_r (device-side) x heads
  const int64_t idx = (int64_t)blockIdx.x * kSplitMergeWarps + warp;
  if (idx >= rows || idx >= p.num_q_tokens * Hq) return;
  const int64_t tok = idx / Hq;
  const int h = (int)(idx - tok * Hq);
  const int64_t part = p.num_q_tokens * Hq;                           // rows per split
  const float l = lane < S ? p.split_lse[lane * part + idx] : -INFINITY;
  const float l0 = __shfl_sync(0xffffffffu, l, 0);
  if (l0 == INFINITY || l0 != l0) return;       // not split: split 0 (+inf) or the unsplit launch (NaN fill) wrote the row
  const float m = warp_max(l);
  QT* orow = reinterpret_cast<QT*>(p.o) + tok * p.o_stride_t + (int64_t)h * p.o_stride_h;
  if (m == -INFINITY) {                                               // no part saw a key (cannot happen under the
    for (int d = lane; d < D; d += kWarp) orow[d] = from_float<QT>(0.f);   // causal rule; kept for masks)
    if (p.lse != nullptr && lane == 0) p.lse[idx] = -INFINITY;
    return;
  }
  const float w = lane < S ? fast_exp2(l - m) : 0.f;                  // 2^(-inf) = 0 for empty parts
  float den = w;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) den += __shfl_xor_sync(0xffffffffu, den, o);
  if (p.sinks != nullptr) den += exp2f(p.sinks[h] * kLog2e - m);
  const float inv = p.v_scale / den;
  for (int d = lane * 4; d < D; d += 4 * kWarp) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    // all parts' loads are issued before the first FMA (ncu: the dependent-load version of this loop took 12 us for
    // 16 parts); parts with lse = -inf hold no data and are not read
#pragma unroll
    for (int s0 = 0; s0 < B200_EXTEND_MAX_SPLITS; s0 += 8) {
      if (s0 >= S) break;
      float4 v[8];
      float ws[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        ws[u] = __shfl_sync(0xffffffffu, w, s0 + u);                  // lanes >= S hold 0
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (ws[u] > 0.f) v[u] = *reinterpret_cast<const float4*>(p.split_o + ((s0 + u) * part + idx) * D + d);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        acc.x = fmaf(ws[u], v[u].x, acc.x); acc.y = fmaf(ws[u], v[u].y, acc.y);
        acc.z = fmaf(ws[u], v[u].z, acc.z); acc.w = fmaf(ws[u], v[u].w, acc.w);
      }
    }
    *reinterpret_cast<uint2*>(orow + d) = make_uint2(pack2<QT>(acc.x * inv, acc.y * inv), pack2<QT>(acc.z * inv, acc.w * inv));
  }
  if (p.lse != nullptr && lane == 0) p.lse[idx] = (m + log2f(den)) * kLn2;
}

int launch_split_merge(const ExtendParams& p, int q_dtype, cudaStream_t stream) {
  const int64_t rows = p.num_q_tokens * p.num_q_heads;
  const unsigned grid = (unsigned)((rows + kSplitMergeWarps - 1) / kSplitMergeWarps);
  if (q_dtype == B200_BF16) extend_split_merge_kernel<__nv_bfloat16><<<grid, kSplitMergeWarps * kWarp, 0, stream>>>(p);
  else extend_split_merge_kernel<__half><<<grid, kSplitMergeWarps * kWarp, 0, stream>>>(p);
  B200_CHECK_CUDA(cudaGetLastError());
  return B200_OK;
}
}  // namespace
}  // namespace b200

extern "C" int64_t b200attn_extend_split_ws_floats(int32_t num_splits, int64_t num_q_tokens, int32_t num_q_heads,
                                                   int32_t head_dim_v) {
  if (num_splits <= 1) return 0;
  return (int64_t)num_splits * num_q_tokens * num_q_heads * ((int64_t)head_dim_v + 1);
}

extern "C" int32_t b200attn_extend_pick_splits(int32_t items, int32_t num_sms, int32_t kv_tokens_hint) {
  if (items <= 0 || num_sms <= 0 || items > 8 * num_sms) return 1;    // >= 8 waves: the quantisation loss is small
  if (items > num_sms) {
    // several waves: only the tail (the items of the last, partly filled wave) is split, see num_tiles_hint
    const int tail = items % num_sms, full = items / num_sms;
    if (tail == 0) return 1;
    const int S = std::min<int>(B200_EXTEND_MAX_SPLITS, num_sms / tail);
    if (S < 2) return 1;
    const double blocks_t = (kv_tokens_hint > 0 ? kv_tokens_hint : 4096) / 128.0 + 1.0;
    const double pc = 4.0 / blocks_t, mc = 3.7 / blocks_t;
    // measured on 24 / 40 / 64-request verify batches at 4 K: the second attention launch (it starts when the first
    // has drained), the NaN fill and the merge cost ~13 kv blocks' worth of time together
    const double unsplit_t = (full + 1) * (1.0 + pc), tail_t = full * (1.0 + pc) + (1.0 / S + pc) + 13.0 / blocks_t;
    (void)mc;
    return tail_t < (kv_tokens_hint > 0 ? 0.95 : 0.90) * unsplit_t ? S : 1;
  }
  // Makespan in units of one unsplit item: ceil(items * S / SMs) waves, each 1 / S of an item plus the fixed cost of a
  // part (TMEM allocation, Q fill, pipeline ramp, fp32 partial epilogue: ~6.5 us = 4 kv blocks of 1.64 us, measured
  // on the 64 x (4 K + 8) verify shape), plus the merge launch (~6 us) once.  A split has to win by 5 %.
  const double blocks = (kv_tokens_hint > 0 ? kv_tokens_hint : 4096) / 128.0 + 1.0;
  const double part_cost = 4.0 / blocks, merge_cost = 3.7 / blocks;
  const int w1 = (items + num_sms - 1) / num_sms;
  const double unsplit = w1 * (1.0 + part_cost);
  int best = 1;
  double best_cost = unsplit * 0.95;
  for (int S = 2; S <= B200_EXTEND_MAX_SPLITS; ++S) {
    const int waves = (items * S + num_sms - 1) / num_sms;
    const double cost = waves * (1.0 / S + part_cost) + merge_cost;
    if (cost < best_cost - 1e-9) { best_cost = cost; best = S; }
  }
  return best;
}

extern "C" int b200attn_extend_fast_supported(int32_t dqk, int32_t dv, int32_t group, int32_t q_dtype,
                                              int32_t kv_dtype) {
  return b200::extend_fast_supported(dqk, dv, group, q_dtype, kv_dtype) ? 1 : 0;
}

extern "C" int b200attn_extend_block_q(int32_t dqk, int32_t dv, int32_t group, int32_t q_dtype,
                                       int32_t kv_dtype, int32_t kernel) {
  if (kernel != B200_KERNEL_GENERIC && b200::extend_fast_supported(dqk, dv, group, q_dtype, kv_dtype))
    return b200::extend_fast_block_q(group, dqk);
  return 64;  // the generic kernel ignores tiles; any value gives a valid plan
}

extern "C" int b200attn_extend(const b200_extend_args* a, b200_stream_t stream_) {
  using namespace b200;
  B200_CHECK_ARG(a != nullptr, "extend: null args");
  if (a->batch_size == 0 || a->num_q_tokens == 0) return B200_OK;
  B200_CHECK_ARG(a->q && a->o && a->k_cache && a->v_cache && a->req_to_token && a->req_pool_indices &&
                     a->extend_prefix_lens && a->extend_seq_lens && a->plan,
                 "extend: null pointer");
  B200_CHECK_ARG(a->num_q_heads > 0 && a->num_kv_heads > 0 && a->num_q_heads % a->num_kv_heads == 0,
                 "extend: Hq=%d must be a multiple of Hkv=%d", a->num_q_heads, a->num_kv_heads);
  B200_CHECK_ARG(a->head_dim_qk > 0 && a->head_dim_qk <= kGenMaxD && a->head_dim_v > 0 &&
                     a->head_dim_v <= kGenMaxD,
                 "extend: head dims (%d, %d) must be in [1, %d]", a->head_dim_qk, a->head_dim_v, kGenMaxD);
  B200_CHECK_ARG(a->q_dtype == B200_BF16 || a->q_dtype == B200_F16, "extend: q dtype %d", a->q_dtype);
  B200_CHECK_ARG(a->batch_size <= a->max_bs && a->num_ctas > 0, "extend: bs %d > max_bs %d or num_ctas",
                 a->batch_size, a->max_bs);
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  ExtendParams p;
  p.q = a->q; p.o = a->o; p.k_cache = a->k_cache; p.v_cache = a->v_cache;
  p.req_to_token = a->req_to_token; p.req_pool_indices = a->req_pool_indices;
  p.extend_prefix_lens = a->extend_prefix_lens; p.extend_seq_lens = a->extend_seq_lens;
  p.plan = a->plan; p.lse = a->lse;
  p.q_stride_t = a->q_stride_t; p.q_stride_h = a->q_stride_h;
  p.o_stride_t = a->o_stride_t; p.o_stride_h = a->o_stride_h;
  p.k_stride_s = a->k_stride_s; p.k_stride_h = a->k_stride_h;
  p.v_stride_s = a->v_stride_s; p.v_stride_h = a->v_stride_h;
  p.req_to_token_stride = a->req_to_token_stride; p.num_q_tokens = a->num_q_tokens;
  p.req_pool_indices_is_i64 = a->req_pool_indices_is_i64; p.lens_is_i64 = a->lens_is_i64;
  p.batch_size = a->batch_size; p.num_q_heads = a->num_q_heads; p.num_kv_heads = a->num_kv_heads;
  p.group = a->num_q_heads / a->num_kv_heads;
  p.head_dim_qk = a->head_dim_qk; p.head_dim_v = a->head_dim_v; p.max_bs = a->max_bs; p.causal = a->causal;
  p.sm_scale = a->sm_scale; p.k_scale = a->k_scale; p.v_scale = a->v_scale; p.logit_cap = a->logit_cap;
  p.window_keys = a->window_keys > 0 ? a->window_keys : 0; p.sinks = a->sinks;
  p.custom_mask = a->custom_mask; p.mask_indptr = a->mask_indptr; p.mask_skip_prefix = a->mask_skip_prefix;
  p.kv_e5m2 = a->kv_dtype == B200_FP8_E5M2;
  p.num_splits = 1; p.split_o = nullptr; p.split_lse = nullptr;
  p.tile_offset = 0; p.tile_count = 0;
  p.d_chunks = a->head_dim_qk / 8;
  p.tma_rows = 0;
  p.pool_slots = a->pool_slots > 0 && a->pool_slots < 0x7fffffff ? (int32_t)a->pool_slots : 0;
  B200_CHECK_ARG(a->custom_mask == nullptr || a->mask_indptr != nullptr, "extend: custom_mask needs mask_indptr");
  B200_CHECK_ARG(!(p.window_keys > 0 && !a->causal), "extend: a sliding window needs causal attention");

  const bool fast_ok = extend_fast_supported(a->head_dim_qk, a->head_dim_v, p.group, a->q_dtype, a->kv_dtype) &&
                       extend_fast_layout_ok(p, a->kv_dtype);
  B200_CHECK_ARG(!(a->kernel == B200_KERNEL_FAST && !fast_ok),
                 "extend: fast kernel does not cover Dk=%d Dv=%d group=%d q=%d kv=%d (or unaligned strides)",
                 a->head_dim_qk, a->head_dim_v, p.group, a->q_dtype, a->kv_dtype);
  const bool use_fast = fast_ok && a->kernel != B200_KERNEL_GENERIC;
  if (a->num_splits > 1) {
    B200_CHECK_ARG(a->num_splits <= B200_EXTEND_MAX_SPLITS, "extend: num_splits %d > %d", a->num_splits,
                   B200_EXTEND_MAX_SPLITS);
    B200_CHECK_ARG(use_fast && a->block_q == extend_fast_block_q_one_tile(p.group),
                   "extend: num_splits %d needs the tcgen05 kernel's one-tile plan (block_q %d, got %d)", a->num_splits,
                   extend_fast_block_q_one_tile(p.group), a->block_q);
    const int64_t need = b200attn_extend_split_ws_floats(a->num_splits, a->num_q_tokens, a->num_q_heads, a->head_dim_v);
    B200_CHECK_ARG(a->split_ws != nullptr && a->split_ws_floats >= need && ((uintptr_t)a->split_ws & 15) == 0,
                   "extend: split workspace of %lld floats (16-byte aligned) needed for %d splits, got %lld",
                   (long long)need, a->num_splits, (long long)a->split_ws_floats);
    B200_CHECK_ARG(a->o_stride_t % 4 == 0 && a->o_stride_h % 4 == 0 && ((uintptr_t)a->o & 7) == 0,
                   "extend: split merge writes 8-byte vectors: o strides must be multiples of 4 elements");
    p.num_splits = a->num_splits;
    p.split_o = a->split_ws;
    // the LSE block starts 16-byte aligned as well: S * T * Hq * Dv floats with Dv a multiple of 64
    p.split_lse = a->split_ws + (int64_t)a->num_splits * a->num_q_tokens * a->num_q_heads * a->head_dim_v;
    // More work items than SMs and an exact tile count: split the tail only.  The tiles that fill whole waves run
    // unsplit; the rest (which would occupy a fraction of the SMs for a whole item's time) is cut into parts.
    if (a->num_tiles_hint > 0) {
      const int sms = sm_count();
      const int64_t items = (int64_t)a->num_tiles_hint * a->num_kv_heads;
      if (items > sms) {
        const int tiles_full = (int)((items / sms) * sms / a->num_kv_heads);
        const int tiles_tail = a->num_tiles_hint - tiles_full;
        const int64_t tail_items = (int64_t)tiles_tail * a->num_kv_heads;
        const int s_tail = tail_items > 0 ? (int)std::min<int64_t>(a->num_splits, sms / tail_items) : 0;
        if (tiles_full <= 0 || s_tail < 2) {              // nothing to gain: one unsplit launch
          p.num_splits = 1;
          return launch_extend_fast(p, *a, stream);
        }
        // rows that no part will write: split 0's LSE = NaN tells the merge to leave them alone
        B200_CHECK_CUDA(cudaMemsetAsync(p.split_lse, 0xff, (size_t)a->num_q_tokens * a->num_q_heads * sizeof(float), stream));
        ExtendParams p1 = p;
        p1.num_splits = 1; p1.tile_offset = 0; p1.tile_count = tiles_full;
        int rc = launch_extend_fast(p1, *a, stream);
        if (rc != B200_OK) return rc;
        p.num_splits = s_tail; p.tile_offset = tiles_full; p.tile_count = tiles_tail;
        rc = launch_extend_fast(p, *a, stream);
        if (rc != B200_OK) return rc;
        return launch_split_merge(p, a->q_dtype, stream);
      }
    }
    const int rc = launch_extend_fast(p, *a, stream);
    if (rc != B200_OK) return rc;
    return launch_split_merge(p, a->q_dtype, stream);
  }
  if (use_fast) return launch_extend_fast(p, *a, stream);
  if (a->q_dtype == B200_BF16) return launch_generic<__nv_bfloat16>(p, a->kv_dtype, a->num_ctas, stream);
  return launch_generic<__half>(p, a->kv_dtype, a->num_ctas, stream);
}

#if defined(B200_DEV) && defined(B200_TRACE)
// dev builds only: copy the traced CTA's clock stamps out (16 slots x 512 kv blocks)
extern "C" int b200attn_dev_trace_read(long long* out) {
  return (int)cudaMemcpyFromSymbol(out, b200::g_ext_trace, sizeof(long long) * 16 * 512);
}
#endif
