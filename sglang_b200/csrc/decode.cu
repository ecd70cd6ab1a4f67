// Paged decode attention for the RadixAttention hot path (sm_100a):
//   planner  -> split-KV attention (page-table gather inside the kernel) -> split merge
//
// Replaces (reference python/sglang/...):
//   srt/layers/attention/torch_native_backend.py:176-277   per-request gather + SDPA (oracle semantics)
//   kernels/ops/attention/decode_attention.py:96-281,383-608,731-805  Triton split-KV stage 1/2
//   kernels/ops/kvcache/kv_indices.py:8-45  + triton_backend.py:304-351  kv_indices gather / num_kv_splits
//
// HBM layout read here: K/V pool rows [slot][Hkv][D] (bf16/f16/fp8), page table
// req_to_token[int32][max_context_len], q [bs][Hq][D].  A work item is (chunk of one
// request's tokens, kv head); all `group = Hq/Hkv` query heads of that kv head share
// the single read of the K/V rows.  Algorithmic bytes per launch:
//   sum_r L_r * Hkv * (Dk + Dv) * s_kv + sum_r L_r * 4 + bs * Hq * (Dk + Dv) * 2.
#include <type_traits>

#include <cstdlib>

#include "common.cuh"
#include "generic_attn.cuh"
#include "sm100_ptx.cuh"

namespace b200 {
namespace {

// ----------------------------------------------------------------------------
// planner
// ----------------------------------------------------------------------------
struct PlanParams {
  const void* seq_lens;
  int32_t* plan;
  int32_t seq_lens_is_i64, batch_size, max_bs, max_chunks;
  int32_t num_kv_heads, num_ctas, waves, min_chunk, fixed_chunk;
  int32_t window_keys;   // sliding window: only the last window_keys tokens of a request are attended (0 = off)
};

// tokens of a request that are attended: all of them, or the last `window_keys`
__device__ __forceinline__ int64_t attended_len(const void* seq_lens, int is_i64, int r, int window_keys) {
  const int64_t L = load_index(seq_lens, is_i64, r);
  return (window_keys > 0 && L > window_keys) ? (int64_t)window_keys : L;
}

constexpr int kPlanThreads = 1024;
constexpr int kPlanCache = 2048;     // requests whose lengths are kept in shared memory by the planner

__device__ __forceinline__ int64_t block_sum_i64(int64_t v, int64_t* s_red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if (lane == 0) s_red[warp] = v;
  __syncthreads();
  if (warp == 0) {
    int64_t t = lane < (int)(blockDim.x >> 5) ? s_red[lane] : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if (lane == 0) s_red[0] = t;
  }
  __syncthreads();
  const int64_t r = s_red[0];
  __syncthreads();
  return r;
}

__global__ void __launch_bounds__(kPlanThreads) decode_plan_kernel(const PlanParams p) {
  __shared__ int64_t s_red[32];
  __shared__ int32_t s_warp_tot[32];
  __shared__ int32_t s_chunk, s_carry;
  __shared__ int64_t s_cost[32], s_cand[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int bs = p.batch_size;
  // The attended lengths are read once into shared memory as 32-bit values (negative -> 0): the search below evaluates
  // items() ~10 times per warp, and 64-bit divisions on re-loaded int64 lengths made this kernel 19 us for 64 requests.
  __shared__ uint32_t s_len[kPlanCache];
  auto len_at = [&](int r) -> uint32_t {
    if (r < kPlanCache) return s_len[r];
    const int64_t L = attended_len(p.seq_lens, p.seq_lens_is_i64, r, p.window_keys);
    return L > 0 ? (uint32_t)(L < 0x7fffffff ? L : 0x7fffffff) : 0u;
  };

  int64_t local = 0;
  for (int r = tid; r < bs; r += kPlanThreads) {
    const int64_t L = attended_len(p.seq_lens, p.seq_lens_is_i64, r, p.window_keys);
    const uint32_t l32 = L > 0 ? (uint32_t)(L < 0x7fffffff ? L : 0x7fffffff) : 0u;
    if (r < kPlanCache) s_len[r] = l32;
    local += l32;
  }
  const int64_t total = block_sum_i64(local, s_red);      // (its barriers also publish s_len)

  // Candidate m = warp + 1 "rounds": the smallest chunk (multiple of 16 tokens, >= min_chunk) whose item
  // count  items(chunk) = G * sum_r ceil(L_r / chunk)  fits m * num_ctas, found by bisection (items() is
  // monotone).  Cost model: rounds actually needed x (chunk + per-item overhead); lowest cost wins, ties go
  // to fewer rounds.  Every warp works alone (no block barrier inside the search).
  {
    const int m = warp + 1;
    uint32_t maxl = 0;
    for (int r = lane; r < bs; r += 32) maxl = max(maxl, len_at(r));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) maxl = max(maxl, __shfl_xor_sync(0xffffffffu, maxl, o));
    auto items = [&](uint32_t chunk) -> int64_t {          // chunk <= 2^31: L + chunk - 1 fits 32 bits
      int64_t n = 0;
      for (int r = lane; r < bs; r += 32) {
        const uint32_t L = len_at(r);
        if (L > 0) n += (L + chunk - 1) / chunk;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
      return n * p.num_kv_heads;
    };
    const uint32_t u_min = (uint32_t)(p.min_chunk + 15) / 16;
    uint32_t u_max = (uint32_t)(((int64_t)maxl + 15) / 16);
    if (u_max < u_min) u_max = u_min;
    const int64_t cap = (int64_t)m * p.num_ctas;
    uint32_t lo = u_min, hi = u_max;           // smallest u in [u_min, u_max] with items(16 u) <= cap, else u_max
    while (lo < hi) {
      const uint32_t mid = (lo + hi) >> 1;
      if (items(16 * mid) <= cap) hi = mid; else lo = mid + 1;
    }
    const int64_t chunk = 16 * (int64_t)lo;
    const int64_t rounds = (items((uint32_t)chunk) + p.num_ctas - 1) / p.num_ctas;
    if (lane == 0) {
      s_cost[warp] = (m <= p.waves) ? rounds * (chunk + p.min_chunk / 4) : (int64_t(1) << 62);
      s_cand[warp] = chunk;
    }
  }
  __syncthreads();
  if (tid == 0) {
    int64_t chunk;
    if (p.fixed_chunk > 0) {
      chunk = p.fixed_chunk;
    } else {
      int best = 0;
      for (int w = 1; w < 32; ++w)
        if (s_cost[w] < s_cost[best]) best = w;
      chunk = s_cand[best];
    }
    // capacity: sum_r ceil(L_r / chunk) <= total / chunk + bs must fit the work list
    while (total / chunk + bs > p.max_chunks && chunk < (int64_t(1) << 30)) chunk *= 2;
    s_chunk = (int32_t)chunk;
    s_carry = 0;
  }
  __syncthreads();
  const int chunk = s_chunk;

  int32_t* chunk_offset = p.plan + kPlanHeader;
  int32_t* chunk_req = p.plan + kPlanHeader + p.max_bs + 1;

  // exclusive scan of chunks-per-request, 1024 requests per pass
  for (int base = 0; base < bs; base += kPlanThreads) {
    const int r = base + tid;
    int32_t n = 0;
    if (r < bs) {
      const uint32_t L = len_at(r);
      n = L > 0 ? (int32_t)((L + (uint32_t)chunk - 1) / (uint32_t)chunk) : 0;
    }
    int32_t inc = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int32_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp_tot[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      int32_t w = s_warp_tot[lane];
      int32_t winc = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int32_t t = __shfl_up_sync(0xffffffffu, winc, o);
        if (lane >= o) winc += t;
      }
      s_warp_tot[lane] = winc - w;  // exclusive warp offsets
      if (lane == 31) s_red[0] = winc;  // pass total
    }
    __syncthreads();
    const int32_t carry = s_carry;
    if (r < bs) chunk_offset[r] = carry + s_warp_tot[warp] + inc - n;
    __syncthreads();
    if (tid == 0) s_carry = carry + (int32_t)s_red[0];
    __syncthreads();
  }
  const int32_t total_chunks = s_carry;
  if (tid == 0) {
    chunk_offset[bs] = total_chunks;
    p.plan[0] = chunk;
    p.plan[1] = total_chunks;
    p.plan[2] = bs;
    p.plan[3] = p.max_chunks;
    p.plan[4] = p.window_keys;
  }
  __syncthreads();
  // owner of each chunk: binary search in chunk_offset (block-visible after the barrier)
  for (int c = tid; c < total_chunks && c < p.max_chunks; c += kPlanThreads) {
    int lo = 0, hi = bs;  // largest r with chunk_offset[r] <= c
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (chunk_offset[mid] <= c) lo = mid; else hi = mid;
    }
    chunk_req[c] = lo;
  }
}

// ----------------------------------------------------------------------------
// shared kernel parameters
// ----------------------------------------------------------------------------
struct DecodeParams {
  const void* q;
  void* o;
  const void* k_cache;
  const void* v_cache;
  const int32_t* req_to_token;
  const void* req_pool_indices;
  const void* seq_lens;
  const int32_t* plan;
  float* ws_o;
  float* ws_lse;
  int64_t q_stride_t, q_stride_h, o_stride_t, o_stride_h;
  int64_t k_stride_s, k_stride_h, v_stride_s, v_stride_h;
  int64_t req_to_token_stride;
  int32_t req_pool_indices_is_i64, seq_lens_is_i64;
  int32_t batch_size, num_q_heads, num_kv_heads, group;
  int32_t head_dim_qk, head_dim_v;
  int32_t max_bs;
  int32_t window_keys;
  float sm_scale, k_scale, v_scale, logit_cap;
  const float* sinks;   // [Hq] natural-log sink logits or nullptr
  float* lse;           // [bs, Hq] natural-log LSE output or nullptr
};

// attention-sink logit of q head h in log2 units (-inf when there are no sinks)
__device__ __forceinline__ float sink_log2(const DecodeParams& p, int h) {
  return p.sinks != nullptr ? p.sinks[h] * kLog2e : -INFINITY;
}

struct ChunkInfo {
  int req;        // batch row
  int start;      // first token of the chunk inside the request
  int n_tok;      // tokens in the chunk
  int n_chunks;   // chunks of this request (1 -> write the final output directly)
  int64_t pool_row;  // row of req_to_token
};

__device__ __forceinline__ ChunkInfo load_chunk(const DecodeParams& p, int c) {
  const int32_t* chunk_offset = p.plan + kPlanHeader;
  const int32_t* chunk_req = p.plan + kPlanHeader + p.max_bs + 1;
  ChunkInfo ci;
  const int chunk = p.plan[0];
  ci.req = chunk_req[c];
  const int off = chunk_offset[ci.req];
  ci.n_chunks = chunk_offset[ci.req + 1] - off;
  const int64_t L_full = load_index(p.seq_lens, p.seq_lens_is_i64, ci.req);
  const int64_t L = (p.window_keys > 0 && L_full > p.window_keys) ? (int64_t)p.window_keys : L_full;
  // equal-sized chunks inside a request: n = ceil(L / chunk) pieces of roundup16(ceil(L / n)) tokens
  // (16 = the kernels' token tile; cs <= chunk because chunk is a multiple of 16, so (n-1) cs < L)
  const int cs = (int)(((L + ci.n_chunks - 1) / ci.n_chunks + 15) / 16 * 16);
  const int rel = (c - off) * cs;                       // first token of the chunk inside the attended range
  const int64_t rem = L - rel;
  ci.n_tok = (int)(rem < cs ? rem : cs);
  ci.start = (int)(L_full - L) + rel;                   // the attended range is the last L tokens of the request
  (void)chunk;
  ci.pool_row = load_index(p.req_pool_indices, p.req_pool_indices_is_i64, ci.req);
  return ci;
}

// ----------------------------------------------------------------------------
// generic kernel: any head_dim <= 512, any dtype pair.  One warp per (chunk, q head).
// ----------------------------------------------------------------------------
constexpr int kGenWarps = 4;

template <typename QT, typename KVT>
__global__ void __launch_bounds__(kGenWarps* kWarp) decode_generic_kernel(const DecodeParams p) {
  extern __shared__ float smem_f[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float* q_s = smem_f + warp * p.head_dim_qk;
  const int Dk = p.head_dim_qk, Dv = p.head_dim_v, Hq = p.num_q_heads;
  const int total_chunks = p.plan[1];
  const int64_t n_items = (int64_t)total_chunks * Hq;
  const QT* q = reinterpret_cast<const QT*>(p.q);
  const KVT* kc = reinterpret_cast<const KVT*>(p.k_cache);
  const KVT* vc = reinterpret_cast<const KVT*>(p.v_cache);
  const float qk_scale = p.sm_scale * p.k_scale;

  for (int64_t item = (int64_t)blockIdx.x * kGenWarps + warp; item < n_items;
       item += (int64_t)gridDim.x * kGenWarps) {
    const int c = (int)(item / Hq), h = (int)(item % Hq);
    const ChunkInfo ci = load_chunk(p, c);
    const int kvh = h / p.group;
    const QT* qrow = q + (int64_t)ci.req * p.q_stride_t + (int64_t)h * p.q_stride_h;
    __syncwarp();
    for (int d = lane; d < Dk; d += kWarp) q_s[d] = to_float(qrow[d]);
    __syncwarp();
    const int32_t* pt = p.req_to_token + ci.pool_row * p.req_to_token_stride + ci.start;

    RowState st;
#pragma unroll
    for (int i = 0; i < kGenAcc; ++i) st.acc[i] = 0.f;
    attend_row_generic<KVT>(st, q_s, pt, ci.n_tok, kc, vc, p.k_stride_s, p.v_stride_s,
                            (int64_t)kvh * p.k_stride_h, (int64_t)kvh * p.v_stride_h, Dk, Dv, qk_scale,
                            p.logit_cap);
    if (ci.n_chunks == 1) {
      const float l_fin = st.l + exp2f(sink_log2(p, h) - st.m);   // + the attention sink's share of the denominator
      const float inv_l = p.v_scale / l_fin;
      QT* orow = reinterpret_cast<QT*>(p.o) + (int64_t)ci.req * p.o_stride_t + (int64_t)h * p.o_stride_h;
#pragma unroll
      for (int i = 0; i < kGenAcc; ++i) {
        const int d = lane + i * kWarp;
        if (d < Dv) orow[d] = from_float<QT>(st.acc[i] * inv_l);
      }
      if (p.lse != nullptr && lane == 0) p.lse[(int64_t)ci.req * Hq + h] = (st.m + log2f(l_fin)) * kLn2;
    } else {
      const float inv_l = 1.f / st.l;
      float* wo = p.ws_o + ((int64_t)c * Hq + h) * Dv;
#pragma unroll
      for (int i = 0; i < kGenAcc; ++i) {
        const int d = lane + i * kWarp;
        if (d < Dv) wo[d] = st.acc[i] * inv_l;
      }
      if (lane == 0) p.ws_lse[(int64_t)c * Hq + h] = st.m + log2f(st.l);
    }
  }
}

// ----------------------------------------------------------------------------
// split merge: one warp per (request, q head); requests with a single chunk were
// already written by the attention kernel.  Also writes zeros for empty requests.
// ----------------------------------------------------------------------------
constexpr int kMergeWarps = 4;

template <typename QT>
__global__ void __launch_bounds__(kMergeWarps* kWarp) decode_merge_kernel(const DecodeParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t w = (int64_t)blockIdx.x * kMergeWarps + (threadIdx.x >> 5);
  const int Hq = p.num_q_heads, Dv = p.head_dim_v;
  if (w >= (int64_t)p.batch_size * Hq) return;
  const int r = (int)(w / Hq), h = (int)(w % Hq);
  const int32_t* chunk_offset = p.plan + kPlanHeader;
  const int c0 = chunk_offset[r], n = chunk_offset[r + 1] - c0;
  if (n == 1) return;
  QT* orow = reinterpret_cast<QT*>(p.o) + (int64_t)r * p.o_stride_t + (int64_t)h * p.o_stride_h;
  if (n == 0) {
    for (int d = lane; d < Dv; d += kWarp) orow[d] = from_float<QT>(0.f);
    if (p.lse != nullptr && lane == 0) p.lse[(int64_t)r * Hq + h] = -INFINITY;
    return;
  }
  float mx = -INFINITY;
  for (int i = lane; i < n; i += kWarp) mx = fmaxf(mx, p.ws_lse[(int64_t)(c0 + i) * Hq + h]);
  mx = warp_max(mx);
  float den = 0.f;
  for (int i = lane; i < n; i += kWarp) den += exp2f(p.ws_lse[(int64_t)(c0 + i) * Hq + h] - mx);
  den = warp_sum(den) + exp2f(sink_log2(p, h) - mx);
  if (p.lse != nullptr && lane == 0) p.lse[(int64_t)r * Hq + h] = (mx + log2f(den)) * kLn2;
  const float inv = p.v_scale / den;
  for (int d0 = 0; d0 < Dv; d0 += kWarp) {
    const int d = d0 + lane;
    float acc = 0.f;
    for (int i = 0; i < n; ++i) {
      const float wgt = exp2f(p.ws_lse[(int64_t)(c0 + i) * Hq + h] - mx);
      if (d < Dv) acc = fmaf(wgt, p.ws_o[((int64_t)(c0 + i) * Hq + h) * Dv + d], acc);
    }
    if (d < Dv) orow[d] = from_float<QT>(acc * inv);
  }
}

}  // namespace

// fast path lives in its own translation unit section
#include "decode_mma.cuh"
#include "decode_hg.cuh"

namespace {

template <typename QT>
int launch_generic(const DecodeParams& p, int kv_dtype, int num_ctas, cudaStream_t stream) {
  const size_t smem = (size_t)kGenWarps * p.head_dim_qk * sizeof(float);
  const int grid = num_ctas * 8;
  switch (kv_dtype) {
    case B200_BF16: decode_generic_kernel<QT, __nv_bfloat16><<<grid, kGenWarps * kWarp, smem, stream>>>(p); break;
    case B200_F16: decode_generic_kernel<QT, __half><<<grid, kGenWarps * kWarp, smem, stream>>>(p); break;
    case B200_FP8_E4M3: decode_generic_kernel<QT, e4m3_t><<<grid, kGenWarps * kWarp, smem, stream>>>(p); break;
    case B200_FP8_E5M2: decode_generic_kernel<QT, e5m2_t><<<grid, kGenWarps * kWarp, smem, stream>>>(p); break;
    default: set_error("decode: unsupported kv dtype %d", kv_dtype); return B200_EINVAL;
  }
  B200_CHECK_CUDA(cudaGetLastError());
  return B200_OK;
}

}  // namespace
}  // namespace b200

extern "C" int64_t b200attn_decode_plan_ints(int32_t max_bs, int32_t max_chunks) {
  return (int64_t)b200::kPlanHeader + (int64_t)max_bs + 1 + (int64_t)max_chunks;
}

extern "C" int b200attn_decode_plan(const b200_decode_plan_args* a, b200_stream_t stream_) {
  using namespace b200;
  B200_CHECK_ARG(a && a->seq_lens && a->plan, "decode_plan: null pointer");
  B200_CHECK_ARG(a->batch_size >= 0 && a->batch_size <= a->max_bs, "decode_plan: bs %d > max_bs %d",
                 a->batch_size, a->max_bs);
  B200_CHECK_ARG(a->max_chunks > a->batch_size, "decode_plan: max_chunks %d must exceed bs %d",
                 a->max_chunks, a->batch_size);
  B200_CHECK_ARG(a->num_kv_heads > 0 && a->num_ctas > 0 && a->waves > 0, "decode_plan: bad grid hints");
  B200_CHECK_ARG(a->min_chunk_tokens > 0 && a->min_chunk_tokens % 16 == 0,
                 "decode_plan: min_chunk_tokens must be a positive multiple of 16");
  B200_CHECK_ARG(a->fixed_chunk_tokens % 16 == 0, "decode_plan: fixed_chunk_tokens must be a multiple of 16");
  B200_CHECK_ARG(a->fixed_chunk_tokens >= 0, "decode_plan: negative fixed_chunk_tokens");
  PlanParams p;
  p.seq_lens = a->seq_lens; p.plan = a->plan; p.seq_lens_is_i64 = a->seq_lens_is_i64;
  p.batch_size = a->batch_size; p.max_bs = a->max_bs; p.max_chunks = a->max_chunks;
  p.num_kv_heads = a->num_kv_heads; p.num_ctas = a->num_ctas; p.waves = a->waves;
  p.min_chunk = a->min_chunk_tokens; p.fixed_chunk = a->fixed_chunk_tokens;
  p.window_keys = a->window_keys > 0 ? a->window_keys : 0;
  decode_plan_kernel<<<1, kPlanThreads, 0, reinterpret_cast<cudaStream_t>(stream_)>>>(p);
  B200_CHECK_CUDA(cudaGetLastError());
  return B200_OK;
}

namespace b200 {
namespace {
constexpr int kDraftTableThreads = 256;

__global__ void __launch_bounds__(kDraftTableThreads) draft_table_kernel(const b200_draft_table_args a) {
  const int row = blockIdx.x;
  const int b = row / a.topk, t = row - b * a.topk;
  const int64_t L64 = load_index(a.seq_lens, a.seq_lens_is_i64, b);
  const int64_t width = a.req_to_token_stride < a.table_stride ? a.req_to_token_stride : a.table_stride;
  const int L = (int)(L64 < 0 ? 0 : (L64 < width ? L64 : width));
  const int32_t* src = a.req_to_token + load_index(a.req_pool_indices, a.req_pool_indices_is_i64, b) * a.req_to_token_stride;
  int32_t* dst = a.table + (int64_t)row * a.table_stride;
  for (int p = threadIdx.x; p < L; p += kDraftTableThreads) dst[p] = __ldg(src + p);
  // first drafted slot of this branch inside the request's row (cache_locs.py:104-122)
  int64_t start;
  if (a.page_size == 1 || a.topk == 1) {
    start = (int64_t)L + (int64_t)t * a.num_steps;
  } else {
    const int last_page_len = L % a.page_size;
    const int new_pages = (last_page_len + a.num_steps + a.page_size - 1) / a.page_size;
    start = (int64_t)(L / a.page_size) * a.page_size + (int64_t)t * new_pages * a.page_size + last_page_len;
  }
  for (int j = threadIdx.x; j < a.num_steps; j += kDraftTableThreads)
    if (L + j < a.table_stride) dst[L + j] = start + j < a.req_to_token_stride ? __ldg(src + start + j) : 0;
}
}  // namespace
}  // namespace b200

extern "C" int b200attn_draft_table(const b200_draft_table_args* a, b200_stream_t stream_) {
  using namespace b200;
  B200_CHECK_ARG(a != nullptr, "draft_table: null args");
  if (a->num_seqs == 0) return B200_OK;
  B200_CHECK_ARG(a->req_to_token && a->req_pool_indices && a->seq_lens && a->table, "draft_table: null pointer");
  B200_CHECK_ARG(a->num_seqs > 0 && a->topk > 0 && a->num_steps > 0 && a->page_size > 0 && a->table_stride > 0 &&
                     a->req_to_token_stride > 0,
                 "draft_table: num_seqs %d topk %d num_steps %d page_size %d", a->num_seqs, a->topk, a->num_steps,
                 a->page_size);
  draft_table_kernel<<<(unsigned)(a->num_seqs * a->topk), kDraftTableThreads, 0, reinterpret_cast<cudaStream_t>(stream_)>>>(*a);
  B200_CHECK_CUDA(cudaGetLastError());
  return B200_OK;
}

extern "C" int b200attn_decode_fast_supported(int32_t dqk, int32_t dv, int32_t group, int32_t q_dtype,
                                              int32_t kv_dtype) {
  return b200::decode_fast_supported(dqk, dv, group, q_dtype, kv_dtype) ? 1 : 0;
}

namespace b200 {
namespace {
// -DB200_DEV builds only: B200ATTN_DECODE_HG=0 forces the per-head kernel (A/B measurements).  The shipped library
// reads no environment variable on a launch path.
bool decode_hg_enabled() {
#ifdef B200_DEV
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200ATTN_DECODE_HG");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v != 0;
#else
  return true;
#endif
}
bool decode_hg_shape_ok(int hq, int hkv, int dqk, int dv, int q_dtype, int kv_dtype) {
  if (!decode_hg_enabled() || hkv <= 0 || hq % hkv != 0 || hkv % kHG != 0) return false;
  if (!decode_hg_dims_supported(dqk, dv, hq / hkv, q_dtype, kv_dtype)) return false;
  return !(dqk == 256 && kv_dtype != B200_FP8_E4M3);
}
}  // namespace
}  // namespace b200

extern "C" int b200attn_decode_ctas_per_sm(int32_t head_dim, int32_t kv_dtype) {
  const int a = b200::decode_fast_ctas_per_sm(head_dim, kv_dtype);
  const int b = b200::decode_hg_enabled() && (head_dim == 64 || head_dim == 128 || head_dim == 256 || head_dim == 80 ||
                                              head_dim == 96 || head_dim == 112)
                    ? b200::decode_hg_ctas_per_sm(head_dim, kv_dtype) : 0;
  return a > b ? a : b;
}

extern "C" int b200attn_decode_items_per_chunk(int32_t num_q_heads, int32_t num_kv_heads, int32_t dqk, int32_t dv,
                                               int32_t q_dtype, int32_t kv_dtype) {
  if (b200::decode_hg_shape_ok(num_q_heads, num_kv_heads, dqk, dv, q_dtype, kv_dtype))
    return num_kv_heads / b200::kHG;
  return num_kv_heads;
}

extern "C" int b200attn_decode(const b200_decode_args* a, b200_stream_t stream_) {
  using namespace b200;
  B200_CHECK_ARG(a != nullptr, "decode: null args");
  if (a->batch_size == 0) return B200_OK;
  B200_CHECK_ARG(a->q && a->o && a->k_cache && a->v_cache && a->req_to_token && a->req_pool_indices &&
                     a->seq_lens && a->plan && a->ws_o && a->ws_lse,
                 "decode: null pointer");
  B200_CHECK_ARG(a->num_q_heads > 0 && a->num_kv_heads > 0 && a->num_q_heads % a->num_kv_heads == 0,
                 "decode: Hq=%d must be a multiple of Hkv=%d", a->num_q_heads, a->num_kv_heads);
  B200_CHECK_ARG(a->head_dim_qk > 0 && a->head_dim_qk <= kGenMaxD && a->head_dim_v > 0 &&
                     a->head_dim_v <= kGenMaxD,
                 "decode: head dims (%d, %d) must be in [1, %d]", a->head_dim_qk, a->head_dim_v, kGenMaxD);
  B200_CHECK_ARG(a->q_dtype == B200_BF16 || a->q_dtype == B200_F16, "decode: q dtype %d", a->q_dtype);
  B200_CHECK_ARG(a->batch_size <= a->max_bs, "decode: bs %d > max_bs %d", a->batch_size, a->max_bs);
  B200_CHECK_ARG(a->num_ctas > 0, "decode: num_ctas");
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);

  DecodeParams p;
  p.q = a->q; p.o = a->o; p.k_cache = a->k_cache; p.v_cache = a->v_cache;
  p.req_to_token = a->req_to_token; p.req_pool_indices = a->req_pool_indices; p.seq_lens = a->seq_lens;
  p.plan = a->plan; p.ws_o = a->ws_o; p.ws_lse = a->ws_lse;
  p.q_stride_t = a->q_stride_t; p.q_stride_h = a->q_stride_h;
  p.o_stride_t = a->o_stride_t; p.o_stride_h = a->o_stride_h;
  p.k_stride_s = a->k_stride_s; p.k_stride_h = a->k_stride_h;
  p.v_stride_s = a->v_stride_s; p.v_stride_h = a->v_stride_h;
  p.req_to_token_stride = a->req_to_token_stride;
  p.req_pool_indices_is_i64 = a->req_pool_indices_is_i64; p.seq_lens_is_i64 = a->seq_lens_is_i64;
  p.batch_size = a->batch_size; p.num_q_heads = a->num_q_heads; p.num_kv_heads = a->num_kv_heads;
  p.group = a->num_q_heads / a->num_kv_heads;
  p.head_dim_qk = a->head_dim_qk; p.head_dim_v = a->head_dim_v; p.max_bs = a->max_bs;
  p.sm_scale = a->sm_scale; p.k_scale = a->k_scale; p.v_scale = a->v_scale; p.logit_cap = a->logit_cap;
  p.window_keys = a->window_keys > 0 ? a->window_keys : 0; p.sinks = a->sinks; p.lse = a->lse;

  // the head-group kernel also covers D = 64 FP8 pools, which the per-head kernel does not
  const bool hg_ok = decode_hg_enabled() && decode_fast_layout_ok(p, a->kv_dtype) &&
                     decode_hg_supported(p, a->q_dtype, a->kv_dtype);
  const bool fast_ok = hg_ok || (decode_fast_supported(a->head_dim_qk, a->head_dim_v, p.group, a->q_dtype, a->kv_dtype) &&
                                 decode_fast_layout_ok(p, a->kv_dtype));
  B200_CHECK_ARG(!(a->kernel == B200_KERNEL_FAST && !fast_ok),
                 "decode: fast kernel does not cover Dk=%d Dv=%d group=%d q=%d kv=%d (or unaligned strides)",
                 a->head_dim_qk, a->head_dim_v, p.group, a->q_dtype, a->kv_dtype);
  const bool use_fast = fast_ok && a->kernel != B200_KERNEL_GENERIC;
  int rc;
  if (a->ev_attn_start) B200_CHECK_CUDA(cudaEventRecord(reinterpret_cast<cudaEvent_t>(a->ev_attn_start), stream));
  const bool use_hg = use_fast && hg_ok;
  if (use_hg) {
    rc = launch_decode_hg(p, a->q_dtype, a->kv_dtype, a->num_ctas, stream);
  } else if (use_fast) {
    rc = launch_decode_fast(p, a->q_dtype, a->kv_dtype, a->num_ctas, stream);
  } else if (a->q_dtype == B200_BF16) {
    rc = launch_generic<__nv_bfloat16>(p, a->kv_dtype, a->num_ctas, stream);
  } else {
    rc = launch_generic<__half>(p, a->kv_dtype, a->num_ctas, stream);
  }
  if (rc != B200_OK) return rc;
  if (a->ev_attn_stop) B200_CHECK_CUDA(cudaEventRecord(reinterpret_cast<cudaEvent_t>(a->ev_attn_stop), stream));
  if (use_fast) return B200_OK;   // the fast path launches its own split merge
  const int64_t rows = (int64_t)a->batch_size * a->num_q_heads;
  const unsigned mgrid = (unsigned)((rows + kMergeWarps - 1) / kMergeWarps);
  if (a->q_dtype == B200_BF16) {
    decode_merge_kernel<__nv_bfloat16><<<mgrid, kMergeWarps * kWarp, 0, stream>>>(p);
  } else {
    decode_merge_kernel<__half><<<mgrid, kMergeWarps * kWarp, 0, stream>>>(p);
  }
  B200_CHECK_CUDA(cudaGetLastError());
  return B200_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// b200attn_decode_static_ws: b200attn_decode for callers that cannot allocate (no counterpart in the reference).
// The split-KV partials go to a file-scope device array instead of
// caller memory.  `g_decode_static_ws` is a HOST-SIDE STAND-IN in host code, not a device address: the address the
// kernels get comes from cudaGetSymbolAddress, resolved once per device with that device current.
// Every call on a device shares the array (overlapping calls race); nothing in it has to start at zero, because
// each launch writes every (chunk, head) partial its merge reads.
// ---------------------------------------------------------------------------------------------------------------
namespace b200 {
namespace {
constexpr int64_t kStaticWsFloats = (int64_t)B200_DECODE_STATIC_WS_BYTES / (int64_t)sizeof(float);
static __device__ __align__(256) float g_decode_static_ws[kStaticWsFloats];

constexpr int kStaticWsMaxDevices = 64;
std::atomic<float*> g_static_ws_addr[kStaticWsMaxDevices];

// device address of the array on the CURRENT device (set-up on first use, no allocation)
int static_ws_address(float** out) {
  int dev = -1;
  B200_CHECK_CUDA(cudaGetDevice(&dev));
  B200_CHECK_ARG(dev >= 0 && dev < kStaticWsMaxDevices, "decode_static_ws: device ordinal %d >= %d", dev,
                 kStaticWsMaxDevices);
  float* ptr = g_static_ws_addr[dev].load(std::memory_order_acquire);
  if (ptr == nullptr) {
    void* sym = nullptr;
    B200_CHECK_CUDA(cudaGetSymbolAddress(&sym, g_decode_static_ws));
    B200_CHECK_ARG(sym != nullptr, "decode_static_ws: cudaGetSymbolAddress returned null on device %d", dev);
    ptr = static_cast<float*>(sym);
    g_static_ws_addr[dev].store(ptr, std::memory_order_release);
  }
  *out = ptr;
  return B200_OK;
}
}  // namespace
}  // namespace b200

extern "C" int b200attn_decode_static_ws(const b200_decode_args* a, b200_stream_t stream_) {
  using namespace b200;
  B200_CHECK_ARG(a != nullptr, "decode_static_ws: null args");
  if (a->batch_size == 0) return B200_OK;
  B200_CHECK_ARG(a->max_chunks > 0 && a->num_q_heads > 0 && a->head_dim_v > 0,
                 "decode_static_ws: max_chunks %d, Hq %d, Dv %d must be positive", a->max_chunks, a->num_q_heads,
                 a->head_dim_v);
  // ws_o first (rounded up so that ws_lse starts 256-byte aligned), then ws_lse
  const int64_t o_floats = ((int64_t)a->max_chunks * a->num_q_heads * a->head_dim_v + 63) / 64 * 64;
  const int64_t lse_floats = (int64_t)a->max_chunks * a->num_q_heads;
  B200_CHECK_ARG(o_floats + lse_floats <= kStaticWsFloats,
                 "decode_static_ws: max_chunks %d x Hq %d x (Dv %d + 1) needs %lld bytes of split workspace, the "
                 "static array holds %lld (B200_DECODE_STATIC_WS_BYTES); plan with fewer chunks or use "
                 "b200attn_decode with caller memory",
                 a->max_chunks, a->num_q_heads, a->head_dim_v, (long long)((o_floats + lse_floats) * 4),
                 (long long)B200_DECODE_STATIC_WS_BYTES);
  float* ws = nullptr;
  const int rc = static_ws_address(&ws);
  if (rc != B200_OK) return rc;
  b200_decode_args b = *a;
  b.ws_o = ws;
  b.ws_lse = ws + o_floats;
  return b200attn_decode(&b, stream_);
}
