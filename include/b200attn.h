/*
 * b200attn.h -- C ABI of the B200-native RadixAttention hot path (libb200attn.so).
 *
 * Plain pointers, sizes and a cudaStream_t; no torch types.  Every entry point
 * returns 0 on success and a negative B200_E* code on failure (the text is
 * available from b200attn_last_error()).  Nothing here allocates device memory:
 * workspaces are sized by the *_ints/_floats helpers and owned by the caller, so
 * every launch is CUDA-graph capturable.
 *
 * Reference interfaces each entry point stands in for (paths relative to the
 * reference tree, python/sglang/...):
 *   b200attn_store_kv        srt/mem_cache/memory_pool.py:141-192 (_set_kv_buffer_impl),
 *                            :2305-2381 (MHATokenToKVPool.set_kv_buffer cast/scale/view rules),
 *                            kernels/ops/kvcache/kvcache.py:57-122 (store_cache),
 *                            kernels/jit/csrc/elementwise/kvcache.cuh:188-219 (store_kvcache),
 *                            kernels/jit/csrc/attention/fused_fp8_qkv_kv_cache.cuh:58-90 (fused FP8 variant)
 *   b200attn_decode_plan     srt/layers/attention/triton_backend.py:304-351 (get_num_kv_splits),
 *                            :430-446 (_fill_kv_indptr_and_indices),
 *                            kernels/ops/kvcache/kv_indices.py:8-45 (page-table gather; done in-kernel here)
 *   b200attn_decode          srt/layers/attention/torch_native_backend.py:176-277 (oracle semantics),
 *                            kernels/ops/attention/decode_attention.py:968-1044 (decode_attention_fwd:
 *                            split-KV stage 1 + stage-2 merge)
 *   b200attn_decode_static_ws  none: no counterpart in the reference, added
 *                            (b200attn_decode with the split workspace in a static device array; see below)
 *   b200attn_draft_table     kernels/ops/speculative/cache_locs.py:56-130 (generate_draft_decode_kv_indices: the kv
 *                            positions branch t of a tree draft attends at step i), triton_backend.py:1912-1955
 *   b200attn_extend_plan     srt/layers/attention/triton_backend.py:860-900 (qo_indptr / kv_indptr for extend)
 *   b200attn_rope_store_kv   srt/layers/rotary_embedding/base.py:234-275,363-385 + set_kv_buffer (see below)
 *   b200attn_extend          srt/layers/attention/torch_native_backend.py:61-174 (oracle semantics),
 *                            kernels/ops/attention/extend_attention.py:664-812 (extend_attention_fwd)
 *   b200attn_merge_state     kernels/aot/csrc/attention/merge_attn_states.cu, kernels/ops/attention/merge_state.py
 *   b200attn_oproj_allreduce srt/layers/linear.py:1563 (RowParallelLinear all-reduce), custom_all_reduce_v2.py
 */
#ifndef B200ATTN_H_
#define B200ATTN_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200ATTN_ABI_VERSION 8

typedef void* b200_stream_t; /* a cudaStream_t */

/* element types */
enum {
  B200_BF16 = 0,
  B200_F16 = 1,
  B200_FP8_E4M3 = 2, /* torch.float8_e4m3fn, stored as uint8 in the pool */
  B200_FP8_E5M2 = 3  /* torch.float8_e5m2 */
};

/* error codes */
enum {
  B200_OK = 0,
  B200_EINVAL = -1,   /* unsupported shape / dtype / null pointer */
  B200_ECUDA = -2,    /* a CUDA runtime call failed */
  B200_ENOTSM100 = -3 /* the device is not compute capability 10.x */
};

/* kernel selection for b200attn_decode / b200attn_extend */
enum {
  B200_KERNEL_AUTO = 0,
  B200_KERNEL_GENERIC = 1, /* any head_dim <= 512, CUDA-core path */
  B200_KERNEL_FAST = 2     /* decode: mma + cp.async split-KV; extend: tcgen05/TMEM */
};

const char* b200attn_last_error(void);
int b200attn_abi_version(void);
/* sm_count / cc of the current device */
int b200attn_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* ------------------------------------------------------------------------ *
 * KV-store scatter:  k_cache[loc[i]] = cast(k[i] / k_scale),  same for V.
 *   - rows whose loc equals skip_index are not written (slot-0 sink convention of
 *     store_cache(); pass -1 to write every row like `k_cache[loc] = k`)
 *   - loc outside [0, size_limit) traps on the device (kvcache.cuh:209)
 *   - src_dtype == dst_dtype is a byte copy; BF16/F16 -> FP8 is torch's
 *     round-to-nearest-even, non-saturating cast (|x| > 464 -> NaN for e4m3)
 *   - k_scale/v_scale != 1: divide in fp32, round to src_dtype, then cast --
 *     the double rounding of `cache_k.div_(k_scale); cache_k.to(fp8)`
 * ------------------------------------------------------------------------ */
typedef struct b200_store_kv_args {
  const void* k;             /* [num_tokens, k_row_elems] src_dtype */
  const void* v;             /* [num_tokens, v_row_elems] src_dtype */
  int64_t k_src_stride;      /* elements between source rows */
  int64_t v_src_stride;
  void* k_cache;             /* [slots, k_row_elems] dst_dtype */
  void* v_cache;
  int64_t k_cache_stride;    /* elements between slots */
  int64_t v_cache_stride;
  const void* loc;           /* [num_tokens] int32 or int64 slot ids */
  int32_t loc_is_i64;
  int64_t num_tokens;
  int32_t k_row_elems;       /* Hkv * Dk */
  int32_t v_row_elems;       /* Hkv * Dv */
  int32_t src_dtype;
  int32_t dst_dtype;
  float k_scale;             /* 1.0f = none */
  float v_scale;
  int64_t size_limit;        /* valid slot ids are [0, size_limit) */
  int64_t skip_index;        /* -1 = none */
} b200_store_kv_args;

int b200attn_store_kv(const b200_store_kv_args* args, b200_stream_t stream);

/* ------------------------------------------------------------------------ *
 * Decode planner.  One tiny launch per forward (graph-capturable): reads
 * seq_lens on the device and writes the split-KV work list into `plan`:
 *   plan[0] = chunk_tokens   plan[1] = total_chunks (over all requests)
 *   plan[2] = batch_size     plan[3] = max_chunks
 *   plan[8 .. 8+bs]                    chunk_offset[r]  (exclusive scan of chunks per request)
 *   plan[8+max_bs+1 .. +max_chunks]    chunk_req[c]     (owning request of chunk c)
 * A work item of the decode kernel is (chunk c, kv head h).
 * ------------------------------------------------------------------------ */
int64_t b200attn_decode_plan_ints(int32_t max_bs, int32_t max_chunks);

typedef struct b200_decode_plan_args {
  const void* seq_lens;      /* [bs] int32 or int64, device */
  int32_t seq_lens_is_i64;
  int32_t batch_size;
  int32_t max_bs;            /* capacity the plan buffer was sized for */
  int32_t max_chunks;
  int32_t num_kv_heads;      /* work multiplicity per chunk */
  int32_t num_ctas;          /* persistent CTAs the decode kernel will run */
  int32_t waves;             /* most rounds (work items per CTA) the planner considers, <= 32 */
  int32_t min_chunk_tokens;  /* smallest chunk the planner may choose, multiple of 16; min_chunk/4 is also its
                                per-item overhead estimate (in tokens) when it ranks candidate chunk sizes */
  int32_t fixed_chunk_tokens;/* > 0 overrides the heuristic (deterministic splits) */
  int32_t* plan;             /* device, b200attn_decode_plan_ints() ints */
  int32_t window_keys;       /* sliding window: only the last window_keys tokens of a request are attended
                                (0 = off).  torch_native's sliding_window_size W is window_keys = W + 1
                                (torch_native_backend.py:36-48: k_pos >= q_pos - W) */
} b200_decode_plan_args;

int b200attn_decode_plan(const b200_decode_plan_args* args, b200_stream_t stream);

/* ------------------------------------------------------------------------ *
 * Paged decode attention (one query token per request), split-KV + merge.
 *   o[r, h, :] = softmax(q[r, h, :] . K[tok(r, 0..L_r)]^T * sm_scale * k_scale) V[...] * v_scale
 *   tok(r, j) = req_to_token[req_pool_indices[r], j]   (gathered inside the kernel)
 * ------------------------------------------------------------------------ */
typedef struct b200_decode_args {
  const void* q;             /* [bs, Hq, Dk] q_dtype */
  void* o;                   /* [bs, Hq, Dv] q_dtype */
  int64_t q_stride_t, q_stride_h;   /* elements */
  int64_t o_stride_t, o_stride_h;
  const void* k_cache;       /* [slots, Hkv, Dk] kv_dtype */
  const void* v_cache;       /* [slots, Hkv, Dv] kv_dtype */
  int64_t k_stride_s, k_stride_h;   /* elements */
  int64_t v_stride_s, v_stride_h;
  const int32_t* req_to_token;      /* [max_reqs+1, max_context_len] */
  int64_t req_to_token_stride;
  const void* req_pool_indices;     /* [bs] int32 or int64 */
  int32_t req_pool_indices_is_i64;
  const void* seq_lens;             /* [bs] int32 or int64 */
  int32_t seq_lens_is_i64;
  int32_t batch_size;
  int32_t num_q_heads, num_kv_heads;
  int32_t head_dim_qk, head_dim_v;
  int32_t q_dtype, kv_dtype;
  float sm_scale;            /* layer.scaling */
  float k_scale, v_scale;    /* FP8 descale factors, 1.0f = none */
  float logit_cap;           /* 0 = off; tanh cap otherwise */
  const int32_t* plan;       /* written by b200attn_decode_plan on the same stream */
  int32_t max_bs, max_chunks;
  float* ws_o;               /* [max_chunks * Hq * Dv] partial outputs */
  float* ws_lse;             /* [max_chunks * Hq] partial log2-sum-exp */
  int32_t num_ctas;          /* persistent grid size (same as in the plan) */
  int32_t kernel;            /* B200_KERNEL_* */
  void* ev_attn_start;       /* optional cudaEvent_t recorded right before the attention kernel */
  void* ev_attn_stop;        /* optional cudaEvent_t recorded right after it (before a separate merge) */
  int32_t window_keys;       /* sliding window, same value the plan was built with (0 = off) */
  const float* sinks;        /* optional [Hq] fp32 attention-sink logits (natural-log units, not scaled):
                                denominator += exp(sink_h - max)  (extend_attention.py:633-635,
                                decode_attention.py stage 2); NULL = none */
  float* lse;                /* optional [bs, Hq] natural-log LSE output (sink included), NULL = none */
} b200_decode_args;

int b200attn_decode(const b200_decode_args* args, b200_stream_t stream);

/* ------------------------------------------------------------------------ *
 * b200attn_decode without a caller-owned split-KV workspace.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments, same kernels and same results as b200attn_decode,
 * except that args->ws_o / args->ws_lse are IGNORED (may be NULL): the split partials live in a file-scope
 * `static __device__` array of B200_DECODE_STATIC_WS_BYTES inside the library (decode.cu), which every GPU that
 * loads the library holds until exit.  The call allocates nothing (no cudaMalloc / cudaFree / allocating library
 * routine); inputs, outputs and the plan stay caller-owned.
 *   limit   max_chunks * Hq * Dv floats (rounded up to 64) + max_chunks * Hq floats <= B200_DECODE_STATIC_WS_BYTES / 4;
 *           beyond it the call returns B200_EINVAL and b200attn_last_error() names both sizes.  Choose max_chunks
 *           (the plan capacity) accordingly, e.g. Hq 32, Dv 128: max_chunks <= 1 016.
 *   set-up  the first call on a device resolves the array's device address (cudaGetSymbolAddress, cached per
 *           device); make that first call with the device current and outside stream capture.
 *   sharing every call on a device uses the SAME array, and it keeps the last call's partials.  Nothing in it has to
 *           start at zero (each launch writes every partial it later reads), so nothing is cleared -- but two
 *           calls that overlap on the GPU (different streams, or graphs replayed concurrently) race on it.
 * ------------------------------------------------------------------------ */
#define B200_DECODE_STATIC_WS_BYTES (16 * 1024 * 1024)
int b200attn_decode_static_ws(const b200_decode_args* args, b200_stream_t stream);

/* ------------------------------------------------------------------------ *
 * EAGLE tree drafting (topk > 1): branch t of request b attends the request's prefix plus the tokens drafted on ITS
 * branch, whose slots the worker keeps at req_to_token[b, seq_len + t * num_steps + j] (page_size 1 or topk 1) or at the
 * page-aligned position of generate_draft_decode_kv_indices (cache_locs.py:104-122) -- not a prefix of the row.  This
 * builds, once per draft, one page-table row per (request, branch):
 *   table[b * topk + t, 0 .. seq_len)            = req_to_token[pool(b), 0 .. seq_len)
 *   table[b * topk + t, seq_len + j], j < num_steps = the branch's j-th drafted slot
 * so that step i of the draft is a plain b200attn_decode over `table` with lengths seq_len + i + 1 and
 * req_pool_indices = 0 .. num_seqs * topk - 1.  Copies only seq_len + num_steps entries per row (read on the device):
 * static launch shape, CUDA-graph recordable.
 * ------------------------------------------------------------------------ */
typedef struct b200_draft_table_args {
  const int32_t* req_to_token;
  int64_t req_to_token_stride;   /* also the readable width of a row */
  const void* req_pool_indices;  /* [num_seqs] */
  int32_t req_pool_indices_is_i64;
  const void* seq_lens;          /* [num_seqs] prefix lengths */
  int32_t seq_lens_is_i64;
  int32_t num_seqs, topk, num_steps, page_size;
  int32_t* table;                /* [num_seqs * topk, table_stride] */
  int64_t table_stride;
} b200_draft_table_args;

int b200attn_draft_table(const b200_draft_table_args* args, b200_stream_t stream);

/* 1 if the per-head tensor-core decode kernel covers this shape, else 0.  The head-group kernel (4 adjacent kv heads
 * per CTA, picked when num_kv_heads % 4 == 0) additionally covers D = 64 with FP8 pools and D = 80 / 96 / 112 with
 * 16-bit pools; b200attn_decode_items_per_chunk() returns num_kv_heads / 4 exactly when it will run. */
int b200attn_decode_fast_supported(int32_t head_dim_qk, int32_t head_dim_v, int32_t group,
                                   int32_t q_dtype, int32_t kv_dtype);
/* persistent CTAs per SM the fast decode kernel is tuned for */
int b200attn_decode_ctas_per_sm(int32_t head_dim, int32_t kv_dtype);
/* Work items the attention kernel creates per split chunk: num_kv_heads for the per-head kernel,
 * num_kv_heads / 4 when the head-group kernel (4 adjacent kv heads per CTA) covers the shape.  Pass it
 * as b200_decode_plan_args.num_kv_heads so that the planner balances the persistent grid. */
int b200attn_decode_items_per_chunk(int32_t num_q_heads, int32_t num_kv_heads, int32_t head_dim_qk,
                                    int32_t head_dim_v, int32_t q_dtype, int32_t kv_dtype);

/* ------------------------------------------------------------------------ *
 * Extend (prefill over a cached prefix) planner: qo_indptr = exclusive scan of
 * extend_seq_lens and the list of (request, q-tile) work items.
 *   plan[0] = num_tiles   plan[1] = block_q   plan[2] = batch_size
 *   plan[8 .. 8+bs]                qo_indptr[r]
 *   plan[8+max_bs+1 + 2*i + {0,1}] tile i: {request, first q row inside the request}
 * ------------------------------------------------------------------------ */
int64_t b200attn_extend_plan_ints(int32_t max_bs, int32_t max_tiles);

typedef struct b200_extend_plan_args {
  const void* extend_seq_lens;  /* [bs] int32 or int64 */
  int32_t lens_is_i64;
  int32_t batch_size;
  int32_t max_bs;
  int32_t max_tiles;
  int32_t block_q;              /* query rows (tokens) per tile */
  int32_t* plan;
  /* Optional page check (req_to_token != NULL and page_size a power of two >= 16): the plan also verifies, on the
   * device, that every request's positions sit on consecutive pool rows inside each page, which lets
   * b200attn_extend(page_size) fetch K/V by TMA page boxes.  A table that fails (or no check) is read row by row --
   * the result is the same either way. */
  const int32_t* req_to_token;
  int64_t req_to_token_stride;
  const void* req_pool_indices;     /* [bs] */
  int32_t req_pool_indices_is_i64;
  const void* extend_prefix_lens;   /* [bs], same integer type as extend_seq_lens */
  int32_t page_size;
} b200_extend_plan_args;

int b200attn_extend_plan(const b200_extend_plan_args* args, b200_stream_t stream);

/* ------------------------------------------------------------------------ *
 * Extend attention.  Query token i of request r (0 <= i < E_r) attends the pool
 * rows tok(r, 0 .. P_r + i) -- the cached prefix plus the causal part of the new
 * tokens, which the caller has already stored (store_kv runs first, exactly as
 * torch_native_backend.py:298-301 does before :312).  causal == 0 attends all
 * P_r + E_r rows (encoder-only / cross attention).
 * ------------------------------------------------------------------------ */
typedef struct b200_extend_args {
  const void* q;             /* [sum E_r, Hq, Dk] */
  void* o;                   /* [sum E_r, Hq, Dv] */
  int64_t q_stride_t, q_stride_h;
  int64_t o_stride_t, o_stride_h;
  const void* k_cache;
  const void* v_cache;
  int64_t k_stride_s, k_stride_h;
  int64_t v_stride_s, v_stride_h;
  const int32_t* req_to_token;
  int64_t req_to_token_stride;
  const void* req_pool_indices;
  int32_t req_pool_indices_is_i64;
  const void* extend_prefix_lens;   /* [bs] P_r */
  const void* extend_seq_lens;      /* [bs] E_r */
  int32_t lens_is_i64;
  int32_t batch_size;
  int64_t num_q_tokens;             /* sum E_r (host-known: forward_batch.extend_num_tokens) */
  int32_t num_q_heads, num_kv_heads;
  int32_t head_dim_qk, head_dim_v;
  int32_t q_dtype, kv_dtype;
  float sm_scale, k_scale, v_scale, logit_cap;
  int32_t causal;
  const int32_t* plan;       /* written by b200attn_extend_plan */
  int32_t max_bs, max_tiles;
  int32_t num_ctas;
  int32_t kernel;
  float* lse;                /* optional [sum E_r, Hq] natural-log LSE output, may be NULL */
  int64_t pool_slots;        /* rows in k_cache / v_cache (bounds the TMA descriptors) */
  int32_t window_keys;       /* sliding window: query at position x attends kv positions (x - window_keys, x]
                                (0 = off; extend_attention.py:391-398 with SLIDING_WINDOW_SIZE = window_keys - 1) */
  const float* sinks;        /* optional [Hq] fp32 attention-sink logits, NULL = none */
  const uint8_t* custom_mask;   /* optional tree / speculative mask (extend_attention.py:378-390,525-539): request r,
                                   query i, kv position x is visible iff custom_mask[mask_indptr[r] + i * (P_r + E_r) + x]
                                   != 0.  Replaces the causal rule on the new tokens (x >= P_r); NULL = none */
  const int64_t* mask_indptr;   /* [bs + 1] offsets into custom_mask */
  int32_t mask_skip_prefix;     /* 1: prefix positions (x < P_r) are always visible (SKIP_PREFIX_CUSTOM_MASK) */
  int32_t block_q;              /* the block_q the plan was built with (checked against the kernel's tile size;
                                   0 = not stated) */
  int32_t page_size;            /* KV page size of the pool (req_to_token entries of one page are consecutive slots,
                                   pages are page_size-aligned); >= 16 lets the fill use one TMA box per page; 0/1 =
                                   token-granular */
  /* Split-KV for one-tile plans (block_q == 128 / group: TARGET_VERIFY / DRAFT_EXTEND / a few new tokens behind a long
   * prefix), where bs x Hkv work items cannot fill 148 SMs: every item's kv range is cut into up to num_splits
   * parts (the device trims that per item so that a part keeps >= 4 kv blocks), each part is its own CTA writing a
   * normalised fp32 partial + base-2 LSE to split_ws, and a merge launch combines them (sinks, v_scale and the LSE
   * output applied there).  The reference's Triton extend kernel has no such split (extend_attention.py:664-812 walks
   * the whole prefix per program); its decode kernel does (decode_attention.py:968-1044).  0 / 1 = off. */
  int32_t num_splits;
  float* split_ws;              /* >= b200attn_extend_split_ws_floats(...) floats; required when num_splits > 1 */
  int64_t split_ws_floats;      /* capacity of split_ws */
  int32_t num_tiles_hint;       /* exact number of plan tiles when the host knows it (speculative batches: one per
                                   request), 0 = unknown.  With it, a batch of MORE work items than SMs is split in
                                   the tail only: the tiles that fill whole waves run unsplit, the remaining ones are
                                   cut into min(num_splits, SMs / tail items) parts so that the last, partly filled
                                   wave ends earlier (two attention launches + the merge) */
} b200_extend_args;

int b200attn_extend(const b200_extend_args* args, b200_stream_t stream);
/* floats of split workspace b200attn_extend needs for num_splits > 1 */
int64_t b200attn_extend_split_ws_floats(int32_t num_splits, int64_t num_q_tokens, int32_t num_q_heads, int32_t head_dim_v);
/* split count that fills `num_sms` SMs best for `items` one-tile work items (bs x Hkv for speculative batches) of about
 * `kv_tokens_hint` kv tokens each (0 = unknown, e.g. under CUDA-graph capture: 4096 is assumed); 1 when the items already
 * make several waves or a split would not pay for its fixed costs */
int32_t b200attn_extend_pick_splits(int32_t items, int32_t num_sms, int32_t kv_tokens_hint);
#define B200_EXTEND_MAX_SPLITS 16
int b200attn_extend_fast_supported(int32_t head_dim_qk, int32_t head_dim_v, int32_t group,
                                   int32_t q_dtype, int32_t kv_dtype);
/* query tokens per tile the kernel that would be picked uses (for the plan) */
int b200attn_extend_block_q(int32_t head_dim_qk, int32_t head_dim_v, int32_t group,
                            int32_t q_dtype, int32_t kv_dtype, int32_t kernel);

/* ------------------------------------------------------------------------ *
 * Fused rotary embedding + KV-store (the step right before RadixAttention.forward).
 * Stands in for  q, k = rotary_emb(positions, q, k)  followed by  set_kv_buffer(layer, loc, k, v):
 *   srt/layers/rotary_embedding/base.py:234-275 (forward_native) / :363-385 (forward_cuda with
 *   fused_set_kv_buffer_arg), srt/layers/rotary_embedding/utils.py:36-60 (apply_rotary_emb),
 *   kernels/ops/kvcache/rope_cache.py:99-545 (fused_qk_rope_reshape_and_cache), srt/models/llama.py:219-223.
 * One launch: reads q, k, v once; writes the rotated q in place, the rotated k in place (write_k != 0)
 * and k_cache[loc] / v_cache[loc] with the cast / scale / skip rules of b200attn_store_kv.
 *   cos_sin_cache[pos] = cos(pos * inv_freq[0..rot/2)) | sin(...)   (rotary_dim values per position)
 *   neox:  pairs (i, i + rot/2);  gpt-j: pairs (2i, 2i + 1);  elements >= rotary_dim pass through.
 *   rounding = B200_ROPE_FP32:   o = round(x1*cos - x2*sin) computed in fp32, one rounding (what the
 *                                reference's CUDA kernel does with its fp32 cache)
 *   rounding = B200_ROPE_NATIVE: cos/sin rounded to the model dtype, every product and the sum rounded to
 *                                the model dtype -- bit-exact with forward_native / apply_rotary_emb
 * ------------------------------------------------------------------------ */
enum { B200_ROPE_FP32 = 0, B200_ROPE_NATIVE = 1 };
enum { B200_CACHE_F32 = 0, B200_CACHE_MODEL_DTYPE = 1 };

typedef struct b200_rope_store_args {
  void* q;                   /* [num_tokens, num_q_heads, head_dim] src_dtype, rotated in place */
  void* k;                   /* [num_tokens, num_kv_heads, head_dim] */
  const void* v;             /* [num_tokens, num_kv_heads, v_head_dim] */
  int64_t q_stride_t;        /* elements between tokens; heads are contiguous (stride head_dim) */
  int64_t k_stride_t;
  int64_t v_stride_t;
  const void* positions;     /* [num_tokens] int32 or int64 */
  int32_t positions_is_i64;
  const void* cos_sin_cache; /* [max_position, rotary_dim] */
  int32_t cache_dtype;       /* B200_CACHE_F32 or B200_CACHE_MODEL_DTYPE (same as src_dtype) */
  int64_t max_position;      /* positions outside [0, max_position) trap */
  int32_t rotary_dim;        /* even, <= head_dim */
  int32_t head_dim;
  int32_t v_head_dim;
  int32_t num_q_heads;
  int32_t num_kv_heads;
  int32_t is_neox;
  int32_t rounding;          /* B200_ROPE_* */
  int32_t write_k;           /* also write the rotated k back to `k` */
  /* store part: same meaning as in b200_store_kv_args; k_cache == NULL skips the store */
  void* k_cache;
  void* v_cache;
  int64_t k_cache_stride;
  int64_t v_cache_stride;
  const void* loc;
  int32_t loc_is_i64;
  int64_t num_tokens;
  int32_t src_dtype;         /* B200_BF16 or B200_F16 */
  int32_t dst_dtype;         /* pool dtype */
  float k_scale;
  float v_scale;
  int64_t size_limit;
  int64_t skip_index;
} b200_rope_store_args;

int b200attn_rope_store_kv(const b200_rope_store_args* args, b200_stream_t stream);

/* ------------------------------------------------------------------------ *
 * Merge two partial attention states (chunked prefix / cascade):
 *   lse = log(exp(lse_a) + exp(lse_b));  o = o_a * exp(lse_a - lse) + o_b * exp(lse_b - lse)
 * Stands in for merge_state_v2 (kernels/aot/csrc/attention/merge_attn_states.cu:60-70) and the Triton
 * merge_state_kernel (kernels/ops/attention/merge_state.py:9-66): an LSE of +inf is treated as -inf
 * (empty state).
 * ------------------------------------------------------------------------ */
typedef struct b200_merge_state_args {
  const void* o_a;           /* [T, H, D] dtype */
  const float* lse_a;        /* [T, H] natural log */
  const void* o_b;
  const float* lse_b;
  void* o;                   /* [T, H, D] dtype (may alias o_a or o_b) */
  float* lse;                /* optional [T, H] */
  int64_t num_tokens;
  int32_t num_heads;
  int32_t head_dim;          /* multiple of 8 */
  int32_t dtype;             /* B200_BF16 / B200_F16 */
  int64_t oa_stride_t, oa_stride_h, ob_stride_t, ob_stride_h, o_stride_t, o_stride_h;  /* elements */
} b200_merge_state_args;

int b200attn_merge_state(const b200_merge_state_args* args, b200_stream_t stream);

/* ------------------------------------------------------------------------ *
 * o_proj GEMM fused with its tensor-parallel all-reduce: on every rank r of the group
 *   out[T, N] = sum_r  x_r[T, K_r] . w_r[N, K_r]^T        (w_r = this rank's column shard of the weight)
 * Stands in for RowParallelLinear.forward -> tensor_model_parallel_all_reduce
 * (srt/layers/linear.py:1563, srt/distributed/communication_op.py:18-20, parallel_state.py:622) and the
 * custom all-reduce it would use (srt/distributed/device_communicators/custom_all_reduce_v2.py,
 * kernels/jit/csrc/distributed/custom_all_reduce.cuh).  One launch per rank, same arguments on every rank
 * except rank / x / w.  The caller owns three SYMMETRIC buffers (same size on every rank, each rank's copy
 * mapped into every other rank: CUDA IPC or a VMM/multicast allocation):
 *   staging  2 x [max_tokens, N] dtype   this rank's partial products (two buffers, alternating by launch)
 *   out      [max_tokens, N] dtype   the reduced result (valid on every rank when the kernel completes)
 *   flags    b200attn_oproj_allreduce_flag_words(max_tiles) uint32, zero-initialised once
 * The call sequence number that separates one launch's flags from the next lives in the flag area and is
 * advanced by the kernel itself, so the launch is CUDA-graph capturable and replayable.  `*_mc` are the multicast
 * (NVLS) addresses of staging / out, or NULL to reduce with peer loads.  world == 1 runs the GEMM alone and
 * writes out_peers[0].  All ranks must launch; a missing peer traps after ~2 s instead of hanging.
 * ------------------------------------------------------------------------ */
#define B200_MAX_WORLD 8
enum { B200_OAR_AUTO = 0, B200_OAR_ONE_SHOT = 1, B200_OAR_TWO_SHOT = 2 };
typedef struct b200_oproj_allreduce_args {
  const void* x;             /* [T, K] dtype, row stride x_stride elements */
  const void* w;             /* [N, K] dtype (torch Linear layout), row stride w_stride elements */
  int64_t x_stride, w_stride;
  int32_t num_tokens;        /* T */
  int32_t out_features;      /* N, multiple of 64 */
  int32_t in_features;       /* K (this rank's shard), multiple of 128 */
  int32_t dtype;             /* B200_BF16 / B200_F16 */
  int32_t rank, world;
  uint32_t reserved0;
  int32_t max_tiles;         /* capacity of the flag area, >= b200attn_oproj_allreduce_tiles(T, N) */
  void* staging_peers[B200_MAX_WORLD];
  void* out_peers[B200_MAX_WORLD];
  void* flags_peers[B200_MAX_WORLD];
  void* staging_mc;
  void* out_mc;
  int64_t staging_bytes;     /* bytes of ONE staging buffer; the staging area holds two (alternating by launch) */
  int32_t mode;              /* B200_OAR_AUTO / ONE_SHOT (every rank reduces every tile itself: shortest chain,
                                world x the output over the fabric) / TWO_SHOT (tile owner reduces + broadcasts) */
  void* debug_times;         /* optional uint64 [tiles][8]: %globaltimer at the epilogue's phase boundaries; written by -DB200_DEV builds only */
} b200_oproj_allreduce_args;

int b200attn_oproj_allreduce(const b200_oproj_allreduce_args* args, b200_stream_t stream);
int64_t b200attn_oproj_allreduce_flag_words(int32_t max_tiles);
int32_t b200attn_oproj_allreduce_tiles(int32_t num_tokens, int32_t out_features);

#ifdef __cplusplus
}
#endif
#endif /* B200ATTN_H_ */
