// Merge of two partial attention states (chunked prefix / cascade attention) for sm_100a.
//
// Replaces merge_state_v2 (reference python/sglang/kernels/aot/csrc/attention/merge_attn_states.cu:60-120)
// and the Triton merge_state_kernel (kernels/ops/attention/merge_state.py:9-66):
//   lse = log(exp(lse_a) + exp(lse_b));  o = o_a * exp(lse_a - lse) + o_b * exp(lse_b - lse)
// with an LSE of +inf read as -inf (an empty state), exactly like both reference kernels.
//
// HBM-bound elementwise work: a thread owns 8 consecutive head-dim elements (one 16-byte load from each
// input, one 16-byte store); the two LSE values of its (token, head) are broadcast loads.
// Algorithmic bytes = T * H * (3 * D * 2 + 12).
#include "common.cuh"

namespace b200 {
namespace {

struct MergeParams {
  const void* o_a;
  const void* o_b;
  void* o;
  const float* lse_a;
  const float* lse_b;
  float* lse;
  int64_t n_units;       // T * H * D / 8
  int32_t num_heads, upr;  // upr = D / 8 units per (token, head) row
  int64_t oa_st, oa_sh, ob_st, ob_sh, o_st, o_sh;
};

template <typename T>
__global__ void __launch_bounds__(256) merge_state_kernel(const MergeParams p) {
  const int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= p.n_units) return;
  const int64_t row = u / p.upr;
  const int c = (int)(u - row * p.upr);
  const int64_t t = row / p.num_heads;
  const int h = (int)(row - t * p.num_heads);
  const uint4 ra = *reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(p.o_a) + t * p.oa_st + h * p.oa_sh + c * 8);
  const uint4 rb = *reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(p.o_b) + t * p.ob_st + h * p.ob_sh + c * 8);
  float la = p.lse_a[row], lb = p.lse_b[row];
  if (la == INFINITY) la = -INFINITY;
  if (lb == INFINITY) lb = -INFINITY;
  const float mx = fmaxf(la, lb);
  // both empty: exp(-inf - -inf) is NaN in the reference too; keep its arithmetic
  const float ea = expf(la - mx), eb = expf(lb - mx);
  const float se = ea + eb;
  const float wa = ea / se, wb = eb / se;
  const uint32_t a[4] = {ra.x, ra.y, ra.z, ra.w}, b[4] = {rb.x, rb.y, rb.z, rb.w};
  uint32_t r[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float a0, a1, b0, b1;
    if constexpr (dtype_of<T>::value == B200_BF16) {
      a0 = __uint_as_float(a[i] << 16); a1 = __uint_as_float(a[i] & 0xffff0000u);
      b0 = __uint_as_float(b[i] << 16); b1 = __uint_as_float(b[i] & 0xffff0000u);
    } else {
      const float2 fa = __half22float2(*reinterpret_cast<const __half2*>(&a[i]));
      const float2 fb = __half22float2(*reinterpret_cast<const __half2*>(&b[i]));
      a0 = fa.x; a1 = fa.y; b0 = fb.x; b1 = fb.y;
    }
    r[i] = pack2<T>(a0 * wa + b0 * wb, a1 * wa + b1 * wb);
  }
  *reinterpret_cast<uint4*>(reinterpret_cast<T*>(p.o) + t * p.o_st + h * p.o_sh + c * 8) = make_uint4(r[0], r[1], r[2], r[3]);
  if (p.lse != nullptr && c == 0) p.lse[row] = logf(se) + mx;
}

}  // namespace
}  // namespace b200

extern "C" int b200attn_merge_state(const b200_merge_state_args* a, b200_stream_t stream_) {
  using namespace b200;
  B200_CHECK_ARG(a != nullptr, "merge_state: null args");
  if (a->num_tokens == 0) return B200_OK;
  B200_CHECK_ARG(a->o_a && a->o_b && a->o && a->lse_a && a->lse_b, "merge_state: null pointer");
  B200_CHECK_ARG(a->num_tokens > 0 && a->num_heads > 0 && a->head_dim > 0 && a->head_dim % 8 == 0,
                 "merge_state: bad sizes T=%lld H=%d D=%d (D must be a multiple of 8)", (long long)a->num_tokens,
                 a->num_heads, a->head_dim);
  B200_CHECK_ARG(a->dtype == B200_BF16 || a->dtype == B200_F16, "merge_state: dtype %d", a->dtype);
  auto ok = [](const void* ptr, int64_t st, int64_t sh) {
    return ((uintptr_t)ptr & 15) == 0 && st % 8 == 0 && sh % 8 == 0;
  };
  B200_CHECK_ARG(ok(a->o_a, a->oa_stride_t, a->oa_stride_h) && ok(a->o_b, a->ob_stride_t, a->ob_stride_h) &&
                     ok(a->o, a->o_stride_t, a->o_stride_h),
                 "merge_state: pointers must be 16-byte aligned and strides multiples of 8 elements");
  MergeParams p;
  p.o_a = a->o_a; p.o_b = a->o_b; p.o = a->o; p.lse_a = a->lse_a; p.lse_b = a->lse_b; p.lse = a->lse;
  p.num_heads = a->num_heads; p.upr = a->head_dim / 8;
  p.n_units = a->num_tokens * a->num_heads * p.upr;
  p.oa_st = a->oa_stride_t; p.oa_sh = a->oa_stride_h; p.ob_st = a->ob_stride_t; p.ob_sh = a->ob_stride_h;
  p.o_st = a->o_stride_t; p.o_sh = a->o_stride_h;
  const int64_t blocks = (p.n_units + 255) / 256;
  B200_CHECK_ARG(blocks < (int64_t)0x7fffffff, "merge_state: too many rows");
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if (a->dtype == B200_BF16) merge_state_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, 0, stream>>>(p);
  else merge_state_kernel<__half><<<(unsigned)blocks, 256, 0, stream>>>(p);
  B200_CHECK_CUDA(cudaGetLastError());
  return B200_OK;
}
