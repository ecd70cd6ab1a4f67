// CUDA-core attention of one query row over a paged KV range: the any-shape path
// (head_dim <= 512, any q/kv dtype pair) shared by the generic decode and extend kernels.
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int kGenMaxD = 512;
constexpr int kGenAcc = kGenMaxD / kWarp;

struct RowState {
  float m = -INFINITY;  // running max (log2 domain)
  float l = 0.f;        // running sum
  float acc[kGenAcc];   // lane owns output dims lane, lane+32, ...
};

__device__ __forceinline__ float capped_logit_log2(float raw_dot, float qk_scale, float logit_cap) {
  float s = raw_dot * qk_scale;
  if (logit_cap > 0.f) s = logit_cap * tanhf(s / logit_cap);
  return s * kLog2e;
}

// One warp.  q_s: Dk floats in shared memory (already converted).  pt: page-table
// entries of the n_tok tokens to attend.  QK: one token per lane; PV: lanes split Dv
// and the probabilities are broadcast by shuffle.  A block of 32 tokens that is entirely masked leaves the state
// untouched (m stays -inf until the first visible token).
template <typename KVT>
__device__ __forceinline__ void attend_row_generic(RowState& st, const float* q_s, const int32_t* pt, int n_tok,
                                                   const KVT* kc, const KVT* vc, int64_t k_stride_s,
                                                   int64_t v_stride_s, int64_t k_head_off, int64_t v_head_off,
                                                   int Dk, int Dv, float qk_scale, float logit_cap,
                                                   const uint8_t* mask = nullptr, int mask_from = 0) {
  const int lane = threadIdx.x & 31;
  for (int t0 = 0; t0 < n_tok; t0 += kWarp) {
    const int tok = t0 + lane;
    // custom (tree) mask: token `tok` of the range is visible iff mask[tok] != 0; tokens before mask_from always are
    const bool valid = tok < n_tok && (mask == nullptr || tok < mask_from || mask[tok] != 0);
    const int64_t slot = valid ? (int64_t)pt[tok] : 0;
    float s = -INFINITY;
    if (valid) {
      const KVT* krow = kc + slot * k_stride_s + k_head_off;
      float dot = 0.f;
      for (int d = 0; d < Dk; ++d) dot = fmaf(q_s[d], to_float(krow[d]), dot);
      s = capped_logit_log2(dot, qk_scale, logit_cap);
    }
    const float m_new = fmaxf(st.m, warp_max(s));
    if (m_new == -INFINITY) continue;                  // nothing visible so far
    const float pr = valid ? exp2f(s - m_new) : 0.f;
    const float corr = exp2f(st.m - m_new);
    st.l = st.l * corr + warp_sum(pr);
#pragma unroll
    for (int i = 0; i < kGenAcc; ++i) st.acc[i] *= corr;
    const int cnt = min(kWarp, n_tok - t0);
    for (int j = 0; j < cnt; ++j) {
      const float pj = __shfl_sync(0xffffffffu, pr, j);
      const int64_t sj = __shfl_sync(0xffffffffu, slot, j);
      const KVT* vrow = vc + sj * v_stride_s + v_head_off;
#pragma unroll
      for (int i = 0; i < kGenAcc; ++i) {
        const int d = lane + i * kWarp;
        if (d < Dv) st.acc[i] = fmaf(pj, to_float(vrow[d]), st.acc[i]);
      }
    }
    st.m = m_new;
  }
}

}  // namespace b200
