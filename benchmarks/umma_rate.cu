// Micro-benchmark: issue rate of tcgen05.mma (kind::f16, M=128) for the operand sources / shapes the
// extend kernel can choose between.  Answers "is an SS-mode N=64 MMA shared-memory bound?".
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o benchmarks/umma_rate benchmarks/umma_rate.cu
//   ./benchmarks/umma_rate
//
// Prints cycles per MMA instruction (K=16) and the implied fraction of the 8192 flop/clk/SM peak.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#include "../sglang_b200/csrc/sm100_ptx.cuh"

using namespace b200::ptx;

__device__ __forceinline__ void umma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// mode: 0 = SS (A K-major, B K-major)   1 = SS (A K-major, B MN-major)
//       2 = TS (A tmem,   B K-major)    3 = TS (A tmem,   B MN-major)
//       4 = attention mix per kv block of 64: 8 x [SS N64 K-major] + 4 x [TS N128 MN-major]
//       5 = attention mix, Q in TMEM:     8 x [TS N64 K-major] + 4 x [TS N128 MN-major]
//       6 = FA4-style mix per kv block 128: 8 x [SS N128 K-major] + 8 x [TS N128 MN-major]
template <int mode, int n>
__global__ void __launch_bounds__(128, 1) umma_rate_kernel(int groups, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base + 160 * 1024, slot = bar + 16;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x)
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(base + 4 * i), "r"(0x3c003c00u));
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    mbar_fence_init();
  }
  if (threadIdx.x < 32) tmem_alloc<512>(slot);
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  uint32_t tmem;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(slot));
  if (threadIdx.x == 0) {
    const uint32_t a_smem = base, b_smem = base + 64 * 1024;
    const uint32_t id64 = umma_idesc_f16(128, 64, 1, 0, 0), id128 = umma_idesc_f16(128, 128, 1, 0, 0);
    const uint32_t id128mn = umma_idesc_f16(128, 128, 1, 0, 1);
    const uint32_t idn = umma_idesc_f16(128, n, 1, 0, 0), idnmn = umma_idesc_f16(128, n, 1, 0, 1);
    long long t0 = clock64();
#pragma unroll 1
    for (int g = 0; g < groups; ++g) {
      if constexpr (mode <= 3) {
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint64_t a = umma_desc_sw128(a_smem + (ks >> 2) * 16384 + (ks & 3) * 32, 16, 1024);
          const uint64_t bk = umma_desc_sw128(b_smem + (ks >> 2) * (n * 128) + (ks & 3) * 32, 16, 1024);
          const uint64_t bm = umma_desc_sw128(b_smem + ks * 2048, n * 64, 1024);
          if constexpr (mode == 0) umma_f16_ss(tmem + 256, a, bk, idn, ks > 0);
          if constexpr (mode == 1) umma_f16_ss(tmem + 256, a, bm, idnmn, ks > 0);
          if constexpr (mode == 2) umma_f16_ts(tmem + 256, tmem + ks * 8, bk, idn, ks > 0);
          if constexpr (mode == 3) umma_f16_ts(tmem + 256, tmem + ks * 8, bm, idnmn, ks > 0);
        }
      } else if constexpr (mode == 4 || mode == 5) {
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint64_t a = umma_desc_sw128(a_smem + (ks >> 2) * 16384 + (ks & 3) * 32, 16, 1024);
          const uint64_t bk = umma_desc_sw128(b_smem + (ks >> 2) * 8192 + (ks & 3) * 32, 16, 1024);
          if constexpr (mode == 4) umma_f16_ss(tmem + 128, a, bk, id64, ks > 0);
          else umma_f16_ts(tmem + 128, tmem + ks * 8, bk, id64, ks > 0);
        }
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t bm = umma_desc_sw128(b_smem + 16384 + ks * 2048, 64 * 128, 1024);
          umma_f16_ts(tmem + 256, tmem + 64 + ks * 8, bm, id128mn, 1);
        }
      } else {
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint64_t a = umma_desc_sw128(a_smem + (ks >> 2) * 16384 + (ks & 3) * 32, 16, 1024);
          const uint64_t bk = umma_desc_sw128(b_smem + (ks >> 2) * 16384 + (ks & 3) * 32, 16, 1024);
          umma_f16_ss(tmem + 0, a, bk, id128, ks > 0);
        }
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) {
          const uint64_t bm = umma_desc_sw128(b_smem + 32768 + ks * 2048, 128 * 128, 1024);
          umma_f16_ts(tmem + 256, tmem + 128 + ks * 8, bm, id128mn, 1);
        }
      }
    }
    umma_commit(bar);
    mbar_wait(bar, 0, 1);
    long long t1 = clock64();
    out[blockIdx.x] = t1 - t0;
  }
  tc_fence_before_sync();
  __syncthreads();
  if (threadIdx.x < 32) {
    tc_fence_after_sync();
    tmem_dealloc<512>(tmem);
  }
}


template <int mode, int n>
void run(int ctas, int groups, long long* out, const char* name, double ideal_per_group) {
  const int smem = 160 * 1024 + 2048;
  cudaFuncSetAttribute(umma_rate_kernel<mode, n>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  umma_rate_kernel<mode, n><<<ctas, 128, smem>>>(groups, out);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("mode %d n %d: %s\n", mode, n, cudaGetErrorString(e)); exit(1); }
  long long mx = 0;
  for (int i = 0; i < ctas; ++i) mx = out[i] > mx ? out[i] : mx;
  const double cyc = (double)mx / groups;
  printf("ctas %3d  %-34s N=%3d : %7.1f cyc/group (ideal %6.1f)  %5.1f%% of peak\n", ctas, name, n, cyc,
         ideal_per_group, 100.0 * ideal_per_group / cyc);
}

int main() {
  long long* out;
  cudaMallocManaged(&out, 148 * sizeof(long long));
  const int groups = 2000;
  for (int ctas : {1, 148}) {
    run<0, 64>(ctas, groups, out, "8x SS B=K-major", 8 * 32.0);
    run<0, 128>(ctas, groups, out, "8x SS B=K-major", 8 * 64.0);
    run<0, 256>(ctas, groups, out, "8x SS B=K-major", 8 * 128.0);
    run<1, 64>(ctas, groups, out, "8x SS B=MN-major", 8 * 32.0);
    run<1, 128>(ctas, groups, out, "8x SS B=MN-major", 8 * 64.0);
    run<2, 64>(ctas, groups, out, "8x TS B=K-major", 8 * 32.0);
    run<2, 128>(ctas, groups, out, "8x TS B=K-major", 8 * 64.0);
    run<2, 256>(ctas, groups, out, "8x TS B=K-major", 8 * 128.0);
    run<3, 64>(ctas, groups, out, "8x TS B=MN-major", 8 * 32.0);
    run<3, 128>(ctas, groups, out, "8x TS B=MN-major", 8 * 64.0);
    run<4, 64>(ctas, groups, out, "mix 8xSS N64 + 4xTS N128mn", 512.0);
    run<5, 64>(ctas, groups, out, "mix 8xTS N64 + 4xTS N128mn", 512.0);
    run<6, 128>(ctas, groups, out, "mix 8xSS N128 + 8xTS N128mn (FA4)", 1024.0);
  }
  return 0;
}
