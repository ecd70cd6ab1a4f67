// Micro-benchmark: issue rate of tcgen05.mma.cta_group::2 (kind::f16, M = 256 over a CTA pair) for the shapes the
// CTA-pair extend kernel uses -- is an M256 x N128 pair MMA (each CTA supplies 64 rows of B) as fast per SM as the
// one-CTA M128 x N128?
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o benchmarks/umma_pair_rate benchmarks/umma_pair_rate.cu
//   ./benchmarks/umma_pair_rate
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>

#include "../sglang_b200/csrc/sm100_ptx.cuh"

using namespace b200::ptx;

// mode 0: 8 x SS (A K-major, B K-major), N = n          mode 1: 8 x TS (A tmem, B MN-major), N = n
// mode 2: attention mix per kv block of n: 8 x SS N=n K-major + (n/16) x TS N=128 MN-major
template <int mode, int n>
__global__ void __launch_bounds__(128, 1) pair_rate_kernel(int groups, long long* out) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base + 160 * 1024, slot = bar + 16;
  const uint32_t rank = cluster_ctarank();
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x)
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(base + 4 * i), "r"(0x3c003c00u));
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    mbar_fence_init();
  }
  if (threadIdx.x < 32) tmem_alloc_pair<512>(slot);
  fence_proxy_async_smem();
  tc_fence_before_sync();
  cluster_sync_all();
  tc_fence_after_sync();
  uint32_t tmem;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(slot));
  if (threadIdx.x == 0) {
    long long t0 = clock64();
    if (rank == 0) {
      const uint32_t a_smem = base, b_smem = base + 64 * 1024;
      const uint32_t idn = umma_idesc_f16(256, n, 1, 0, 0), id128mn = umma_idesc_f16(256, 128, 1, 0, 1);
      const uint32_t idnmn = umma_idesc_f16(256, n, 1, 0, 1);
#pragma unroll 1
      for (int g = 0; g < groups; ++g) {
        if constexpr (mode == 0 || mode == 2) {
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint64_t a = umma_desc_sw128(a_smem + (ks >> 2) * 16384 + (ks & 3) * 32, 16, 1024);
            const uint64_t bk = umma_desc_sw128(b_smem + (ks >> 2) * (n / 2 * 128) + (ks & 3) * 32, 16, 1024);
            umma_f16_ss_pair(tmem + 0, a, bk, idn, ks > 0);
          }
        }
        if constexpr (mode == 1) {
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {
            const uint64_t bm = umma_desc_sw128(b_smem + ks * 2048, 16384, 1024);
            umma_f16_ts_pair(tmem + 256, tmem + ks * 8, bm, idnmn, 1);
          }
        }
        if constexpr (mode == 2) {
#pragma unroll
          for (int ks = 0; ks < n / 16; ++ks) {
            const uint64_t bm = umma_desc_sw128(b_smem + 32768 + ks * 2048, 16384, 1024);
            umma_f16_ts_pair(tmem + 256, tmem + (ks & 7) * 8, bm, id128mn, 1);
          }
        }
      }
      umma_commit_pair(bar);
    }
    mbar_wait(bar, 0, 1);
    long long t1 = clock64();
    out[blockIdx.x] = t1 - t0;
  }
  tc_fence_before_sync();
  cluster_sync_all();
  if (threadIdx.x < 32) {
    tc_fence_after_sync();
    tmem_dealloc_pair<512>(tmem);
  }
}

template <int mode, int n>
void run(int ctas, int groups, long long* out, const char* name, double ideal_per_group) {
  const int smem = 160 * 1024 + 2048;
  auto kern = pair_rate_kernel<mode, n>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(ctas, 1, 1);
  cfg.blockDim = dim3(128, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, groups, out);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("mode %d n %d: %s\n", mode, n, cudaGetErrorString(e)); exit(1); }
  long long mx = 0;
  for (int i = 0; i < ctas; ++i) mx = out[i] > mx ? out[i] : mx;
  const double cyc = (double)mx / groups;
  printf("ctas %3d  %-40s N=%3d : %7.1f cyc/group (ideal %6.1f)  %5.1f%% of peak\n", ctas, name, n, cyc,
         ideal_per_group, 100.0 * ideal_per_group / cyc);
}

int main() {
  long long* out;
  cudaMallocManaged(&out, 148 * sizeof(long long));
  const int groups = 2000;
  for (int ctas : {2, 148}) {
    run<0, 64>(ctas, groups, out, "pair 8x SS M256 B=K-major", 8 * 32.0);
    run<0, 128>(ctas, groups, out, "pair 8x SS M256 B=K-major", 8 * 64.0);
    run<0, 256>(ctas, groups, out, "pair 8x SS M256 B=K-major", 8 * 128.0);
    run<1, 128>(ctas, groups, out, "pair 8x TS M256 B=MN-major", 8 * 64.0);
    run<2, 128>(ctas, groups, out, "pair mix 8xSS N128 + 8xTS N128mn", 1024.0);
    run<2, 256>(ctas, groups, out, "pair mix 8xSS N256 + 16xTS N128mn", 2048.0);
  }
  return 0;
}
