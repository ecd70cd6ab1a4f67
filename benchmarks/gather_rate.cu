// Micro-benchmark: HBM bandwidth of a random gather as a function of the contiguous segment size.
// This is the memory-side roofline of paged decode at page_size 1: one (token, kv head) row of K or V is
// a contiguous segment of head_dim * elem_size bytes (128 B for FP8 D=128, 256 B for bf16 D=128) at a
// random pool slot.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o benchmarks/gather_rate benchmarks/gather_rate.cu
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>
#include <cuda_runtime.h>

// Each group of (seg/16) lanes reads one segment with 16-byte loads; kUnroll segments in flight per lane group.
template <int kSeg, int kUnroll>
__global__ void __launch_bounds__(256) gather_kernel(const uint4* __restrict__ buf, const uint32_t* __restrict__ idx,
                                                     long long n_seg, long long stride16, uint4* out) {
  constexpr int kLanes = kSeg / 16;               // lanes per segment
  constexpr int kPerWarp = 32 / kLanes;           // segments per warp load
  const int lane = threadIdx.x & 31;
  const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const long long n_warps = ((long long)gridDim.x * blockDim.x) >> 5;
  uint4 acc = make_uint4(0, 0, 0, 0);
  for (long long s0 = warp * kPerWarp * kUnroll; s0 < n_seg; s0 += n_warps * kPerWarp * kUnroll) {
    uint4 v[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long s = s0 + u * kPerWarp + lane / kLanes;
      const uint32_t slot = s < n_seg ? idx[s] : 0;
      v[u] = __ldcs(buf + slot * stride16 + (lane % kLanes));
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) { acc.x ^= v[u].x; acc.y ^= v[u].y; acc.z ^= v[u].z; acc.w ^= v[u].w; }
  }
  if (acc.x == 0x12345u) out[threadIdx.x] = acc;
}

template <int kSeg>
void run(const uint4* buf, uint32_t* d_idx, long long pool_bytes, long long row_bytes, bool random, uint4* out) {
  // rows of row_bytes; the segment is the first kSeg bytes of a row (other heads of the token are skipped)
  const long long n_rows = pool_bytes / row_bytes;
  std::vector<uint32_t> idx(n_rows);
  for (long long i = 0; i < n_rows; ++i) idx[i] = (uint32_t)i;
  if (random) { std::mt19937 g(1234); std::shuffle(idx.begin(), idx.end(), g); }
  cudaMemcpy(d_idx, idx.data(), n_rows * 4, cudaMemcpyHostToDevice);
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  float best = 1e30f;
  for (int it = 0; it < 4; ++it) {
    cudaEventRecord(a);
    gather_kernel<kSeg, 8><<<148 * 8, 256>>>(buf, d_idx, n_rows, row_bytes / 16, out);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    best = std::min(best, ms);
  }
  const double gbs = (double)n_rows * (kSeg + 4) / best / 1e6;
  printf("segment %4d B  row stride %5lld B  %s : %8.1f GB/s\n", kSeg, row_bytes, random ? "random slots " : "sequential   ", gbs);
}

int main() {
  const long long pool = 4ll << 30;
  uint4 *buf, *out;
  uint32_t* d_idx;
  cudaMalloc(&buf, pool);
  cudaMemset(buf, 1, pool);
  cudaMalloc(&out, 4096 * 16);
  cudaMalloc(&d_idx, (pool / 128) * 4);
  for (bool random : {false, true}) {
    run<128>(buf, d_idx, pool, 128, random, out);     // dense 128 B segments (Hkv = 1, FP8 D = 128)
    run<128>(buf, d_idx, pool, 1024, random, out);    // FP8 D=128, 8 kv heads: one head of a 1 KiB token row
    run<256>(buf, d_idx, pool, 256, random, out);
    run<256>(buf, d_idx, pool, 2048, random, out);    // bf16 D=128, 8 kv heads
    run<512>(buf, d_idx, pool, 512, random, out);
    run<512>(buf, d_idx, pool, 2048, random, out);
  }
  // whole token rows (all 8 heads) gathered by one warp: what a "all heads per CTA" mapping would see
  run<512>(buf, d_idx, pool, 1024, true, out);        // half of a 1 KiB row
  return 0;
}
