// Micro-benchmark: per-scheduler issue cost (cycles per warp instruction) of the instructions the decode
// kernel leans on: legacy HMMA m16n8k16, e4m3->f16 unpack, LOP3/SHF, ldmatrix, MUFU.EX2.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o benchmarks/pipe_rate benchmarks/pipe_rate.cu
// One CTA on one SM; W warps per scheduler (W = 1, 2, 4); independent chains so latency is hidden.
#include <cstdint>
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

constexpr int kIters = 4096;

template <int OP>
__global__ void rate_kernel(long long* out, uint32_t* sink, uint32_t seed) {
  __shared__ __align__(128) uint8_t sm[8192];
  for (int i = threadIdx.x; i < 2048; i += blockDim.x) reinterpret_cast<uint32_t*>(sm)[i] = i * 2654435761u;
  __syncthreads();
  uint32_t r[16];
  float f[8][4];
#pragma unroll
  for (int i = 0; i < 16; ++i) r[i] = seed * (i + 1) + threadIdx.x;
#pragma unroll
  for (int i = 0; i < 8; ++i) f[i][0] = f[i][1] = f[i][2] = f[i][3] = 0.f;
  const uint32_t saddr = (uint32_t)__cvta_generic_to_shared(sm) + (threadIdx.x & 31) * 16;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < kIters; ++it) {
    if constexpr (OP == 0) {  // 8 independent HMMA
#pragma unroll
      for (int i = 0; i < 8; ++i)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(f[i][0]), "+f"(f[i][1]), "+f"(f[i][2]), "+f"(f[i][3])
                     : "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4 + (i & 3)]), "r"(r[8 + (i & 3)]));
    } else if constexpr (OP == 1) {  // 8 independent e4m3x2 -> f16x2
#pragma unroll
      for (int i = 0; i < 8; ++i)
        asm volatile("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(r[8 + i]) : "h"((uint16_t)r[i]));
#pragma unroll
      for (int i = 0; i < 8; ++i) r[i] ^= r[8 + i];
    } else if constexpr (OP == 2) {  // 8 independent LOP3
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(r[i]) : "r"(r[8 + i]), "r"(seed));
    } else if constexpr (OP == 3) {  // 8 independent shifts
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("shf.r.wrap.b32 %0, %0, %1, 3;" : "+r"(r[i]) : "r"(r[8 + i]));
    } else if constexpr (OP == 4) {  // 4 ldmatrix.x4
#pragma unroll
      for (int i = 0; i < 4; ++i)
        asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                     : "=r"(r[4 * i]), "=r"(r[4 * i + 1]), "=r"(r[4 * i + 2]), "=r"(r[4 * i + 3])
                     : "r"(saddr + 512 * i));
    } else if constexpr (OP == 5) {  // 8 MUFU.EX2
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(f[i][0]));
    } else if constexpr (OP == 6) {  // 8 PRMT
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("prmt.b32 %0, %0, %1, 0x3120;" : "+r"(r[i]) : "r"(r[8 + i]));
    } else if constexpr (OP == 7) {  // 8 FP8 HMMA m16n8k32 e4m3
#pragma unroll
      for (int i = 0; i < 8; ++i)
        asm volatile("mma.sync.aligned.m16n8k32.row.col.f32.e4m3.e4m3.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(f[i][0]), "+f"(f[i][1]), "+f"(f[i][2]), "+f"(f[i][3])
                     : "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4 + (i & 3)]), "r"(r[8 + (i & 3)]));
    } else if constexpr (OP == 9) {  // 8 packed-half exp2 (two results per lane)
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(r[i]));
    } else if constexpr (OP == 10) {  // 8 packed-bf16 exp2
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(r[i]));
    } else if constexpr (OP == 11) {  // 8 FFMA2 (fma.rn.f32x2)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        uint64_t a, b;
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(f[i][0]), "f"(f[i][1]));
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(b) : "f"(f[i][2]), "f"(f[i][3]));
        asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(a) : "l"(b));
        asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(f[i][0]), "=f"(f[i][1]) : "l"(a));
      }
    } else if constexpr (OP == 12) {  // 8 FADD2 (add.rn.f32x2)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        uint64_t a, b;
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(f[i][0]), "f"(f[i][1]));
        asm volatile("mov.b64 %0, {%1, %2};" : "=l"(b) : "f"(f[i][2]), "f"(f[i][3]));
        asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(a) : "l"(b));
        asm volatile("mov.b64 {%0, %1}, %2;" : "=f"(f[i][0]), "=f"(f[i][1]) : "l"(a));
      }
    } else if constexpr (OP == 13) {  // 8 F2FP f32,f32 -> bf16x2
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r[8 + i]) : "f"(f[i][0]), "f"(f[i][1]));
        f[i][0] += 1.f;   // FADD keeps the chain alive (counted below as one of two instructions)
      }
    } else if constexpr (OP == 14) {  // 8 FMNMX3
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(f[i][0]) : "f"(f[i][1]), "f"(f[i][2]));
    } else if constexpr (OP == 15) {  // 8 FFMA
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i][0]) : "f"(f[i][1]), "f"(f[i][2]));
    } else if constexpr (OP == 16) {  // 8 F2FP f32,f32 -> f16x2 feeding 8 packed-half exp2
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        asm volatile("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r[i]) : "f"(f[i][0]), "f"(f[i][1]));
        asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(r[i]));
        f[i][0] = __uint_as_float(r[i]);
      }
    } else if constexpr (OP == 8) {  // 8 IMAD (shift-left stand-in)
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("mad.lo.u32 %0, %0, 256, %1;" : "+r"(r[i]) : "r"(r[8 + i]));
    }
  }
  const long long t1 = clock64();
  uint32_t acc = 0;
  float facc = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) acc ^= r[i];
#pragma unroll
  for (int i = 0; i < 8; ++i) facc += f[i][0] + f[i][1] + f[i][2] + f[i][3];
  sink[threadIdx.x] = acc + __float_as_uint(facc);
  if (threadIdx.x == 0) out[0] = t1 - t0;
}

uint32_t* g_sink;
template <int OP>
void run(const char* name, int per_iter, long long* out) {
  for (int warps_per_sched : {1, 2, 4}) {
    rate_kernel<OP><<<1, 128 * warps_per_sched>>>(out, g_sink, 12345u);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
    const double cyc = (double)out[0] / kIters;
    printf("%-28s warps/sched %d : %7.2f cyc per warp-instr per scheduler (raw %lld)\n", name, warps_per_sched,
           cyc / (per_iter * warps_per_sched), out[0]);
  }
}

int main() {
  long long* out;
  cudaMallocManaged(&out, 16);
  cudaMalloc(&g_sink, 4096);
  run<0>("HMMA.16816 f16 (f32 acc)", 8, out);
  run<7>("HMMA.16832 e4m3 (f32 acc)", 8, out);
  run<1>("F2FP e4m3x2->f16x2 (+xor)", 8, out);
  run<2>("LOP3", 8, out);
  run<3>("SHF", 8, out);
  run<8>("IMAD", 8, out);
  run<6>("PRMT", 8, out);
  run<4>("LDSM.x4", 4, out);
  run<5>("MUFU.EX2", 8, out);
  run<9>("MUFU.EX2 f16x2", 8, out);
  run<10>("MUFU.EX2 bf16x2", 8, out);
  run<11>("FFMA2 (fma.rn.f32x2)", 8, out);
  run<12>("FADD2 (add.rn.f32x2)", 8, out);
  run<15>("FFMA", 8, out);
  run<14>("FMNMX3", 8, out);
  run<13>("F2FP bf16x2 + FADD", 16, out);
  run<16>("F2FP f16x2 + EX2 f16x2", 16, out);
  return 0;
}
