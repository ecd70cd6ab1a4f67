// Micro-benchmark: cycles one SM sub-partition needs for the S -> P pass of the extend kernel (row max, scale,
// exp2 -- MUFU or polynomial --, bf16 pack, row sum) on register data, without TMEM / barriers / MMAs:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I sglang_b200/csrc -o benchmarks/softmax_rate benchmarks/softmax_rate.cu
// One CTA on one SM, W warps per scheduler; each thread owns COLS columns of one row per pass.  Reports cycles per
// pass and per element, for several polynomial fractions and code shapes.  This is the floor the real kernel's
// softmax warps can reach; the tensor pipe needs 1 024 cycles per (tile, kv block).
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#include "common.cuh"
#include "sm100_ptx.cuh"
namespace b200 { void set_error(const char*, ...) {} }
using namespace b200;
using namespace b200::ptx;

constexpr int kIters = 2048;

// exp2 of a pair on the FMA/ALU pipes, exponent patched with shift + integer add (LEA, ALU pipe) instead of IMAD
__device__ __forceinline__ float2 exp2_poly2_lea(float2 x) {
  x.x = fmaxf(x.x, -126.f);
  x.y = fmaxf(x.y, -126.f);
  const float2 t = fadd2(x, make_float2(12582912.f, 12582912.f));
  const float2 nf = fadd2(t, make_float2(-12582912.f, -12582912.f));
  const float2 f = ffma2(nf, make_float2(-1.f, -1.f), x);
  float2 p = ffma2(make_float2(0.055171407759189606f, 0.055171407759189606f), f,
                   make_float2(0.24261075258255005f, 0.24261075258255005f));
  p = ffma2(p, f, make_float2(0.6932609677314758f, 0.6932609677314758f));
  p = ffma2(p, f, make_float2(0.9999281167984009f, 0.9999281167984009f));
  float2 r;
  int a, b;
  asm("shl.b32 %0, %1, 23;" : "=r"(a) : "r"(__float_as_int(t.x)));
  asm("shl.b32 %0, %1, 23;" : "=r"(b) : "r"(__float_as_int(t.y)));
  asm("add.s32 %0, %1, %2;" : "=r"(a) : "r"(a), "r"(__float_as_int(p.x)));
  asm("add.s32 %0, %1, %2;" : "=r"(b) : "r"(b), "r"(__float_as_int(p.y)));
  r.x = __int_as_float(a);
  r.y = __int_as_float(b);
  return r;
}

// MODE 0: the kernel's pass (max, then exp/pack with the sum in the same loop)
// MODE 1: sum after the pack loop
// MODE 2: like 1, polynomial with LEA
// MODE 3: no max pass (exp/pack/sum only)
// MODE 4: scalar FFMA/FADD instead of packed
template <int COLS, int EMU, int MODE>
__global__ void __launch_bounds__(COLS == 128 ? 256 : 512, 1) pass_kernel(long long* out, uint32_t* sink, float scale) {
  uint32_t sv[COLS];
#pragma unroll
  for (int i = 0; i < COLS; ++i) sv[i] = __float_as_uint((float)((threadIdx.x * 31 + i * 17) % 97) * 0.01f);
  float m_run = -INFINITY, l_run = 0.f;
  uint32_t keep = 0;
  __syncthreads();
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < kIters; ++it) {
    float m_new = m_run;
    if (MODE != 3) {
      float mx[COLS / 32];
#pragma unroll
      for (int c = 0; c < COLS / 32; ++c) {
        mx[c] = __uint_as_float(sv[32 * c]);
#pragma unroll
        for (int i = 1; i < 31; i += 2)
          mx[c] = fmaxf(mx[c], fmaxf(__uint_as_float(sv[32 * c + i]), __uint_as_float(sv[32 * c + i + 1])));
        mx[c] = fmaxf(mx[c], __uint_as_float(sv[32 * c + 31]));
      }
      float mp = mx[0];
#pragma unroll
      for (int c = 1; c < COLS / 32; ++c) mp = fmaxf(mp, mx[c]);
      m_new = fmaxf(m_run, mp);
      if ((m_new - m_run) * scale <= 8.f) m_new = m_run;
    }
    const float corr = fast_exp2((m_run - m_new) * scale);
    const float neg = -m_new * scale;
    float2 sum2 = make_float2(0.f, 0.f);
    constexpr int CH = COLS < 64 ? COLS : 64;
#pragma unroll
    for (int c = 0; c < COLS / CH; ++c) {
      uint32_t pk[CH / 2];
#pragma unroll
      for (int i = 0; i < CH / 2; ++i) {
        const int c0 = CH * c + 2 * i, c1 = c0 + 1;
        float2 e;
        if (MODE == 4) {
          e.x = fmaf(__uint_as_float(sv[c0]), scale, neg);
          e.y = fmaf(__uint_as_float(sv[c1]), scale, neg);
        } else {
          e = ffma2(make_float2(__uint_as_float(sv[c0]), __uint_as_float(sv[c1])), make_float2(scale, scale),
                    make_float2(neg, neg));
        }
        if ((i & 7) < EMU) {
          e = MODE == 2 ? exp2_poly2_lea(e) : exp2_poly2(e);
        } else {
          e.x = fast_exp2(e.x);
          e.y = fast_exp2(e.y);
        }
        if (MODE == 0 || MODE == 3) sum2 = fadd2(sum2, e);
        if (MODE == 4) { sum2.x += e.x; sum2.y += e.y; }
        sv[c0] = __float_as_uint(e.x);
        sv[c1] = __float_as_uint(e.y);
        pk[i] = pack2<__nv_bfloat16>(e.x, e.y);
      }
#pragma unroll
      for (int i = 0; i < CH / 2; ++i) asm volatile("" ::"r"(pk[i]));     // P leaves through tcgen05.st in the kernel
      keep ^= pk[0] ^ pk[CH / 2 - 1];
    }
    if (MODE == 1 || MODE == 2) {
#pragma unroll
      for (int i = 0; i < COLS; i += 2)
        sum2 = fadd2(sum2, make_float2(__uint_as_float(sv[i]), __uint_as_float(sv[i + 1])));
    }
    l_run = l_run * corr + (sum2.x + sum2.y);
    m_run = m_new;
    // next pass: the values move a little so nothing is loop invariant
#pragma unroll
    for (int i = 0; i < COLS; i += 16) sv[i] ^= (uint32_t)it & 0xffu;
  }
  const long long t1 = clock64();
  sink[threadIdx.x] = keep + __float_as_uint(l_run + m_run);
  if (threadIdx.x == 0) out[0] = t1 - t0;
}

uint32_t* g_sink;
template <int COLS, int EMU, int MODE>
void run(const char* name, long long* out) {
  for (int w : {1, 2, 4}) {
    if (COLS == 128 && w == 4) continue;   // 128 live values x 512 threads do not fit the register file
    pass_kernel<COLS, EMU, MODE><<<1, 128 * w>>>(out, g_sink, 0.1275f);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
    const double cyc = (double)out[0] / kIters;
    printf("%-44s cols %3d emu %d/8 warps/sched %d : %7.1f cyc/pass  %5.2f cyc per element per scheduler\n", name,
           COLS, EMU, w, cyc, cyc / (COLS * w));
  }
}

int main() {
  long long* out;
  cudaMallocManaged(&out, 16);
  cudaMalloc(&g_sink, 4096);
  run<128, 3, 0>("kernel pass r1 (row per thread)", out);
  run<64, 3, 0>("kernel pass, sum in loop", out);
  run<64, 0, 0>("kernel pass, sum in loop", out);
  run<64, 1, 0>("kernel pass, sum in loop", out);
  run<64, 2, 0>("kernel pass, sum in loop", out);
  run<64, 4, 0>("kernel pass, sum in loop", out);
  run<64, 8, 0>("kernel pass, sum in loop", out);
  run<64, 0, 1>("sum after", out);
  run<64, 2, 1>("sum after", out);
  run<64, 3, 1>("sum after", out);
  run<64, 2, 2>("sum after, LEA poly", out);
  run<64, 3, 2>("sum after, LEA poly", out);
  run<64, 4, 2>("sum after, LEA poly", out);
  run<64, 0, 3>("no max pass", out);
  run<64, 3, 3>("no max pass", out);
  run<64, 0, 4>("scalar fma/add", out);
  run<64, 3, 4>("scalar fma/add", out);
  run<32, 0, 0>("kernel pass, sum in loop", out);
  run<32, 3, 0>("kernel pass, sum in loop", out);
  return 0;
}
