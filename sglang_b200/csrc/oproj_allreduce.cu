// o_proj GEMM fused with its tensor-parallel all-reduce over NVLink / NVSwitch peer memory (sm_100a).
//
// The step right after the attention path.  Replaces, for the attention block's output
// projection, the reference's  RowParallelLinear.forward  ->  tensor_model_parallel_all_reduce
//   srt/layers/linear.py:1563 (all-reduce of the [T, hidden] partial products),
//   srt/distributed/communication_op.py:18-20, srt/distributed/parallel_state.py:622,
//   srt/distributed/device_communicators/custom_all_reduce_v2.py + kernels/jit/csrc/distributed/custom_all_reduce.cuh
//   (one-shot / two-shot peer-memory all-reduce), srt/layers/flashinfer_comm_fusion.py (fused variants)
// with ONE kernel:  out = sum_r  x_r [T, K_r] . w_r[N, K_r]^T   on every rank r of the TP group.
//
// Per output tile (128 tokens x 64 hidden columns), on every rank:
//   1. tcgen05.mma (M128 x N64 x K16, bf16/f16 -> fp32 in TMEM), operands staged by cp.async producers into a
//      6-stage ring of 128B-swizzled K-major tiles; x is re-read from L2 by every CTA, w streams from HBM once.
//   2. the fp32 accumulators are rounded to the model dtype (what the reference's GEMM output would hold) and
//      written to this rank's symmetric staging buffer; a release flag tells the tile's OWNER rank (tile % world).
//   3. the owner waits for all `world` flags, sums the tile across ranks -- in the NVSwitch with
//      multimem.ld_reduce (fp32 accumulation) when a multicast mapping exists, else by peer loads in rank order --
//      and writes the result into every rank's output buffer (multimem.st / peer stores), then flags "done".
//   4. every rank waits for the tile's done flag, so kernel completion == output complete on that rank.
// The transfer of one tile overlaps the GEMM of the others; wire traffic per rank is 2 (world-1)/world of the
// output (two-shot), not (world-1) x output (one-shot pull).
// All waits are bounded (trap after ~2 s): a missing peer ends in a launch failure, never in a hung GPU.
//
// Roofline: decode (T <= 128) is HBM-bound on w: bytes = N K s + T K s + T N s;  prefill is tensor-bound.
#include <cstdlib>

#include "common.cuh"
#include "sm100_ptx.cuh"

namespace b200 {
namespace {
namespace oar {

constexpr int BM = 128, BN = 64, BK = 128;         // BK * 2 bytes = two 128-byte swizzle rows per tile row
// Ring of up to kMaxStages k-blocks.  The w tiles (16 KiB each) sit back to back; the x tiles are packed at
// a_rows * 256 bytes each, a_rows = min(T, 128) rounded up to 8: the MMA still reads 128 rows from a swizzle
// block's base, i.e. it runs on into the following x rows (or the 32 KiB tail), which only produces output rows
// >= T that nobody stores (MMA rows are independent).  A decode batch of 64 tokens gets 5 stages = 80 KiB of w
// in flight per SM (128-token batches: 4).
constexpr int kMaxStages = 12;
constexpr int kATile = BM * BK * 2, kBTile = BN * BK * 2;
constexpr int kSmemBudget = 227 * 1024 - 1024 /*align*/ - 512 /*barriers*/;
constexpr int kEpiWarps = 4, kMmaWarp = 4, kProdWarp0 = 5, kProdWarps = 4;
constexpr int kProdThreads = kProdWarps * 32;
constexpr int kThreads = (kProdWarp0 + kProdWarps) * 32;   // 288
constexpr int kTmemCols = 64;
constexpr int kMaxWorld = 8;
constexpr int kExBatch = 4;            // tile pieces a thread keeps in flight during the exchange
// barrier slots
constexpr int kBarFull = 0, kBarEmpty = kMaxStages, kBarAccFull = 2 * kMaxStages, kBarAccEmpty = 2 * kMaxStages + 1;
constexpr int kNumBars = 2 * kMaxStages + 2;
static_assert(kNumBars * 8 + 8 <= 512, "barrier area");

struct Params {
  const void* x;
  const void* w;
  int64_t x_stride, w_stride;
  int32_t T, N, K;
  int32_t rank, world;
  int32_t tiles_m, tiles_n;
  int32_t max_tiles;
  int32_t stages, a_rows;           // ring depth and x rows per stage (see kMaxStages)
  int32_t one_shot;                 // 1: every rank reduces every tile itself; 0: owner reduces + broadcasts
  int64_t staging_bytes;            // one of the two staging buffers (alternating by launch parity)
  uint32_t* epoch_state;            // local: {epoch of the previous launch, CTAs that have finished this launch}
  uint8_t* staging_local;
  uint8_t* staging_mc;              // multicast address of the staging buffers (nullptr: peer loads)
  uint8_t* out_mc;                  // multicast address of the output buffers
  uint8_t* staging_peers[kMaxWorld];
  uint8_t* out_peers[kMaxWorld];    // out_peers[rank] is the local output
  uint32_t* flags_peers[kMaxWorld]; // per rank: arrive[max_tiles][8] then done[max_tiles]
  unsigned long long* debug_times;  // optional [tiles][8] globaltimer stamps of the epilogue's phases
};

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// Bounded spin on a flag written by a peer GPU.  Flags are launch epochs and only ever grow, so the wait is
// "at least `want`" (wrap-safe): a fast peer may already have posted epoch want + 1 for the same tile before this
// rank polled -- its epoch-`want` data is then complete and still intact, because the staging buffers alternate
// with the epoch's parity.
__device__ __forceinline__ void wait_flag(const uint32_t* p, uint32_t want, int tag) {
  long long t0 = 0;
  uint32_t polls = 0;
  while ((int32_t)(ld_acquire_sys(p) - want) < 0) {
    if ((++polls & 1023u) == 0) {
      const long long now = clock64();
      if (t0 == 0) t0 = now;
      if (now - t0 > 4000000000LL) {
        printf("b200attn oproj_allreduce: peer flag wait timed out (tag %d, block %d, want %u, have %u)\n", tag,
               blockIdx.x, want, ld_acquire_sys(p));
        __trap();
      }
    }
  }
}
__device__ __forceinline__ uint4 ld_sys_16(const void* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void st_sys_16(void* p, uint4 v) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
template <typename T> __device__ __forceinline__ uint4 multimem_ld_reduce_16(const void* p);
template <> __device__ __forceinline__ uint4 multimem_ld_reduce_16<__nv_bfloat16>(const void* p) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
template <> __device__ __forceinline__ uint4 multimem_ld_reduce_16<__half>(const void* p) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void multimem_st_16(void* p, uint4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
               : "memory");
}
__device__ __forceinline__ void epi_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

template <typename T> __device__ __forceinline__ float2 unpack2(uint32_t w);
template <> __device__ __forceinline__ float2 unpack2<__nv_bfloat16>(uint32_t w) {
  return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
}
template <> __device__ __forceinline__ float2 unpack2<__half>(uint32_t w) {
  return __half22float2(*reinterpret_cast<const __half2*>(&w));
}

}  // namespace oar

template <typename T>
__global__ void __launch_bounds__(oar::kThreads, 1) oproj_allreduce_kernel(const oar::Params p) {
  using namespace oar;
  using namespace ptx;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kStages = p.stages;
  const uint32_t a_stage = (uint32_t)p.a_rows * 256u;
  const uint32_t smem_a = smem_base + kStages * kBTile;                  // x ring (+ 16 KiB tail) after the w ring
  const uint32_t smem_bar = smem_a + kStages * a_stage + kATile;
  auto bar = [&](int i) -> uint32_t { return smem_bar + 8 * i; };
  const uint32_t tmem_slot = smem_bar + 8 * kNumBars;
  pdl_launch_dependents();

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(bar(kBarFull + s), kProdThreads);
      mbar_init(bar(kBarEmpty + s), 1);
    }
    mbar_init(bar(kBarAccFull), 1);
    mbar_init(bar(kBarAccEmpty), kEpiWarps);
    mbar_fence_init();
  }
  if (warp == kMmaWarp) tmem_alloc<kTmemCols>(tmem_slot);
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

  const int n_tiles = p.tiles_m * p.tiles_n;
  const int n_kb = p.K / BK;
  const int64_t row_bytes = (int64_t)p.N * 2;            // staging / output rows: [T, N] in the model dtype

  // Programmatic dependent launch: this kernel may become resident while the attention kernels before it still run.
  // The weights do not depend on them, so the producers fill the w half of the first ring stages right away; x
  // (the attention output), the epoch and every flag are read only after griddepcontrol.wait.
  int w_prefetched = 0;                                     // ring steps whose w tile is already in flight
  if (warp >= kProdWarp0) {
    const int pt = threadIdx.x - kProdWarp0 * 32;
    const int chunk = pt & 15, r0 = pt >> 4;
    const uint32_t swz_lo = (uint32_t)(chunk & 7), blk = (uint32_t)(chunk >> 3);
    for (int tile = blockIdx.x; tile < n_tiles && w_prefetched < kStages; tile += gridDim.x) {
      const int n0 = (tile % p.tiles_n) * BN;
      const uint8_t* wb = reinterpret_cast<const uint8_t*>(p.w) + ((int64_t)n0 * p.w_stride) * 2 + chunk * 16;
      for (int kb = 0; kb < n_kb && w_prefetched < kStages; ++kb, ++w_prefetched) {
        const uint32_t b_dst = smem_base + w_prefetched * kBTile + blk * (kBTile / 2);
#pragma unroll
        for (int i = 0; i < BN / 8; ++i) {
          const int row = i * 8 + r0;
          cp_async16_full(b_dst + row * 128 + ((swz_lo ^ (row & 7)) << 4),
                          wb + (int64_t)row * p.w_stride * 2 + (int64_t)kb * BK * 2);
        }
      }
    }
  }
  pdl_wait();

  // The epoch lives in device memory so that a CUDA-graph replay (same kernel arguments) still gets a fresh one:
  // every CTA reads it at start, the last CTA to finish bumps it.  Every rank launches the same sequence, so the
  // epochs agree across ranks without any host involvement.
  uint32_t epoch = 0;
  if (p.world > 1) epoch = *reinterpret_cast<volatile uint32_t*>(p.epoch_state) + 1;
  // Two staging buffers, alternating by launch: a rank may still be reading launch n's partials of a faster peer
  // that has already started launch n + 1 (one-shot mode has no "done" handshake); it cannot still be reading
  // them when that peer starts launch n + 2, because launch n + 1 needed this rank's arrive flags.
  const int64_t stage_off = (epoch & 1u) ? p.staging_bytes : 0;

  if (warp >= kProdWarp0) {
    // ------------------------------------------------------------------ producers: x and w tiles -> smem ring
    // A k-block is BK = 128 elements = 256 bytes per row = two 128-byte swizzle blocks side by side: 16 lanes
    // cover one row (two whole lines of HBM per w row and k-block), 8 rows per round.
    const int pt = threadIdx.x - kProdWarp0 * 32;
    const int chunk = pt & 15, r0 = pt >> 4;
    const uint32_t a_half = a_stage >> 1;                    // bytes of one swizzle block of an x tile
    const uint32_t swz_lo = (uint32_t)(chunk & 7), blk = (uint32_t)(chunk >> 3);
    uint32_t s = 0, round = 0;                               // ring slot and how often the ring has wrapped
    int step = 0;                                            // k-blocks issued so far (the first w_prefetched have their w)
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const int tm = tile / p.tiles_n, tn = tile - tm * p.tiles_n;
      const int m0 = tm * BM, n0 = tn * BN;
      const uint8_t* xb = reinterpret_cast<const uint8_t*>(p.x) + chunk * 16;
      const uint8_t* wb = reinterpret_cast<const uint8_t*>(p.w) + ((int64_t)n0 * p.w_stride) * 2 + chunk * 16;
      for (int kb = 0; kb < n_kb; ++kb) {
        if (round > 0) mbar_wait(bar(kBarEmpty + s), (round - 1) & 1, 10 + s);
        const uint32_t a_dst = smem_a + s * a_stage + blk * a_half, b_dst = smem_base + s * kBTile + blk * (kBTile / 2);
        const int64_t koff = (int64_t)kb * BK * 2;
        if (step++ >= w_prefetched) {
#pragma unroll
          for (int i = 0; i < BN / 8; ++i) {
            const int row = i * 8 + r0;
            cp_async16_full(b_dst + row * 128 + ((swz_lo ^ (row & 7)) << 4), wb + (int64_t)row * p.w_stride * 2 + koff);
          }
        }
#pragma unroll
        for (int i = 0; i < BM / 8; ++i) {
          const int row = i * 8 + r0;
          if (row < p.a_rows) {
            const bool valid = m0 + row < p.T;
            const uint8_t* src = xb + (valid ? (int64_t)(m0 + row) * p.x_stride * 2 + koff : 0);
            cp_async16(a_dst + row * 128 + ((swz_lo ^ (row & 7)) << 4), src, valid);
          }
        }
        cp_async_mbar_arrive_noinc(bar(kBarFull + s));
        if (++s == (uint32_t)kStages) { s = 0; ++round; }
      }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
  } else if (warp == kMmaWarp) {
    // ------------------------------------------------------------------ MMA issue (one elected lane)
    // This warp's loop is the serial spine of the kernel (wait -> fence -> 8 MMAs -> commit per k-block): no
    // divisions, stage and phase counted incrementally.
    constexpr int kFmt = dtype_of<T>::value == B200_BF16 ? 1 : 0;
    constexpr uint32_t idesc = umma_idesc_f16(BM, BN, kFmt, 0, 0);
    constexpr uint32_t kHi = (1024u >> 4) | (1u << 14) | (2u << 29);     // SBO 1024 B | version 1 | SWIZZLE_128B
    constexpr uint32_t kLbo = (16u >> 4) << 16;
    auto mk = [](uint32_t lo, uint32_t hi) -> uint64_t { return ((uint64_t)hi << 32) | lo; };
    const bool leader = elect_one();
    const uint32_t a_half16 = a_stage >> 5;                  // (a_stage / 2) >> 4
    uint32_t s = 0, phase = 0, it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      if (it > 0) {                                         // the epilogue has drained the accumulator
        mbar_wait(bar(kBarAccEmpty), (it - 1) & 1, 20);
        tc_fence_after_sync();
      }
      for (int kb = 0; kb < n_kb; ++kb) {
        mbar_wait(bar(kBarFull + s), phase, 30 + s);
        fence_proxy_async_smem();
        tc_fence_after_sync();
        const uint32_t a_lo = ((smem_a + s * a_stage) >> 4) | kLbo;
        const uint32_t b_lo = ((smem_base + s * kBTile) >> 4) | kLbo;
#pragma unroll
        for (int ks = 0; ks < BK / 16; ++ks) {
          const uint32_t a_off = (ks >> 2) * a_half16 + (ks & 3) * 2, b_off = (ks >> 2) * (kBTile >> 5) + (ks & 3) * 2;
          if (leader) umma_f16_ss(tmem_base, mk(a_lo + a_off, kHi), mk(b_lo + b_off, kHi), idesc, (kb | ks) ? 1u : 0u);
        }
        if (leader) umma_commit(bar(kBarEmpty + s));
        if (++s == (uint32_t)kStages) { s = 0; phase ^= 1u; }
      }
      if (leader) umma_commit(bar(kBarAccFull));
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ epilogue + cross-GPU reduction
    const int et = threadIdx.x;                              // 0..127 == MMA row == TMEM lane
    const uint32_t t_addr = tmem_base + ((uint32_t)(warp * 32) << 16);
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const int tm = tile / p.tiles_n, tn = tile - tm * p.tiles_n;
      const int m0 = tm * BM, n0 = tn * BN;
      const int rows = min(BM, p.T - m0);
#ifdef B200_DEV
      unsigned long long* dbg = (p.debug_times != nullptr && et == 0) ? p.debug_times + (int64_t)tile * 8 : nullptr;
#else
      unsigned long long* const dbg = nullptr;             // phase stamps exist in -DB200_DEV builds only
#endif
      if (dbg) dbg[0] = globaltimer_ns();
      mbar_wait(bar(kBarAccFull), it & 1, 40);
      if (dbg) dbg[1] = globaltimer_ns();
      tc_fence_after_sync();
      uint32_t acc[2][32];
      tmem_ld32(t_addr, acc[0]);
      tmem_ld32(t_addr + 32, acc[1]);
      tmem_ld_wait();
      tc_fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(kBarAccEmpty));
      // this rank's partial tile, rounded to the model dtype: straight to the output when there is no peer
      uint8_t* dst_base = p.world == 1 ? p.out_peers[0] : p.staging_local + stage_off;
      if (et < rows) {
        uint8_t* dst = dst_base + (int64_t)(m0 + et) * row_bytes + (int64_t)n0 * 2;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          uint4 v;
          const uint32_t* a = &acc[c >> 2][(c & 3) * 8];
          v.x = pack2<T>(__uint_as_float(a[0]), __uint_as_float(a[1]));
          v.y = pack2<T>(__uint_as_float(a[2]), __uint_as_float(a[3]));
          v.z = pack2<T>(__uint_as_float(a[4]), __uint_as_float(a[5]));
          v.w = pack2<T>(__uint_as_float(a[6]), __uint_as_float(a[7]));
          *reinterpret_cast<uint4*>(dst + c * 16) = v;
        }
      }
      if (p.world == 1) continue;

      // ---- exchange.  Ordering: every thread's staging stores -> bar.sync -> ONE release at system scope
      // (cumulative) per flag; on the other side one acquire per flag -> bar.sync -> everybody's loads.
      const int owner = tile % p.world;
      uint32_t* my_flags = p.flags_peers[p.rank];
      uint32_t* arrive = my_flags + (int64_t)tile * kMaxWorld;
      const int n_units = rows * 8;                            // 16-byte pieces of the tile
      auto unit_off = [&](int u) -> int64_t {
        return (int64_t)(m0 + (u >> 3)) * row_bytes + (int64_t)n0 * 2 + (u & 7) * 16;
      };
      // sum of the tile piece at `off` over all ranks: in the switch, or by peer loads in rank order
      auto peer_sum = [&](int64_t off) -> uint4 {
        float2 acc2[4] = {{0.f, 0.f}, {0.f, 0.f}, {0.f, 0.f}, {0.f, 0.f}};
        uint4 v[kMaxWorld];
#pragma unroll
        for (int r = 0; r < kMaxWorld; ++r)
          if (r < p.world) v[r] = ld_sys_16(p.staging_peers[r] + stage_off + off);
#pragma unroll
        for (int r = 0; r < kMaxWorld; ++r) {
          if (r < p.world) {
            const uint32_t w4[4] = {v[r].x, v[r].y, v[r].z, v[r].w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float2 f = unpack2<T>(w4[i]);
              acc2[i].x += f.x; acc2[i].y += f.y;
            }
          }
        }
        return make_uint4(pack2<T>(acc2[0].x, acc2[0].y), pack2<T>(acc2[1].x, acc2[1].y),
                          pack2<T>(acc2[2].x, acc2[2].y), pack2<T>(acc2[3].x, acc2[3].y));
      };
      epi_sync();
      if (dbg) dbg[7] = globaltimer_ns();
      if (p.one_shot) {
        // one-shot: every rank tells every rank, then reduces the whole tile for itself (latency-optimal)
        if (et < p.world) st_release_sys(p.flags_peers[et] + (int64_t)tile * kMaxWorld + p.rank, epoch);
        if (dbg) dbg[2] = globaltimer_ns();
        if (et < p.world) wait_flag(arrive + et, epoch, 50 + et);
        epi_sync();
        if (dbg) dbg[3] = globaltimer_ns();
        uint8_t* out = p.out_peers[p.rank];
        for (int u0 = et; u0 < n_units; u0 += 128 * kExBatch) {
          uint4 v[kExBatch];
#pragma unroll
          for (int b = 0; b < kExBatch; ++b) {
            const int u = u0 + b * 128;
            if (u < n_units)
              v[b] = p.staging_mc != nullptr ? multimem_ld_reduce_16<T>(p.staging_mc + stage_off + unit_off(u))
                                             : peer_sum(unit_off(u));
          }
#pragma unroll
          for (int b = 0; b < kExBatch; ++b) {
            const int u = u0 + b * 128;
            if (u < n_units) *reinterpret_cast<uint4*>(out + unit_off(u)) = v[b];
          }
        }
        if (dbg) dbg[6] = globaltimer_ns();
        continue;
      }
      // two-shot: the tile's owner reduces it and writes the result into every rank's output
      if (et == 0) st_release_sys(p.flags_peers[owner] + (int64_t)tile * kMaxWorld + p.rank, epoch);
      if (dbg) dbg[2] = globaltimer_ns();
      if (owner == p.rank) {
        if (et < p.world) wait_flag(arrive + et, epoch, 50 + et);
        epi_sync();
        if (dbg) dbg[3] = globaltimer_ns();
        for (int u0 = et; u0 < n_units; u0 += 128 * kExBatch) {
          uint4 v[kExBatch];
#pragma unroll
          for (int b = 0; b < kExBatch; ++b) {
            const int u = u0 + b * 128;
            if (u < n_units)
              v[b] = p.staging_mc != nullptr ? multimem_ld_reduce_16<T>(p.staging_mc + stage_off + unit_off(u))
                                             : peer_sum(unit_off(u));
          }
#pragma unroll
          for (int b = 0; b < kExBatch; ++b) {
            const int u = u0 + b * 128;
            if (u < n_units) {
              if (p.staging_mc != nullptr) {
                multimem_st_16(p.out_mc + unit_off(u), v[b]);
              } else {
                for (int r = 0; r < p.world; ++r) st_sys_16(p.out_peers[r] + unit_off(u), v[b]);
              }
            }
          }
        }
        if (dbg) dbg[4] = globaltimer_ns();
        epi_sync();
        if (dbg) dbg[5] = globaltimer_ns();
        if (et < p.world) st_release_sys(p.flags_peers[et] + (int64_t)p.max_tiles * kMaxWorld + tile, epoch);
      }
      if (et == 0) wait_flag(my_flags + (int64_t)p.max_tiles * kMaxWorld + tile, epoch, 60);
      if (dbg) dbg[6] = globaltimer_ns();
      epi_sync();
    }
  }

  __syncthreads();
  if (warp == kMmaWarp) {
    tc_fence_after_sync();
    tmem_dealloc<kTmemCols>(tmem_base);
  }
  if (p.world > 1 && threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(p.epoch_state + 1, 1u) == gridDim.x - 1) {   // last CTA of this launch
      p.epoch_state[1] = 0;
      __threadfence();
      *reinterpret_cast<volatile uint32_t*>(p.epoch_state) = epoch;
    }
  }
}

// -DB200_DEV builds only: B200ATTN_OAR_MULTICAST=0 forces the peer-load exchange (A/B measurements)
bool oar_mc_enabled() {
#ifdef B200_DEV
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200ATTN_OAR_MULTICAST");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v != 0;
#else
  return true;
#endif
}

}  // namespace
}  // namespace b200

extern "C" int64_t b200attn_oproj_allreduce_flag_words(int32_t max_tiles) {
  return (int64_t)max_tiles * (b200::oar::kMaxWorld + 1) + 2;   // arrive | done | {epoch, exit counter}
}

extern "C" int32_t b200attn_oproj_allreduce_tiles(int32_t T, int32_t N) {
  using namespace b200::oar;
  return ((T + BM - 1) / BM) * ((N + BN - 1) / BN);
}

extern "C" int b200attn_oproj_allreduce(const b200_oproj_allreduce_args* a, b200_stream_t stream_) {
  using namespace b200;
  using namespace b200::oar;
  B200_CHECK_ARG(a != nullptr, "oproj_allreduce: null args");
  if (a->num_tokens == 0) return B200_OK;
  B200_CHECK_ARG(a->x && a->w, "oproj_allreduce: null pointer");
  B200_CHECK_ARG(a->dtype == B200_BF16 || a->dtype == B200_F16, "oproj_allreduce: dtype %d", a->dtype);
  B200_CHECK_ARG(a->num_tokens > 0 && a->out_features > 0 && a->in_features > 0 && a->out_features % BN == 0 &&
                     a->in_features % BK == 0,
                 "oproj_allreduce: T=%d N=%d K=%d (N must be a multiple of 64, K of 128)", a->num_tokens, a->out_features,
                 a->in_features);
  B200_CHECK_ARG(a->world >= 1 && a->world <= kMaxWorld && a->rank >= 0 && a->rank < a->world,
                 "oproj_allreduce: rank %d / world %d", a->rank, a->world);
  auto a16 = [](const void* ptr) { return ((uintptr_t)ptr & 15) == 0; };
  B200_CHECK_ARG(a16(a->x) && a16(a->w) && a->x_stride % 8 == 0 && a->w_stride % 8 == 0,
                 "oproj_allreduce: x / w must be 16-byte aligned with strides multiple of 8 elements");
  Params p;
  p.x = a->x; p.w = a->w; p.x_stride = a->x_stride; p.w_stride = a->w_stride;
  p.T = a->num_tokens; p.N = a->out_features; p.K = a->in_features;
  p.rank = a->rank; p.world = a->world;
  p.tiles_m = (p.T + BM - 1) / BM; p.tiles_n = p.N / BN;
  p.max_tiles = a->max_tiles; p.epoch_state = nullptr; p.one_shot = 0;
  p.staging_bytes = a->staging_bytes;
  p.debug_times = reinterpret_cast<unsigned long long*>(a->debug_times);
  const int n_tiles = p.tiles_m * p.tiles_n;
  for (int r = 0; r < kMaxWorld; ++r) {
    p.staging_peers[r] = nullptr; p.out_peers[r] = nullptr; p.flags_peers[r] = nullptr;
  }
  if (a->world == 1) {
    B200_CHECK_ARG(a->out_peers[0] != nullptr && a16(a->out_peers[0]), "oproj_allreduce: null / unaligned output");
    p.out_peers[0] = reinterpret_cast<uint8_t*>(a->out_peers[0]);
    p.staging_local = nullptr; p.staging_mc = nullptr; p.out_mc = nullptr;
  } else {
    B200_CHECK_ARG(n_tiles <= a->max_tiles, "oproj_allreduce: %d tiles exceed the flag area (%d)", n_tiles, a->max_tiles);
    for (int r = 0; r < a->world; ++r) {
      B200_CHECK_ARG(a->staging_peers[r] && a->out_peers[r] && a->flags_peers[r] && a16(a->staging_peers[r]) &&
                         a16(a->out_peers[r]),
                     "oproj_allreduce: null / unaligned peer pointer for rank %d", r);
      p.staging_peers[r] = reinterpret_cast<uint8_t*>(a->staging_peers[r]);
      p.out_peers[r] = reinterpret_cast<uint8_t*>(a->out_peers[r]);
      p.flags_peers[r] = reinterpret_cast<uint32_t*>(a->flags_peers[r]);
    }
    p.staging_local = p.staging_peers[a->rank];
    B200_CHECK_ARG(a->staging_bytes >= (int64_t)a->num_tokens * a->out_features * 2 && a->staging_bytes % 16 == 0,
                   "oproj_allreduce: staging_bytes %lld too small for [%d, %d]", (long long)a->staging_bytes,
                   a->num_tokens, a->out_features);
    // one-shot moves world x the output through the fabric but has the shorter dependency chain: use it while
    // the whole exchange stays small (decode batches); B200_OAR_* force a mode
    const int64_t fabric_bytes = (int64_t)a->num_tokens * a->out_features * 2 * a->world;
    // measured on 8 x B200 (profiles/r1_oproj_allreduce_tp8.json): hidden 8192, T = 32 one-shot 21 us / two-shot
    // 23 us, T = 64: 27 / 24 us; with two ranks both modes move the same bytes and one-shot always wins
    p.one_shot = a->mode == B200_OAR_ONE_SHOT ||
                 (a->mode == B200_OAR_AUTO && (a->world <= 2 || fabric_bytes <= (int64_t)(4 << 20)));
    p.epoch_state = p.flags_peers[a->rank] + (int64_t)a->max_tiles * (kMaxWorld + 1);
    const bool mc = oar_mc_enabled() && a->staging_mc != nullptr && a->out_mc != nullptr;
    p.staging_mc = mc ? reinterpret_cast<uint8_t*>(a->staging_mc) : nullptr;
    p.out_mc = mc ? reinterpret_cast<uint8_t*>(a->out_mc) : nullptr;
  }
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  p.a_rows = ((p.T < BM ? p.T : BM) + 7) / 8 * 8;
  p.stages = (kSmemBudget - kATile) / (kBTile + p.a_rows * 256);
  if (p.stages > kMaxStages) p.stages = kMaxStages;
#ifdef B200_DEV
  {
    static int cap = -1;                                   // B200ATTN_OAR_STAGES: tuning aid of dev builds
    if (cap < 0) { const char* e = getenv("B200ATTN_OAR_STAGES"); cap = e ? atoi(e) : 0; }
    if (cap >= 2 && cap < p.stages) p.stages = cap;
  }
#endif
  const int smem_bytes = p.stages * (kBTile + p.a_rows * 256) + kATile + 512 + 1024;
  const int kSmemMax = 227 * 1024;
  const int grid = n_tiles < sm_count() ? n_tiles : sm_count();   // every CTA resident: peers wait on each other
  B200_CHECK_ARG(grid > 0, "oproj_allreduce: no SMs?");
  if (a->dtype == B200_BF16) {
    static std::atomic<uint64_t> configured{0};
    auto kern = oproj_allreduce_kernel<__nv_bfloat16>;
    B200_CHECK_CUDA(set_max_dynamic_smem(kern, kSmemMax, configured));
    B200_CHECK_CUDA(launch_pdl(kern, dim3(grid), dim3(kThreads), smem_bytes, stream, p));
  } else {
    static std::atomic<uint64_t> configured{0};
    auto kern = oproj_allreduce_kernel<__half>;
    B200_CHECK_CUDA(set_max_dynamic_smem(kern, kSmemMax, configured));
    B200_CHECK_CUDA(launch_pdl(kern, dim3(grid), dim3(kThreads), smem_bytes, stream, p));
  }
  B200_CHECK_CUDA(cudaGetLastError());
  return B200_OK;
}
