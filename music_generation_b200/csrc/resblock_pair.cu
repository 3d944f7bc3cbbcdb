// resblock_pair: one ResBlock1 pair of the Generator's 32-channel stage in ONE kernel,
//     out = x + conv2(lrelu(conv1(lrelu(x))))            (VISinger/modules.py:210-223)
// on the tcgen05 tensor cores with the intermediate tile in shared memory: 4 bytes of HBM traffic per element (read the
// activated input, write the result; the residual re-read of the same rows is served by L2) instead of the 10 the
// two conv1d_tc launches move.  Same arithmetic in the same order as conv1d_tc -> conv1d_tc (bit-identical results).
//
// Data: channels-last [B, L, 32] 16-bit tensors holding lrelu(x) ("activated stream"); a row pair
// (2u, 2u+1) is one 128-byte "folded" row of 64 channels.  conv2 (dilation 1) always runs time-pair folded
// (packer.fold_time_pairs: 64 -> 64 channels, (k+1)/2+1 taps); conv1 runs folded when its dilation is 1 and unfolded
// (32 -> 32 channels, k taps, dilation d, K = 32) otherwise.
//
// Tile = 128 folded rows of the INTERMEDIATE t = lrelu(conv1(.)) starting at folded row s_mid; conv2 is evaluated on the
// same 128 rows and its outputs are valid for rows [p2, 128 - p2) (p2 = conv2's folded padding), so consecutive tiles
// advance by R_out = 128 - 2*p2 rows and recompute 2*p2 of 128 intermediate rows (1.6-4.7 %).
//
// Warp roles (384 threads, 1 CTA/SM, persistent over tiles); everything meets at mbarriers only:
//   warp 0 lane 0 : TMA producer of the conv1 input tile (rows with halo; zero padding = TMA out-of-bounds fill)
//   warp 1 lane 0 : tcgen05.mma issuer of every conv1
//   warp 3 lane 0 : loads both weight sets once (resident for the whole kernel), then tcgen05.mma issuer of every conv2
//   warp 2        : TMEM alloc / dealloc; lane 0 = TMA store thread
//   warps 4..11   : both epilogues, in the MMA thread's order (epilogue 1 of tile i+1, then epilogue 2 of tile i); warp w
//                   and w + 4 share a TMEM lane quadrant and split the 64 folded columns 32 / 32:
//                   epilogue 1: acc1 (TMEM) + bias -> lrelu -> fp16 -> intermediate tile in shared memory in the UMMA
//                   K-major SWIZZLE_128B layout (rows outside the sequence are written as zeros = conv2's padding)
//                   epilogue 2: acc2 + bias + residual (un-lrelu'd input row, read from global / L2) [+ 16-bit MRF sum]
//                   -> [scale] -> [lrelu] -> fp16 -> staging slot -> TMA store
#include <cstdlib>
#include <string.h>
#include "common.cuh"
#include "api_internal.h"
#include "../../include/b200svs_debug.h"

namespace b200svs {

#ifndef B200SVS_PAIR_POLL
#define B200SVS_PAIR_POLL 1
#endif
#define pwait(bar, parity) do { if (p.poll) mbar_wait_poll(bar, parity); else mbar_wait(bar, parity); } while (0)

constexpr int kPairThreads = 640;        // 4 control warps + 2 epilogue groups of 8 warps (alternate tiles)
constexpr int kMidGuard = 8;                                  // guard rows above and below the 128 intermediate rows
constexpr int kMidBytes = (128 + 2 * kMidGuard) * 128;        // 18 KB, 1024-byte aligned
constexpr int kPairSlotBytes = 128 * 128;                     // 16 KB output staging slot

struct PairParams {
  int B, Lh;                    // batch, folded rows per utterance (L / 2)
  int n_t, total_tiles;         // tiles per utterance, in all
  int r_out, p2, taps2;         // conv2 (folded)
  int c1_folded, taps1, dil1, pad1;      // conv1: folded (taps1, pad1 in folded rows) or unfolded (k taps, dilation, pad in time steps)
  int a_rows, a_boxes, a_box_rows, a_stage_bytes, na;        // conv1 input tile
  int nm;                       // intermediate buffers (1 or 2)
  int w1_tile_bytes, w2_tile_bytes;
  const float* bias1; const float* bias2;
  const void* x;                // [B, L, 32] activated 16-bit input = residual source
  const void* mrf;              // nullable: 16-bit MRF sum so far, same shape
  float res_unlrelu, scale, post_slope;
  int out_lrelu;                // 1: out = lrelu(v, post_slope); 0: raw v
  int plain;                    // scale == 1
  uint32_t idesc1, idesc2;
  int poll;                     // 1: polling mbarrier waits, 0: waits with the suspend-time hint
  long long* trace;             // diagnostic: [grid][trace_cap][16] clock64 stamps (NULL: off)
  int trace_cap;
};

struct __align__(8) PairBars {
  uint64_t w_full;
  uint64_t a_full[2], a_empty[2];
  uint64_t acc1_full[2], acc1_empty[2];
  uint64_t mid_full[2], mid_empty[2];
  uint64_t acc2_full[2], acc2_empty[2];
  uint64_t slot_full[2], slot_empty[2];
  uint32_t tmem_base;
};

__device__ __forceinline__ void pstamp(const PairParams& p, int it, int slot) {
  if (p.trace && it < p.trace_cap) p.trace[(static_cast<size_t>(blockIdx.x) * p.trace_cap + it) * 16 + slot] = clock64();
}
__device__ __forceinline__ uint4 ldg_nc128(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void sts128_u4(uint32_t a, uint4 v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ float4 lds128(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
// packed fp32x2 arithmetic (FADD2 / FMUL2): same IEEE results as two scalar operations
__device__ __forceinline__ float2 add2(float2 a, float2 b) {
  unsigned long long r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)));
  return *reinterpret_cast<float2*>(&r);
}
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
  unsigned long long r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)));
  return *reinterpret_cast<float2*>(&r);
}
__device__ __forceinline__ float2 h2f2(uint32_t u) { return __half22float2(*reinterpret_cast<const __half2*>(&u)); }

__global__ void __launch_bounds__(kPairThreads, 1)
resblock_pair_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW1,
                     const __grid_constant__ CUtensorMap tmW2, const __grid_constant__ CUtensorMap tmO, const PairParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // pointer arithmetic keeps the shared address space (LDS/STS, not generic LD/ST)
  uint8_t* smA = smem;                                                 // [na][a_stage_bytes]
  uint8_t* smM = smA + p.na * p.a_stage_bytes;                         // [nm][kMidBytes]
  uint8_t* smS = smM + p.nm * kMidBytes;                               // [2][kPairSlotBytes]
  uint8_t* smW1 = smS + 2 * kPairSlotBytes;                            // [taps1][w1_tile_bytes]
  uint8_t* smW2 = smW1 + p.taps1 * p.w1_tile_bytes;                    // [taps2][w2_tile_bytes]
  PairBars* bars = reinterpret_cast<PairBars*>(smW2 + p.taps2 * p.w2_tile_bytes);
  float* bias1_s = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 256);   // [64]
  float* bias2_s = bias1_s + 64;                                                       // [64]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmX); prefetch_tmap(&tmW1); prefetch_tmap(&tmW2); prefetch_tmap(&tmO);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(&bars->w_full, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&bars->a_full[s], 1); mbar_init(&bars->a_empty[s], 1);
      mbar_init(&bars->acc1_full[s], 1); mbar_init(&bars->acc1_empty[s], 8);
      mbar_init(&bars->mid_full[s], 8); mbar_init(&bars->mid_empty[s], 1);
      mbar_init(&bars->acc2_full[s], 1); mbar_init(&bars->acc2_empty[s], 8);
      mbar_init(&bars->slot_full[s], 8); mbar_init(&bars->slot_empty[s], 1);
    }
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(&bars->tmem_base, 256); tmem_relinquish(); }
  if (threadIdx.x < 64) {
    const int n1 = p.c1_folded ? 64 : 32;
    bias1_s[threadIdx.x] = (p.bias1 && static_cast<int>(threadIdx.x) < n1) ? __ldg(p.bias1 + threadIdx.x) : 0.0f;
    bias2_s[threadIdx.x] = p.bias2 ? __ldg(p.bias2 + threadIdx.x) : 0.0f;
  }
  // the guard rows of the intermediate buffers are read by conv2's edge taps (their outputs are discarded): keep them finite
  for (int i = threadIdx.x; i < p.nm * kMidBytes / 16; i += kPairThreads)
    reinterpret_cast<uint4*>(smM)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;       // acc1: columns [0,64) + 128*buf ; acc2: [64,128) + 128*buf
  griddep_launch_dependents();

  const int n_my = (p.total_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);

  if (warp == 0) {
    if (lane == 0) {
      griddep_wait();
      // ------------------------------------------------------------ conv1 input tiles
      for (int it = 0; it < n_my; ++it) {
        const int tile = blockIdx.x + it * gridDim.x;
        const int b = tile / p.n_t, tt = tile - b * p.n_t;
        const int s_mid = tt * p.r_out - p.p2;
        const int ab = p.na == 2 ? (it & 1) : 0;
        const int use = p.na == 2 ? (it >> 1) : it;
        pwait(&bars->a_empty[ab], (use & 1) ^ 1);
        pstamp(p, it, 0);
        uint8_t* dst = smA + ab * p.a_stage_bytes;
        mbar_arrive_expect_tx(&bars->a_full[ab], static_cast<uint32_t>(p.a_boxes) * p.a_box_rows * 128);
        const int row0 = p.c1_folded ? s_mid - p.pad1 : 2 * s_mid - p.pad1;
        for (int bx = 0; bx < p.a_boxes; ++bx)
          tma_load_3d(dst + bx * p.a_box_rows * 128, &tmX, &bars->a_full[ab], 0, row0 + bx * p.a_box_rows, b);
      }
    }
  } else if (warp == 1 || warp == 3) {
    if (warp == 3 && lane == 0) {
      // ------------------------------------------------------------ both weight sets, once
      mbar_arrive_expect_tx(&bars->w_full, static_cast<uint32_t>(p.taps1 * p.w1_tile_bytes + p.taps2 * p.w2_tile_bytes));
      const int n1 = p.c1_folded ? 64 : 32;
      for (int v = 0; v < p.taps1; ++v) tma_load_2d(smW1 + v * p.w1_tile_bytes, &tmW1, &bars->w_full, 0, v * n1);
      for (int v = 0; v < p.taps2; ++v) tma_load_2d(smW2 + v * p.w2_tile_bytes, &tmW2, &bars->w_full, 0, v * 64);
    }
    __syncwarp();
    if (elect_one()) {
      // ------------------------------------------------------------ MMA issuers: warp 1 issues every conv1, warp 3 every
      // conv2.  A barrier round trip costs an issuing thread ~230-450 cycles during which the tensor pipe drains its
      // queue; with one thread per conv the two sets of waits overlap and the pipe stays fed (tcgen05.commit tracks the
      // MMAs of the executing thread only, so each thread's commits cover exactly its own conv).
      const uint64_t dA0 = make_desc_k128(smem_u32(smA)), dM0 = make_desc_k128(smem_u32(smM));
      const uint64_t dW1 = make_desc_k128(smem_u32(smW1)), dW2 = make_desc_k128(smem_u32(smW2));
      const uint32_t a_stage_d = static_cast<uint32_t>(p.a_stage_bytes) >> 4, w1_d = static_cast<uint32_t>(p.w1_tile_bytes) >> 4;
      const uint32_t w2_d = static_cast<uint32_t>(p.w2_tile_bytes) >> 4;
      constexpr uint32_t kRowD = 128 >> 4;                           // one 128-byte row in descriptor units
      auto conv1 = [&](int it) {
        const int ab = p.na == 2 ? (it & 1) : 0;
        const int use = p.na == 2 ? (it >> 1) : it;
        const int buf = it & 1;
        pstamp(p, it, 15);
        pwait(&bars->a_full[ab], use & 1);
        pwait(&bars->acc1_empty[buf], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        pstamp(p, it, 1);
        const uint32_t d = tmem_base + buf * 128;
        const uint64_t da = dA0 + ab * a_stage_d;
        if (p.c1_folded) {
          uint32_t accum = 0;
          for (int v = 0; v < p.taps1; ++v) {
#pragma unroll
            for (int k = 0; k < 4; ++k) { umma_f16(d, da + static_cast<uint32_t>(v * p.dil1) * kRowD + 2 * k, dW1 + v * w1_d + 2 * k, p.idesc1, accum); accum = 1; }
          }
        } else {
#pragma unroll
          for (int ms = 0; ms < 2; ++ms) {
            uint32_t accum = 0;
            for (int j = 0; j < p.taps1; ++j) {
              const uint64_t a = da + static_cast<uint32_t>(ms * 128 + j * p.dil1) * kRowD;
#pragma unroll
              for (int k = 0; k < 2; ++k) { umma_f16(d + ms * 32, a + 2 * k, dW1 + j * w1_d + 2 * k, p.idesc1, accum); accum = 1; }
            }
          }
        }
        umma_commit(&bars->a_empty[ab]);
        umma_commit(&bars->acc1_full[buf]);
        pstamp(p, it, 2);
      };
      auto conv2 = [&](int it) {
        const int mb = p.nm == 2 ? (it & 1) : 0;
        const int use = p.nm == 2 ? (it >> 1) : it;
        const int buf = it & 1;
        pstamp(p, it, 13);
        pwait(&bars->mid_full[mb], use & 1);
        // acc2[buf] is free as well: its last reader, epilogue 2 of tile it-2, precedes epilogue 1 of tile it in the same
        // warp group, whose arrival on mid_full(it) was just observed.
        tc_fence_after();
        pstamp(p, it, 3);
        const uint32_t d = tmem_base + buf * 128 + 64;
        const uint64_t dm = dM0 + static_cast<uint32_t>(mb * kMidBytes >> 4) + static_cast<uint32_t>(kMidGuard - p.p2) * kRowD;
        uint32_t accum = 0;
        for (int v = 0; v < p.taps2; ++v) {
#pragma unroll
          for (int k = 0; k < 4; ++k) { umma_f16(d, dm + v * kRowD + 2 * k, dW2 + v * w2_d + 2 * k, p.idesc2, accum); accum = 1; }
        }
        umma_commit(&bars->mid_empty[mb]);
        umma_commit(&bars->acc2_full[buf]);
        pstamp(p, it, 14);
      };
      pwait(&bars->w_full, 0);
      tc_fence_after();
      if (warp == 1) {
        for (int it = 0; it < n_my; ++it) conv1(it);
      } else {
        for (int it = 0; it < n_my; ++it) conv2(it);
      }
    }
    __syncwarp();
  } else if (warp == 2) {
    if (lane == 0) {
      // ------------------------------------------------------------ TMA store thread
      griddep_wait();
      for (int it = 0; it < n_my; ++it) {
        const int tile = blockIdx.x + it * gridDim.x;
        const int b = tile / p.n_t, tt = tile - b * p.n_t;
        const int s = it & 1;
        pwait(&bars->slot_full[s], (it >> 1) & 1);
        pstamp(p, it, 10);
        tma_store_3d(&tmO, smS + s * kPairSlotBytes, 0, tt * p.r_out, b);
        tma_store_commit();
        tma_store_wait_read<0>();
        mbar_arrive(&bars->slot_empty[s]);
        pstamp(p, it, 11);
      }
      tma_store_wait_all<0>();
    }
    __syncwarp();
  } else if (warp >= 4) {
    // -------------------------------------------------------------- epilogue warps 4..11: warp w and w + 4 share TMEM
    // lane quadrant w % 4 (row r of the tile) and split the 64 folded columns 32 / 32 (half = time step of the row pair).
    // Both phases run on all eight warps, in the MMA thread's order: epilogue 1 of tile i+1, then epilogue 2 of tile i.
    griddep_wait();
    const int grp = (warp - 4) >> 3;                    // group g owns tiles g, g + 2, ... of this CTA
    const int q = warp & 3, half = ((warp - 4) >> 2) & 1, r = q * 32 + lane;
    const bool tracer = ((threadIdx.x - 128) & 255) == 0;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    const uint32_t smM_u = smem_u32(smM), smS_u = smem_u32(smS);
    const uint32_t b1_u = smem_u32(bias1_s) + (p.c1_folded ? half * 128 : 0), b2_u = smem_u32(bias2_s) + half * 128;
    const float rinv = p.res_unlrelu > 1.0f ? p.res_unlrelu : 1.0f;
    const float2 slope1_2 = make_float2(0.1f, 0.1f), rinv2 = make_float2(rinv, rinv);
    const float2 scale2 = make_float2(p.scale, p.scale), pslope2 = make_float2(p.post_slope, p.post_slope);

    // t = fp16(lrelu(acc1 + bias1)) for this thread's 32 columns -> 4 cells of the intermediate tile
    auto epi1 = [&](int it) {
      const int tile = blockIdx.x + it * gridDim.x;
      const int tt = tile % p.n_t;
      const int s_mid = tt * p.r_out - p.p2;
      const int buf = it & 1;
      const int mb = p.nm == 2 ? (it & 1) : 0;
      const int use = p.nm == 2 ? (it >> 1) : it;
      pwait(&bars->acc1_full[buf], (it >> 1) & 1);
      tc_fence_after();
      if (tracer) pstamp(p, it, 4);
      uint32_t a[32];
      tmem_ld32_nowait(tmem_base + buf * 128 + lane_off + half * 32, a);
      tmem_ld_wait32(a);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->acc1_empty[buf]);
      pwait(&bars->mid_empty[mb], (use & 1) ^ 1);          // conv2 of the tile that used this buffer has read it
      if (tracer) pstamp(p, it, 5);
      // folded conv1: row r of the tile, columns [32*half, +32) = time step `half` of the pair -> cells 4*half ..
      // plain conv1 : sub-tile `half`, time step half*128 + r, its 32 channels -> row (half*128 + r)/2, cells 4*(r&1) ..
      int R, cell0;
      bool ok;
      if (p.c1_folded) {
        const int g = s_mid + r;
        ok = g >= 0 && g < p.Lh;
        R = r; cell0 = half * 4;
      } else {
        const int tl = half * 128 + r, t = 2 * s_mid + tl;
        ok = t >= 0 && t < 2 * p.Lh;
        R = tl >> 1; cell0 = (tl & 1) * 4;
      }
      const uint32_t rowa = smM_u + mb * kMidBytes + (kMidGuard + R) * 128;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float4 bq0 = lds128(b1_u + c * 32), bq1 = lds128(b1_u + c * 32 + 16);
        float2 v[4];
        v[0] = add2(make_float2(__uint_as_float(a[8 * c]), __uint_as_float(a[8 * c + 1])), make_float2(bq0.x, bq0.y));
        v[1] = add2(make_float2(__uint_as_float(a[8 * c + 2]), __uint_as_float(a[8 * c + 3])), make_float2(bq0.z, bq0.w));
        v[2] = add2(make_float2(__uint_as_float(a[8 * c + 4]), __uint_as_float(a[8 * c + 5])), make_float2(bq1.x, bq1.y));
        v[3] = add2(make_float2(__uint_as_float(a[8 * c + 6]), __uint_as_float(a[8 * c + 7])), make_float2(bq1.z, bq1.w));
        uint32_t pk[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 m = mul2(v[e], slope1_2);
          pk[e] = ok ? cvt_f16x2_sat(fmaxf(v[e].x, m.x), fmaxf(v[e].y, m.y)) : 0u;
        }
        sts128_u4(rowa + (((cell0 + c) ^ (R & 7)) << 4), make_uint4(pk[0], pk[1], pk[2], pk[3]));
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->mid_full[mb]);
      if (tracer) pstamp(p, it, 6);
    };

    // out = acc2 + bias2 + x [+ mrf] -> [scale] -> [lrelu] -> fp16 for this thread's 32 columns -> 4 cells of the slot
    auto epi2 = [&](int it) {
      const int tile = blockIdx.x + it * gridDim.x;
      const int b = tile / p.n_t, tt = tile - b * p.n_t;
      const int s_mid = tt * p.r_out - p.p2;
      const int buf = it & 1, s = it & 1;
      const int g = s_mid + r;
      const bool ok = r >= p.p2 && r < 128 - p.p2 && g < p.Lh;
      if (tracer) pstamp(p, it, 12);
      // residual (and MRF) half rows first: their latency hides behind the wait for conv2
      uint4 xr[4] = {};
      const size_t row_off = (static_cast<size_t>(b) * p.Lh + (ok ? g : 0)) * 128 + half * 64;
      if (ok) {
#pragma unroll
        for (int c = 0; c < 4; ++c) xr[c] = ldg_nc128(static_cast<const uint8_t*>(p.x) + row_off + c * 16);
      }
      pwait(&bars->acc2_full[buf], (it >> 1) & 1);
      tc_fence_after();
      if (tracer) pstamp(p, it, 7);
      uint32_t a[32];
      tmem_ld32_nowait(tmem_base + buf * 128 + 64 + lane_off + half * 32, a);
      tmem_ld_wait32(a);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->acc2_empty[buf]);
      pwait(&bars->slot_empty[s], ((it >> 1) & 1) ^ 1);
      if (tracer) pstamp(p, it, 8);
      if (ok) {
        const int ro = r - p.p2;
        const uint32_t rowa = smS_u + s * kPairSlotBytes + ro * 128;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const float4 bq0 = lds128(b2_u + c * 32), bq1 = lds128(b2_u + c * 32 + 16);
          const float2 bb[4] = {make_float2(bq0.x, bq0.y), make_float2(bq0.z, bq0.w), make_float2(bq1.x, bq1.y), make_float2(bq1.z, bq1.w)};
          const uint32_t xw[4] = {xr[c].x, xr[c].y, xr[c].z, xr[c].w};
          uint4 mrc = make_uint4(0, 0, 0, 0);      // plain load: these rows are overwritten by this tile's own store when out aliases mrf
          if (p.mrf) mrc = *reinterpret_cast<const uint4*>(static_cast<const uint8_t*>(p.mrf) + row_off + c * 16);
          const uint32_t mw[4] = {mrc.x, mrc.y, mrc.z, mrc.w};
          uint32_t pk[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float2 v = add2(make_float2(__uint_as_float(a[8 * c + 2 * e]), __uint_as_float(a[8 * c + 2 * e + 1])), bb[e]);
            const float2 f = h2f2(xw[e]);
            const float2 m = mul2(f, rinv2);
            v = add2(v, make_float2(fminf(f.x, m.x), fminf(f.y, m.y)));
            if (p.mrf) v = add2(v, h2f2(mw[e]));
            if (!p.plain) v = mul2(v, scale2);
            if (p.out_lrelu) {
              const float2 l = mul2(v, pslope2);
              v = make_float2(fmaxf(v.x, l.x), fmaxf(v.y, l.y));
            }
            pk[e] = cvt_f16x2_sat(v.x, v.y);
          }
          sts128_u4(rowa + (((half * 4 + c) ^ (ro & 7)) << 4), make_uint4(pk[0], pk[1], pk[2], pk[3]));
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->slot_full[s]);
      if (tracer) pstamp(p, it, 9);
    };

    // conv2's result of tile i is drained, then conv1's of tile i + 2 (the order in which the tensor pipe produces them)
    if (grp < n_my) epi1(grp);
    for (int it = grp; it < n_my; it += 2) {
      epi2(it);
      if (it + 2 < n_my) epi1(it + 2);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

}  // namespace b200svs

using namespace b200svs;

static long long* g_pair_trace = nullptr;
static int g_pair_trace_cap = 0;
extern "C" int b200svs_resblock_pair_set_trace(void* dev_buf, int32_t tiles_per_cta) {
  g_pair_trace = static_cast<long long*>(dev_buf);
  g_pair_trace_cap = dev_buf ? tiles_per_cta : 0;
  return 0;
}

// x / out / mrf: [B, L, 32] fp16 channels-last, dense (row stride 32), 16-byte aligned; L even.  out must not alias x
// (tiles read halo rows of x that neighbouring tiles would overwrite); out may alias mrf (row-wise in place).
extern "C" int b200svs_resblock_pair(const b200svs_conv_w* c1, const b200svs_conv_w* c2, const void* x, int32_t B, int32_t L,
                                     const void* mrf, void* out, int32_t out_lrelu, float post_slope, float scale,
                                     float res_unlrelu, void* stream) {
  B200SVS_REQUIRE(c1 && c2 && x && out && B > 0 && L > 0, "resblock_pair: bad argument");
  B200SVS_REQUIRE(out != x, "resblock_pair: out must not alias x");
  B200SVS_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(mrf)) & 15) == 0,
                  "resblock_pair: tensors must be 16-byte aligned");
  B200SVS_REQUIRE(!c1->f32 && !c1->gate && !c2->f32 && !c2->gate && c1->cin_pad == 64 && c2->cin_pad == 64 && c2->dil == 1,
                  "resblock_pair: 16-bit tensor-core convs with K <= 64 expected, conv2 with dilation 1");
  // 64-channel stage: rows are time steps (64 channels = 128 bytes), both convs plain 64 -> 64.
  // 32-channel stage: rows are time-step PAIRS; conv2 folded (64 -> 64), conv1 folded or plain 32 -> 32.
  const bool wide = c2->fold <= 1;
  if (wide) {
    B200SVS_REQUIRE(c2->cin == 64 && c2->cout_pad == 64 && c1->cin == 64 && c1->cout_pad == 64 && c1->fold <= 1,
                    "resblock_pair: 64-channel pair needs two plain 64 -> 64 convs");
  } else {
    B200SVS_REQUIRE(L % 2 == 0, "resblock_pair: L=%d must be even (time-pair folding)", L);
    B200SVS_REQUIRE(c2->fold == 2 && c2->cin == 64 && c2->cout_pad == 64, "resblock_pair: conv2 must be the folded 32-channel conv");
    B200SVS_REQUIRE(c1->fold == 2 ? (c1->cin == 64 && c1->cout_pad == 64 && c1->dil == 1) : (c1->cin == 32 && c1->cout_pad == 32),
                    "resblock_pair: conv1 must be the folded (64 -> 64) or plain (32 -> 32) 32-channel conv");
  }
  const bool f1 = wide || c1->fold == 2;            // conv1 works on 128-byte rows of 64 channels
  PairParams p;
  memset(&p, 0, sizeof(p));
  p.B = B; p.Lh = wide ? L : L / 2;
  p.taps2 = c2->ntaps; p.p2 = c2->pad_left;
  B200SVS_REQUIRE(p.p2 >= 0 && p.p2 <= kMidGuard && p.taps2 - 1 - p.p2 <= kMidGuard && p.taps2 <= 8, "resblock_pair: conv2 taps=%d pad=%d",
                  p.taps2, p.p2);
  p.r_out = 128 - 2 * (p.p2 > p.taps2 - 1 - p.p2 ? p.p2 : p.taps2 - 1 - p.p2);
  B200SVS_REQUIRE(p.taps2 - 1 - p.p2 == p.p2, "resblock_pair: conv2 must be centred (taps=%d pad=%d)", p.taps2, p.p2);
  p.n_t = (p.Lh + p.r_out - 1) / p.r_out;
  p.total_tiles = p.n_t * B;
  p.c1_folded = f1 ? 1 : 0;
  p.taps1 = c1->ntaps; p.dil1 = c1->dil; p.pad1 = c1->pad_left;
  B200SVS_REQUIRE(p.taps1 >= 1 && p.taps1 <= 16, "resblock_pair: conv1 taps=%d", p.taps1);
  p.a_rows = f1 ? 128 + (p.taps1 - 1) * p.dil1 : 256 + (p.taps1 - 1) * p.dil1;
  p.a_boxes = p.a_rows > 256 ? 2 : 1;
  p.a_box_rows = ((p.a_rows + p.a_boxes - 1) / p.a_boxes + 7) / 8 * 8;
  p.a_stage_bytes = p.a_boxes * p.a_box_rows * 128;
  p.w1_tile_bytes = (f1 ? 64 : 32) * 128;
  p.w2_tile_bytes = 64 * 128;
  p.bias1 = c1->bias; p.bias2 = c2->bias;
  p.x = x; p.mrf = mrf;
  p.res_unlrelu = res_unlrelu; p.scale = scale; p.post_slope = post_slope; p.out_lrelu = out_lrelu; p.plain = scale == 1.0f;
  p.idesc1 = make_idesc_f16(128, f1 ? 64 : 32, false);
  p.idesc2 = make_idesc_f16(128, 64, false);
  p.trace = g_pair_trace; p.trace_cap = g_pair_trace_cap;
  static const int poll_on = [] { const char* e = getenv("B200SVS_PAIR_POLL"); return e ? atoi(e) : 1; }();
  p.poll = poll_on;
  // shared memory: weights (resident) + 2 output slots + barriers / biases; then as many intermediate buffers and input
  // stages as fit (2 + 2, else 1 intermediate buffer, else 1 input stage)
  const int fixed = p.taps1 * p.w1_tile_bytes + p.taps2 * p.w2_tile_bytes + 2 * kPairSlotBytes + 1024 + 1024;
  const int budget = max_smem_optin() - fixed;
  const int combos[4][2] = {{2, 2}, {1, 2}, {2, 1}, {1, 1}};
  for (int i = 0; i < 4; ++i) {
    p.na = combos[i][0]; p.nm = combos[i][1];
    if (budget >= p.na * p.a_stage_bytes + p.nm * kMidBytes) break;
  }
  B200SVS_REQUIRE(budget >= p.na * p.a_stage_bytes + p.nm * kMidBytes, "resblock_pair: shapes need %d bytes of shared memory",
                  fixed + p.a_stage_bytes + kMidBytes);
  const size_t smem = static_cast<size_t>(fixed) + p.na * p.a_stage_bytes + p.nm * kMidBytes;

  CUtensorMap tmX, tmW1, tmW2, tmO;
  {
    uint64_t dims[3], strides[2];
    if (f1) { dims[0] = 64; dims[1] = p.Lh; dims[2] = B; strides[0] = 128; strides[1] = static_cast<uint64_t>(p.Lh) * 128; }
    else { dims[0] = 32; dims[1] = L; dims[2] = B; strides[0] = 64; strides[1] = static_cast<uint64_t>(L) * 64; }
    uint32_t box[3] = {64, static_cast<uint32_t>(p.a_box_rows), 1};
    int r = make_tmap(&tmX, 0, x, 3, dims, strides, box, 128);
    if (r) return r;
  }
  {
    uint64_t dims[2] = {64, static_cast<uint64_t>(p.taps1) * c1->cout_pad}, strides[1] = {128};
    uint32_t box[2] = {64, static_cast<uint32_t>(f1 ? 64 : 32)};
    int r = make_tmap(&tmW1, 0, c1->w, 2, dims, strides, box, 128);
    if (r) return r;
  }
  {
    uint64_t dims[2] = {64, static_cast<uint64_t>(p.taps2) * 64}, strides[1] = {128};
    uint32_t box[2] = {64, 64};
    int r = make_tmap(&tmW2, 0, c2->w, 2, dims, strides, box, 128);
    if (r) return r;
  }
  {
    uint64_t dims[3] = {64, static_cast<uint64_t>(p.Lh), static_cast<uint64_t>(B)};
    uint64_t strides[2] = {128, static_cast<uint64_t>(p.Lh) * 128};
    uint32_t box[3] = {64, static_cast<uint32_t>(p.r_out), 1};
    int r = make_tmap(&tmO, 0, out, 3, dims, strides, box, 128);
    if (r) return r;
  }
  B200SVS_SMEM_OPTIN(resblock_pair_kernel, "cudaFuncSetAttribute(resblock_pair)");
  const int grid = p.total_tiles < num_sms() ? p.total_tiles : num_sms();
  static const int pdl_on = [] { const char* e = getenv("B200SVS_PDL"); return e ? atoi(e) : 1; }();
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(kPairThreads); cfg.dynamicSmemBytes = smem; cfg.stream = as_stream(stream);
  cudaLaunchAttribute attr[1];
  if (pdl_on) {
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
  }
  const int r = check_cuda(cudaLaunchKernelEx(&cfg, resblock_pair_kernel, tmX, tmW1, tmW2, tmO, p), "resblock_pair_kernel launch");
  if (r == B200SVS_OK) count_launch();
  return r;
}
