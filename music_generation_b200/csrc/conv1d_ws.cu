// conv1d_ws: weight-stationary ("transposed") implicit-GEMM conv for the 64-channel, dilation-1 ResBlock convs of the
// Generator (VISinger/modules.py:210-223, k = 3 / 7 / 11).
//
// Why.  conv1d_tc computes D[time, Cout] = X[time, K] . W^T with both operands in shared memory; at N = Cout = 64 an
// M = 128 MMA fetches 4 KB of activations + 2 KB of weights for 128 x 64 x 16 MACs and takes ~75 cycles - 2.3 x its
// math time (32 cycles): these convs are bound by the tensor core's shared-memory operand fetch.
// attention_ts showed that an MMA whose A operand comes from TENSOR MEMORY runs at the math floor.  The activations
// cannot be that operand (every tap needs a row-shifted view, TMEM lanes do not shift) - but the weights can, if the
// product is transposed:
//     D^T[(s, co), g] = sum_{j, ci} A[(s, co), (j, ci)] . X[2g + j, ci]          g = time PAIR, s = step within the pair
// A = the weights as a banded [128 x (k+1)*64] matrix (row (s, co) holds W[co, ci, j - s + pad], zero outside 0..k-1):
// M = 128 rows = 2 time steps x 64 output channels, stationary in TMEM for the whole kernel.  B = the activations as
// they lie in memory: a [groups x 64 channels] K-major tile per step parity (TMA with the strided view [T/2][2][64]), and
// chunk j of K is that tile read from a start address shifted by (j - parity) / 2 rows.  Per MMA the tensor core now
// fetches only N x 32 B of activations; 11 of the 12 chunks of a k = 11 row are non-zero (density k / (k+1)).
// The price is a transposed epilogue: a thread owns ONE output channel of one step parity for 32 time pairs, so bias /
// residual / activation are scalar per element and go to the output tile in shared memory as 2-byte stores (a warp
// writes 64 contiguous bytes: conflict-free), from where TMA stores them with the same strided view.
//
// Tile = 64 time pairs (128 steps) of one utterance; persistent CTAs; roles: warp 0 TMA loads (4 stages), warp 1 MMA
// issuer, warp 2 TMEM alloc + TMA store thread, warps 4-11 epilogue (two per TMEM lane quadrant, 32 columns each).
// Accumulator: 64 columns x 2 buffers; weights: (k+1) x 32 columns (384 for k = 11).
#include "common.cuh"
#include "api_internal.h"

namespace b200svs {

constexpr int kWsC = 64;                     // channels in = out
constexpr int kWsNT = 64;                    // time pairs per tile
constexpr int kWsXTileBytes = 9 * 1024;      // one parity tile of the input: up to 64 + 2*3 rows of 128 B, 1024-aligned
constexpr int kWsOTileBytes = kWsNT * 128;   // 8 KB: one parity tile of the output / residual / MRF staging
// Staging / output stage counts are EVEN: the two epilogue groups take alternate tiles, so with an even count a stage always
// belongs to the same group and every waiter sees every phase of its barrier (a parity wait cannot tell phase k from k - 2).
constexpr int kWsXStages = 3, kWsRStages = 4, kWsOStages = 2;      // R: 4 stages of 16 KB (residual only) or 32 KB (+ MRF)
constexpr int kWsEpiWarps = 8;              // per group; two groups work on alternate tiles
constexpr int kWsThreads = 128 + 2 * kWsEpiWarps * 32;

struct WsParams {
  const uint32_t* w_img;      // [128][nj * 32] packed fp16 pairs (the TMEM image of A)
  const float* bias;          // [64]
  int nj, pad, halo;          // chunks = taps + 1; left pad (k-1)/2; halo pairs per side (pad+1)/2
  int B, T, tiles_per_row, total_tiles;
  int has_res, has_mrf, out_lrelu;
  float res_unlrelu, post_slope, scale;
  int rst, r_stage_bytes;     // residual / MRF staging: stages in use and bytes per stage
  int debug;                  // timing experiments only (B200SVS_WS_DEBUG): 1 skip the MMAs, 2 skip the epilogue arithmetic, 4 skip the stores
};

struct __align__(8) WsBars {
  uint64_t x_full[kWsXStages], x_empty[kWsXStages];
  uint64_t r_full[kWsRStages], r_empty[kWsRStages];
  uint64_t acc_full[2], acc_empty[2];
  uint64_t o_ready[kWsOStages], o_free[kWsOStages];
  uint64_t w_ready;
  uint32_t tmem_base;
};

__device__ __forceinline__ void tma_load_4d(void* smem, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(smem)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* m, const void* smem, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(smem)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}

template <int NJ>
__global__ void __launch_bounds__(kWsThreads, 1)
conv1d_ws_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmR,
                 const __grid_constant__ CUtensorMap tmM, const __grid_constant__ CUtensorMap tmO, const WsParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* smX = smem;                                              // [stages][2 parities][9 KB]
  uint8_t* smR = smX + kWsXStages * 2 * kWsXTileBytes;              // [stages][res | mrf][2 parities][8 KB]
  uint8_t* smO = smR + kWsRStages * p.r_stage_bytes;                // [stages][2 parities][8 KB]
  float* bias_s = reinterpret_cast<float*>(smO + kWsOStages * 2 * kWsOTileBytes);
  WsBars* bars = reinterpret_cast<WsBars*>(bias_s + kWsC);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int i = 0; i < kWsXStages; ++i) { mbar_init(&bars->x_full[i], 1); mbar_init(&bars->x_empty[i], 1); }
    for (int i = 0; i < kWsRStages; ++i) { mbar_init(&bars->r_full[i], 1); mbar_init(&bars->r_empty[i], kWsEpiWarps); }
    for (int i = 0; i < 2; ++i) { mbar_init(&bars->acc_full[i], 1); mbar_init(&bars->acc_empty[i], kWsEpiWarps); }
    for (int i = 0; i < kWsOStages; ++i) { mbar_init(&bars->o_ready[i], kWsEpiWarps); mbar_init(&bars->o_free[i], 1); }
    mbar_init(&bars->w_ready, 4);
    fence_barrier_init();
    prefetch_tmap(&tmX); prefetch_tmap(&tmO);
    if (p.has_res) prefetch_tmap(&tmR);
    if (p.has_mrf) prefetch_tmap(&tmM);
  }
  if (warp == 2) {
    tmem_alloc(&bars->tmem_base, 512);
    tmem_relinquish();
  }
  if (tid < kWsC) bias_s[tid] = __ldg(p.bias + tid);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_acc = bars->tmem_base;           // 2 x 64 columns
  const uint32_t tmem_w = tmem_acc + 128;              // nj x 32 columns
  constexpr int kPad = (NJ - 2) / 2, kHalo = (kPad + 1) / 2;      // taps = NJ - 1
  constexpr int xrows = kWsNT + 2 * kHalo;

  if (warp == 0) {
    if (lane == 0) {
      // ------------------------------------------------------------ TMA loads: input parity tiles, residual / MRF staging
      int it = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
        const int b = tile / p.tiles_per_row, g0 = (tile - b * p.tiles_per_row) * kWsNT;
        const int st = it % kWsXStages;
        mbar_wait(&bars->x_empty[st], ((it / kWsXStages) & 1) ^ 1);
        mbar_arrive_expect_tx(&bars->x_full[st], 2 * xrows * 128);
        uint8_t* dst = smX + st * 2 * kWsXTileBytes;
        tma_load_4d(dst, &tmX, &bars->x_full[st], 0, 0, g0 - kHalo, b);
        tma_load_4d(dst + kWsXTileBytes, &tmX, &bars->x_full[st], 0, 1, g0 - kHalo, b);
        if (p.has_res) {
          const int rs = it % p.rst;
          mbar_wait(&bars->r_empty[rs], ((it / p.rst) & 1) ^ 1);
          mbar_arrive_expect_tx(&bars->r_full[rs], p.r_stage_bytes);
          uint8_t* rd = smR + rs * p.r_stage_bytes;
          tma_load_4d(rd, &tmR, &bars->r_full[rs], 0, 0, g0, b);
          tma_load_4d(rd + kWsOTileBytes, &tmR, &bars->r_full[rs], 0, 1, g0, b);
          if (p.has_mrf) {
            tma_load_4d(rd + 2 * kWsOTileBytes, &tmM, &bars->r_full[rs], 0, 0, g0, b);
            tma_load_4d(rd + 3 * kWsOTileBytes, &tmM, &bars->r_full[rs], 0, 1, g0, b);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      // ------------------------------------------------------------ MMA issuer
      const uint32_t idesc = make_idesc_f16(128, kWsNT, false);
      mbar_wait(&bars->w_ready, 0);
      tc_fence_after();
      int it = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
        const int st = it % kWsXStages, ab = it & 1;
        mbar_wait(&bars->acc_empty[ab], ((it >> 1) & 1) ^ 1);
        mbar_wait(&bars->x_full[st], (it / kWsXStages) & 1);
        tc_fence_after();
        const uint32_t x_u = smem_u32(smX + st * 2 * kWsXTileBytes);
        const uint64_t d0 = make_desc_k128(x_u);
        if (!(p.debug & 1))
#pragma unroll
        for (int ji = 0; ji < NJ; ++ji) {
              const int j = ji - kPad;                      // input step 2g + j
          const int s = j & 1;                          // its parity (two's complement: -5 & 1 = 1)
          const int shift = (j - s) >> 1;               // pair offset, exact
          const uint64_t db = d0 + static_cast<uint64_t>((s * kWsXTileBytes + (kHalo + shift) * 128) >> 4);   // compile-time offsets
#pragma unroll
          for (int k = 0; k < 4; ++k)
            umma_f16_ts(tmem_acc + ab * kWsNT, tmem_w + ji * 32 + k * 8, db + 2 * k, idesc, (ji | k) ? 1u : 0u);
        }
        umma_commit(&bars->acc_full[ab]);
        umma_commit(&bars->x_empty[st]);
      }
    }
    __syncwarp();
  } else if (warp == 2) {
    if (lane == 0) {
      // ------------------------------------------------------------ TMA store thread
      int it = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
        const int b = tile / p.tiles_per_row, g0 = (tile - b * p.tiles_per_row) * kWsNT;
        const int ob = it % kWsOStages;
        mbar_wait(&bars->o_ready[ob], (it / kWsOStages) & 1);
        const uint8_t* src = smO + ob * 2 * kWsOTileBytes;
        if (!(p.debug & 4)) {
          tma_store_4d(&tmO, src, 0, 0, g0, b);
          tma_store_4d(&tmO, src + kWsOTileBytes, 0, 1, g0, b);
        }
        tma_store_commit();
        tma_store_wait_read<0>();
        mbar_arrive(&bars->o_free[ob]);
      }
      tma_store_wait_all<0>();
    }
  } else if (warp >= 4) {
    // ---------------------------------------------------------------- epilogue warps 4-11
    const int q = warp & 3, ch = ((warp - 4) >> 2) & 1, grp = (warp - 4) >> 3;    // TMEM lane quadrant, column half, tile parity
    const int m = q * 32 + lane;                            // accumulator row: (step parity, output channel)
    const int s_out = m >> 6, co = m & 63;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    if (ch == 0 && grp == 0) {     // the weights' TMEM image: this thread's row of A, 16 columns at a time
      const uint32_t* src = p.w_img + static_cast<size_t>(m) * (NJ * 32);
      for (int c = 0; c < NJ * 2; ++c) {
        uint32_t r[16];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const uint4 v = __ldg(reinterpret_cast<const uint4*>(src + c * 16) + e);
          r[4 * e] = v.x; r[4 * e + 1] = v.y; r[4 * e + 2] = v.z; r[4 * e + 3] = v.w;
        }
        tmem_st16(tmem_w + lane_off + c * 16, r);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->w_ready);
    }
    const float bias = bias_s[co];
    const uint32_t cell = static_cast<uint32_t>(co >> 3), sub = static_cast<uint32_t>(co & 7) * 2;
    // group `grp` takes the tiles with it & 1 == grp (its own accumulator buffer); staging / output stages rotate over all tiles
    for (int it = grp; blockIdx.x + static_cast<long long>(it) * gridDim.x < p.total_tiles; it += 2) {
      const int ab = grp, ob = it % kWsOStages, rs = it % p.rst;
      mbar_wait(&bars->acc_full[ab], (it >> 1) & 1);
      tc_fence_after();
      uint32_t v[32];
      tmem_ld32_nowait(tmem_acc + lane_off + ab * kWsNT + ch * 32, v);
      tmem_ld_wait32(v);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->acc_empty[ab]);
      if (p.has_res) mbar_wait(&bars->r_full[rs], (it / p.rst) & 1);
      mbar_wait(&bars->o_free[ob], ((it / kWsOStages) & 1) ^ 1);
      const uint8_t* rt = smR + rs * p.r_stage_bytes + s_out * kWsOTileBytes;
      const uint8_t* mt = rt + 2 * kWsOTileBytes;
      uint8_t* ot = smO + (ob * 2 + s_out) * kWsOTileBytes;
      const uint32_t base = static_cast<uint32_t>(ch) * 32 * 128 + sub;
      // row n = ch * 32 + e, n & 7 = e & 7: the swizzled 16-byte cell only depends on e & 7 (hoisted by unrolling).  Loads of
      // a batch of 16 elements are issued before the first store: the compiler cannot prove that the staging tiles do not
      // alias, and element-by-element code exposed one shared-memory round trip per element (3 900 cycles per tile).
      if (!(p.debug & 2)) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          __half rr[16], mm[16];
          if (p.has_res) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const int e = h * 16 + i;
              const uint32_t off = base + e * 128 + ((cell ^ static_cast<uint32_t>(e & 7)) << 4);
              rr[i] = *reinterpret_cast<const __half*>(rt + off);
              if (p.has_mrf) mm[i] = *reinterpret_cast<const __half*>(mt + off);
            }
          }
          __half yy[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float y = __uint_as_float(v[h * 16 + i]) + bias;
            if (p.has_res) {
              const float r = __half2float(rr[i]);
              y += fminf(r, r * p.res_unlrelu);                  // x recovered from lrelu(x)
              if (p.has_mrf) y += __half2float(mm[i]);
              y *= p.scale;
            }
            if (p.out_lrelu) y = fmaxf(y, y * p.post_slope);      // 0 <= slope < 1
            yy[i] = cvt_f16_sat(y);
          }
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int e = h * 16 + i;
            const uint32_t off = base + e * 128 + ((cell ^ static_cast<uint32_t>(e & 7)) << 4);
            *reinterpret_cast<__half*>(ot + off) = yy[i];
          }
        }
      }
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&bars->o_ready[ob]);
        if (p.has_res) mbar_arrive(&bars->r_empty[rs]);
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(bars->tmem_base, 512);
  }
}

}  // namespace b200svs

using namespace b200svs;

extern "C" int b200svs_conv1d_ws(const void* w_img, const float* bias, int32_t ntaps, const void* x, int32_t B, int32_t T,
                                 const void* res, float res_unlrelu, const void* mrf, void* out, int32_t out_lrelu,
                                 float post_slope, float scale, void* stream) {
  B200SVS_REQUIRE(w_img && bias && x && out, "conv1d_ws: null argument");
  B200SVS_REQUIRE(ntaps == 3 || ntaps == 7 || ntaps == 11, "conv1d_ws: kernel size %d (3, 7 or 11)", ntaps);
  B200SVS_REQUIRE(B > 0 && T > 0 && T % 2 == 0, "conv1d_ws: T=%d must be even (time pairs)", T);
  B200SVS_REQUIRE(!mrf || res, "conv1d_ws: an MRF input needs the residual form");
  B200SVS_REQUIRE(out != x && out != res, "conv1d_ws: the output must not alias an input (tiles read a halo)");
  int r = b200svs_check_device();
  if (r) return r;
  WsParams p;
  p.w_img = static_cast<const uint32_t*>(w_img); p.bias = bias;
  p.nj = ntaps + 1; p.pad = (ntaps - 1) / 2; p.halo = (p.pad + 1) / 2;
  p.B = B; p.T = T;
  const int G = T / 2;
  p.tiles_per_row = (G + kWsNT - 1) / kWsNT; p.total_tiles = p.tiles_per_row * B;
  p.has_res = res != nullptr; p.has_mrf = mrf != nullptr; p.out_lrelu = out_lrelu;
  p.res_unlrelu = res_unlrelu; p.post_slope = post_slope; p.scale = scale;
  static const int dbg = [] { const char* e = getenv("B200SVS_WS_DEBUG"); return e ? atoi(e) : 0; }();
  p.debug = dbg;

  CUtensorMap tmX, tmR, tmM, tmO;
  uint64_t dims[4] = {(uint64_t)kWsC, 2, (uint64_t)G, (uint64_t)B};
  uint64_t strides[3] = {(uint64_t)kWsC * 2, (uint64_t)kWsC * 4, (uint64_t)T * kWsC * 2};
  uint32_t box_in[4] = {(uint32_t)kWsC, 1, (uint32_t)(kWsNT + 2 * p.halo), 1}, box_t[4] = {(uint32_t)kWsC, 1, (uint32_t)kWsNT, 1};
  r = make_tmap(&tmX, 0, x, 4, dims, strides, box_in, 128);
  if (r) return r;
  r = make_tmap(&tmO, 0, out, 4, dims, strides, box_t, 128);
  if (r) return r;
  r = make_tmap(&tmR, 0, res ? res : x, 4, dims, strides, box_t, 128);
  if (r) return r;
  r = make_tmap(&tmM, 0, mrf ? mrf : x, 4, dims, strides, box_t, 128);
  if (r) return r;

  p.rst = kWsRStages; p.r_stage_bytes = (mrf ? 4 : 2) * kWsOTileBytes;
  const size_t smem = (size_t)kWsXStages * 2 * kWsXTileBytes + (size_t)kWsRStages * p.r_stage_bytes +
                      (size_t)kWsOStages * 2 * kWsOTileBytes + kWsC * sizeof(float) + sizeof(WsBars) + 1024;
  const int grid = p.total_tiles < num_sms() ? p.total_tiles : num_sms();
  if (ntaps == 3) {
    B200SVS_SMEM_OPTIN(conv1d_ws_kernel<4>, "cudaFuncSetAttribute(conv1d_ws<3 taps>)");
    conv1d_ws_kernel<4><<<grid, kWsThreads, smem, as_stream(stream)>>>(tmX, tmR, tmM, tmO, p);
  } else if (ntaps == 7) {
    B200SVS_SMEM_OPTIN(conv1d_ws_kernel<8>, "cudaFuncSetAttribute(conv1d_ws<7 taps>)");
    conv1d_ws_kernel<8><<<grid, kWsThreads, smem, as_stream(stream)>>>(tmX, tmR, tmM, tmO, p);
  } else {
    B200SVS_SMEM_OPTIN(conv1d_ws_kernel<12>, "cudaFuncSetAttribute(conv1d_ws<11 taps>)");
    conv1d_ws_kernel<12><<<grid, kWsThreads, smem, as_stream(stream)>>>(tmX, tmR, tmM, tmO, p);
  }
  return check_launch("conv1d_ws_kernel");
}
