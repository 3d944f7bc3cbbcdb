// conv1d_tc (v2): implicit-GEMM 1-D convolution on the 5th-gen tensor cores (tcgen05 + TMEM), every
// global<->shared transfer done by TMA, fp16/bf16 operands, fp32 accumulate, fused epilogue.
//
//   D[t, co] = sum_tap sum_ci  X[b, t + tap*dil - pad, ci] * W[tap][co][ci]
//
// Tile: msub*128 time rows (msub in {1,2}; one 128-lane TMEM accumulator per sub-tile) x block_n
// output channels.  Per 64-channel K chunk the activation rows [t0-pad, t0+msub*128+(ntaps-1)*dil-pad)
// are loaded ONCE (rank-3 map {C,T,B}: conv zero padding and batch edges are TMA out-of-bounds fill)
// and every tap reads them through a UMMA descriptor whose start address is shifted by tap*dil rows
// (SWIZZLE_128B is a function of the absolute smem address, so a row shift keeps the pattern; checked
// on hardware by b200svs_probe_rowshift).  Each weight tile [block_n x 64] feeds msub MMAs.
//
// Epilogue I/O goes through shared memory: the residual tile is TMA-loaded into a staging slot ahead
// of time, the epilogue warps (one TMEM lane = one time row per thread) combine it with the
// accumulator in place, and TMA stores write the fp32 and the 16-bit result (the latter with the next
// layer's leaky-ReLU applied) fully coalesced; the tensor maps clip rows >= T and columns >= n_valid.
//
// Warp roles (416 threads, 1 CTA/SM, persistent over tiles):
//   warp 0 lane 0 : TMA producer of the activation ring
//   warp 1 lane 0 : tcgen05.mma issuer (accumulators double-buffered in TMEM)
//   warp 2        : TMEM alloc / dealloc; lane 0 = TMA store thread (drains finished epilogue chunks)
//   warp 3 lane 0 : TMA producer of the epilogue staging slots (residual tiles)
//   warps 4..11   : epilogue; warp w and w+4 share a TMEM lane quadrant and split each 32-column chunk 16/16
//   warp 12 lane 0: TMA producer of the weight ring
//
// CTA pairs (conv1d_tc_kernel<true>): for tensor-bound single-N-tile convs two CTAs of a cluster run the same roles on
// adjacent tiles while the leader's MMA thread issues tcgen05.mma.cta_group::2 (M = 256) over both CTAs' activation
// rows and weight-tile halves; see the PAIR branches and DESIGN.md 4.1.
//
// Replaces: nn.Conv1d/ConvTranspose1d + surrounding pointwise ops of the reference
// (VISinger/models_ae.py:327-346, modules.py:148-176,210-223,329-349, attentions.py:138-146,277-285).
#include <cstdlib>
#include "common.cuh"
#include "api_internal.h"

namespace b200svs {

constexpr int kBlockK = 64;
constexpr int kRowBytes = kBlockK * 2;     // one K-chunk row of a 16-bit operand tile: 128 B
constexpr int kMaxAStages = 4;
constexpr int kMaxWStages = 8;
constexpr int kMaxSlots = 8;
constexpr int kMaxAcc = 4;               // TMEM accumulator ring (2 or 4 buffers)
constexpr int kEpiCols = 32;               // output columns per epilogue chunk
constexpr int kSlot32Bytes = 128 * kEpiCols * 4;   // 16 KB: [128 rows][32 fp32], SWIZZLE_128B
constexpr int kSlot16Bytes = 128 * kEpiCols * 2;   //  8 KB: [128 rows][32 x 16-bit], SWIZZLE_64B
constexpr int kEpiThreads = 256;            // 8 epilogue warps
constexpr int kTcThreads = 128 + kEpiThreads + 32;   // + warp 12: weight-ring producer

struct ConvTcParams {
  int T, batch, tiles_t, m_tiles, n_tiles, total_tiles;
  int ntaps, nchunks, dil, pad_left, cout_pad, block_n, msub;
  int a_boxes, a_box_rows, a_stage_bytes;  // activation tile = a_boxes TMA boxes of a_box_rows rows
  int a_stages, w_stages, w_stage_bytes;
  int n_slots, slot_bytes, slot16_off;     // epilogue staging slot: [fp32 area][res2 fp32 area][16-bit area]
  int slot_r2_off;                         // offset of the res2 area inside a slot (0 = none)
  int slot_raw_off;                        // offset of the raw 16-bit area (res16 in / out16r out) inside a slot
  int stores_in_flight;                    // TMA store groups left pending before a slot is recycled
  int acc_shift;                           // log2 of the accumulator ring depth (1 or 2)
  int w_resident;                          // all weight tiles stay in shared memory (w_stages = taps x K chunks)
  int epi_alt;                             // epilogue warp pairs take alternate chunks (32 columns each)
  int pair;                                // CTA pairs (cluster of 2, tcgen05 cta_group::2): M = 256 MMAs, each CTA holds half of B
  int reverse;                             // walk the tiles from the last one down (L2 reuse across back-to-back launches)
  int cw;                                  // 32-column units per epilogue chunk: 2 = [128 x 64] 16-bit chunks (128-byte rows, one
                                           // slot wait / fence / arrive / TMA store per 64 columns); residual-free lean launches only
  int lean16;                              // process_o16 variant: 1 res16 -> out16 (or conv1: no residual), 2 res16 -> out16r,
                                           // 3 res16 + 16-bit res2 -> out16r, 4 res16 + 16-bit res2 -> scaled out16; 0: generic
  int xpre;                                // epilogue prefetches the next tile's first TMEM chunk
  int peek;                                // MMA thread peeks the next tile's barriers before the last tap
  int chunks_per_sub, n_out_tile;          // output chunks per 128-row sub-tile; output columns per N tile
  int tmem_cols;
  int ksub;                                // K=16 MMA steps per K chunk (4, or fewer when Cin < 64)
  uint32_t idesc;
  unsigned int* sat_events;                // overflow canary: += 1 per epilogue thread that stored a saturated / NaN fp16 value
  long long* trace;                        // diagnostic: [grid][trace_cap][16] clock64 stamps (NULL: off)
  int trace_cap;
  b200svs_epilogue ep;
};

struct __align__(8) TcBarriers {
  uint64_t a_full[kMaxAStages], a_empty[kMaxAStages];
  uint64_t w_full[kMaxWStages], w_empty[kMaxWStages];
  uint64_t tmem_full[kMaxAcc], tmem_empty[kMaxAcc];
  uint64_t e_full[kMaxSlots], e_empty[kMaxSlots];
  uint64_t s_ready[kMaxSlots];             // epilogue warps -> store thread: slot holds a finished chunk
  uint32_t tmem_base;
};

template <int ACT>
__device__ __forceinline__ float act_fn(float v) {
  if (ACT == B200SVS_ACT_RELU) return fmaxf(v, 0.f);
  if (ACT == B200SVS_ACT_GELU) return gelu_erf_fast(v);
  if (ACT == B200SVS_ACT_SIGMOID) return 1.0f / (1.0f + expf(-v));
  if (ACT == B200SVS_ACT_TANH) return tanhf(v);
  return v;
}

__device__ __forceinline__ float4 lds128(uint32_t a) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t a, float x, float y, float z, float w) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}
__device__ __forceinline__ void sts128u(uint32_t a, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}
// saturating conversions (common.cuh): an fp16 overflow clips to +-65504 instead of becoming inf
template <bool BF>
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  return BF ? cvt_bf16x2_sat(a, b) : cvt_f16x2_sat(a, b);
}

// packed fp32x2 arithmetic (sm_100: FADD2 / FMUL2, one issue slot for two lanes' worth of fp32 work)
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

static_assert(sizeof(TcBarriers) <= 512, "barrier block must fit its 512-byte reservation");

__device__ __forceinline__ void trace_stamp(const ConvTcParams& p, int it, int slot, long long v) {
  if (p.trace && it < p.trace_cap) p.trace[(static_cast<size_t>(blockIdx.x) * p.trace_cap + it) * 16 + slot] = v;
}

// Walks this CTA's tiles (tile = blockIdx.x + i * gridDim.x, ordered [n_idx][b][time tile]) without the two
// integer divisions per tile a direct decode costs every role.
struct TileIter {
  int rank, tile, n_idx, b, tt, d_b, d_t;     // rank: position in this CTA's walk (validity); tile: the tile it maps to
  __device__ __forceinline__ explicit TileIter(const ConvTcParams& p) {
    rank = blockIdx.x;
    tile = p.reverse ? p.total_tiles - 1 - rank : rank;
    n_idx = tile / p.m_tiles;
    const int m = tile - n_idx * p.m_tiles;
    b = m / p.tiles_t;
    tt = m - b * p.tiles_t;
    d_b = gridDim.x / p.tiles_t;
    d_t = gridDim.x - d_b * p.tiles_t;
    if (p.pair && n_idx >= p.n_tiles) { n_idx = 0; b = p.batch; }
  }
  // CTA pairs walk in lock step: the peer (odd CTA) runs exactly the leader's iterations; if the tile count is odd its
  // last tile is a phantom (b == batch: every load is zero fill, every store clipped)
  __device__ __forceinline__ int lead_rank(const ConvTcParams& p) const { return p.pair ? rank - static_cast<int>(blockIdx.x & 1) : rank; }
  __device__ __forceinline__ bool valid(const ConvTcParams& p) const { return lead_rank(p) < p.total_tiles; }
  // true when this CTA has another tile after the current one
  __device__ __forceinline__ bool has_next(const ConvTcParams& p) const { return lead_rank(p) + static_cast<int>(gridDim.x) < p.total_tiles; }
  __device__ __forceinline__ int t0(const ConvTcParams& p) const { return tt * (128 * p.msub); }
  __device__ __forceinline__ void next(const ConvTcParams& p) {
    rank += gridDim.x;
    if (!p.reverse) {
      tile += gridDim.x;
      tt += d_t;
      b += d_b;
      if (tt >= p.tiles_t) { tt -= p.tiles_t; ++b; }
      while (b >= p.batch) { b -= p.batch; ++n_idx; }
      if (p.pair && n_idx >= p.n_tiles) { n_idx = 0; b = p.batch; }
    } else {               // same walk from the last tile down (ranks past the end are never decoded: valid() is false)
      tile -= gridDim.x;
      tt -= d_t;
      b -= d_b;
      if (tt < 0) { tt += p.tiles_t; --b; }
      while (b < 0 && n_idx > 0) { b += p.batch; --n_idx; }
    }
  }
};

// The epilogue of one CTA: 8 warps; warp w and w + 4 share TMEM lane quadrant w % 4 and split each 32-column
// chunk 16 / 16 (two warps per scheduler).  One thread owns 16 output columns of one time row per chunk.
// Everything constant for the launch is a template parameter or lives in a register (swizzled staging
// offsets, slopes), so a chunk is straight-line code; the accumulators of chunk i + 1 are fetched from TMEM
// while chunk i is processed.  Finished chunks are handed to the store thread through s_ready (no CTA barrier).
//   RES: 0 none, 1 res1 only, 2 res2 (+ res1 if present).  PLAIN: alpha == scale == 1, no sequence mask, no
//   speaker bias (the Generator's case).  BF: 16-bit output is bf16.  SPLIT: fp16 [hi | lo | hi] output.
//   ALT: warp w and w + 4 take alternate 32-column chunks (even / odd) and do all 32 columns of theirs, instead of
//   splitting every chunk 16 / 16: half as many slot waits, fences, barrier arrivals and loop trips per byte.
//   OUTS != 0 (act NONE, RES == 0, no alpha / mask / speaker bias): the Generator's launches.  The only output is
//   the activated 16-bit tensor (OUTS == 1) or the un-activated one (OUTS == 2, the MRF stream), the residuals are
//   res16 (R16) and optionally the 16-bit MRF sum (R2H), `scale` applies when !PLAIN.  Every output / residual flag
//   is a compile-time constant and the arithmetic runs on packed fp32x2 instructions (process_o16).
template <int ACT, int RES, bool PLAIN, bool BF, bool SPLIT, bool R16 = false, bool ALT = false, int OUTS = 0, bool R2H = false>
__device__ __forceinline__ void epilogue_loop(const ConvTcParams& p, TcBarriers* bars, uint32_t tmem_base,
                                              uint32_t smE_u, uint32_t bias_u, bool remote_te) {
  constexpr bool GATE = ACT == B200SVS_ACT_GATE;
  const b200svs_epilogue& ep = p.ep;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int q = warp & 3;
  const int row = q * 32 + lane;
  const int half = (warp - 4) >> 2;
  const int block_n = p.block_n, half_n = p.block_n >> 1;
  const int acc_cols = p.msub * block_n;
  const int acc_shift = p.acc_shift, acc_mask = (1 << p.acc_shift) - 1;
  const int cps = p.chunks_per_sub;
  const int n_chunks = p.msub * cps;
  const int n_slots = p.n_slots, slot_bytes = p.slot_bytes;
  const bool need32 = ep.out32 != nullptr, has16 = ep.out16 != nullptr;
  const bool has_res1 = ep.res1 != nullptr;
  const float slope = ep.post_lrelu ? ep.post_slope : 1.0f;     // max(v, v * 1) == v: no branch for "no lrelu"
  const float alpha = ep.alpha;
  const float* gbias = PLAIN ? nullptr : ep.gbias;
  // this thread's 16-byte cells inside a staging slot (fp32 area SWIZZLE_128B, 16-bit area SWIZZLE_64B)
  constexpr int NH = ALT ? 2 : 1;      // column halves this thread handles
  uint32_t off32[NH][4], off16[NH][2];
#pragma unroll
  for (int h = 0; h < NH; ++h) {
    const int hh = ALT ? h : half;
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) off32[h][jj] = row * 128 + (((hh * 4 + jj) ^ (row & 7)) << 4);
#pragma unroll
    for (int jj = 0; jj < 2; ++jj) off16[h][jj] = p.slot16_off + row * 64 + (((hh * 2 + jj) ^ ((row >> 1) & 3)) << 4);
  }
  const uint32_t r2_off = p.slot_r2_off;
  // raw 16-bit area (R16): residual in, un-activated result out, same cells as the activated 16-bit copy
  const uint32_t raw_delta = static_cast<uint32_t>(p.slot_raw_off - p.slot16_off);
  const bool has_res16 = R16 && ep.res16 != nullptr, has_out16r = R16 && ep.out16r != nullptr;
  const bool has_r2h = R16 && ep.res2 != nullptr && ep.res2_is16;      // 16-bit res2 (MRF stream), own slot area
  const uint32_t r2h_delta = static_cast<uint32_t>(p.slot_r2_off - p.slot16_off);
  const float rinv = (R16 && ep.res16_unlrelu > 1.0f) ? ep.res16_unlrelu : 1.0f;   // min(r, r * 1) == r: no branch

  uint32_t sat = 0;          // fp16 overflow canary (sat_probe_f16x2 of every packed pair this thread stores)

  auto process = [&](const uint32_t* a0, const uint32_t* a1, uint32_t slot_u, int pcol, int b, float s, int h) {
    float v[16];
    const uint32_t ba = bias_u + pcol * 4;
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      const float4 bq = lds128(ba + 16 * jj);
      v[4 * jj] = __uint_as_float(a0[4 * jj]) + bq.x; v[4 * jj + 1] = __uint_as_float(a0[4 * jj + 1]) + bq.y;
      v[4 * jj + 2] = __uint_as_float(a0[4 * jj + 2]) + bq.z; v[4 * jj + 3] = __uint_as_float(a0[4 * jj + 3]) + bq.w;
    }
    if constexpr (!PLAIN) {
      if (gbias) {
        const float* gb = gbias + static_cast<size_t>(b) * ep.gbias_ld + pcol;
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] += __ldg(gb + i);
      }
    }
    if constexpr (GATE) {
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        const float4 bq = lds128(ba + half_n * 4 + 16 * jj);
        float y[4] = {__uint_as_float(a1[4 * jj]) + bq.x, __uint_as_float(a1[4 * jj + 1]) + bq.y,
                      __uint_as_float(a1[4 * jj + 2]) + bq.z, __uint_as_float(a1[4 * jj + 3]) + bq.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          if (gbias) y[i] += __ldg(gbias + static_cast<size_t>(b) * ep.gbias_ld + pcol + half_n + 4 * jj + i);
          v[4 * jj + i] = gate_fast(v[4 * jj + i], y[i]);
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = act_fn<ACT>(v[i]);
    }
    float r16v[R16 ? 16 : 1];
    if constexpr (R16) {
#pragma unroll
      for (int i = 0; i < 16; ++i) r16v[i] = 0.f;
      if (has_r2h) {
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
          const float4 q = lds128(slot_u + r2h_delta + off16[h][jj]);
          const uint32_t u[4] = {__float_as_uint(q.x), __float_as_uint(q.y), __float_as_uint(q.z), __float_as_uint(q.w)};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            if (BF) {
              r16v[8 * jj + 2 * e] = __uint_as_float(u[e] << 16); r16v[8 * jj + 2 * e + 1] = __uint_as_float(u[e] & 0xFFFF0000u);
            } else {
              const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&u[e]));
              r16v[8 * jj + 2 * e] = f.x; r16v[8 * jj + 2 * e + 1] = f.y;
            }
          }
        }
      }
      if (has_res16) {      // rinv > 1: the residual tensor holds lrelu(x, 1 / rinv); x = min(r, r * rinv)
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
          const float4 q = lds128(slot_u + raw_delta + off16[h][jj]);
          const uint32_t u[4] = {__float_as_uint(q.x), __float_as_uint(q.y), __float_as_uint(q.z), __float_as_uint(q.w)};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float2 f;
            if (BF) {
              f.x = __uint_as_float(u[e] << 16); f.y = __uint_as_float(u[e] & 0xFFFF0000u);
            } else {
              f = __half22float2(*reinterpret_cast<const __half2*>(&u[e]));
            }
            r16v[8 * jj + 2 * e] += fminf(f.x, f.x * rinv); r16v[8 * jj + 2 * e + 1] += fminf(f.y, f.y * rinv);
          }
        }
      }
    }
    if constexpr (RES != 0) {
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
        if (RES == 1 || has_res1) r = lds128(slot_u + off32[h][jj]);
        if constexpr (R16) { r.x += r16v[4 * jj]; r.y += r16v[4 * jj + 1]; r.z += r16v[4 * jj + 2]; r.w += r16v[4 * jj + 3]; }
        if constexpr (RES == 2) {
          const float4 q2 = lds128(slot_u + r2_off + off32[h][jj]);
          r.x += q2.x; r.y += q2.y; r.z += q2.z; r.w += q2.w;
        }
        if constexpr (PLAIN) {
          v[4 * jj] += r.x; v[4 * jj + 1] += r.y; v[4 * jj + 2] += r.z; v[4 * jj + 3] += r.w;
        } else {   // (alpha * v + res1 + res2) * scale * mask
          v[4 * jj] = fmaf(v[4 * jj], alpha, r.x) * s; v[4 * jj + 1] = fmaf(v[4 * jj + 1], alpha, r.y) * s;
          v[4 * jj + 2] = fmaf(v[4 * jj + 2], alpha, r.z) * s; v[4 * jj + 3] = fmaf(v[4 * jj + 3], alpha, r.w) * s;
        }
      }
    } else if constexpr (R16) {
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = PLAIN ? v[i] + r16v[i] : fmaf(v[i], alpha, r16v[i]) * s;
    } else if constexpr (!PLAIN) {
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = fmaf(v[i], alpha, 0.0f) * s;
    }
    if (need32) {
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) sts128(slot_u + off32[h][jj], v[4 * jj], v[4 * jj + 1], v[4 * jj + 2], v[4 * jj + 3]);
    }
    if constexpr (R16) {
      if (has_out16r) {
        uint32_t p16[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) { p16[i] = pack2<BF>(v[2 * i], v[2 * i + 1]); if (!BF) sat |= sat_probe_f16x2(p16[i]); }
        sts128u(slot_u + raw_delta + off16[h][0], p16[0], p16[1], p16[2], p16[3]);
        sts128u(slot_u + raw_delta + off16[h][1], p16[4], p16[5], p16[6], p16[7]);
      }
    }
    if (has16) {
      uint32_t p16[8];
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], v[i] * slope);     // leaky-ReLU, 0 < slope <= 1
#pragma unroll
      for (int i = 0; i < 8; ++i) { p16[i] = pack2<BF>(v[2 * i], v[2 * i + 1]); if (!BF) sat |= sat_probe_f16x2(p16[i]); }
      sts128u(slot_u + off16[h][0], p16[0], p16[1], p16[2], p16[3]);
      sts128u(slot_u + off16[h][1], p16[4], p16[5], p16[6], p16[7]);
      if constexpr (SPLIT) {   // lo = v - hi, so that hi + lo == v to ~2^-22
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] -= __half2float(cvt_f16_sat(v[i]));
#pragma unroll
        for (int i = 0; i < 8; ++i) p16[i] = pack2<false>(v[2 * i], v[2 * i + 1]);
        sts128u(slot_u + kSlot16Bytes + off16[h][0], p16[0], p16[1], p16[2], p16[3]);
        sts128u(slot_u + kSlot16Bytes + off16[h][1], p16[4], p16[5], p16[6], p16[7]);
      }
    }
  };

  // OUTS == 1: v = acc + bias [+ unlrelu(res16)]; out16 = cvt16(lrelu(v)).  16 columns of one row.
  auto process_o16 = [&](const uint32_t* a0, uint32_t slot_u, int pcol, int h, const uint32_t* o16) {
    const uint32_t ba = bias_u + pcol * 4;
    float2 v2[8];
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      const float4 bq = lds128(ba + 16 * jj);
      v2[2 * jj] = add2(make_float2(__uint_as_float(a0[4 * jj]), __uint_as_float(a0[4 * jj + 1])), make_float2(bq.x, bq.y));
      v2[2 * jj + 1] = add2(make_float2(__uint_as_float(a0[4 * jj + 2]), __uint_as_float(a0[4 * jj + 3])), make_float2(bq.z, bq.w));
    }
    if constexpr (R16) {
      const float2 rinv2 = make_float2(rinv, rinv);
#pragma unroll
      for (int jj = 0; jj < 2; ++jj) {
        const float4 q = lds128(slot_u + raw_delta + o16[jj]);
        const uint32_t u[4] = {__float_as_uint(q.x), __float_as_uint(q.y), __float_as_uint(q.z), __float_as_uint(q.w)};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float2 f;
          if (BF) {
            f.x = __uint_as_float(u[e] << 16); f.y = __uint_as_float(u[e] & 0xFFFF0000u);
          } else {
            f = __half22float2(*reinterpret_cast<const __half2*>(&u[e]));
          }
          const float2 m = mul2(f, rinv2);
          v2[4 * jj + e] = add2(v2[4 * jj + e], make_float2(fminf(f.x, m.x), fminf(f.y, m.y)));
        }
      }
    }
    if constexpr (R2H) {       // + the 16-bit MRF sum so far
#pragma unroll
      for (int jj = 0; jj < 2; ++jj) {
        const float4 q = lds128(slot_u + r2h_delta + o16[jj]);
        const uint32_t u[4] = {__float_as_uint(q.x), __float_as_uint(q.y), __float_as_uint(q.z), __float_as_uint(q.w)};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float2 f;
          if (BF) {
            f.x = __uint_as_float(u[e] << 16); f.y = __uint_as_float(u[e] & 0xFFFF0000u);
          } else {
            f = __half22float2(*reinterpret_cast<const __half2*>(&u[e]));
          }
          v2[4 * jj + e] = add2(v2[4 * jj + e], f);
        }
      }
    }
    if constexpr (!PLAIN) {
      const float2 scale2 = make_float2(ep.scale, ep.scale);
#pragma unroll
      for (int i = 0; i < 8; ++i) v2[i] = mul2(v2[i], scale2);
    }
    uint32_t p16[8];
    if constexpr (OUTS == 2) {   // un-activated 16-bit result (out16r area of the slot)
#pragma unroll
      for (int i = 0; i < 8; ++i) { p16[i] = pack2<BF>(v2[i].x, v2[i].y); if (!BF) sat |= sat_probe_f16x2(p16[i]); }
      sts128u(slot_u + raw_delta + o16[0], p16[0], p16[1], p16[2], p16[3]);
      sts128u(slot_u + raw_delta + o16[1], p16[4], p16[5], p16[6], p16[7]);
    } else {
      const float2 slope2 = make_float2(slope, slope);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float2 m = mul2(v2[i], slope2);
        p16[i] = pack2<BF>(fmaxf(v2[i].x, m.x), fmaxf(v2[i].y, m.y));
        if (!BF) sat |= sat_probe_f16x2(p16[i]);
      }
      sts128u(slot_u + o16[0], p16[0], p16[1], p16[2], p16[3]);
      sts128u(slot_u + o16[1], p16[4], p16[5], p16[6], p16[7]);
    }
  };

  if constexpr (ALT) {                 // n_chunks is even (host): this warp's chunks are grp, grp + 2, ...
    const int grp = half;
    const int cw = p.cw, ccols = kEpiCols * cw;
    // wide chunks (cw == 2): the slot is [128 rows][64 x 16-bit] = 128-byte rows, SWIZZLE_128B; 32-column unit j of the
    // chunk fills cells 4j .. 4j+3 of a row
    uint32_t offw[2][NH][2];
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int h = 0; h < NH; ++h)
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) offw[j][h][jj] = p.slot16_off + row * 128 + (((j * 4 + h * 2 + jj) ^ (row & 7)) << 4);
    int it = 0, sl = grp;
    uint32_t eph = 0;
    const uint32_t taddr0 = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    uint32_t ra[32];
    bool pre = false;                  // ra already holds the in-flight load of this tile's first chunk
    for (TileIter ti(p); ti.valid(p); ti.next(p), ++it) {
      const int acc = it & acc_mask;
      const int tile_b = ti.b, tile_t0 = ti.t0(p);
      const int len_b = ep.lens ? __ldg(ep.lens + tile_b) : p.T;
      const int pcol0 = ti.n_idx * block_n;
      const uint32_t taddr = taddr0 + acc * acc_cols;
      if (!pre) {
        mbar_wait(&bars->tmem_full[acc], (it >> acc_shift) & 1);
        tc_fence_after();
        tmem_ld16_nowait(taddr + grp * ccols, ra);
        tmem_ld16_nowait(taddr + grp * ccols + 16, ra + 16);
      }
      pre = false;
      if (threadIdx.x == 128) trace_stamp(p, it, 5, clock64());
      for (int ci = grp; ci < n_chunks; ci += 2) {
        const int ms = ci >= cps ? 1 : 0;
        const int ch = ci - (ms ? cps : 0);
        const int tsub = tile_t0 + ms * 128;
        const float s = tsub + row < len_b ? ep.scale : 0.0f;
        mbar_wait(&bars->e_full[sl], eph);
        const uint32_t slot_u = smE_u + sl * slot_bytes;
        for (int j = 0; j < cw; ++j) {
          tmem_ld_wait16(ra);
          tmem_ld_wait16(ra + 16);
          if (tsub < p.T) {
            const int pc = pcol0 + ch * ccols + j * kEpiCols;
            if constexpr (OUTS != 0) {
              process_o16(ra, slot_u, pc, 0, cw == 2 ? offw[j][0] : off16[0]);
              process_o16(ra + 16, slot_u, pc + 16, 1, cw == 2 ? offw[j][NH - 1] : off16[NH - 1]);
            } else {
              process(ra, ra, slot_u, pc, tile_b, s, 0);
              process(ra + 16, ra, slot_u, pc + 16, tile_b, s, 1);
            }
          }
          // the accumulators of this warp's next 32 columns are fetched under the arithmetic / fence / barrier traffic
          if (j + 1 < cw) {
            tmem_ld16_nowait(taddr + ci * ccols + (j + 1) * kEpiCols, ra);
            tmem_ld16_nowait(taddr + ci * ccols + (j + 1) * kEpiCols + 16, ra + 16);
          } else if (ci + 2 < n_chunks) {
            tmem_ld16_nowait(taddr + (ci + 2) * ccols, ra);
            tmem_ld16_nowait(taddr + (ci + 2) * ccols + 16, ra + 16);
          } else if (p.xpre && ti.has_next(p) &&
                     mbar_test(&bars->tmem_full[(it + 1) & acc_mask], ((it + 1) >> acc_shift) & 1)) {
            tc_fence_after();
            const uint32_t tn = taddr0 + ((it + 1) & acc_mask) * acc_cols + grp * ccols;
            tmem_ld16_nowait(tn, ra);
            tmem_ld16_nowait(tn + 16, ra + 16);
            pre = true;
          }
        }
        fence_proxy_async();             // generic-proxy smem writes -> visible to the TMA store
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->s_ready[sl]);
        sl += 2;
        if (sl >= n_slots) { sl -= n_slots; eph ^= 1; }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { if (remote_te) mbar_arrive_remote(&bars->tmem_empty[acc], 0); else mbar_arrive(&bars->tmem_empty[acc]); }
      if (threadIdx.x == 128) trace_stamp(p, it, 6, clock64());
    }
    if ((sat & 0x80008000u) && p.sat_events) atomicAdd(p.sat_events, 1u);
    return;
  }

  // non-ALT: both warps of a pair work on every 32-column unit (16 columns each); with wide chunks (cw == 2, lean
  // launches only) two units share one staging slot of 128-byte rows and one wait / fence / arrive / store
  const int cwn = p.cw, cps32 = cps * cwn, n_units = p.msub * cps32;
  uint32_t offn[2][2];                 // [unit j of the wide chunk][jj]: this thread's 16-byte cells, 128-byte rows
#pragma unroll
  for (int j = 0; j < 2; ++j)
#pragma unroll
    for (int jj = 0; jj < 2; ++jj) offn[j][jj] = p.slot16_off + row * 128 + (((j * 4 + half * 2 + jj) ^ (row & 7)) << 4);
  int it = 0, sl = 0;
  uint32_t eph = 0;
  const uint32_t taddr0 = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + half * 16;
  uint32_t ra0[16], rb0[GATE ? 1 : 16], ra1[GATE ? 16 : 1];
  bool pre = false;                    // ra0 already holds the in-flight load of this tile's first chunk
  for (TileIter ti(p); ti.valid(p); ti.next(p), ++it) {
    const int acc = it & acc_mask;
    const int tile_b = ti.b, tile_t0 = ti.t0(p);
    const int len_b = ep.lens ? __ldg(ep.lens + tile_b) : p.T;
    const int pcol0 = ti.n_idx * block_n + half * 16;
    if (!pre) {
      mbar_wait(&bars->tmem_full[acc], (it >> acc_shift) & 1);
      tc_fence_after();
    }
    if (threadIdx.x == 128) trace_stamp(p, it, 5, clock64());
    const uint32_t taddr = taddr0 + acc * acc_cols;
    // chunk ci: sub-tile ms = ci / cps, column chunk ch = ci % cps (msub <= 2)
    auto finish = [&](int ci, const uint32_t* r0, const uint32_t* r1) {      // ci: 32-column unit of the tile
      const int ms = ci >= cps32 ? 1 : 0;
      const int ch = ci - (ms ? cps32 : 0);
      const int j = cwn == 2 ? (ch & 1) : 0;
      const int tsub = tile_t0 + ms * 128;
      const float s = tsub + row < len_b ? ep.scale : 0.0f;
      if (j == 0) mbar_wait(&bars->e_full[sl], eph);
      if (tsub < p.T) {
        if constexpr (OUTS != 0) process_o16(r0, smE_u + sl * slot_bytes, pcol0 + ch * kEpiCols, 0, cwn == 2 ? offn[j] : off16[0]);
        else process(r0, r1, smE_u + sl * slot_bytes, pcol0 + ch * kEpiCols, tile_b, s, 0);
      }
      if (j == cwn - 1) {
        fence_proxy_async();             // generic-proxy smem writes -> visible to the TMA store
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->s_ready[sl]);
        if (++sl == n_slots) { sl = 0; eph ^= 1; }
      }
    };
    if constexpr (GATE) {              // two accumulator halves per chunk: no register room for a prefetch
      for (int ci = 0; ci < n_chunks; ++ci) {
        const int ms = ci >= cps ? 1 : 0;
        const uint32_t col = ms * block_n + (ci - (ms ? cps : 0)) * kEpiCols;
        tmem_ld16_nowait(taddr + col, ra0);
        tmem_ld16_nowait(taddr + half_n + col, ra1);
        tmem_ld_wait16(ra0);
        tmem_ld_wait16(ra1);
        finish(ci, ra0, ra1);
      }
    } else {                           // block_n == cps * 32: chunk ci sits at TMEM column ci * 32
      if (!pre) tmem_ld16_nowait(taddr, ra0);
      pre = false;
      for (int ci = 0; ci < n_units; ci += 2) {
        tmem_ld_wait16(ra0);
        if (ci + 1 < n_units) tmem_ld16_nowait(taddr + (ci + 1) * kEpiCols, rb0);
        finish(ci, ra0, ra1);
        if (ci + 1 < n_units) {
          tmem_ld_wait16(rb0);
          if (ci + 2 < n_units) {
            tmem_ld16_nowait(taddr + (ci + 2) * kEpiCols, ra0);
          } else if (p.xpre && ti.has_next(p) &&
                     mbar_test(&bars->tmem_full[(it + 1) & acc_mask], ((it + 1) >> acc_shift) & 1)) {
            // the next tile's accumulator is already complete (the usual case when the epilogue is the slower
            // side): start fetching its first chunk under the last chunk of this tile
            tc_fence_after();
            tmem_ld16_nowait(taddr0 + ((it + 1) & acc_mask) * acc_cols, ra0);
            pre = true;
          }
          finish(ci + 1, rb0, ra1);
        }
      }
    }
    tc_fence_before();
    __syncwarp();
    if (lane == 0) { if (remote_te) mbar_arrive_remote(&bars->tmem_empty[acc], 0); else mbar_arrive(&bars->tmem_empty[acc]); }
    if (threadIdx.x == 128) trace_stamp(p, it, 6, clock64());
  }
  if ((sat & 0x80008000u) && p.sat_events) atomicAdd(p.sat_events, 1u);
}

template <int RES, bool BF>
__device__ __forceinline__ void epilogue_dispatch2(const ConvTcParams& p, TcBarriers* bars, uint32_t tmem_base,
                                                   uint32_t smE_u, uint32_t bias_s, bool plain, bool split, bool remote_te) {
  const int act = p.ep.act;
  if (p.ep.res16 || p.ep.out16r || (p.ep.res2 && p.ep.res2_is16)) {   // 16-bit residual stream: host restricts it to act NONE, non-split, RES in {0, 2}
    if (RES == 1) return;
    constexpr int R = RES == 1 ? 0 : RES;
    if (R == 0 && p.epi_alt) {       // no residual to stage (the upsampling convs): alternate-chunk epilogue
      if (plain) epilogue_loop<B200SVS_ACT_NONE, 0, true, BF, false, true, true>(p, bars, tmem_base, smE_u, bias_s, remote_te);
      else epilogue_loop<B200SVS_ACT_NONE, 0, false, BF, false, true, true>(p, bars, tmem_base, smE_u, bias_s, remote_te);
      return;
    }
    if (R == 0 && p.lean16) {   // the Generator's conv2 launches: res16 (+ 16-bit MRF sum) in, one 16-bit tensor out
      switch (p.lean16) {
        case 1: epilogue_loop<B200SVS_ACT_NONE, 0, true, BF, false, true, false, 1, false>(p, bars, tmem_base, smE_u, bias_s, remote_te); return;
        case 2: epilogue_loop<B200SVS_ACT_NONE, 0, true, BF, false, true, false, 2, false>(p, bars, tmem_base, smE_u, bias_s, remote_te); return;
        case 3: epilogue_loop<B200SVS_ACT_NONE, 0, true, BF, false, true, false, 2, true>(p, bars, tmem_base, smE_u, bias_s, remote_te); return;
        case 4: epilogue_loop<B200SVS_ACT_NONE, 0, false, BF, false, true, false, 1, true>(p, bars, tmem_base, smE_u, bias_s, remote_te); return;
        default: break;
      }
    }
    if (plain) epilogue_loop<B200SVS_ACT_NONE, R, true, BF, false, true>(p, bars, tmem_base, smE_u, bias_s, remote_te);
    else epilogue_loop<B200SVS_ACT_NONE, R, false, BF, false, true>(p, bars, tmem_base, smE_u, bias_s, remote_te);
    return;
  }
  if (split) {   // host restricts split outputs to NONE / RELU, fp16
    if (act == B200SVS_ACT_RELU) epilogue_loop<B200SVS_ACT_RELU, RES, false, false, true>(p, bars, tmem_base, smE_u, bias_s, remote_te);
    else epilogue_loop<B200SVS_ACT_NONE, RES, false, false, true>(p, bars, tmem_base, smE_u, bias_s, remote_te);
    return;
  }
  switch (act) {
    case B200SVS_ACT_GATE: epilogue_loop<B200SVS_ACT_GATE, RES, false, BF, false>(p, bars, tmem_base, smE_u, bias_s, remote_te); break;
    case B200SVS_ACT_RELU: epilogue_loop<B200SVS_ACT_RELU, RES, false, BF, false>(p, bars, tmem_base, smE_u, bias_s, remote_te); break;
    case B200SVS_ACT_GELU: epilogue_loop<B200SVS_ACT_GELU, RES, false, BF, false>(p, bars, tmem_base, smE_u, bias_s, remote_te); break;
    case B200SVS_ACT_SIGMOID: epilogue_loop<B200SVS_ACT_SIGMOID, RES, false, BF, false>(p, bars, tmem_base, smE_u, bias_s, remote_te); break;
    case B200SVS_ACT_TANH: epilogue_loop<B200SVS_ACT_TANH, RES, false, BF, false>(p, bars, tmem_base, smE_u, bias_s, remote_te); break;
    default:
      if (RES == 0 && plain && p.epi_alt && p.lean16) epilogue_loop<B200SVS_ACT_NONE, 0, true, BF, false, false, true, 1>(p, bars, tmem_base, smE_u, bias_s, remote_te);
      else if (RES == 0 && plain && p.epi_alt) epilogue_loop<B200SVS_ACT_NONE, 0, true, BF, false, false, true>(p, bars, tmem_base, smE_u, bias_s, remote_te);
      else if (plain) epilogue_loop<B200SVS_ACT_NONE, RES, true, BF, false>(p, bars, tmem_base, smE_u, bias_s, remote_te);
      else epilogue_loop<B200SVS_ACT_NONE, RES, false, BF, false>(p, bars, tmem_base, smE_u, bias_s, remote_te);
      break;
  }
}
// remote_te: this CTA is the peer of a CTA pair - its epilogue frees the accumulator on the LEADER's tmem_empty barrier
__device__ __forceinline__ void epilogue_dispatch(const ConvTcParams& p, TcBarriers* bars, uint32_t tmem_base,
                                                  uint32_t smE_u, uint32_t bias_s, bool remote_te) {
  const b200svs_epilogue& ep = p.ep;
  const int res = (ep.res2 != nullptr && !ep.res2_is16) ? 2 : (ep.res1 != nullptr ? 1 : 0);   // a 16-bit res2 rides the R16 path
  const bool plain = ep.alpha == 1.0f && ep.scale == 1.0f && ep.lens == nullptr && ep.gbias == nullptr;
  const bool split = ep.out16 && ep.out16_dtype == B200SVS_DT_F16_SPLIT;
  const bool bf = ep.out16 && ep.out16_dtype == B200SVS_DT_BF16;
  if (res == 2) {
    if (bf) epilogue_dispatch2<2, true>(p, bars, tmem_base, smE_u, bias_s, plain, split, remote_te);
    else epilogue_dispatch2<2, false>(p, bars, tmem_base, smE_u, bias_s, plain, split, remote_te);
  } else if (res == 1) {
    if (bf) epilogue_dispatch2<1, true>(p, bars, tmem_base, smE_u, bias_s, plain, split, remote_te);
    else epilogue_dispatch2<1, false>(p, bars, tmem_base, smE_u, bias_s, plain, split, remote_te);
  } else {
    if (bf) epilogue_dispatch2<0, true>(p, bars, tmem_base, smE_u, bias_s, plain, split, remote_te);
    else epilogue_dispatch2<0, false>(p, bars, tmem_base, smE_u, bias_s, plain, split, remote_te);
  }
}

// The tcgen05.mma issue loop of one CTA, run by a single thread.  KSUB = K=16 steps per 64-channel chunk
// (0: runtime p.ksub); MSUB = 128-row sub-tiles per tile.  Descriptor arithmetic is done on the 14-bit
// (address >> 4) field directly: every operand tile lies below 256 KB, so the adds never carry out of it.
template <int MSUB, int KSUB, bool WRES, bool PAIR = false>
__device__ __forceinline__ void mma_issue_loop(const ConvTcParams& p, TcBarriers* bars, uint32_t tmem_base,
                                               uint32_t smA_u, uint32_t smW_u) {
  const uint64_t descA0 = make_desc_k128(smA_u), descW0 = make_desc_k128(smW_u);
  const uint32_t a_stage_d = static_cast<uint32_t>(p.a_stage_bytes) >> 4;
  const uint32_t w_stage_d = static_cast<uint32_t>(p.w_stage_bytes) >> 4;
  const uint32_t tap_d = static_cast<uint32_t>(p.dil * kRowBytes) >> 4;
  constexpr uint32_t kSubD = (128 * kRowBytes) >> 4;       // next 128-row sub-tile of the activation stage
  const int ksub = KSUB ? KSUB : p.ksub;
  const int acc_cols = p.msub * p.block_n;
  const int acc_shift = p.acc_shift, acc_mask = (1 << p.acc_shift) - 1;
  const uint32_t idesc = p.idesc;
  int as = 0, ws = 0, it = 0;
  uint32_t aph = 0, wph = 0;
  if (WRES) {            // every weight tile of the conv is resident: one wait for the whole kernel
    mbar_wait(&bars->w_full[0], 0);
    tc_fence_after();
  }
  // The waits of tile i+1 (accumulator drained, first activation chunk landed) are peeked before the last tap
  // of tile i is issued, while the tensor pipe still has a tap's worth of MMAs queued: in a tensor-bound conv
  // both barriers completed long ago and only their latency is left, which otherwise drains the MMA queue at
  // every tile boundary.  A failed peek falls back to the blocking wait in its usual place.
  bool peeked = false;
  for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
    const int acc = it & acc_mask;
    if (!peeked) mbar_wait(&bars->tmem_empty[acc], ((it >> acc_shift) & 1) ^ 1);
    tc_fence_after();
    trace_stamp(p, it, 2, clock64());
    const uint32_t d_tmem = tmem_base + acc * acc_cols;
    uint32_t accum = 0;
    for (int ch = 0; ch < p.nchunks; ++ch) {
      if (!(peeked && ch == 0)) mbar_wait(&bars->a_full[as], aph);
      if (ch == 0) trace_stamp(p, it, 3, clock64());
      uint64_t da = descA0 + as * a_stage_d;
      const bool last_ch = ch == p.nchunks - 1;
      for (int tap = 0; tap < p.ntaps; ++tap) {
        if (!WRES) {
          mbar_wait(&bars->w_full[ws], wph);
          tc_fence_after();
        }
        if (last_ch && tap == p.ntaps - 1) {
          const int as_n = as + 1 == p.a_stages ? 0 : as + 1;
          const uint32_t aph_n = as + 1 == p.a_stages ? aph ^ 1 : aph;
          peeked = p.peek && tile + static_cast<int>(gridDim.x) < p.total_tiles &&
                   mbar_test(&bars->tmem_empty[(it + 1) & acc_mask], (((it + 1) >> acc_shift) & 1) ^ 1) && mbar_test(&bars->a_full[as_n], aph_n);
        }
        const uint64_t db = descW0 + (WRES ? static_cast<uint32_t>(ch * p.ntaps + tap) : static_cast<uint32_t>(ws)) * w_stage_d;
#pragma unroll
        for (int ms = 0; ms < MSUB; ++ms) {
          if (PAIR) {
            umma_f16_pair(d_tmem + ms * p.block_n, da + ms * kSubD, db, idesc, accum);
#pragma unroll
            for (int k = 1; k < KSUB; ++k) umma_f16_pair(d_tmem + ms * p.block_n, da + ms * kSubD + 2 * k, db + 2 * k, idesc, 1u);
          } else {
            umma_f16(d_tmem + ms * p.block_n, da + ms * kSubD, db, idesc, accum);
            if (KSUB) {
#pragma unroll
              for (int k = 1; k < KSUB; ++k) umma_f16(d_tmem + ms * p.block_n, da + ms * kSubD + 2 * k, db + 2 * k, idesc, 1u);
            } else {
              for (int k = 1; k < ksub; ++k) umma_f16(d_tmem + ms * p.block_n, da + ms * kSubD + 2 * k, db + 2 * k, idesc, 1u);
            }
          }
        }
        accum = 1;
        da += tap_d;
        if (!WRES) {
          if (PAIR) umma_commit_pair(&bars->w_empty[ws]); else umma_commit(&bars->w_empty[ws]);
          if (++ws == p.w_stages) { ws = 0; wph ^= 1; }
        }
      }
      if (PAIR) umma_commit_pair(&bars->a_empty[as]); else umma_commit(&bars->a_empty[as]);
      if (++as == p.a_stages) { as = 0; aph ^= 1; }
    }
    if (PAIR) umma_commit_pair(&bars->tmem_full[acc]); else umma_commit(&bars->tmem_full[acc]);
    trace_stamp(p, it, 4, clock64());
  }
}

template <bool PAIR>
__global__ void __launch_bounds__(kTcThreads, 1)
conv1d_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                 const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmR2,
                 const __grid_constant__ CUtensorMap tmO32,
                 const __grid_constant__ CUtensorMap tmO16, const __grid_constant__ CUtensorMap tmR16,
                 const __grid_constant__ CUtensorMap tmO16r, const ConvTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // pointer arithmetic keeps the shared address space (LDS/STS, not generic LD/ST)
  uint8_t* smA = smem;
  uint8_t* smW = smA + p.a_stages * p.a_stage_bytes;
  uint8_t* smE = smW + p.w_stages * p.w_stage_bytes;
  TcBarriers* bars = reinterpret_cast<TcBarriers*>(smE + p.n_slots * p.slot_bytes);
  float* bias_s = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 512);   // [cout_pad]

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const b200svs_epilogue& ep = p.ep;
  const bool has_res1 = ep.res1 != nullptr;
  const bool has_res2 = ep.res2 != nullptr;
  // PAIR is a separate kernel instantiation: a kernel that contains cta_group::2 instructions can only be launched
  // as a cluster of two
  constexpr bool pair = PAIR;
  uint32_t crank = 0;                                       // 0: leader (issues the MMAs of the pair)
  if constexpr (PAIR) crank = cluster_ctarank();

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmW);
    if (has_res1) prefetch_tmap(&tmR);
    if (has_res2) prefetch_tmap(&tmR2);
    if (ep.out32) prefetch_tmap(&tmO32);
    if (ep.out16) prefetch_tmap(&tmO16);
    if (ep.res16) prefetch_tmap(&tmR16);
    if (ep.out16r) prefetch_tmap(&tmO16r);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.a_stages; ++s) { mbar_init(&bars->a_full[s], 1); mbar_init(&bars->a_empty[s], 1); }
    for (int s = 0; s < (p.w_resident ? 1 : p.w_stages); ++s) { mbar_init(&bars->w_full[s], 1); mbar_init(&bars->w_empty[s], 1); }
    for (int s = 0; s < kMaxAcc; ++s) { mbar_init(&bars->tmem_full[s], 1); mbar_init(&bars->tmem_empty[s], (pair ? 2 : 1) * (kEpiThreads / 32)); }
    for (int s = 0; s < p.n_slots; ++s) {
      mbar_init(&bars->e_full[s], 1); mbar_init(&bars->e_empty[s], 1);
      mbar_init(&bars->s_ready[s], p.epi_alt ? kEpiThreads / 64 : kEpiThreads / 32);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    if constexpr (PAIR) { tmem_alloc_pair(&bars->tmem_base, p.tmem_cols); tmem_relinquish_pair(); }
    else { tmem_alloc(&bars->tmem_base, p.tmem_cols); tmem_relinquish(); }
  }
  for (int i = threadIdx.x; i < p.cout_pad; i += kTcThreads) bias_s[i] = p.ep.bias ? __ldg(p.ep.bias + i) : 0.0f;
  tc_fence_before();
  if constexpr (PAIR) cluster_sync_all();      // the peer's barriers exist before anything of this CTA signals them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;
  // Programmatic dependent launch: this CTA's setup (barriers, TMEM, bias, tensor-map prefetch, and below the
  // resident-weight loads, none of which read what the previous kernel writes) ran while the previous kernel of
  // the stream was still draining; let the next kernel's CTAs be scheduled as soon as ours exit.  Every thread
  // that reads or writes activations waits for the previous grid to complete and flush first.
  griddep_launch_dependents();

  if (warp == 0) {
    if (lane == 0) {
      griddep_wait();
      // ------------------------------------------------------------ activation-ring TMA producer (its own
      // thread: the weight ring's back-pressure must not delay the prefetch of the next tiles' activations)
      int as = 0, it = 0;
      uint32_t aph = 0;
      const uint32_t a_tx = static_cast<uint32_t>(p.a_boxes) * p.a_box_rows * kRowBytes;
      for (TileIter ti(p); ti.valid(p); ti.next(p), ++it) {
        const int tile_t0 = ti.t0(p);
        for (int ch = 0; ch < p.nchunks; ++ch) {
          mbar_wait(&bars->a_empty[as], aph ^ 1);
          if (ch == 0) trace_stamp(p, it, 0, clock64());
          uint8_t* sa = smA + as * p.a_stage_bytes;
          if constexpr (PAIR) {          // both CTAs' bytes land on the leader's barrier
            if (crank == 0) mbar_arrive_expect_tx(&bars->a_full[as], 2 * a_tx);
            for (int bx = 0; bx < p.a_boxes; ++bx)
              tma_load_3d_pair(sa + bx * p.a_box_rows * kRowBytes, &tmA, &bars->a_full[as], ch * kBlockK,
                               tile_t0 - p.pad_left + bx * p.a_box_rows, ti.b);
          } else {
            mbar_arrive_expect_tx(&bars->a_full[as], a_tx);
            for (int bx = 0; bx < p.a_boxes; ++bx)
              tma_load_3d(sa + bx * p.a_box_rows * kRowBytes, &tmA, &bars->a_full[as], ch * kBlockK,
                          tile_t0 - p.pad_left + bx * p.a_box_rows, ti.b);
          }
          if (++as == p.a_stages) { as = 0; aph ^= 1; }
        }
      }
    }
  } else if (warp == 12) {
    if (lane == 0) {
      // ------------------------------------------------------------ weight-ring TMA producer
      int ws = 0, it = 0;
      uint32_t wph = 0;
      const int wrow0 = pair ? static_cast<int>(crank) * (p.block_n / 2) : 0;      // a pair's CTA holds half of every B tile
      if (p.w_resident) {        // the whole conv's weights (one N tile) fit: loaded once, never recycled
        const int ksteps = p.nchunks * p.ntaps;
        if (!pair || crank == 0) mbar_arrive_expect_tx(&bars->w_full[0], static_cast<uint32_t>(ksteps) * p.w_stage_bytes * (pair ? 2 : 1));
        for (int ch = 0; ch < p.nchunks; ++ch)
          for (int tap = 0; tap < p.ntaps; ++tap) {
            if constexpr (PAIR) tma_load_2d_pair(smW + (ch * p.ntaps + tap) * p.w_stage_bytes, &tmW, &bars->w_full[0], ch * kBlockK, tap * p.cout_pad + wrow0);
            else tma_load_2d(smW + (ch * p.ntaps + tap) * p.w_stage_bytes, &tmW, &bars->w_full[0], ch * kBlockK, tap * p.cout_pad);
          }
      }
      for (TileIter ti(p); !p.w_resident && ti.valid(p); ti.next(p), ++it) {
        const int n_idx = ti.n_idx;
        for (int ch = 0; ch < p.nchunks; ++ch) {
          for (int tap = 0; tap < p.ntaps; ++tap) {
            mbar_wait(&bars->w_empty[ws], wph ^ 1);
            if constexpr (PAIR) {
              if (crank == 0) mbar_arrive_expect_tx(&bars->w_full[ws], 2 * p.w_stage_bytes);
              tma_load_2d_pair(smW + ws * p.w_stage_bytes, &tmW, &bars->w_full[ws], ch * kBlockK,
                               tap * p.cout_pad + n_idx * p.block_n + wrow0);
            } else {
              mbar_arrive_expect_tx(&bars->w_full[ws], p.w_stage_bytes);
              tma_load_2d(smW + ws * p.w_stage_bytes, &tmW, &bars->w_full[ws], ch * kBlockK,
                          tap * p.cout_pad + n_idx * p.block_n);
            }
            if (++ws == p.w_stages) { ws = 0; wph ^= 1; }
          }
        }
        trace_stamp(p, it, 1, clock64());
      }
    }
  } else if (warp == 1) {
    // -------------------------------------------------------------- MMA issuer: one elected thread runs the
    // whole loop (descriptors stay in uniform registers; MSUB / KSUB are compile-time so a tap is a straight
    // run of UTCHMMAs with immediate descriptor offsets)
    if constexpr (PAIR) {          // host: ksub == 4, msub == 2 (N = 128) or 1 (N = 256); only the leader issues
      if (crank == 0 && elect_one()) {
        const uint32_t smA_u = smem_u32(smA), smW_u = smem_u32(smW);
        if (p.msub == 2) {
          if (p.w_resident) mma_issue_loop<2, 4, true, true>(p, bars, tmem_base, smA_u, smW_u);
          else mma_issue_loop<2, 4, false, true>(p, bars, tmem_base, smA_u, smW_u);
        } else {
          if (p.w_resident) mma_issue_loop<1, 4, true, true>(p, bars, tmem_base, smA_u, smW_u);
          else mma_issue_loop<1, 4, false, true>(p, bars, tmem_base, smA_u, smW_u);
        }
      }
    } else if (elect_one()) {
      const uint32_t smA_u = smem_u32(smA), smW_u = smem_u32(smW);
      if (p.w_resident) {
        if (p.msub == 2) {
          if (p.ksub == 4) mma_issue_loop<2, 4, true>(p, bars, tmem_base, smA_u, smW_u);
          else if (p.ksub == 2) mma_issue_loop<2, 2, true>(p, bars, tmem_base, smA_u, smW_u);
          else mma_issue_loop<2, 0, true>(p, bars, tmem_base, smA_u, smW_u);
        } else {
          if (p.ksub == 4) mma_issue_loop<1, 4, true>(p, bars, tmem_base, smA_u, smW_u);
          else mma_issue_loop<1, 0, true>(p, bars, tmem_base, smA_u, smW_u);
        }
      } else if (p.msub == 2) {
        if (p.ksub == 4) mma_issue_loop<2, 4, false>(p, bars, tmem_base, smA_u, smW_u);
        else if (p.ksub == 2) mma_issue_loop<2, 2, false>(p, bars, tmem_base, smA_u, smW_u);
        else mma_issue_loop<2, 0, false>(p, bars, tmem_base, smA_u, smW_u);
      } else {
        if (p.ksub == 4) mma_issue_loop<1, 4, false>(p, bars, tmem_base, smA_u, smW_u);
        else mma_issue_loop<1, 0, false>(p, bars, tmem_base, smA_u, smW_u);
      }
    }
    __syncwarp();
  } else if (warp == 2) {
    if (lane == 0) {
      // ------------------------------------------------------------ TMA store thread: drains finished chunks so
      // the epilogue warps never meet at a CTA-wide barrier; keeps `stores_in_flight` bulk groups pending and
      // recycles the slot of the oldest completed one
      griddep_wait();
      int sl = 0, n_done = 0, free_sl = 0;
      uint32_t sph = 0;
      const bool split16 = ep.out16 && ep.out16_dtype == B200SVS_DT_F16_SPLIT;
      int it = 0;
      for (TileIter ti(p); ti.valid(p); ti.next(p), ++it) {
        const TileIter& tc = ti;
        for (int ms = 0; ms < p.msub; ++ms) {
          const int tsub = ti.t0(p) + ms * 128;
          for (int ch = 0; ch < p.chunks_per_sub; ++ch) {
            mbar_wait(&bars->s_ready[sl], sph);
            if (ms == 0 && ch == 0) trace_stamp(p, it, 8, clock64());
            if (tsub < p.T) {
              const int ocol = tc.n_idx * p.n_out_tile + ch * kEpiCols * p.cw;
              uint8_t* slot = smE + sl * p.slot_bytes;
              if (ep.out32) tma_store_3d(&tmO32, slot, ocol, tsub, tc.b);
              if (ep.out16) tma_store_3d(&tmO16, slot + p.slot16_off, ocol, tsub, tc.b);
              if (ep.out16r) tma_store_3d(&tmO16r, slot + p.slot_raw_off, ocol, tsub, tc.b);
              if (split16) {   // [hi | lo | hi]
                tma_store_3d(&tmO16, slot + p.slot16_off + kSlot16Bytes, ep.n_valid + ocol, tsub, tc.b);
                tma_store_3d(&tmO16, slot + p.slot16_off, 2 * ep.n_valid + ocol, tsub, tc.b);
              }
            }
            tma_store_commit();
            switch (p.stores_in_flight) {
              case 0: tma_store_wait_read<0>(); break;
              case 1: tma_store_wait_read<1>(); break;
              case 2: tma_store_wait_read<2>(); break;
              case 3: tma_store_wait_read<3>(); break;
              case 4: tma_store_wait_read<4>(); break;
              case 5: tma_store_wait_read<5>(); break;
              case 6: tma_store_wait_read<6>(); break;
              default: tma_store_wait_read<7>(); break;
            }
            if (++n_done > p.stores_in_flight) {
              mbar_arrive(&bars->e_empty[free_sl]);
              if (++free_sl == p.n_slots) free_sl = 0;
            }
            if (++sl == p.n_slots) { sl = 0; sph ^= 1; }
          }
        }
        trace_stamp(p, it, 9, clock64());
      }
      tma_store_wait_all<0>();
    }
    __syncwarp();
  } else if (warp == 3) {
    if (lane == 0) {
      // ------------------------------------------------------------ epilogue staging producer
      griddep_wait();
      int sl = 0;
      uint32_t eph = 0;
      for (TileIter ti(p); ti.valid(p); ti.next(p)) {
        const TileIter& tc = ti;
        for (int ms = 0; ms < p.msub; ++ms) {
          const int t = ti.t0(p) + ms * 128;
          for (int c = 0; c < p.chunks_per_sub; ++c) {
            mbar_wait(&bars->e_empty[sl], eph ^ 1);
            if ((has_res1 || has_res2 || ep.res16) && t < p.T) {
              const int col = tc.n_idx * p.n_out_tile + c * kEpiCols * p.cw;
              mbar_arrive_expect_tx(&bars->e_full[sl], (has_res1 ? kSlot32Bytes : 0) + (has_res2 ? (ep.res2_is16 ? kSlot16Bytes * p.cw : kSlot32Bytes) : 0) +
                                                           (ep.res16 ? kSlot16Bytes * p.cw : 0));
              if (ep.res16) tma_load_3d(smE + sl * p.slot_bytes + p.slot_raw_off, &tmR16, &bars->e_full[sl], col, t, tc.b);
              if (has_res1) tma_load_3d(smE + sl * p.slot_bytes, &tmR, &bars->e_full[sl], col, t, tc.b);
              if (has_res2) tma_load_3d(smE + sl * p.slot_bytes + p.slot_r2_off, &tmR2, &bars->e_full[sl], col, t, tc.b);
            } else {
              mbar_arrive(&bars->e_full[sl]);
            }
            if (++sl == p.n_slots) { sl = 0; eph ^= 1; }
          }
        }
      }
    }
  } else if (warp >= 4 && warp < 12) {
    griddep_wait();
    epilogue_dispatch(p, bars, tmem_base, smem_u32(smE), smem_u32(bias_s), PAIR && crank != 0);
  }

  tc_fence_before();
  if constexpr (PAIR) cluster_sync_all();      // neither CTA may exit (or free TMEM) while the pair's MMAs can still touch it
  else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    if constexpr (PAIR) tmem_dealloc_pair(tmem_base, p.tmem_cols); else tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

static int next_pow2_cols(int c) {
  int v = 32;
  while (v < c) v <<= 1;
  return v;
}
static int round_up_i(int x, int m) { return (x + m - 1) / m * m; }

}  // namespace b200svs

using namespace b200svs;

static long long* g_trace = nullptr;
static int g_trace_cap = 0;
extern "C" int b200svs_conv1d_tc_set_trace(void* dev_buf, int tiles_per_cta) {
  g_trace = static_cast<long long*>(dev_buf);
  g_trace_cap = dev_buf ? tiles_per_cta : 0;
  return 0;
}

extern "C" int b200svs_conv1d_tc(const b200svs_conv_args* a, void* stream) {
  B200SVS_REQUIRE(a && a->x && a->w, "conv1d_tc: null argument");
  B200SVS_REQUIRE(a->B > 0 && a->T > 0 && a->Cin > 0, "conv1d_tc: empty shape B=%d T=%d Cin=%d", a->B, a->T, a->Cin);
  B200SVS_REQUIRE(a->Cin_pad % kBlockK == 0 && a->Cin_pad >= a->Cin, "conv1d_tc: Cin_pad=%d must be a multiple of 64 >= Cin=%d", a->Cin_pad, a->Cin);
  B200SVS_REQUIRE(a->block_n >= 32 && a->block_n <= 256 && a->block_n % 32 == 0, "conv1d_tc: block_n=%d must be a multiple of 32 in [32,256]", a->block_n);
  B200SVS_REQUIRE(a->Cout_pad % a->block_n == 0, "conv1d_tc: Cout_pad=%d not a multiple of block_n=%d", a->Cout_pad, a->block_n);
  B200SVS_REQUIRE(a->x_ld % 8 == 0 && (reinterpret_cast<uintptr_t>(a->x) & 15) == 0, "conv1d_tc: x must be 16-B aligned with ld %% 8 == 0 (ld=%d)", a->x_ld);
  B200SVS_REQUIRE((reinterpret_cast<uintptr_t>(a->w) & 15) == 0, "conv1d_tc: w must be 16-B aligned");
  B200SVS_REQUIRE(a->ntaps >= 1 && a->dil >= 1, "conv1d_tc: ntaps=%d dil=%d", a->ntaps, a->dil);
  B200SVS_REQUIRE((a->ntaps - 1) * a->dil <= 120, "conv1d_tc: receptive field (ntaps-1)*dil=%d exceeds 120 rows", (a->ntaps - 1) * a->dil);
  const b200svs_epilogue& ep = a->ep;
  const bool gate = ep.act == B200SVS_ACT_GATE;
  B200SVS_REQUIRE(!gate || a->block_n % 64 == 0, "conv1d_tc: GATE needs block_n %% 64 == 0");
  B200SVS_REQUIRE(ep.out32 || ep.out16 || ep.out16r, "conv1d_tc: no output");
  B200SVS_REQUIRE(!ep.out32 || (ep.out32_ld % 4 == 0 && (reinterpret_cast<uintptr_t>(ep.out32) & 15) == 0), "conv1d_tc: out32 alignment");
  B200SVS_REQUIRE(!ep.out16 || (ep.out16_ld % 8 == 0 && (reinterpret_cast<uintptr_t>(ep.out16) & 15) == 0), "conv1d_tc: out16 alignment");
  B200SVS_REQUIRE(!ep.res1 || (ep.res1_ld % 4 == 0 && (reinterpret_cast<uintptr_t>(ep.res1) & 15) == 0), "conv1d_tc: res1 alignment");
  B200SVS_REQUIRE(!ep.res2 || (ep.res2_ld % (ep.res2_is16 ? 8 : 4) == 0 && (reinterpret_cast<uintptr_t>(ep.res2) & 15) == 0), "conv1d_tc: res2 alignment");
  const int n_tiles = a->Cout_pad / a->block_n;
  const int n_out_tile = gate ? a->block_n / 2 : a->block_n;
  B200SVS_REQUIRE(ep.n_valid > 0 && ep.n_valid <= n_tiles * n_out_tile, "conv1d_tc: n_valid=%d", ep.n_valid);

  ConvTcParams p;
  p.T = a->T; p.batch = a->B;
  p.ntaps = a->ntaps; p.nchunks = (a->Cin + kBlockK - 1) / kBlockK;
  p.dil = a->dil; p.pad_left = a->pad_left; p.cout_pad = a->Cout_pad; p.block_n = a->block_n;
  p.n_tiles = n_tiles; p.n_out_tile = n_out_tile;
  p.chunks_per_sub = n_out_tile / kEpiCols;
  // 256-row tiles (each weight tile feeds two MMAs) when TMEM holds 2 x 2 accumulators and the grid stays full
  p.msub = 1;
  if (4 * a->block_n <= 512) {
    const long long tiles2 = static_cast<long long>((a->T + 255) / 256) * a->B * n_tiles;
    if (tiles2 >= 2LL * num_sms()) p.msub = 2;
  }
  p.tiles_t = (a->T + 128 * p.msub - 1) / (128 * p.msub);
  p.m_tiles = p.tiles_t * a->B;
  p.total_tiles = p.m_tiles * n_tiles;
  const int ra = 128 * p.msub + (a->ntaps - 1) * a->dil;
  p.a_boxes = ra > 256 ? 2 : 1;
  p.a_box_rows = round_up_i((ra + p.a_boxes - 1) / p.a_boxes, 8);
  p.a_stage_bytes = p.a_boxes * p.a_box_rows * kRowBytes;
  {   // CTA pairs (cta_group::2) for single-N-tile N = 128 convs (the Generator's 128-channel stage): an N = 128 MMA
      // reads 4 KB of A and 4 KB of B per 64 cycles - all of an SM's shared-memory bandwidth, 79 cycles measured with
      // the weight ring's TMA writes next to it; a pair reads 4 KB + 2 KB per SM and halves the weight footprint
    // Only for tensor-bound launches (>= pair_min K steps of 64 per tile: k >= 7 at 128 channels): measured -9..-15 %
    // there, while the HBM-bound k = 3 launches lose 6-17 % to the lock step of the two CTAs.
    static const int pair_on = [] { const char* e = getenv("B200SVS_TC_PAIR"); return e ? atoi(e) : 1; }();
    static const int pair_min = [] { const char* e = getenv("B200SVS_TC_PAIR_MIN"); return e ? atoi(e) : 10; }();
    static const int pair256 = [] { const char* e = getenv("B200SVS_TC_PAIR256"); return e ? atoi(e) : 1; }();
    static const int pair_any = [] { const char* e = getenv("B200SVS_TC_PAIR_ANY"); return e ? atoi(e) : 1; }();
    static const int pair_tiles = [] { const char* e = getenv("B200SVS_TC_PAIR_TILES"); return e ? atoi(e) : 0; }();   // tests: pair small launches too
    // generator shapes: one N tile of 128 (256-row tiles) or 256; pair_any: every other wide shape whose pairs never
    // straddle two N tiles (m_tiles even) - the WN gate convs (N = 192 x 2: the weight ring re-streamed per 128-row tile
    // is what bounds them, a pair loads half of it per SM), the encoders' FFN / projection convs
    const bool gen_shape = n_tiles == 1 && ((a->block_n == 128 && p.msub == 2) || (pair256 && a->block_n == 256 && p.msub == 1));
    const bool any_shape = pair_any && a->block_n >= 128 && a->block_n % 32 == 0 && (n_tiles == 1 || p.m_tiles % 2 == 0);
    // a phantom tile (odd tile count) must not index lens / gbias
    const bool phantom_ok = p.total_tiles % 2 == 0 || (!ep.lens && !ep.gbias);
    const int min_tiles = pair_tiles > 0 ? pair_tiles : 2 * num_sms();
    p.pair = pair_on && (gen_shape || any_shape) && phantom_ok && a->Cin >= kBlockK && a->Cin % kBlockK == 0 &&
             a->ntaps * ((a->Cin + kBlockK - 1) / kBlockK) >= pair_min && p.total_tiles >= min_tiles && p.total_tiles >= 2 &&
             num_sms() % 2 == 0;
  }
  {   // 64-column epilogue chunks for the residual-free launches whose only output is the activated 16-bit tensor (the
      // Generator's conv1 / ConvTranspose launches): a chunk is a latency chain (slot wait, TMEM load, proxy fence,
      // barrier arrive, store hand-off) of ~600-680 cycles whatever its width; [128 x 64] 16-bit chunks are one
      // 128-byte-row TMA box and halve the number of chains per byte (k=3 conv1 at 64 channels: 52.6 -> 42.4 us = 6.3 TB/s).
      // Needs the alternate-chunk lean epilogue; not for CTA pairs (tensor-bound, measured +10 % there).
    static const int wide_on = [] { const char* e = getenv("B200SVS_TC_WIDE16"); return e ? atoi(e) : 1; }();
    static const int wide_res = [] { const char* e = getenv("B200SVS_TC_WIDE16_RES"); return e ? atoi(e) : 0; }();
    static const int alt_on0 = [] { const char* e = getenv("B200SVS_TC_ALT"); return e ? atoi(e) : 1; }();
    static const int lean_on0 = [] { const char* e = getenv("B200SVS_TC_LEAN"); return e ? atoi(e) : 1; }();
    const bool base0 = wide_on && lean_on0 && !p.pair && !gate && ep.act == B200SVS_ACT_NONE && ep.alpha == 1.0f && !ep.lens &&
                       !ep.gbias && !(ep.out16 && ep.out16_dtype == B200SVS_DT_F16_SPLIT) && !ep.out32 && !ep.res1 &&
                       !(ep.res2 && !ep.res2_is16) && n_out_tile % 64 == 0;
    bool wide = false;
    if (base0 && !ep.res16) {          // conv1 / ConvTranspose: alternate-chunk lean epilogue
      wide = alt_on0 && ep.scale == 1.0f && ep.out16 && !ep.out16r && !ep.res2 && (p.msub * (n_out_tile / 64)) % 2 == 0;
    } else if (base0 && wide_res) {    // conv2: the lean16 variants 1-4 (residual staged by TMA, both warps of a pair per unit).
      // Off by default: measured slower (C=64 k=7 75.5 -> 92.8 us, k=11 97.2 -> 127.0 us; k=3 unchanged) - the 32 KB
      // slots cost the resident weights / slot depth that these launches live on.
      wide = (ep.scale == 1.0f && ep.out16 && !ep.out16r && !ep.res2) || (ep.scale == 1.0f && ep.out16r && !ep.out16) ||
             (ep.res2 && ep.out16 && !ep.out16r);
    }
    p.cw = wide ? 2 : 1;
    p.chunks_per_sub = n_out_tile / (kEpiCols * p.cw);
  }
  p.w_stage_bytes = (p.pair ? a->block_n / 2 : a->block_n) * kRowBytes;
  const bool area32 = ep.res1 || ep.out32;
  const bool res2_f32 = ep.res2 && !ep.res2_is16, res2_h = ep.res2 && ep.res2_is16;
  p.slot_r2_off = res2_f32 ? (area32 ? kSlot32Bytes : 0) : 0;
  p.slot16_off = (area32 ? kSlot32Bytes : 0) + (res2_f32 ? kSlot32Bytes : 0);
  const bool split16 = ep.out16 && ep.out16_dtype == B200SVS_DT_F16_SPLIT;
  B200SVS_REQUIRE(!split16 || ep.act == B200SVS_ACT_NONE || ep.act == B200SVS_ACT_RELU, "conv1d_tc: split 16-bit output supports act NONE / RELU");
  B200SVS_REQUIRE(!split16 || ep.n_valid % kEpiCols == 0, "conv1d_tc: split 16-bit output needs n_valid %% 32 == 0 (n_valid=%d)", ep.n_valid);
  const bool raw16 = ep.res16 || ep.out16r || res2_h;
  B200SVS_REQUIRE(!raw16 || (ep.act == B200SVS_ACT_NONE && !split16 && !ep.res1),
                  "conv1d_tc: res16 / out16r / 16-bit res2 need act NONE, a non-split output and no fp32 res1");
  B200SVS_REQUIRE(!ep.res16 || (ep.res16_ld % 8 == 0 && (reinterpret_cast<uintptr_t>(ep.res16) & 15) == 0), "conv1d_tc: res16 alignment");
  B200SVS_REQUIRE(!ep.out16r || (ep.out16r_ld % 8 == 0 && (reinterpret_cast<uintptr_t>(ep.out16r) & 15) == 0), "conv1d_tc: out16r alignment");
  p.slot_raw_off = p.slot16_off + (ep.out16 ? kSlot16Bytes * p.cw : 0) + (split16 ? kSlot16Bytes : 0);
  p.slot_bytes = p.slot_raw_off + ((ep.res16 || ep.out16r) ? kSlot16Bytes * p.cw : 0);
  if (res2_h) { p.slot_r2_off = p.slot_bytes; p.slot_bytes += kSlot16Bytes * p.cw; }     // 16-bit res2 tile: its own area
  // shared memory: 2 activation stages, >= 3 weight stages (or all k-steps of a tile if fewer), then as many
  // staging slots as fit (stores in flight + residual prefetch), the rest to weight stages
  const int bias_bytes = a->Cout_pad * 4;     // staged bias (zeros when there is none)
  const int budget = max_smem_optin() - 1024 - 1024 - bias_bytes;   // alignment slack + barriers + bias
  const int ksteps = a->ntaps * p.nchunks;
  int w_min = ksteps < 3 ? ksteps : 3;
  if (w_min < 2) w_min = 2;
  // as many activation stages (loads in flight) as leave room for the minimum of everything else
  const bool has_res = ep.res1 || ep.res2 || ep.res16;
  const int min_slots = has_res ? 4 : 3;
  // Resident weights: with one N tile and taps x K-chunks x block_n x 128 B small enough (the 32- and 64-channel
  // Generator stages, k = 3 at 128 channels) the weights are loaded once per CTA instead of once per tile, and the
  // MMA thread's tap loop loses its per-tap wait and commit.
  static const int wres_on = [] { const char* e = getenv("B200SVS_TC_WRES"); return e ? atoi(e) : 1; }();
  static const int peek_on = [] { const char* e = getenv("B200SVS_TC_PEEK"); return e ? atoi(e) : 1; }();
  p.peek = peek_on;
  static const int xpre_on = [] { const char* e = getenv("B200SVS_TC_XPRE"); return e ? atoi(e) : 1; }();
  p.xpre = xpre_on;
  p.w_resident = wres_on && n_tiles == 1 &&
                 budget - ksteps * p.w_stage_bytes - 2 * p.a_stage_bytes >= min_slots * p.slot_bytes;
  if (p.w_resident) w_min = ksteps;
  int a_stages = kMaxAStages;
  if (const char* e = getenv("B200SVS_TC_ASTAGES")) { int v = atoi(e); if (v >= 2 && v <= kMaxAStages) a_stages = v; }   // tuning knob
  // a conv with a staged residual wants depth in the slot ring first (the residual of chunk c is requested when
  // slot c mod n_slots is recycled: 4 slots give it a lead of ~1 chunk, under the HBM latency): 6 slots before a
  // third / fourth activation stage (measured: C=32 k=3 conv2 99 -> 92 us, C=64 91.5 vs 93.8, C=128 97 vs 101)
  static const int slots_pref = [] { const char* e = getenv("B200SVS_TC_SLOTS"); return e ? atoi(e) : 6; }();
  const int want_slots = has_res && slots_pref > min_slots ? slots_pref : min_slots;
  while (a_stages > 2 && budget - a_stages * p.a_stage_bytes - w_min * p.w_stage_bytes < want_slots * p.slot_bytes) --a_stages;
  if (p.nchunks * ((p.total_tiles + num_sms() - 1) / num_sms()) < a_stages) a_stages = 2;
  p.a_stages = a_stages;
  const int avail = budget - a_stages * p.a_stage_bytes;
  int n_slots = (avail - w_min * p.w_stage_bytes) / p.slot_bytes;
  if (n_slots > kMaxSlots) n_slots = kMaxSlots;
  if (n_slots < 2) { n_slots = 2; }
  int w_stages = (avail - n_slots * p.slot_bytes) / p.w_stage_bytes;
  if (w_stages > kMaxWStages) w_stages = kMaxWStages;
  if (p.w_resident) w_stages = ksteps;
  B200SVS_REQUIRE(w_stages >= 2 || p.w_resident, "conv1d_tc: not enough shared memory (block_n=%d)", a->block_n);
  p.n_slots = n_slots;
  p.w_stages = w_stages;
  // Slot recycling: after committing the store of chunk i the store thread waits until at most S bulk groups
  // are still reading shared memory and frees the slot of chunk i - S.  wait_group.read only covers the smem
  // read (fast), so S = 1 is enough to keep the store thread from blocking; a small S frees slots early, which
  // gives the residual prefetch (and the epilogue's next write) a lead of n_slots - S - 1 chunks.
  p.stores_in_flight = 1;
  if (const char* e = getenv("B200SVS_TC_SIF")) { int v = atoi(e); if (v >= 0 && v <= n_slots - 2) p.stores_in_flight = v; }   // tuning knob
  // accumulator ring: 4 buffers when TMEM has room (block_n <= 64 with 256-row tiles) - the MMA of tile i+1 then
  // no longer waits for the epilogue of tile i-1, which left the epilogue idle ~25 % of a short (k = 3) tile
  static const int acc4_on = [] { const char* e = getenv("B200SVS_TC_ACC4"); return e ? atoi(e) : 1; }();
  p.acc_shift = (acc4_on && 4 * p.msub * a->block_n <= 512) ? 2 : 1;
  p.tmem_cols = next_pow2_cols((1 << p.acc_shift) * p.msub * a->block_n);
  p.ksub = a->Cin >= kBlockK ? 4 : (a->Cin + 15) / 16;
  {   // alternate-chunk epilogue for convs without a residual input (act NONE): measured -4..-7 % on the k=3 conv1
      // launches; with a staged residual it is slower (two slots held at once halve the residual prefetch lead)
    static const int alt_on = [] { const char* e = getenv("B200SVS_TC_ALT"); return e ? atoi(e) : 1; }();
    const bool plain = ep.alpha == 1.0f && ep.scale == 1.0f && ep.lens == nullptr && ep.gbias == nullptr;
    p.epi_alt = alt_on && ep.act == B200SVS_ACT_NONE && !split16 && (p.msub * p.chunks_per_sub) % 2 == 0 &&
                !has_res && (raw16 || plain);
  }
  {   // the Generator's two dominant launch types get the compile-time-specialised, packed-fp32x2 epilogue
    static const int lean_on = [] { const char* e = getenv("B200SVS_TC_LEAN"); return e ? atoi(e) : 1; }();
    const bool plain = ep.alpha == 1.0f && ep.scale == 1.0f && ep.lens == nullptr && ep.gbias == nullptr;
    const bool base = lean_on && ep.alpha == 1.0f && ep.lens == nullptr && ep.gbias == nullptr && ep.act == B200SVS_ACT_NONE &&
                      !split16 && !ep.out32 && !ep.res1 && !(ep.res2 && !ep.res2_is16);
    p.lean16 = 0;
    if (base && plain && ep.out16 && !ep.out16r && !ep.res2) p.lean16 = (ep.res16 || p.epi_alt) ? 1 : 0;
    else if (base && plain && ep.res16 && ep.out16r && !ep.out16) p.lean16 = ep.res2 ? 3 : 2;
    else if (base && ep.res16 && ep.res2 && ep.out16 && !ep.out16r) p.lean16 = 4;
  }
  {   // Zigzag: consecutive launches walk their tiles in opposite directions, so a launch starts on the rows its
      // predecessor touched last - still in the 126 MB L2 - instead of on rows evicted long ago (the Generator's
      // launches are chains conv1 -> conv2 -> conv1 over the same rows).  Tile order never changes a result.
    static const int zig_on = [] { const char* e = getenv("B200SVS_TC_ZIGZAG"); return e ? atoi(e) : 1; }();
    static unsigned launch_parity = 0;
    p.reverse = zig_on ? static_cast<int>(__atomic_fetch_add(&launch_parity, 1u, __ATOMIC_RELAXED) & 1u) : 0;
  }
  B200SVS_REQUIRE(p.cw == 1 || (p.lean16 != 0 && (p.epi_alt || ep.res16)), "conv1d_tc: internal: wide chunks without a lean epilogue");
  if (p.pair) p.reverse = 0;      // the phantom tile of an odd tile count is defined for the forward walk only
  p.idesc = make_idesc_f16(p.pair ? 256 : 128, a->block_n, a->x_dtype == B200SVS_DT_BF16);
  p.ep = ep;
  p.trace = g_trace; p.trace_cap = g_trace_cap;
  p.sat_events = sat_counter();

  const int el16 = a->x_dtype == B200SVS_DT_BF16 ? 1 : 0;
  CUtensorMap tmA, tmW, tmR, tmR2, tmO32, tmO16, tmR16, tmO16r;
  {
    uint64_t dims[3] = {(uint64_t)a->Cin, (uint64_t)a->T, (uint64_t)a->B};
    uint64_t strides[2] = {(uint64_t)a->x_ld * 2, (uint64_t)a->T * a->x_ld * 2};
    uint32_t box[3] = {(uint32_t)kBlockK, (uint32_t)p.a_box_rows, 1};
    int r = make_tmap(&tmA, el16, a->x, 3, dims, strides, box, 128);
    if (r) return r;
  }
  {
    uint64_t dims[2] = {(uint64_t)a->Cin_pad, (uint64_t)a->ntaps * a->Cout_pad};
    uint64_t strides[1] = {(uint64_t)a->Cin_pad * 2};
    uint32_t box[2] = {(uint32_t)kBlockK, (uint32_t)(p.pair ? a->block_n / 2 : a->block_n)};
    int r = make_tmap(&tmW, el16, a->w, 2, dims, strides, box, 128);
    if (r) return r;
  }
  tmR = tmA; tmR2 = tmA; tmO32 = tmA; tmO16 = tmA; tmR16 = tmA; tmO16r = tmA;
  const uint32_t ebox[3] = {(uint32_t)kEpiCols, 128, 1};
  const uint32_t hbox[3] = {(uint32_t)(kEpiCols * p.cw), 128, 1};      // 16-bit tiles: 64 columns (128-byte rows) with wide chunks
  const uint64_t edims[3] = {(uint64_t)ep.n_valid, (uint64_t)a->T, (uint64_t)a->B};
  if (ep.res1) {
    uint64_t strides[2] = {(uint64_t)ep.res1_ld * 4, (uint64_t)a->T * ep.res1_ld * 4};
    int r = make_tmap(&tmR, 2, ep.res1, 3, edims, strides, ebox, 128);
    if (r) return r;
  }
  if (ep.res2 && ep.res2_is16) {
    uint64_t strides[2] = {(uint64_t)ep.res2_ld * 2, (uint64_t)a->T * ep.res2_ld * 2};
    int r = make_tmap(&tmR2, el16, ep.res2, 3, edims, strides, hbox, p.cw == 2 ? 128 : 64);
    if (r) return r;
  } else if (ep.res2) {
    uint64_t strides[2] = {(uint64_t)ep.res2_ld * 4, (uint64_t)a->T * ep.res2_ld * 4};
    int r = make_tmap(&tmR2, 2, ep.res2, 3, edims, strides, ebox, 128);
    if (r) return r;
  }
  if (ep.out32) {
    uint64_t strides[2] = {(uint64_t)ep.out32_ld * 4, (uint64_t)a->T * ep.out32_ld * 4};
    int r = make_tmap(&tmO32, 2, ep.out32, 3, edims, strides, ebox, 128);
    if (r) return r;
  }
  if (ep.out16) {
    uint64_t strides[2] = {(uint64_t)ep.out16_ld * 2, (uint64_t)a->T * ep.out16_ld * 2};
    uint64_t odims[3] = {(uint64_t)ep.n_valid * (ep.out16_dtype == B200SVS_DT_F16_SPLIT ? 3 : 1), (uint64_t)a->T, (uint64_t)a->B};
    const uint32_t obox[3] = {(uint32_t)(kEpiCols * p.cw), 128, 1};
    int r = make_tmap(&tmO16, ep.out16_dtype == B200SVS_DT_BF16 ? 1 : 0, ep.out16, 3, odims, strides, obox, p.cw == 2 ? 128 : 64);
    if (r) return r;
  }

  if (ep.res16) {
    uint64_t strides[2] = {(uint64_t)ep.res16_ld * 2, (uint64_t)a->T * ep.res16_ld * 2};
    int r = make_tmap(&tmR16, el16, ep.res16, 3, edims, strides, hbox, p.cw == 2 ? 128 : 64);
    if (r) return r;
  }
  if (ep.out16r) {
    uint64_t strides[2] = {(uint64_t)ep.out16r_ld * 2, (uint64_t)a->T * ep.out16r_ld * 2};
    int r = make_tmap(&tmO16r, el16, ep.out16r, 3, edims, strides, hbox, p.cw == 2 ? 128 : 64);
    if (r) return r;
  }

  const size_t smem = (size_t)p.a_stages * p.a_stage_bytes + (size_t)p.w_stages * p.w_stage_bytes +
                      (size_t)p.n_slots * p.slot_bytes + 512 + bias_bytes + 1024;
  if (p.pair) B200SVS_SMEM_OPTIN(conv1d_tc_kernel<true>, "cudaFuncSetAttribute(conv1d_tc<pair>)");
  else B200SVS_SMEM_OPTIN(conv1d_tc_kernel<false>, "cudaFuncSetAttribute(conv1d_tc)");
  int grid = p.total_tiles < num_sms() ? p.total_tiles : num_sms();
  if (p.pair) grid = (grid + 1) & ~1;      // whole clusters; a CTA without a tile of its own runs the phantom
  // launched with programmatic stream serialization (PDL): the kernel may start while its predecessor in the
  // stream drains; it orders itself with griddepcontrol.wait (B200SVS_PDL=0 restores full serialization)
  static const int pdl_on = [] { const char* e = getenv("B200SVS_PDL"); return e ? atoi(e) : 1; }();
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(kTcThreads); cfg.dynamicSmemBytes = smem; cfg.stream = as_stream(stream);
  cudaLaunchAttribute attr[2];
  int n_attr = 0;
  if (pdl_on) {
    attr[n_attr].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n_attr].val.programmaticStreamSerializationAllowed = 1;
    ++n_attr;
  }
  if (p.pair) {                   // grid == num_sms (even): clusters of two CTAs on the SM pair of one TPC
    attr[n_attr].id = cudaLaunchAttributeClusterDimension;
    attr[n_attr].val.clusterDim.x = 2; attr[n_attr].val.clusterDim.y = 1; attr[n_attr].val.clusterDim.z = 1;
    ++n_attr;
  }
  cfg.attrs = attr; cfg.numAttrs = n_attr;
  const int rl = p.pair
      ? check_cuda(cudaLaunchKernelEx(&cfg, conv1d_tc_kernel<true>, tmA, tmW, tmR, tmR2, tmO32, tmO16, tmR16, tmO16r, p),
                   "conv1d_tc_kernel<pair> launch")
      : check_cuda(cudaLaunchKernelEx(&cfg, conv1d_tc_kernel<false>, tmA, tmW, tmR, tmR2, tmO32, tmO16, tmR16, tmO16r, p),
                   "conv1d_tc_kernel launch");
  if (rl == B200SVS_OK) count_launch();
  return rl;
}
