// attention_ts: the attention_tc contract (VISinger/attentions.py:148-179, :199-243; see attention_tc.cu) with the
// tensor-core A operands in TENSOR MEMORY instead of shared memory.
//
// attention_tc's split mode moves 272 KB through shared memory per 64-key tile (MMA operand reads 192 KB, TMA writes
// 48 KB, P writes 32 KB) - at the ~90 B/clk the tensor core sustains on operand fetches that is the whole tile time.
// Here
//   * Q (the stationary A operand of S = Q.K^T) is copied once per CTA from its TMA tile into TMEM by the softmax
//     threads (tcgen05.st), and S MMAs are issued in the `[d], [a_tmem], b_desc` form: only K is read from smem;
//   * P (the A operand of O += P.V) is written by tcgen05.st over the S tile it was computed from (fp32 S_j occupies
//     64 columns, fp16 P_j hi 32 + lo 32): no generic-proxy stores, no fence.proxy.async, no shared-memory P reads;
//   * the PV accumulator is double-buffered in TMEM (WIDE), so publishing P_{j+1} no longer waits for PV_j: the
//     softmax warps fold PV_{j-1} into their register accumulator AFTER handing P_j to the tensor core.
// WIDE = one CTA per SM with all 512 TMEM columns (S 2 x 64 | O 2 x 96 | Q hi 48 | Q lo 48) and two V stages;
// !WIDE = two CTAs per SM with 256 columns each (S 2 x 64 | O 96), Q stays a shared-memory operand (plain mode only).
// Roles as in attention_tc: warps 0-7 softmax (warps q and q+4 share query rows = TMEM lanes 32q.., split the 64 key
// columns and the 96 O columns), warp 8 MMA issuer, warp 9 TMA loader.
#include <type_traits>
#include "attention_common.cuh"

#ifndef B200SVS_ATTN_ACC
#define B200SVS_ATTN_ACC 1
#endif

namespace b200svs {

constexpr int kRelKBytes = 2 * 2 * 16 * 128;     // 8 KB

struct __align__(8) AttnTsBars {
  uint64_t qbar, q_tmem, band_full;
  uint64_t k_full[2], v_full[2];     // stage j & 1: K_j / V_j landed
  uint64_t k_empty[2], v_empty[2];   // S_j / PV_j complete: the stage's K / V buffer may be overwritten
  uint64_t s_full[2];
  uint64_t p_full[2];                // P_j stored in TMEM (count = softmax warps)
  // PV_j complete: barrier j & 1, phase j >> 1.  TWO barriers although ACC has one accumulator: with O in tensor memory the
  // softmax warps no longer wait for every PV, and a parity wait on a single barrier cannot tell phase j from phase j - 2 -
  // the final wait returned while PV_{n-2} and PV_{n-1} were still running whenever the last tile was a quick one (found by
  // tools/fuzz_attn.py at T % 128 == 0).  S_j complete implies PV_{j-2} complete, so per parity a waiter is never a phase behind.
  uint64_t o_full[2];
  uint32_t tmem_base;
};

// PARTS = softmax warps per TMEM lane quadrant (2 or 4): a thread owns 64 / PARTS key columns and 96 / PARTS O columns
template <bool SPLIT, bool WIDE, int PARTS>
__global__ void __launch_bounds__(PARTS * 128 + 64, WIDE ? 1 : 2)
attention_ts_kernel(const __grid_constant__ CUtensorMap tmQh, const __grid_constant__ CUtensorMap tmKh,
                    const __grid_constant__ CUtensorMap tmVh, const __grid_constant__ CUtensorMap tmQl,
                    const __grid_constant__ CUtensorMap tmKl, const __grid_constant__ CUtensorMap tmVl,
                    const AttnTcParams p) {
  static_assert(WIDE || !SPLIT, "split mode needs the 512-column layout");
  static_assert(PARTS == 2 || (PARTS == 4 && WIDE), "4 parts only with one CTA per SM (registers)");
  constexpr int kSW = 4 * PARTS;                 // softmax warps
  constexpr int kThreads = kSW * 32 + 64;
  constexpr int kMmaW = kSW, kTmaW = kSW + 1;
  // ACC: O accumulates in TENSOR MEMORY over all key tiles (every PV MMA after the first adds in place) instead of being
  // folded into registers tile by tile.  The row maximum is then LAZY: it only moves when a tile's maximum exceeds it by
  // more than kLazy (2^8: P <= 256, far inside fp16 / fp32-accumulate range), and only then the warp rescales its rows of O
  // in TMEM (ld, multiply, st) before it publishes P_j - after the first tiles that is practically never, so the per-tile
  // chain loses the PV wait and the fold (310 of ~1550 cycles) and the loop keeps 48 registers less.
  constexpr bool ACC = WIDE && (B200SVS_ATTN_ACC != 0);
  constexpr float kLazy = 8.0f;
  constexpr int KC = kTK / PARTS;                // key columns per thread
  constexpr int kOC = kHD / PARTS;               // O columns per thread
  constexpr int NP = SPLIT ? 2 : 1;
  constexpr int VST = 2;
  constexpr int kKStage = NP * 2 * kKChunkBytes;
  constexpr uint32_t kTmemCols = WIDE ? 512 : 256;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // pointer arithmetic keeps the shared address space (LDS/STS, not generic LD/ST)
  uint8_t* smQ = smem;                                   // [NP][2 chunks][16 KB]
  uint8_t* smK = smQ + NP * 2 * kQChunkBytes;            // [2 stages][NP][2 chunks][8 KB]
  uint8_t* smV = smK + 2 * kKStage;                      // [VST][NP][12 KB]
  uint8_t* smRK = smV + VST * NP * kVBytes;              // rel_k as an MMA B operand: [hi | lo][2 chunks][16 rows x 128 B]
  float* qrel = reinterpret_cast<float*>(smRK + kRelKBytes);    // [128][9]  (q/sqrt(d)).rel_k * log2(e)
  float* band = qrel + kTQ * kNR;                               // [128][9]  band scores (log2 domain)
  float* relv_s = band + kTQ * kNR;                             // [9][96]
  float* pmax = relv_s + kNR * kHD;                             // [2 S buffers][PARTS][128] partial row maxima
  float* lsum = pmax + 2 * PARTS * kTQ;                         // [PARTS][128] partial row sums
  AttnTsBars* bars = reinterpret_cast<AttnTsBars*>(lsum + PARTS * kTQ);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q0 = blockIdx.x * kTQ, h = blockIdx.y, b = blockIdx.z;
  const int len = p.lens ? min(__ldg(p.lens + b), p.T) : p.T;
  const int dt = SPLIT ? B200SVS_DT_F16 : p.dtype;
  const int oc = h * kHD;

  if (q0 >= len) {   // whole tile is padding: zeros (the reference's values there are masked by every consumer)
    const int t = q0 + tid;
    const size_t orow = static_cast<size_t>(b) * p.T + t;
    if (tid < kTQ && t < p.T) {
      for (int c = 0; c < kHD; ++c) {
        if (p.out32) p.out32[orow * p.out32_ld + oc + c] = 0.f;
        if (p.out16) {
          uint16_t* o = reinterpret_cast<uint16_t*>(p.out16) + orow * p.out16_ld + oc + c;
          o[0] = 0;
          if (p.out16_split) { o[p.C] = 0; o[2 * p.C] = 0; }
        }
      }
    }
    return;
  }

  if (warp == kTmaW && lane == 0) {      // the loader thread: Q, K_0 and K_1 are in flight before the CTA-wide sync below
    mbar_init(&bars->qbar, 1);
    mbar_init(&bars->q_tmem, kSW);
    mbar_init(&bars->band_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->k_full[i], 1); mbar_init(&bars->v_full[i], 1);
      mbar_init(&bars->k_empty[i], 1); mbar_init(&bars->v_empty[i], 1);
      mbar_init(&bars->s_full[i], 1); mbar_init(&bars->p_full[i], kSW);
      mbar_init(&bars->o_full[i], 1);
    }
    fence_barrier_init();
    const int nkt0 = (len + kTK - 1) / kTK;
    mbar_arrive_expect_tx(&bars->qbar, NP * 2 * kQChunkBytes);
    for (int c = 0; c < 2; ++c) {
      tma_load_3d(smQ + c * kQChunkBytes, &tmQh, &bars->qbar, oc + c * 64, q0, b);
      if (SPLIT) tma_load_3d(smQ + (2 + c) * kQChunkBytes, &tmQl, &bars->qbar, oc + c * 64, q0, b);
    }
    for (int j = 0; j < 2 && j < nkt0; ++j) {
      mbar_arrive_expect_tx(&bars->k_full[j], kKStage);
      uint8_t* dst = smK + j * kKStage;
      for (int c = 0; c < 2; ++c) {
        tma_load_3d(dst + c * kKChunkBytes, &tmKh, &bars->k_full[j], p.C + oc + c * 64, j * kTK, b);
        if (SPLIT) tma_load_3d(dst + (2 + c) * kKChunkBytes, &tmKl, &bars->k_full[j], p.C + oc + c * 64, j * kTK, b);
      }
    }
    prefetch_tmap(&tmVh);
  }
  if (warp == kMmaW) {
    tmem_alloc(&bars->tmem_base, kTmemCols);
    tmem_relinquish();
  }
  for (int i = tid; i < kNR * kHD; i += kThreads) relv_s[i] = __ldg(p.rel_v + i);
  // rel_k [9][96] fp32 -> 16-bit (hi, lo) rows 0..8 of a 16-row K-major SWIZZLE_128B operand tile (rows 9..15 zero): the
  // band dot products q . rel_k[w] are ONE N = 16 accumulator of MMAs (q.hi + q.lo [+ q_lo.hi]) instead of 430 FMAs per thread
  for (int i = tid; i < 16 * kHD; i += kThreads) {
    const int w = i / kHD, d = i - w * kHD, dd = d & 63;
    const float v = w < kNR ? __ldg(p.rel_k + w * kHD + d) : 0.f;
    const float vh = rt16(v, dt);
    const uint32_t off = (d >> 6) * 2048 + (w >> 3) * 1024 + (w & 7) * 128 + (((dd >> 3) ^ (w & 7)) << 4) + (dd & 7) * 2;
    *reinterpret_cast<uint16_t*>(smRK + off) = static_cast<uint16_t>(pk16(vh, 0.f, dt) & 0xffffu);
    *reinterpret_cast<uint16_t*>(smRK + 4096 + off) = static_cast<uint16_t>(pk16(v - vh, 0.f, dt) & 0xffffu);
  }
  fence_proxy_async();               // generic-proxy writes of the rel_k tile -> visible to the tensor core
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_s = bars->tmem_base;          // S : 2 x 64 columns (P_j overwrites S_j: hi 32 columns, lo 32)
  const uint32_t tmem_o = tmem_s + 128;             // PV: 96 columns, second buffer at +128 (WIDE)
  const uint32_t tmem_q = tmem_s + 384;             // Q hi 48 columns, Q lo at +64 (WIDE)
  const uint32_t tmem_band = tmem_s + (WIDE ? 496 : 224);     // band dot products: 16 columns
  const int n_kt = (len + kTK - 1) / kTK;
#ifdef B200SVS_ATTN_TRACE      // clock64 stamps of CTA (1,0,0) for tools/trace_attn.py (a -D build: the stamps cost ~5 %)
  const bool tr_cta = p.trace && blockIdx.x == 1 && blockIdx.y == 0 && blockIdx.z == 0;
  auto stamp = [&](int j, int slot) {
    if (tr_cta && j < p.trace_cap) p.trace[j * 16 + slot] = clock64();
  };
#else
  auto stamp = [](int, int) {};
#endif
  if (tid == 0) stamp(0, 14);      // setup done

  if (warp == kTmaW) {
    if (lane == 0) {
      // ------------------------------------------------------------ TMA loader (Q, K_0, K_1 went out at setup): V_j, K_{j+2}
      auto load_k = [&](int j, uint64_t* bar) {
        uint8_t* dst = smK + (j & 1) * kKStage;
        for (int c = 0; c < 2; ++c) {
          tma_load_3d(dst + c * kKChunkBytes, &tmKh, bar, p.C + oc + c * 64, j * kTK, b);
          if (SPLIT) tma_load_3d(dst + (2 + c) * kKChunkBytes, &tmKl, bar, p.C + oc + c * 64, j * kTK, b);
        }
      };
      for (int j = 0; j < n_kt; ++j) {
        const int st = j & 1;
        attn_wait(&bars->v_empty[st], ((j >> 1) & 1) ^ 1);          // PV_{j-2} complete (free on a fresh barrier)
        uint8_t* dst = smV + st * NP * kVBytes;
        mbar_arrive_expect_tx(&bars->v_full[st], NP * kVBytes);
        tma_load_3d(dst, &tmVh, &bars->v_full[st], j * kTK, oc, b);
        if (SPLIT) tma_load_3d(dst + kVBytes, &tmVl, &bars->v_full[st], j * kTK, oc, b);
        if (j + 2 < n_kt) {
          attn_wait(&bars->k_empty[st], (j >> 1) & 1);              // S_j complete
          mbar_arrive_expect_tx(&bars->k_full[st], kKStage);
          load_k(j + 2, &bars->k_full[st]);
        }
      }
    }
  } else if (warp == kMmaW) {
    if (elect_one()) {
      // ------------------------------------------------------------ MMA issuer
      const uint32_t idesc_s = make_idesc_f16(128, kTK, dt == B200SVS_DT_BF16);
      const uint32_t idesc_o = make_idesc_f16(128, kHD, dt == B200SVS_DT_BF16);
      const uint32_t q_u = smem_u32(smQ), k_u = smem_u32(smK), v_u = smem_u32(smV);
      // S_j = Q . K_j^T (d = 96 = 64 + 32).  It overwrites S_{j-2} / P_{j-2}: the PV_{j-2} MMAs that read them precede it in
      // issue order, and the softmax warps read S_{j-2} before they published P_{j-2}.
      auto issue_s = [&](int j, int c) {        // K chunk c of S_j: c = 0 is d 0..63 (4 K steps), c = 1 is d 64..95 (2)
        const int st = j & 1;
#pragma unroll
        for (int pr = 0; pr < (SPLIT ? 3 : 1); ++pr) {
          const int pa = pr == 2 ? 1 : 0, pb = pr == 1 ? 1 : 0;       // (hi,hi) (hi,lo) (lo,hi)
          const uint64_t da = make_desc_k128(q_u + (pa * 2 + c) * kQChunkBytes);
          const uint64_t db = make_desc_k128(k_u + st * kKStage + (pb * 2 + c) * kKChunkBytes);
#pragma unroll
          for (int k = 0; k < (c == 0 ? 4 : 2); ++k) {
            const uint32_t accum = (c | pr | k) ? 1u : 0u;
            if (WIDE) umma_f16_ts(tmem_s + st * kTK, tmem_q + pa * 64 + (c * 4 + k) * 8, db + 2 * k, idesc_s, accum);
            else umma_f16(tmem_s + st * kTK, da + 2 * k, db + 2 * k, idesc_s, accum);
          }
        }
        if (c == 1) {
          umma_commit(&bars->s_full[st]);
          umma_commit(&bars->k_empty[st]);
        }
      };
      if (WIDE) attn_wait(&bars->q_tmem, 0); else attn_wait(&bars->qbar, 0);
      tc_fence_after();
      {   // band dot products  D[128 x 16] = Q . rel_k^T  (rel_k always as hi + lo: fp32-class, like the FMAs it replaces)
        const uint32_t idesc_b = make_idesc_f16(128, 16, dt == B200SVS_DT_BF16);
        const uint32_t rk_u = smem_u32(smRK);
#pragma unroll
        for (int c = 0; c < 2; ++c) {
#pragma unroll
          for (int pr = 0; pr < (SPLIT ? 3 : 2); ++pr) {
            const int pa = pr == 2 ? 1 : 0, pb = pr == 1 ? 1 : 0;       // (hi,hi) (hi,lo) [(lo,hi)]
            const uint64_t da = make_desc_k128(q_u + (pa * 2 + c) * kQChunkBytes);
            const uint64_t db = make_desc_k128(rk_u + pb * 4096 + c * 2048);
#pragma unroll
            for (int k = 0; k < (c == 0 ? 4 : 2); ++k) {
              const uint32_t accum = (c | pr | k) ? 1u : 0u;
              if (WIDE) umma_f16_ts(tmem_band, tmem_q + pa * 64 + (c * 4 + k) * 8, db + 2 * k, idesc_b, accum);
              else umma_f16(tmem_band, da + 2 * k, db + 2 * k, idesc_b, accum);
            }
          }
        }
        umma_commit(&bars->band_full);
      }
      for (int j = 0; j < 2 && j < n_kt; ++j) {
        attn_wait(&bars->k_full[j], 0);
        tc_fence_after();
        issue_s(j, 0);
        issue_s(j, 1);
      }
      attn_wait(&bars->v_full[0], 0);
      // Steady state: the only barrier between two bursts of MMAs is P_j.  A barrier check costs this thread ~100-200
      // cycles even when the phase is complete, so K_{j+2} and V_{j+1} (landed a tile ago) are checked between PV_j
      // and the two halves of S_{j+2}, where MMAs already queued cover the checks.  PV_j's accumulator is free because the
      // softmax warps fold PV_{j-2} (WIDE; PV_{j-1} otherwise) before they publish P_j.
      for (int j = 0; j < n_kt; ++j) {
        const int sb = j & 1, ob = (WIDE && !ACC) ? sb : 0;
        stamp(j, 10);
        attn_wait(&bars->p_full[sb], (j >> 1) & 1);
        tc_fence_after();
        stamp(j, 11);
        uint32_t accum = (ACC && j > 0) ? 1u : 0u;
#pragma unroll
        for (int pr = 0; pr < (SPLIT ? 3 : 1); ++pr) {
          const int pa = pr == 2 ? 1 : 0, pb = pr == 1 ? 1 : 0;
          const uint64_t db = make_desc_k128(v_u + (sb * NP + pb) * kVBytes);
#pragma unroll
          for (int k = 0; k < kTK / 16; ++k) {
            umma_f16_ts(tmem_o + ob * 128, tmem_s + sb * kTK + pa * 32 + k * 8, db + 2 * k, idesc_o, accum);
            accum = 1;
          }
        }
        umma_commit(&bars->o_full[sb]);
        umma_commit(&bars->v_empty[sb]);
        stamp(j, 12);
        if (j + 2 < n_kt) {
          attn_wait(&bars->k_full[sb], ((j + 2) >> 1) & 1);
          tc_fence_after();
          issue_s(j + 2, 0);
        }
        if (j + 1 < n_kt) attn_wait(&bars->v_full[sb ^ 1], ((j + 1) >> 1) & 1);
        if (j + 2 < n_kt) issue_s(j + 2, 1);
        stamp(j, 9);
      }
    }
    __syncwarp();
  } else {
    // ---------------------------------------------------------------- softmax warps (PARTS per lane quadrant)
    // Thread (row, part): query row `row` = TMEM lane, key columns [KC*part, KC*part+KC) of every S tile and O columns
    // [kOC*part, +kOC).  Warps q, q+4, .. own the same rows; they exchange the row maximum once per tile through
    // shared memory (named barrier 1+q, 32*PARTS threads) and the row sums once at the end.
    const int q = warp & 3, part = warp >> 2;
    const int row = q * 32 + lane;
    const int t = q0 + row;
    const float scale2 = rsqrtf(static_cast<float>(kHD)) * kLog2e;     // scores live in the log2 domain
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    {   // Q row -> TMEM (WIDE): part c stores cells [12/PARTS * c, ..) - a cell = 8 channels = 4 columns; then the band
        // dot products (q_i / sqrt(d)) . rel_k[w] come back from the tensor core (16 columns, 9 used)
      const int w0 = (kNR * part + PARTS - 1) / PARTS, w1 = (kNR * (part + 1) + PARTS - 1) / PARTS;
      if (WIDE) {
        attn_wait(&bars->qbar, 0);
        if (tid == 0) stamp(1, 14);      // Q landed
#pragma unroll
        for (int i = 0; i < 12 / PARTS; ++i) {      // chunk 0 holds d 0..63 (cells 0..7), chunk 1 d 64..95
          const int cell = part * (12 / PARTS) + i;
          const int c = cell >> 3, jc = cell & 7;
          const uint8_t* src = smQ + c * kQChunkBytes + row * 128 + ((jc ^ (row & 7)) << 4);
          const uint4 qh = *reinterpret_cast<const uint4*>(src);
          const uint32_t uh[4] = {qh.x, qh.y, qh.z, qh.w};
          tmem_st4(tmem_q + lane_off + cell * 4, uh);
          if (SPLIT) {
            const uint4 ql = *reinterpret_cast<const uint4*>(src + 2 * kQChunkBytes);
            const uint32_t ul[4] = {ql.x, ql.y, ql.z, ql.w};
            tmem_st4(tmem_q + 64 + lane_off + cell * 4, ul);
          }
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->q_tmem);
      }
      attn_wait(&bars->band_full, 0);
      tc_fence_after();
      if (tid == 0) stamp(1, 15);        // band dot products done
      uint32_t bv[16];
      tmem_ld16_nowait(tmem_band + lane_off, bv);
      tmem_ld_wait16(bv);
#pragma unroll
      for (int w = 0; w < kNR; ++w) {
        if (w >= w0 && w < w1) {
          qrel[row * kNR + w] = (t < p.T) ? __uint_as_float(bv[w]) * scale2 : 0.f;
          band[row * kNR + w] = -INFINITY;
        }
      }
      named_bar_sync(1 + q, 32 * PARTS);         // every part's band offsets are in place
    }
    float m_run = -INFINITY, l_run = 0.f, alpha_prev = 0.f;    // l_run: partial row sum over this thread's key columns
    float o[kOC];
#pragma unroll
    for (int c = 0; c < kOC; ++c) o[c] = 0.f;
    auto fold = [&](int ob, float alpha) {       // O = O * alpha + PV  (this thread's 48 of the 96 PV columns)
      const uint32_t ta = tmem_o + ob * 128 + lane_off + part * kOC;
      uint32_t r0[16];                           // all loads in flight before the first use
      tmem_ld16_nowait(ta, r0);
      if constexpr (PARTS == 2) {
        uint32_t r1[16], r2[16];
        tmem_ld16_nowait(ta + 16, r1);
        tmem_ld16_nowait(ta + 32, r2);
        tmem_ld_wait16(r0);
#pragma unroll
        for (int e = 0; e < 16; ++e) o[e] = fmaf(o[e], alpha, __uint_as_float(r0[e]));
        tmem_ld_wait16(r1);
#pragma unroll
        for (int e = 0; e < 16; ++e) o[16 + e] = fmaf(o[16 + e], alpha, __uint_as_float(r1[e]));
        tmem_ld_wait16(r2);
#pragma unroll
        for (int e = 0; e < 16; ++e) o[32 + e] = fmaf(o[32 + e], alpha, __uint_as_float(r2[e]));
      } else {
        uint32_t r1[8];
        tmem_ld8_nowait(ta + 16, r1);
        tmem_ld_wait16(r0);
#pragma unroll
        for (int e = 0; e < 16; ++e) o[e] = fmaf(o[e], alpha, __uint_as_float(r0[e]));
        tmem_ld_wait8(r1);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[16 + e] = fmaf(o[16 + e], alpha, __uint_as_float(r1[e]));
      }
    };
    auto o_wait = [&](int i) {                   // PV_i complete
      attn_wait(&bars->o_full[i & 1], (i >> 1) & 1);
      tc_fence_after();
    };
    auto rescale = [&](float alpha) {            // ACC: this thread's 48 (24) columns of its row of O, in place in TMEM
      const uint32_t ta = tmem_o + lane_off + part * kOC;
      uint32_t r0[16];
      tmem_ld16_nowait(ta, r0);
      if constexpr (PARTS == 2) {
        uint32_t r1[16], r2[16];
        tmem_ld16_nowait(ta + 16, r1);
        tmem_ld16_nowait(ta + 32, r2);
        tmem_ld_wait16(r0);
#pragma unroll
        for (int e = 0; e < 16; ++e) r0[e] = __float_as_uint(__uint_as_float(r0[e]) * alpha);
        tmem_st16(ta, r0);
        tmem_ld_wait16(r1);
#pragma unroll
        for (int e = 0; e < 16; ++e) r1[e] = __float_as_uint(__uint_as_float(r1[e]) * alpha);
        tmem_st16(ta + 16, r1);
        tmem_ld_wait16(r2);
#pragma unroll
        for (int e = 0; e < 16; ++e) r2[e] = __float_as_uint(__uint_as_float(r2[e]) * alpha);
        tmem_st16(ta + 32, r2);
      } else {
        uint32_t r1[8];
        tmem_ld8_nowait(ta + 16, r1);
        tmem_ld_wait16(r0);
#pragma unroll
        for (int e = 0; e < 16; ++e) r0[e] = __float_as_uint(__uint_as_float(r0[e]) * alpha);
        tmem_st16(ta, r0);
        tmem_ld_wait8(r1);
#pragma unroll
        for (int e = 0; e < 8; ++e) r1[e] = __float_as_uint(__uint_as_float(r1[e]) * alpha);
        tmem_st8(ta + 16, r1);
      }
    };

    for (int j = 0; j < n_kt; ++j) {
      const int k0 = j * kTK;
      const int sb = j & 1;
      const bool tracer = tid == 0;
      if (tracer) stamp(j, 0);
      attn_wait(&bars->s_full[sb], (j >> 1) & 1);
      tc_fence_after();
      if (tracer) stamp(j, 1);
      const uint32_t ts = tmem_s + sb * kTK + lane_off;                    // S_j / P_j, this thread's lane
      const int kb = k0 + part * KC;                                       // first key of this thread's columns
      const bool near_diag = (k0 <= q0 + kTQ - 1 + kWin) && (k0 + kTK - 1 >= q0 - kWin);
      const bool tail = k0 + kTK > len;
      const bool slow = near_diag || tail;             // CTA-uniform
      // score of key kb + e with the band bias and the masks applied (slow tiles only)
      auto adjust = [&](uint32_t raw, int e) -> float {
        float x = __uint_as_float(raw) * scale2;
        const int key = kb + e;
        const int d = key - t + kWin;
        if (d >= 0 && d < kNR) x += qrel[row * kNR + d];
        if (key >= len) x = -1e4f * kLog2e;            // masked_fill(mask == 0, -1e4)
        if (key >= p.T) x = -INFINITY;                 // beyond the tensor: not a key at all
        return x;
      };
      // ---- S_j: this thread's 32 scores are read from TMEM once and stay in registers for both passes
      uint32_t sv[KC];
      if constexpr (KC == 32) {
        tmem_ld32_nowait(ts + part * KC, sv);
        tmem_ld_wait32(sv);
      } else {
        tmem_ld16_nowait(ts + part * KC, sv);
        tmem_ld_wait16(sv);
      }
      if (tracer) stamp(j, 2);
      // ---- pass 1: row maximum over this thread's keys, then over the quadrant's warps.  Slow tiles keep the adjusted score.
      float mx = -INFINITY;
      if (slow) {
#pragma unroll
        for (int e = 0; e < KC; ++e) {
          const float x = adjust(sv[e], e);
          sv[e] = __float_as_uint(x);
          mx = fmaxf(mx, x);
        }
      } else {
#pragma unroll
        for (int e = 0; e < KC; ++e) mx = fmaxf(mx, __uint_as_float(sv[e]));
      }
      if (!slow) mx *= scale2;                         // scale2 > 0 commutes with max
      pmax[(sb * PARTS + part) * kTQ + row] = mx;
      named_bar_sync(1 + q, 32 * PARTS);               // also: ALL warps of the quadrant have read S_j - P_j may overwrite it
#pragma unroll
      for (int c = 0; c < PARTS; ++c) mx = fmaxf(mx, pmax[(sb * PARTS + c) * kTQ + row]);
      // ACC: lazy maximum - keep m_run unless this tile exceeds it by more than kLazy (identical in every warp of the
      // quadrant: mx is the exchanged row maximum)
      const float m_new = ACC ? ((m_run == -INFINITY || mx > m_run + kLazy) ? mx : m_run) : fmaxf(m_run, mx);
      const float alpha = (m_run == -INFINITY) ? 0.f : ex2f(m_run - m_new);
      if (ACC && j > 0 && __any_sync(0xffffffffu, m_new != m_run)) {      // rare: rescale this warp's rows of O in place
        o_wait(j - 1);                   // no PV MMA in flight on O: PV_{j-1} complete, PV_j not issued before P_j is published
        rescale(alpha);                  // alpha == 1 exactly for the rows whose maximum stays
      }
      const float a_fold = alpha_prev;                 // rescale factor that belongs to PV_{j-1}
      if (tracer) stamp(j, 3);
      // ---- one accumulator (!WIDE): PV_{j-1} has to be folded into O before PV_j may be issued
      if (!WIDE && j > 0) {
        o_wait(j - 1);
        fold(0, a_fold);
      }
      // ---- pass 2: P = 2^(S - m), partial row sum; P goes to TMEM over S_j as the A operand of the PV MMA
      float sum = 0.f;
      uint32_t ph[KC / 2], pl[SPLIT ? KC / 2 : 1];
      auto emit = [&](const uint32_t* r, int g, auto slow_c) {      // 8 keys -> 4 packed columns
        constexpr bool kSlow = decltype(slow_c)::value;
        float pv[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          float x;
          if constexpr (kSlow) {
            const int ee = g * 8 + e;
            x = __uint_as_float(r[e]);                    // adjusted in pass 1
            const int d = kb + ee - t + kWin;
            if (near_diag && d >= 0 && d < kNR && kb + ee < p.T) band[row * kNR + d] = x;
            x -= m_new;
          } else {
            x = fmaf(__uint_as_float(r[e]), scale2, -m_new);
          }
          pv[e] = ex2f(x);                              // ex2(-inf) = 0
          sum += pv[e];
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) ph[4 * g + e] = pk16(pv[2 * e], pv[2 * e + 1], dt);
        if (SPLIT) {
#pragma unroll
          for (int e = 0; e < 8; ++e) pv[e] -= rt16(pv[e], dt);
#pragma unroll
          for (int e = 0; e < 4; ++e) pl[4 * g + e] = pk16(pv[2 * e], pv[2 * e + 1], dt);
        }
      };
      if (slow) {
#pragma unroll
        for (int g = 0; g < KC / 8; ++g) emit(sv + 8 * g, g, std::true_type{});
      } else {
#pragma unroll
        for (int g = 0; g < KC / 8; ++g) emit(sv + 8 * g, g, std::false_type{});
      }
      if (tracer) stamp(j, 4);
      if constexpr (KC == 32) {
        tmem_st16(ts + part * 16, ph);
        if (SPLIT) tmem_st16(ts + 32 + part * 16, pl);
      } else {
        tmem_st8(ts + part * 8, ph);
        if (SPLIT) tmem_st8(ts + 32 + part * 8, pl);
      }
      l_run = l_run * alpha + sum;
      m_run = m_new;
      alpha_prev = alpha;
      tmem_st_wait();
      tc_fence_before();                 // P_j stored (and, !WIDE, PV_{j-1} read out)
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->p_full[sb]);
      if (tracer) stamp(j, 5);
      // ---- two accumulators (WIDE): fold PV_{j-1} now, off the path that feeds the tensor core
      if (WIDE && !ACC && j > 0) {
        o_wait(j - 1);
        if (tracer) stamp(j, 6);
        fold((j - 1) & 1, a_fold);
        if (tracer) stamp(j, 7);
      }
    }
    o_wait(n_kt - 1);
    if (ACC) fold(0, 0.f);               // o = the accumulated O (o starts at 0: 0 * 0 + O)
    else fold(WIDE ? ((n_kt - 1) & 1) : 0, alpha_prev);
    if (tid == 0) stamp(0, 13);    // main loop done

    // ---- row sum over both halves, relative-value band term, normalisation, stores
    lsum[part * kTQ + row] = l_run;
    named_bar_sync(1 + q, 32 * PARTS);
    float l_tot = 0.f;
#pragma unroll
    for (int c = 0; c < PARTS; ++c) l_tot += lsum[c * kTQ + row];
    const bool live = t < len;
    if (t < p.T) {
      if (live) {
#pragma unroll 3
        for (int w = 0; w < kNR; ++w) {
          const float sv = band[row * kNR + w];
          if (sv == -INFINITY) continue;
          const float pw = ex2f(sv - m_run);
#pragma unroll
          for (int c = 0; c < kOC; c += 4) {
            const float4 rv = *reinterpret_cast<const float4*>(relv_s + w * kHD + part * kOC + c);
            o[c] = fmaf(pw, rv.x, o[c]); o[c + 1] = fmaf(pw, rv.y, o[c + 1]);
            o[c + 2] = fmaf(pw, rv.z, o[c + 2]); o[c + 3] = fmaf(pw, rv.w, o[c + 3]);
          }
        }
      }
      const float invl = live ? 1.0f / l_tot : 0.f;
#pragma unroll
      for (int c = 0; c < kOC; ++c) o[c] = live ? o[c] * invl : 0.f;
    }
    // ---- stores.  A thread holds kOC columns of ONE row: stored directly, every instruction would touch 32 rows (32 half-used
    // sectors).  The quadrant's 32 x 96 tile goes through shared memory instead (the Q / K buffers: every MMA and TMA load of
    // this CTA has completed; row stride 100 floats keeps the float4 writes conflict-free) and is written out row by row,
    // 24 lanes x 4 columns = one contiguous 384-byte (fp32) / 192-byte (16-bit) segment per instruction.
    float* stage = reinterpret_cast<float*>(smem);
    constexpr int kStageLd = kHD + 4;
    static_assert(kTQ * kStageLd * sizeof(float) <= 2 * kQChunkBytes + 4 * kKChunkBytes, "output staging must fit in the Q + K buffers");
#pragma unroll
    for (int c = 0; c < kOC; c += 4)
      *reinterpret_cast<float4*>(stage + row * kStageLd + part * kOC + c) = make_float4(o[c], o[c + 1], o[c + 2], o[c + 3]);
    named_bar_sync(1 + q, 32 * PARTS);
    if (lane < kHD / 4) {
      const int odt = p.out16_split ? B200SVS_DT_F16 : p.out16_dtype;
#pragma unroll 4
      for (int i = part; i < 32; i += PARTS) {
        const int r = q * 32 + i;
        if (q0 + r >= p.T) break;
        const size_t gro = static_cast<size_t>(b) * p.T + q0 + r;
        const float4 v = *reinterpret_cast<const float4*>(stage + r * kStageLd + 4 * lane);
        if (p.out32) *reinterpret_cast<float4*>(p.out32 + gro * p.out32_ld + h * kHD + 4 * lane) = v;
        if (p.out16) {
          uint16_t* op = reinterpret_cast<uint16_t*>(p.out16) + gro * p.out16_ld + h * kHD + 4 * lane;
          const uint2 hi = make_uint2(pk16(v.x, v.y, odt), pk16(v.z, v.w, odt));
          *reinterpret_cast<uint2*>(op) = hi;
          if (p.out16_split) {
            *reinterpret_cast<uint2*>(op + p.C) = make_uint2(pk16(v.x - rt16(v.x, odt), v.y - rt16(v.y, odt), odt),
                                                              pk16(v.z - rt16(v.z, odt), v.w - rt16(v.w, odt), odt));
            *reinterpret_cast<uint2*>(op + 2 * p.C) = hi;
          }
        }
      }
    }
    tc_fence_before();
  }

  __syncthreads();
  if (tid == 0) stamp(0, 15);      // CTA done
  if (warp == kMmaW) {
    tc_fence_after();
    tmem_dealloc(bars->tmem_base, kTmemCols);
  }
}

static size_t ts_smem_bytes(bool split, bool wide, int parts) {
  const int np = split ? 2 : 1, vst = 2;
  size_t s = (size_t)np * (2 * kQChunkBytes + 4 * kKChunkBytes + vst * kVBytes) + 2 * kTQ * kNR * sizeof(float) +
             kRelKBytes + kNR * kHD * sizeof(float) + 3 * parts * kTQ * sizeof(float) + sizeof(AttnTsBars) + 1024;
  // WIDE allocates all 512 TMEM columns: a second co-resident CTA would only block in tcgen05.alloc - ask for more than
  // half of the SM's shared memory so that it is not scheduled there in the first place
  if (wide && s < 120 * 1024) s = 120 * 1024;
  return s;
}

template <bool SPLIT, bool WIDE, int PARTS>
static int launch_ts(const CUtensorMap* tmQ, const CUtensorMap* tmK, const CUtensorMap* tmV, const AttnTcParams& p,
                     cudaStream_t stream) {
  dim3 grid((p.T + kTQ - 1) / kTQ, p.H, p.B);
  B200SVS_SMEM_OPTIN((attention_ts_kernel<SPLIT, WIDE, PARTS>), "cudaFuncSetAttribute(attention_ts)");
  attention_ts_kernel<SPLIT, WIDE, PARTS><<<grid, PARTS * 128 + 64, ts_smem_bytes(SPLIT, WIDE, PARTS), stream>>>(
      tmQ[0], tmK[0], tmV[0], tmQ[1], tmK[1], tmV[1], p);
  return check_launch("attention_ts_kernel");
}

int launch_attention_ts(const CUtensorMap* tmQ, const CUtensorMap* tmK, const CUtensorMap* tmV, const AttnTcParams& p,
                        bool split, bool wide, int parts, cudaStream_t stream) {
  if (split) return parts == 4 ? launch_ts<true, true, 4>(tmQ, tmK, tmV, p, stream) : launch_ts<true, true, 2>(tmQ, tmK, tmV, p, stream);
  if (!wide) return launch_ts<false, false, 2>(tmQ, tmK, tmV, p, stream);
  return parts == 4 ? launch_ts<false, true, 4>(tmQ, tmK, tmV, p, stream) : launch_ts<false, true, 2>(tmQ, tmK, tmV, p, stream);
}

}  // namespace b200svs
