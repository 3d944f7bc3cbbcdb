// attention_tc: windowed relative-position multi-head self-attention on the 5th-gen tensor cores.
//
// Restates VISinger/attentions.py:148-179 (+ helpers :199-243) flash style, never materialising the
// [B,H,T,T] scores:
//   S = (Q K^T)/sqrt(d) + band(q.rel_k) ; masked_fill(key >= len, -1e4) ; P = softmax(S)
//   O = P V + sum_{|r|<=w} P[i,i+r] rel_v[r]
// One CTA = 128 query rows of one (batch, head), warp-specialised and software-pipelined over 64-key tiles:
//   warps 0-7 : softmax (warps q and q+4 own query rows 32q..32q+31 = TMEM lanes, and split the 64 key columns and
//               the 96 O columns between them): S_j from TMEM -> online max/sum (fp32,
//               log2 domain, ex2.approx) -> P_j to shared memory as the PV MMA's A operand (SWIZZLE_128B,
//               K-major); folds PV_{j-1} into the register accumulator O (O = O*alpha + PV)
//   warp 8    : tcgen05.mma issuer (one elected thread): S_{j+1} = Q.K_{j+1}^T is issued while the softmax
//               warps work on S_j (S double-buffered in TMEM), PV_j = P_j.V_j as soon as P_j is published
//   warp 9    : TMA loader (K double-buffered, V single-buffered)
// V is consumed K-major from a transposed copy V^T [B,C,T] written by attn_prep_kernel.
//
// SPLIT mode (phoneme encoder and F0 predictor, whose outputs cross float->int boundaries):
// every operand is an fp16 (hi, lo) pair with x = hi + lo to ~2^-22, and every product is three MMAs
// (hi.hi + hi.lo + lo.hi) accumulated in fp32 - tensor-core speed at fp32-class accuracy.
#include <type_traits>
#include "attention_common.cuh"

namespace b200svs {

template <bool SPLIT>
#ifndef B200SVS_ATTN_MINB_PLAIN
#define B200SVS_ATTN_MINB_PLAIN 2
#endif
__global__ void __launch_bounds__(kAttnThreads, SPLIT ? 1 : B200SVS_ATTN_MINB_PLAIN)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmQh, const __grid_constant__ CUtensorMap tmKh,
                    const __grid_constant__ CUtensorMap tmVh, const __grid_constant__ CUtensorMap tmQl,
                    const __grid_constant__ CUtensorMap tmKl, const __grid_constant__ CUtensorMap tmVl,
                    const AttnTcParams p) {
  constexpr int NP = SPLIT ? 2 : 1;
  constexpr int kKStage = NP * 2 * kKChunkBytes;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // pointer arithmetic keeps the shared address space (LDS/STS, not generic LD/ST)
  uint8_t* smQ = smem;                                   // [NP][2 chunks][16 KB]
  uint8_t* smK = smQ + NP * 2 * kQChunkBytes;            // [2 stages][NP][2 chunks][8 KB]
  uint8_t* smV = smK + 2 * kKStage;                      // [NP][12 KB]
  uint8_t* smP = smV + NP * kVBytes;                     // [NP][16 KB]
  float* qrel = reinterpret_cast<float*>(smP + NP * kPBytes);   // [128][9]  (q/sqrt(d)).rel_k * log2(e)
  float* band = qrel + kTQ * kNR;                               // [128][9]  band scores (log2 domain)
  float* relk_s = band + kTQ * kNR;                             // [9][96]
  float* relv_s = relk_s + kNR * kHD;                           // [9][96]
  float* pmax = relv_s + kNR * kHD;                             // [2 S buffers][2 halves][128] partial row maxima
  float* lsum = pmax + 4 * kTQ;                                 // [2 halves][128] partial row sums
  AttnBars* bars = reinterpret_cast<AttnBars*>(lsum + 2 * kTQ);

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

  if (tid == 0) {
    mbar_init(&bars->qbar, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->k_full[i], 1); mbar_init(&bars->k_empty[i], 1);
      mbar_init(&bars->s_full[i], 1); mbar_init(&bars->s_free[i], kSoftWarps);
    }
    mbar_init(&bars->v_full, 1); mbar_init(&bars->v_empty, 1);
    mbar_init(&bars->p_full, kSoftWarps); mbar_init(&bars->o_full, 1);
    fence_barrier_init();
    prefetch_tmap(&tmQh); prefetch_tmap(&tmKh); prefetch_tmap(&tmVh);
  }
  if (warp == kMmaWarp) {
    tmem_alloc(&bars->tmem_base, 256);
    tmem_relinquish();
  }
  for (int i = tid; i < kNR * kHD; i += kAttnThreads) {
    relk_s[i] = __ldg(p.rel_k + i);
    relv_s[i] = __ldg(p.rel_v + i);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_s = bars->tmem_base;          // S : 2 x 64 columns
  const uint32_t tmem_o = tmem_s + 2 * kTK;         // PV: 96 columns
  const int n_kt = (len + kTK - 1) / kTK;

  if (warp == kTmaWarp) {
    if (lane == 0) {
      // ------------------------------------------------------------ TMA loader: Q, K0, then (K_{j+1}, V_j)
      auto load_k = [&](int j) {
        const int st = j & 1;
        attn_wait(&bars->k_empty[st], ((j >> 1) & 1) ^ 1);
        uint8_t* dst = smK + st * kKStage;
        mbar_arrive_expect_tx(&bars->k_full[st], kKStage);
        for (int c = 0; c < 2; ++c) {
          tma_load_3d(dst + c * kKChunkBytes, &tmKh, &bars->k_full[st], p.C + oc + c * 64, j * kTK, b);
          if (SPLIT) tma_load_3d(dst + (2 + c) * kKChunkBytes, &tmKl, &bars->k_full[st], p.C + oc + c * 64, j * kTK, b);
        }
      };
      mbar_arrive_expect_tx(&bars->qbar, NP * 2 * kQChunkBytes);
      for (int c = 0; c < 2; ++c) {
        tma_load_3d(smQ + c * kQChunkBytes, &tmQh, &bars->qbar, oc + c * 64, q0, b);
        if (SPLIT) tma_load_3d(smQ + (2 + c) * kQChunkBytes, &tmQl, &bars->qbar, oc + c * 64, q0, b);
      }
      load_k(0);
      for (int j = 0; j < n_kt; ++j) {
        if (j + 1 < n_kt) load_k(j + 1);
        attn_wait(&bars->v_empty, (j & 1) ^ 1);
        mbar_arrive_expect_tx(&bars->v_full, NP * kVBytes);
        tma_load_3d(smV, &tmVh, &bars->v_full, j * kTK, oc, b);
        if (SPLIT) tma_load_3d(smV + kVBytes, &tmVl, &bars->v_full, j * kTK, oc, b);
      }
    }
  } else if (warp == kMmaWarp) {
    if (elect_one()) {
      // ------------------------------------------------------------ MMA issuer
      const uint32_t idesc_s = make_idesc_f16(128, kTK, dt == B200SVS_DT_BF16);
      const uint32_t idesc_o = make_idesc_f16(128, kHD, dt == B200SVS_DT_BF16);
      const uint32_t q_u = smem_u32(smQ), k_u = smem_u32(smK), v_u = smem_u32(smV), p_u = smem_u32(smP);
      auto issue_s = [&](int j) {      // S_j = Q . K_j^T  (d = 96 = 64 + 32)
        const int st = j & 1;
        attn_wait(&bars->k_full[st], (j >> 1) & 1);
        attn_wait(&bars->s_free[st], ((j >> 1) & 1) ^ 1);
        tc_fence_after();
        uint32_t accum = 0;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
#pragma unroll
          for (int pr = 0; pr < (SPLIT ? 3 : 1); ++pr) {
            const int pa = pr == 2 ? 1 : 0, pb = pr == 1 ? 1 : 0;       // (hi,hi) (hi,lo) (lo,hi)
            const uint64_t da = make_desc_k128(q_u + (pa * 2 + c) * kQChunkBytes);
            const uint64_t db = make_desc_k128(k_u + st * kKStage + (pb * 2 + c) * kKChunkBytes);
#pragma unroll
            for (int k = 0; k < (c == 0 ? 4 : 2); ++k) {
              umma_f16(tmem_s + st * kTK, da + 2 * k, db + 2 * k, idesc_s, accum);
              accum = 1;
            }
          }
        }
        umma_commit(&bars->s_full[st]);
        umma_commit(&bars->k_empty[st]);
      };
      attn_wait(&bars->qbar, 0);
      issue_s(0);
      for (int j = 0; j < n_kt; ++j) {
        if (j + 1 < n_kt) issue_s(j + 1);
        attn_wait(&bars->p_full, j & 1);
        attn_wait(&bars->v_full, j & 1);
        tc_fence_after();
        uint32_t accum = 0;
#pragma unroll
        for (int pr = 0; pr < (SPLIT ? 3 : 1); ++pr) {
          const int pa = pr == 2 ? 1 : 0, pb = pr == 1 ? 1 : 0;
          const uint64_t da = make_desc_k128(p_u + pa * kPBytes);
          const uint64_t db = make_desc_k128(v_u + pb * kVBytes);
#pragma unroll
          for (int k = 0; k < kTK / 16; ++k) {
            umma_f16(tmem_o, da + 2 * k, db + 2 * k, idesc_o, accum);
            accum = 1;
          }
        }
        umma_commit(&bars->o_full);
        umma_commit(&bars->v_empty);
      }
    }
    __syncwarp();
  } else {
    // ---------------------------------------------------------------- softmax warps 0-7 (256 threads)
    // Thread (row, half): query row `row` = TMEM lane, key columns [32*half, 32*half+32) of every S tile and
    // O columns [48*half, 48*half+48).  Warps q and q+4 own the same rows; they exchange the row maximum once
    // per tile through shared memory (named barrier 1+q, 64 threads) and the row sums once at the end.
    const int q = warp & 3, half = warp >> 2;
    const int row = q * 32 + lane;
    const int t = q0 + row;
    const size_t orow = static_cast<size_t>(b) * p.T + t;
    const float scale2 = rsqrtf(static_cast<float>(kHD)) * kLog2e;     // scores live in the log2 domain
    {   // (q_i / sqrt(d)) . rel_k[w]: half 0 computes band offsets 0..4, half 1 offsets 5..8
      const int w0 = half ? 5 : 0, nw = half ? 4 : 5;
      float a[5];
#pragma unroll
      for (int w = 0; w < 5; ++w) a[w] = 0.f;
      attn_wait(&bars->qbar, 0);
#pragma unroll 1
      for (int cell = 0; cell < kHD / 8; ++cell) {      // 12 cells of 8 channels: chunk 0 holds d 0..63, chunk 1 d 64..95
        const int c = cell >> 3, jc = cell & 7;
        const uint8_t* src = smQ + c * kQChunkBytes + row * 128 + ((jc ^ (row & 7)) << 4);
        const uint4 qh = *reinterpret_cast<const uint4*>(src);
        const uint32_t uh[4] = {qh.x, qh.y, qh.z, qh.w};
        float qv[8];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          if (dt == B200SVS_DT_BF16) {
            qv[2 * e] = __uint_as_float(uh[e] << 16); qv[2 * e + 1] = __uint_as_float(uh[e] & 0xFFFF0000u);
          } else {
            const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&uh[e]));
            qv[2 * e] = f.x; qv[2 * e + 1] = f.y;
          }
        }
        if (SPLIT) {
          const uint4 ql = *reinterpret_cast<const uint4*>(src + 2 * kQChunkBytes);
          const uint32_t ul[4] = {ql.x, ql.y, ql.z, ql.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&ul[e]));
            qv[2 * e] += f.x; qv[2 * e + 1] += f.y;
          }
        }
#pragma unroll
        for (int w = 0; w < 5; ++w) {
          if (w < nw) {
            const float4 k0v = *reinterpret_cast<const float4*>(relk_s + (w0 + w) * kHD + cell * 8);
            const float4 k1v = *reinterpret_cast<const float4*>(relk_s + (w0 + w) * kHD + cell * 8 + 4);
            float x = a[w];
            x = fmaf(qv[0], k0v.x, x); x = fmaf(qv[1], k0v.y, x); x = fmaf(qv[2], k0v.z, x); x = fmaf(qv[3], k0v.w, x);
            x = fmaf(qv[4], k1v.x, x); x = fmaf(qv[5], k1v.y, x); x = fmaf(qv[6], k1v.z, x); x = fmaf(qv[7], k1v.w, x);
            a[w] = x;
          }
        }
      }
#pragma unroll
      for (int w = 0; w < 5; ++w) {
        if (w < nw) {
          qrel[row * kNR + w0 + w] = (t < p.T) ? a[w] * scale2 : 0.f;
          band[row * kNR + w0 + w] = -INFINITY;
        }
      }
      named_bar_sync(1 + q, 64);         // both halves' band offsets are in place
    }
    float m_run = -INFINITY, l_run = 0.f, alpha_prev = 0.f;    // l_run: partial row sum over this thread's key columns
    constexpr int kOC = kHD / 2;                               // 48 O columns per thread
    float o[kOC];
#pragma unroll
    for (int c = 0; c < kOC; ++c) o[c] = 0.f;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    auto fold = [&](float alpha) {       // O = O * alpha + PV  (this thread's 48 of the 96 PV columns)
      uint32_t r0[16], r1[16], r2[16];   // all three loads in flight before the first use
      tmem_ld16_nowait(tmem_o + lane_off + half * kOC, r0);
      tmem_ld16_nowait(tmem_o + lane_off + half * kOC + 16, r1);
      tmem_ld16_nowait(tmem_o + lane_off + half * kOC + 32, r2);
      tmem_ld_wait16(r0);
#pragma unroll
      for (int e = 0; e < 16; ++e) o[e] = fmaf(o[e], alpha, __uint_as_float(r0[e]));
      tmem_ld_wait16(r1);
#pragma unroll
      for (int e = 0; e < 16; ++e) o[16 + e] = fmaf(o[16 + e], alpha, __uint_as_float(r1[e]));
      tmem_ld_wait16(r2);
#pragma unroll
      for (int e = 0; e < 16; ++e) o[32 + e] = fmaf(o[32 + e], alpha, __uint_as_float(r2[e]));
    };

    for (int j = 0; j < n_kt; ++j) {
      const int k0 = j * kTK;
      const int sb = j & 1;
      attn_wait(&bars->s_full[sb], (j >> 1) & 1);
      tc_fence_after();
      const uint32_t ta = tmem_s + sb * kTK + lane_off + half * 32;      // this thread's 32 key columns
      const int kb = k0 + half * 32;                                       // first key of those columns
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
      // ---- S_j: this thread's 32 scores are read from TMEM ONCE and stay in registers for both passes; the S buffer
      // is handed back to the MMA warp right away (S_{j+2} can be issued while this tile's exponentials run)
      uint32_t sv[32];
      tmem_ld32_nowait(ta, sv);
      tmem_ld_wait32(sv);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->s_free[sb]);
      // ---- pass 1: row maximum over this thread's 32 keys, then over the pair.  Slow tiles keep the adjusted score.
      float mx = -INFINITY;
      if (slow) {
#pragma unroll
        for (int e = 0; e < 32; ++e) {
          const float x = adjust(sv[e], e);
          sv[e] = __float_as_uint(x);
          mx = fmaxf(mx, x);
        }
      } else {
#pragma unroll
        for (int e = 0; e < 32; ++e) mx = fmaxf(mx, __uint_as_float(sv[e]));
      }
      if (!slow) mx *= scale2;                         // scale2 > 0 commutes with max
      pmax[(sb * 2 + half) * kTQ + row] = mx;
      named_bar_sync(1 + q, 64);
      mx = fmaxf(mx, pmax[(sb * 2 + (half ^ 1)) * kTQ + row]);
      const float m_new = fmaxf(m_run, mx);
      const float alpha = (m_run == -INFINITY) ? 0.f : ex2f(m_run - m_new);
      // ---- PV_{j-1} has to be folded into O (and the P buffer released by its MMAs) before P_j is written
      if (j > 0) {
        attn_wait(&bars->o_full, (j - 1) & 1);
        tc_fence_after();
        fold(alpha_prev);
      }
      // ---- pass 2: P = 2^(S - m), partial row sum, publish P as the A operand of the PV MMA
      float sum = 0.f;
      uint8_t* prow = smP + row * 128;
      auto emit = [&](const uint32_t* r, int g, auto slow_c) {      // 8 keys -> one 16-byte cell of the P row
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
        const int phys = ((half * 4 + g) ^ (row & 7)) << 4;
        *reinterpret_cast<uint4*>(prow + phys) =
            make_uint4(pk16(pv[0], pv[1], dt), pk16(pv[2], pv[3], dt), pk16(pv[4], pv[5], dt), pk16(pv[6], pv[7], dt));
        if (SPLIT) {
#pragma unroll
          for (int e = 0; e < 8; ++e) pv[e] -= rt16(pv[e], dt);
          *reinterpret_cast<uint4*>(prow + kPBytes + phys) =
              make_uint4(pk16(pv[0], pv[1], dt), pk16(pv[2], pv[3], dt), pk16(pv[4], pv[5], dt), pk16(pv[6], pv[7], dt));
        }
      };
      if (slow) {
#pragma unroll
        for (int g = 0; g < 4; ++g) emit(sv + 8 * g, g, std::true_type{});
      } else {
#pragma unroll
        for (int g = 0; g < 4; ++g) emit(sv + 8 * g, g, std::false_type{});
      }
      l_run = l_run * alpha + sum;
      m_run = m_new;
      alpha_prev = alpha;
      fence_proxy_async();               // P (generic-proxy stores) -> visible to the tensor core
      tc_fence_before();                 // PV_{j-1} reads are complete
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->p_full);
    }
    attn_wait(&bars->o_full, (n_kt - 1) & 1);
    tc_fence_after();
    fold(alpha_prev);

    // ---- row sum over both halves, relative-value band term, normalisation, stores
    lsum[half * kTQ + row] = l_run;
    named_bar_sync(1 + q, 64);
    const float l_tot = l_run + lsum[(half ^ 1) * kTQ + row];
    const bool live = t < len;
    const int oc = h * kHD + half * kOC;
    if (t < p.T) {
      if (live) {
#pragma unroll 1
        for (int w = 0; w < kNR; ++w) {
          const float sv = band[row * kNR + w];
          if (sv == -INFINITY) continue;
          const float pw = ex2f(sv - m_run);
#pragma unroll
          for (int c = 0; c < kOC; c += 4) {
            const float4 rv = *reinterpret_cast<const float4*>(relv_s + w * kHD + half * kOC + c);
            o[c] = fmaf(pw, rv.x, o[c]); o[c + 1] = fmaf(pw, rv.y, o[c + 1]);
            o[c + 2] = fmaf(pw, rv.z, o[c + 2]); o[c + 3] = fmaf(pw, rv.w, o[c + 3]);
          }
        }
      }
      const float invl = live ? 1.0f / l_tot : 0.f;
#pragma unroll
      for (int c = 0; c < kOC; ++c) o[c] = live ? o[c] * invl : 0.f;
      if (p.out32) {
        float* op = p.out32 + orow * p.out32_ld + oc;
#pragma unroll
        for (int c = 0; c < kOC; c += 4) *reinterpret_cast<float4*>(op + c) = make_float4(o[c], o[c + 1], o[c + 2], o[c + 3]);
      }
      if (p.out16) {
        uint16_t* op = reinterpret_cast<uint16_t*>(p.out16) + orow * p.out16_ld + oc;
        const int odt = p.out16_split ? B200SVS_DT_F16 : p.out16_dtype;
#pragma unroll
        for (int c = 0; c < kOC; c += 8) {
          const uint4 hi = make_uint4(pk16(o[c], o[c + 1], odt), pk16(o[c + 2], o[c + 3], odt), pk16(o[c + 4], o[c + 5], odt),
                                      pk16(o[c + 6], o[c + 7], odt));
          *reinterpret_cast<uint4*>(op + c) = hi;
          if (p.out16_split) {
            float lo[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) lo[e] = o[c + e] - rt16(o[c + e], odt);
            *reinterpret_cast<uint4*>(op + p.C + c) =
                make_uint4(pk16(lo[0], lo[1], odt), pk16(lo[2], lo[3], odt), pk16(lo[4], lo[5], odt), pk16(lo[6], lo[7], odt));
            *reinterpret_cast<uint4*>(op + 2 * p.C + c) = hi;
          }
        }
      }
    }
    tc_fence_before();
  }

  __syncthreads();
  if (warp == kMmaWarp) {
    tc_fence_after();
    tmem_dealloc(bars->tmem_base, 256);
  }
}

// qkv fp32 [B,T,3C] -> (q|k) 16-bit [B,T,2C] (hi [+ lo]) and V^T 16-bit [B,C,Tp] (hi [+ lo]).
// One block = 64 time rows x (128 q|k columns, or 32 V channels).  q|k: a thread converts 4 consecutive columns (16-byte
// load, 8-byte stores, a warp covers 512 contiguous bytes of a row).  V: 64 x 32 tile transposed through shared memory,
// a warp writes 64 time steps of one channel as half2 (128 contiguous bytes).
constexpr int kPrepRows = 64;
template <bool BF>
__device__ __forceinline__ void prep_cvt2(float a, float b, uint32_t& hi, uint32_t& lo) {
  if (BF) {
    hi = cvt_bf16x2_sat(a, b);
    lo = 0;
  } else {
    hi = cvt_f16x2_sat(a, b);
    const float2 hf = __half22float2(*reinterpret_cast<const __half2*>(&hi));
    lo = cvt_f16x2_sat(a - hf.x, b - hf.y);
  }
}
template <bool BF>
__global__ void __launch_bounds__(256) attn_prep_kernel(const float* qkv, int qkv_ld, void* qk_hi, void* qk_lo, void* vt_hi,
                                                        void* vt_lo, int n_qk_blocks, int B, int T, int Tp, int C) {
  __shared__ float tile[kPrepRows][33];
  const int b = blockIdx.z, t0 = blockIdx.x * kPrepRows;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  if (static_cast<int>(blockIdx.y) < n_qk_blocks) {
    const int c = blockIdx.y * 128 + tx * 4;
    if (c >= 2 * C) return;
    for (int j = ty; j < kPrepRows; j += 8) {
      const int t = t0 + j;
      if (t >= T) break;
      const size_t row = static_cast<size_t>(b) * T + t;
      const float4 v = __ldg(reinterpret_cast<const float4*>(qkv + row * qkv_ld + c));
      uint32_t h0, l0, h1, l1;
      prep_cvt2<BF>(v.x, v.y, h0, l0);
      prep_cvt2<BF>(v.z, v.w, h1, l1);
      const size_t o = row * (2 * C) + c;
      *reinterpret_cast<uint2*>(reinterpret_cast<uint16_t*>(qk_hi) + o) = make_uint2(h0, h1);
      if (!BF && qk_lo) *reinterpret_cast<uint2*>(reinterpret_cast<uint16_t*>(qk_lo) + o) = make_uint2(l0, l1);
    }
    return;
  }
  const int c0 = (blockIdx.y - n_qk_blocks) * 32;
  for (int j = ty; j < kPrepRows; j += 8) {
    const int t = t0 + j;
    tile[j][tx] = t < T ? __ldg(qkv + (static_cast<size_t>(b) * T + t) * qkv_ld + 2 * C + c0 + tx) : 0.f;
  }
  __syncthreads();
  const int t = t0 + 2 * tx;                 // this lane's two time steps (t0 and Tp are even)
  if (t >= Tp) return;
  for (int j = ty; j < 32; j += 8) {
    uint32_t h, l;
    prep_cvt2<BF>(tile[2 * tx][j], tile[2 * tx + 1][j], h, l);      // rows >= T hold zeros: the pad [T, Tp) is written as 0
    const size_t o = (static_cast<size_t>(b) * C + c0 + j) * Tp + t;
    *reinterpret_cast<uint32_t*>(reinterpret_cast<uint16_t*>(vt_hi) + o) = h;
    if (!BF && vt_lo) *reinterpret_cast<uint32_t*>(reinterpret_cast<uint16_t*>(vt_lo) + o) = l;
  }
}

}  // namespace b200svs

using namespace b200svs;

extern "C" int b200svs_attn_prep(const float* qkv, int32_t qkv_ld, void* qk_hi, void* qk_lo, void* vt_hi, void* vt_lo,
                                 int32_t dtype, int32_t B, int32_t T, int32_t Tp, int32_t C, void* stream) {
  B200SVS_REQUIRE(qkv && qk_hi && vt_hi && B > 0 && T > 0 && C > 0, "attn_prep: bad argument");
  B200SVS_REQUIRE(C % 32 == 0 && Tp >= T && Tp % 8 == 0, "attn_prep: C=%d must be a multiple of 32, Tp=%d a multiple of 8 >= T", C, Tp);
  B200SVS_REQUIRE((!qk_lo && !vt_lo) || (qk_lo && vt_lo && dtype == B200SVS_DT_F16), "attn_prep: split mode is fp16 (hi, lo)");
  B200SVS_REQUIRE(qkv_ld % 4 == 0 && (reinterpret_cast<uintptr_t>(qkv) & 15) == 0 && C % 2 == 0,
                  "attn_prep: qkv must be 16-byte aligned with ld %% 4 == 0 (ld=%d)", qkv_ld);
  const int n_qk = (2 * C + 127) / 128;
  dim3 grid((T + kPrepRows - 1) / kPrepRows, n_qk + C / 32, B);
  if (dtype == B200SVS_DT_BF16)
    attn_prep_kernel<true><<<grid, 256, 0, as_stream(stream)>>>(qkv, qkv_ld, qk_hi, qk_lo, vt_hi, vt_lo, n_qk, B, T, Tp, C);
  else
    attn_prep_kernel<false><<<grid, 256, 0, as_stream(stream)>>>(qkv, qkv_ld, qk_hi, qk_lo, vt_hi, vt_lo, n_qk, B, T, Tp, C);
  return check_launch("attn_prep_kernel");
}

static long long* g_attn_trace = nullptr;
static int g_attn_trace_cap = 0;
extern "C" int b200svs_attention_set_trace(void* dev_buf, int32_t tiles) {
  g_attn_trace = static_cast<long long*>(dev_buf);
  g_attn_trace_cap = dev_buf ? tiles : 0;
  return 0;
}

extern "C" int b200svs_attention_tc(const void* qk_hi, const void* qk_lo, const void* vt_hi, const void* vt_lo,
                                    int32_t dtype, int32_t Tp, float* out32, int32_t out32_ld, void* out16,
                                    int32_t out16_ld, int32_t out16_dtype, int32_t out16_split, const float* rel_k,
                                    const float* rel_v, const int32_t* lens, int32_t B, int32_t T, int32_t C, int32_t H,
                                    int32_t window, void* stream) {
  B200SVS_REQUIRE(qk_hi && vt_hi && rel_k && rel_v && (out32 || out16), "attention_tc: null argument");
  B200SVS_REQUIRE(B > 0 && T > 0, "attention_tc: empty shape B=%d T=%d", B, T);
  B200SVS_REQUIRE(C == H * kHD && window == kWin, "attention_tc: built for head dim 96, window 4 (C=%d H=%d window=%d)", C, H, window);
  const bool split = qk_lo != nullptr;
  B200SVS_REQUIRE(split == (vt_lo != nullptr), "attention_tc: qk_lo and vt_lo must be given together");
  B200SVS_REQUIRE(!split || dtype == B200SVS_DT_F16, "attention_tc: split mode is fp16 (hi, lo)");
  B200SVS_REQUIRE(Tp >= T && Tp % 8 == 0, "attention_tc: Tp=%d must be a multiple of 8 >= T=%d", Tp, T);
  B200SVS_REQUIRE(!out32 || (out32_ld % 4 == 0 && (reinterpret_cast<uintptr_t>(out32) & 15) == 0), "attention_tc: out32 alignment");
  B200SVS_REQUIRE(!out16 || (out16_ld % 8 == 0 && (reinterpret_cast<uintptr_t>(out16) & 15) == 0), "attention_tc: out16 alignment");
  B200SVS_REQUIRE(!out16_split || out16, "attention_tc: out16_split without out16");

  const int el = dtype == B200SVS_DT_BF16 ? 1 : 0;
  CUtensorMap tmQ[2], tmK[2], tmV[2];
  for (int part = 0; part < (split ? 2 : 1); ++part) {
    const void* qk = part ? qk_lo : qk_hi;
    const void* vt = part ? vt_lo : vt_hi;
    uint64_t dims[3] = {(uint64_t)2 * C, (uint64_t)T, (uint64_t)B};
    uint64_t strides[2] = {(uint64_t)2 * C * 2, (uint64_t)T * 2 * C * 2};
    uint32_t boxq[3] = {64, (uint32_t)kTQ, 1}, boxk[3] = {64, (uint32_t)kTK, 1};
    int r = make_tmap(&tmQ[part], el, qk, 3, dims, strides, boxq, 128);
    if (r) return r;
    r = make_tmap(&tmK[part], el, qk, 3, dims, strides, boxk, 128);
    if (r) return r;
    uint64_t vdims[3] = {(uint64_t)T, (uint64_t)C, (uint64_t)B};
    uint64_t vstrides[2] = {(uint64_t)Tp * 2, (uint64_t)C * Tp * 2};
    uint32_t boxv[3] = {(uint32_t)kTK, (uint32_t)kHD, 1};
    r = make_tmap(&tmV[part], el, vt, 3, vdims, vstrides, boxv, 128);
    if (r) return r;
  }
  if (!split) { tmQ[1] = tmQ[0]; tmK[1] = tmK[0]; tmV[1] = tmV[0]; }

  AttnTcParams p;
  p.qk_hi = qk_hi; p.qk_lo = qk_lo; p.qk_ld = 2 * C;
  p.out32 = out32; p.out32_ld = out32_ld; p.out16 = out16; p.out16_ld = out16_ld; p.out16_dtype = out16_dtype;
  p.out16_split = out16_split; p.rel_k = rel_k; p.rel_v = rel_v; p.lens = lens;
  p.B = B; p.T = T; p.C = C; p.H = H; p.dtype = dtype;
  p.trace = g_attn_trace; p.trace_cap = g_attn_trace_cap;
  // B200SVS_ATTN_TS: 0 = operands from shared memory (this file); 1 = attention_ts for split launches; 2 (default) = also
  // for plain launches, one CTA per SM (Q in TMEM, two PV accumulators); 3 = plain launches with two CTAs per SM (P in
  // TMEM only).  Measured at 64 x 2584 frames: split 1298 -> 1000 us, plain 946 -> 861 (mode 2) / 920 (mode 3).
  static const int ts_mode = [] { const char* e = getenv("B200SVS_ATTN_TS"); return e ? atoi(e) : 2; }();
  static const int ts_parts = [] { const char* e = getenv("B200SVS_ATTN_TS_PARTS"); return e ? atoi(e) : 2; }();
  if (ts_mode >= 1 && (split || ts_mode >= 2))
    return launch_attention_ts(tmQ, tmK, tmV, p, split, ts_mode == 2, ts_parts, as_stream(stream));
  const int np = split ? 2 : 1;
  const size_t smem = (size_t)np * (2 * kQChunkBytes + 4 * kKChunkBytes + kVBytes + kPBytes) + 2 * kTQ * kNR * sizeof(float) + 2 * kNR * kHD * sizeof(float) + 6 * kTQ * sizeof(float) +
                      sizeof(AttnBars) + 1024;
  dim3 grid((T + kTQ - 1) / kTQ, H, B);
  if (split) {
    B200SVS_SMEM_OPTIN(attention_tc_kernel<true>, "cudaFuncSetAttribute(attention_tc<split>)");
    attention_tc_kernel<true><<<grid, kAttnThreads, smem, as_stream(stream)>>>(tmQ[0], tmK[0], tmV[0], tmQ[1], tmK[1], tmV[1], p);
  } else {
    B200SVS_SMEM_OPTIN(attention_tc_kernel<false>, "cudaFuncSetAttribute(attention_tc)");
    attention_tc_kernel<false><<<grid, kAttnThreads, smem, as_stream(stream)>>>(tmQ[0], tmK[0], tmV[0], tmQ[1], tmK[1], tmV[1], p);
  }
  return check_launch("attention_tc_kernel");
}
