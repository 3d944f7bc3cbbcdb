// attention_tc: windowed relative-position multi-head self-attention on the 5th-gen tensor cores.
//
// Restates VISinger/attentions.py:148-179 (+ helpers :199-243) flash style, never materialising the
// [B,H,T,T] scores:
//   S = (Q K^T)/sqrt(d) + band(q.rel_k) ; masked_fill(key >= len, -1e4) ; P = softmax(S)
//   O = P V + sum_{|r|<=w} P[i,i+r] rel_v[r]
// One CTA = 128 query rows of one (batch, head); thread r owns query row r = TMEM lane r.
// Per 64-key tile:  TMA(K_j) -> tcgen05.mma S = Q.K_j^T (TMEM, 64 cols) -> softmax in registers
// (online max/sum, fp32) -> P written to shared memory as the next MMA's A operand (SWIZZLE_128B,
// K-major) -> tcgen05.mma PV = P.V_j (TMEM, 96 cols) -> O = O*alpha + PV in registers.
// V is consumed K-major from a transposed copy V^T [B,C,T] written by attn_prep_kernel.
//
// SPLIT mode (F0 predictor, whose output is quantised by f0_to_coarse: SURVEY.md 7.3-C): every
// operand is an fp16 (hi, lo) pair with x = hi + lo to ~2^-22, and every product is three MMAs
// (hi.hi + hi.lo + lo.hi) accumulated in fp32 - tensor-core speed at fp32-class accuracy.
#include "common.cuh"
#include "api_internal.h"

namespace b200svs {

constexpr int kTQ = 128;      // queries per CTA
constexpr int kTK = 64;       // keys per tile
constexpr int kHD = 96;       // head dim
constexpr int kWin = 4;
constexpr int kNR = 2 * kWin + 1;
constexpr int kQChunkBytes = kTQ * 128;   // 16 KB  [128 rows][64 x 16-bit]
constexpr int kKChunkBytes = kTK * 128;   //  8 KB
constexpr int kVBytes = kHD * 128;        // 12 KB  [96 d rows][64 keys]
constexpr int kPBytes = kTQ * 128;        // 16 KB  [128 q rows][64 keys]

struct AttnTcParams {
  const void* qk_hi; const void* qk_lo; int qk_ld;   // [B,T,2C] 16-bit (q | k)
  float* out32; int out32_ld;
  void* out16; int out16_ld; int out16_dtype; int out16_split;
  const float* rel_k; const float* rel_v;
  const int32_t* lens;
  int B, T, C, H, dtype;
};

struct __align__(8) AttnBars {
  uint64_t qbar, kbar, vbar, sbar, obar;
  uint32_t tmem_base;
};

__device__ __forceinline__ float ld16f(const void* base, size_t idx, int dtype) {
  if (dtype == B200SVS_DT_BF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[idx]);
  return __half2float(reinterpret_cast<const __half*>(base)[idx]);
}
__device__ __forceinline__ uint32_t pk16(float a, float b, int dtype) {
  if (dtype == B200SVS_DT_BF16) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ float rt16(float a, int dtype) {   // round trip through the 16-bit type
  if (dtype == B200SVS_DT_BF16) return __bfloat162float(__float2bfloat16_rn(a));
  return __half2float(__float2half_rn(a));
}

template <bool SPLIT>
__global__ void __launch_bounds__(kTQ, SPLIT ? 1 : 2)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmQh, const __grid_constant__ CUtensorMap tmKh,
                    const __grid_constant__ CUtensorMap tmVh, const __grid_constant__ CUtensorMap tmQl,
                    const __grid_constant__ CUtensorMap tmKl, const __grid_constant__ CUtensorMap tmVl,
                    const AttnTcParams p) {
  constexpr int NP = SPLIT ? 2 : 1;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smQ = smem;                                   // [NP][2 chunks][16 KB]
  uint8_t* smK = smQ + NP * 2 * kQChunkBytes;            // [NP][2 chunks][8 KB]
  uint8_t* smV = smK + NP * 2 * kKChunkBytes;            // [NP][12 KB]
  uint8_t* smP = smV + NP * kVBytes;                     // [NP][16 KB]
  float* qrel = reinterpret_cast<float*>(smP + NP * kPBytes);   // [128][9]
  float* band = qrel + kTQ * kNR;                               // [128][9]
  AttnBars* bars = reinterpret_cast<AttnBars*>(band + kTQ * kNR);

  const int tid = threadIdx.x, warp = tid >> 5;
  const int q0 = blockIdx.x * kTQ, h = blockIdx.y, b = blockIdx.z;
  const int len = p.lens ? min(__ldg(p.lens + b), p.T) : p.T;
  const int dt = SPLIT ? B200SVS_DT_F16 : p.dtype;
  const int t = q0 + tid;
  const size_t orow = static_cast<size_t>(b) * p.T + t;
  const int oc = h * kHD;

  if (q0 >= len) {   // whole tile is padding: zeros (the reference's values there are masked by every consumer)
    if (t < p.T) {
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
    mbar_init(&bars->qbar, 1); mbar_init(&bars->kbar, 1); mbar_init(&bars->vbar, 1);
    mbar_init(&bars->sbar, 1); mbar_init(&bars->obar, 1);
    fence_barrier_init();
    prefetch_tmap(&tmQh); prefetch_tmap(&tmKh); prefetch_tmap(&tmVh);
  }
  if (warp == 0) {
    tmem_alloc(&bars->tmem_base, 256);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_s = bars->tmem_base;          // S : 64 columns
  const uint32_t tmem_o = tmem_s + 64;              // PV: 96 columns
  const int n_kt = (len + kTK - 1) / kTK;

  auto load_k = [&](int j) {
    mbar_arrive_expect_tx(&bars->kbar, NP * 2 * kKChunkBytes);
    for (int c = 0; c < 2; ++c) {
      tma_load_3d(smK + c * kKChunkBytes, &tmKh, &bars->kbar, p.C + oc + c * 64, j * kTK, b);
      if (SPLIT) tma_load_3d(smK + (2 + c) * kKChunkBytes, &tmKl, &bars->kbar, p.C + oc + c * 64, j * kTK, b);
    }
  };
  auto load_v = [&](int j) {
    mbar_arrive_expect_tx(&bars->vbar, NP * kVBytes);
    tma_load_3d(smV, &tmVh, &bars->vbar, j * kTK, oc, b);
    if (SPLIT) tma_load_3d(smV + kVBytes, &tmVl, &bars->vbar, j * kTK, oc, b);
  };
  if (tid == 0) {
    mbar_arrive_expect_tx(&bars->qbar, NP * 2 * kQChunkBytes);
    for (int c = 0; c < 2; ++c) {
      tma_load_3d(smQ + c * kQChunkBytes, &tmQh, &bars->qbar, oc + c * 64, q0, b);
      if (SPLIT) tma_load_3d(smQ + (2 + c) * kQChunkBytes, &tmQl, &bars->qbar, oc + c * 64, q0, b);
    }
    load_k(0);
    load_v(0);
  }

  // (q_i / sqrt(d)) . rel_k[w] for the 9 band offsets; q read back from global (16-bit hi [+ lo])
  const float scale = rsqrtf(static_cast<float>(kHD));
  {
    float a[kNR];
#pragma unroll
    for (int w = 0; w < kNR; ++w) a[w] = 0.f;
    if (t < p.T) {
      const size_t qi = orow * p.qk_ld + oc;
      for (int c = 0; c < kHD; ++c) {
        float qv = ld16f(p.qk_hi, qi + c, dt);
        if (SPLIT) qv += ld16f(p.qk_lo, qi + c, dt);
#pragma unroll
        for (int w = 0; w < kNR; ++w) a[w] = fmaf(qv, __ldg(p.rel_k + w * kHD + c), a[w]);
      }
    }
#pragma unroll
    for (int w = 0; w < kNR; ++w) {
      qrel[tid * kNR + w] = a[w] * scale;
      band[tid * kNR + w] = -INFINITY;
    }
  }

  const uint32_t idesc_s = make_idesc_f16(128, kTK, dt == B200SVS_DT_BF16);
  const uint32_t idesc_o = make_idesc_f16(128, kHD, dt == B200SVS_DT_BF16);
  float m_run = -INFINITY, l_run = 0.f;
  float o[kHD];
#pragma unroll
  for (int c = 0; c < kHD; ++c) o[c] = 0.f;

  for (int j = 0; j < n_kt; ++j) {
    const int k0 = j * kTK;
    const uint32_t ph = j & 1;
    if (tid == 0) {
      if (j == 0) mbar_wait(&bars->qbar, 0);
      mbar_wait(&bars->kbar, ph);
      tc_fence_after();
      bool first = true;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const int ks = c == 0 ? 4 : 2;     // d = 96 = 64 + 32
        for (int pr = 0; pr < (SPLIT ? 3 : 1); ++pr) {
          const int pa = pr == 2 ? 1 : 0, pb = pr == 1 ? 1 : 0;       // (hi,hi) (hi,lo) (lo,hi)
          const uint64_t da = make_desc_k128(smem_u32(smQ + (pa * 2 + c) * kQChunkBytes));
          const uint64_t db = make_desc_k128(smem_u32(smK + (pb * 2 + c) * kKChunkBytes));
          for (int k = 0; k < ks; ++k) {
            umma_f16(tmem_s, da + 2 * k, db + 2 * k, idesc_s, first ? 0u : 1u);
            first = false;
          }
        }
      }
      umma_commit(&bars->sbar);
    }
    mbar_wait(&bars->sbar, ph);
    tc_fence_after();
    if (tid == 0 && j + 1 < n_kt) load_k(j + 1);      // K buffer is free: S MMAs have completed

    // ---- softmax over this tile's 64 keys (row = thread)
    float s[kTK];
    {
      uint32_t r[32];
      const uint32_t ta = tmem_s + (static_cast<uint32_t>(warp * 32) << 16);
      tmem_ld32_nowait(ta, r);
      tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < 32; ++e) s[e] = __uint_as_float(r[e]) * scale;
      tmem_ld32_nowait(ta + 32, r);
      tmem_ld_wait();
#pragma unroll
      for (int e = 0; e < 32; ++e) s[32 + e] = __uint_as_float(r[e]) * scale;
    }
    const bool near_diag = (k0 <= q0 + kTQ - 1 + kWin) && (k0 + kTK - 1 >= q0 - kWin);
    const bool tail = k0 + kTK > len;
    if (near_diag) {
#pragma unroll
      for (int e = 0; e < kTK; ++e) {
        const int d = k0 + e - t + kWin;
        if (d >= 0 && d < kNR) s[e] += qrel[tid * kNR + d];
      }
    }
    if (tail) {
#pragma unroll
      for (int e = 0; e < kTK; ++e) {
        const int key = k0 + e;
        if (key >= len) s[e] = -1e4f;             // masked_fill(mask == 0, -1e4)
        if (key >= p.T) s[e] = -INFINITY;          // beyond the tensor: not a key at all
      }
    }
    if (near_diag) {
#pragma unroll
      for (int e = 0; e < kTK; ++e) {
        const int d = k0 + e - t + kWin;
        if (d >= 0 && d < kNR && k0 + e < p.T) band[tid * kNR + d] = s[e];
      }
    }
    float mx = s[0];
#pragma unroll
    for (int e = 1; e < kTK; ++e) mx = fmaxf(mx, s[e]);
    const float m_new = fmaxf(m_run, mx);
    const float alpha = (m_run == -INFINITY) ? 0.f : expf(m_run - m_new);
    float sum = 0.f;
    uint8_t* prow = smP + tid * 128;
#pragma unroll
    for (int c8 = 0; c8 < 8; ++c8) {
      float pv[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const float x = s[c8 * 8 + e];
        pv[e] = (x == -INFINITY) ? 0.f : (SPLIT ? expf(x - m_new) : __expf(x - m_new));
        sum += pv[e];
      }
      const int phys = (c8 ^ (tid & 7)) << 4;
      *reinterpret_cast<uint4*>(prow + phys) =
          make_uint4(pk16(pv[0], pv[1], dt), pk16(pv[2], pv[3], dt), pk16(pv[4], pv[5], dt), pk16(pv[6], pv[7], dt));
      if (SPLIT) {
#pragma unroll
        for (int e = 0; e < 8; ++e) pv[e] -= rt16(pv[e], dt);
        *reinterpret_cast<uint4*>(prow + kPBytes + phys) =
            make_uint4(pk16(pv[0], pv[1], dt), pk16(pv[2], pv[3], dt), pk16(pv[4], pv[5], dt), pk16(pv[6], pv[7], dt));
      }
    }
    l_run = l_run * alpha + sum;
    m_run = m_new;
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      mbar_wait(&bars->vbar, ph);
      tc_fence_after();
      bool first = true;
      for (int pr = 0; pr < (SPLIT ? 3 : 1); ++pr) {
        const int pa = pr == 2 ? 1 : 0, pb = pr == 1 ? 1 : 0;
        const uint64_t da = make_desc_k128(smem_u32(smP + pa * kPBytes));
        const uint64_t db = make_desc_k128(smem_u32(smV + pb * kVBytes));
        for (int k = 0; k < kTK / 16; ++k) {
          umma_f16(tmem_o, da + 2 * k, db + 2 * k, idesc_o, first ? 0u : 1u);
          first = false;
        }
      }
      umma_commit(&bars->obar);
    }
    mbar_wait(&bars->obar, ph);
    tc_fence_after();
    if (tid == 0 && j + 1 < n_kt) load_v(j + 1);      // V and P buffers are free: PV MMAs have completed
    {
      const uint32_t ta = tmem_o + (static_cast<uint32_t>(warp * 32) << 16);
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        uint32_t r[32];
        tmem_ld32_nowait(ta + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; ++e) o[c * 32 + e] = fmaf(o[c * 32 + e], alpha, __uint_as_float(r[e]));
      }
    }
    tc_fence_before();
  }

  // ---- relative-value band term, normalisation, stores
  const bool live = t < len;
  if (t < p.T) {
    if (live) {
#pragma unroll
      for (int w = 0; w < kNR; ++w) {
        const float sv = band[tid * kNR + w];
        if (sv == -INFINITY) continue;
        const float pw = SPLIT ? expf(sv - m_run) : __expf(sv - m_run);
#pragma unroll
        for (int c = 0; c < kHD; ++c) o[c] = fmaf(pw, __ldg(p.rel_v + w * kHD + c), o[c]);
      }
    }
    const float invl = live ? 1.0f / l_run : 0.f;
#pragma unroll
    for (int c = 0; c < kHD; ++c) o[c] = live ? o[c] * invl : 0.f;
    if (p.out32) {
      float* op = p.out32 + orow * p.out32_ld + oc;
#pragma unroll
      for (int c = 0; c < kHD; c += 4) *reinterpret_cast<float4*>(op + c) = make_float4(o[c], o[c + 1], o[c + 2], o[c + 3]);
    }
    if (p.out16) {
      uint16_t* op = reinterpret_cast<uint16_t*>(p.out16) + orow * p.out16_ld + oc;
      const int odt = p.out16_split ? B200SVS_DT_F16 : p.out16_dtype;
#pragma unroll
      for (int c = 0; c < kHD; c += 8) {
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
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(bars->tmem_base, 256);
  }
}

// qkv fp32 [B,T,3C] -> (q|k) 16-bit [B,T,2C] (hi [+ lo]) and V^T 16-bit [B,C,Tp] (hi [+ lo]).
__global__ void __launch_bounds__(256) attn_prep_kernel(const float* qkv, int qkv_ld, void* qk_hi, void* qk_lo, void* vt_hi,
                                                        void* vt_lo, int dt, int B, int T, int Tp, int C) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z, t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;   // c0 over 3C columns
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const bool is_v = c0 >= 2 * C;
  for (int j = ty; j < 32; j += 8) {
    const int t = t0 + j, c = c0 + tx;
    if (t >= T) continue;
    const float v = qkv[(static_cast<size_t>(b) * T + t) * qkv_ld + c];
    if (is_v) {
      tile[j][tx] = v;
    } else {
      const size_t o = (static_cast<size_t>(b) * T + t) * (2 * C) + c;
      if (dt == B200SVS_DT_BF16) {
        reinterpret_cast<__nv_bfloat16*>(qk_hi)[o] = __float2bfloat16_rn(v);
      } else {
        const __half hi = __float2half_rn(v);
        reinterpret_cast<__half*>(qk_hi)[o] = hi;
        if (qk_lo) reinterpret_cast<__half*>(qk_lo)[o] = __float2half_rn(v - __half2float(hi));
      }
    }
  }
  if (!is_v) return;
  __syncthreads();
  for (int j = ty; j < 32; j += 8) {
    const int c = c0 - 2 * C + j, t = t0 + tx;
    if (t >= T) continue;
    const float v = tile[tx][j];
    const size_t o = (static_cast<size_t>(b) * C + c) * Tp + t;
    if (dt == B200SVS_DT_BF16) {
      reinterpret_cast<__nv_bfloat16*>(vt_hi)[o] = __float2bfloat16_rn(v);
    } else {
      const __half hi = __float2half_rn(v);
      reinterpret_cast<__half*>(vt_hi)[o] = hi;
      if (vt_lo) reinterpret_cast<__half*>(vt_lo)[o] = __float2half_rn(v - __half2float(hi));
    }
  }
}

}  // namespace b200svs

using namespace b200svs;

extern "C" int b200svs_attn_prep(const float* qkv, int32_t qkv_ld, void* qk_hi, void* qk_lo, void* vt_hi, void* vt_lo,
                                 int32_t dtype, int32_t B, int32_t T, int32_t Tp, int32_t C, void* stream) {
  B200SVS_REQUIRE(qkv && qk_hi && vt_hi && B > 0 && T > 0 && C > 0, "attn_prep: bad argument");
  B200SVS_REQUIRE(C % 32 == 0 && Tp >= T && Tp % 8 == 0, "attn_prep: C=%d must be a multiple of 32, Tp=%d a multiple of 8 >= T", C, Tp);
  B200SVS_REQUIRE((!qk_lo && !vt_lo) || (qk_lo && vt_lo && dtype == B200SVS_DT_F16), "attn_prep: split mode is fp16 (hi, lo)");
  dim3 grid((T + 31) / 32, 3 * C / 32, B);
  attn_prep_kernel<<<grid, 256, 0, as_stream(stream)>>>(qkv, qkv_ld, qk_hi, qk_lo, vt_hi, vt_lo, dtype, B, T, Tp, C);
  return check_launch("attn_prep_kernel");
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
  const int np = split ? 2 : 1;
  const size_t smem = (size_t)np * (2 * kQChunkBytes + 2 * kKChunkBytes + kVBytes + kPBytes) + 2 * kTQ * kNR * sizeof(float) +
                      sizeof(AttnBars) + 1024;
  dim3 grid((T + kTQ - 1) / kTQ, H, B);
  if (split) {
    static bool cfg = false;
    if (!cfg) {
      int r = check_cuda(cudaFuncSetAttribute(attention_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "cudaFuncSetAttribute(attention_tc<split>)");
      if (r) return r;
      cfg = true;
    }
    attention_tc_kernel<true><<<grid, kTQ, smem, as_stream(stream)>>>(tmQ[0], tmK[0], tmV[0], tmQ[1], tmK[1], tmV[1], p);
  } else {
    static bool cfg = false;
    if (!cfg) {
      int r = check_cuda(cudaFuncSetAttribute(attention_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "cudaFuncSetAttribute(attention_tc)");
      if (r) return r;
      cfg = true;
    }
    attention_tc_kernel<false><<<grid, kTQ, smem, as_stream(stream)>>>(tmQ[0], tmK[0], tmV[0], tmQ[1], tmK[1], tmV[1], p);
  }
  return check_launch("attention_tc_kernel");
}
