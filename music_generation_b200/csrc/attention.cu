// attention_f32: windowed relative-position multi-head self-attention, flash style, fp32 SIMT.
//
// Restates VISinger/attentions.py:148-179 (+ helpers :199-243) without ever materialising the
// [B,H,T,T] score tensor or the pad/view/slice copies of the relative<->absolute conversions:
//   score[i,j] = (q_i/sqrt(d)) . k_j  +  [|j-i| <= w] (q_i/sqrt(d)) . rel_k[j-i+w]
//   score[i,j] = -1e4 where key j is padding          (masked_fill, :166)
//   p = softmax_j(score);  out_i = sum_j p[i,j] v_j + sum_{|r|<=w} p[i,i+r] rel_v[r+w]
// heads_share=True: one [2w+1, d] table for all heads (:125-128).
//
// One CTA = 64 query rows of one (batch, head); keys/values stream through shared memory in tiles of
// 64 with an online softmax.  Key tiles at or beyond lens[b] are skipped: for a valid query their
// probabilities are exp(-1e4 - max) == 0 in fp32, exactly what the reference computes.  Query rows
// >= lens[b] are written as zeros (the reference leaves finite garbage there that every consumer
// multiplies by the mask, attentions.py:46,278,285).
#include "common.cuh"
#include "api_internal.h"

namespace b200svs {

constexpr int kAQ = 64;    // query rows per CTA
constexpr int kAK = 64;    // keys per tile
constexpr int kAD = 96;    // head dim (hidden 192 / 2 heads)
constexpr int kAW = 4;     // window
constexpr int kAR = 2 * kAW + 1;
constexpr int kKtLd = kAK + 4;

struct AttnSmem {
  float Qs[kAQ][kAD];
  float Kt[kAD][kKtLd];
  float Vs[kAK][kAD];
  float Ps[kAQ][kAK];
  float qrel[kAQ][kAR];   // (q_i/sqrt d) . rel_k[r]
  float band[kAQ][kAR];   // raw score of key i+r-w (or -inf)
  float relk[kAR][kAD];
  float relv[kAR][kAD];
};

struct AttnParams {
  const float* qkv; int qkv_ld;
  float* out; int out_ld;
  void* out16; int out16_ld; int out16_dtype;
  const float* rel_k; const float* rel_v;
  const int32_t* lens;
  int B, T, C, H;
};

__global__ void __launch_bounds__(256) attention_f32_kernel(const AttnParams p) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  AttnSmem& s = *reinterpret_cast<AttnSmem*>(smem_raw);
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int b = blockIdx.z, h = blockIdx.y, q0 = blockIdx.x * kAQ;
  const int len = p.lens ? min(p.lens[b], p.T) : p.T;
  const float* base = p.qkv + static_cast<size_t>(b) * p.T * p.qkv_ld;
  const int qc = h * kAD, kc = p.C + h * kAD, vc = 2 * p.C + h * kAD;
  const float inv = sqrtf(static_cast<float>(kAD));

  // output rows of this CTA that are padding: write zeros and leave early if the whole tile is padding
  if (q0 >= len) {
    for (int i = tid; i < kAQ * kAD; i += 256) {
      const int r = i / kAD, c = i - r * kAD, t = q0 + r;
      if (t < p.T) {
        const size_t row = static_cast<size_t>(b) * p.T + t;
        p.out[row * p.out_ld + qc + c] = 0.f;
        if (p.out16) {
          if (p.out16_dtype == B200SVS_DT_BF16) reinterpret_cast<__nv_bfloat16*>(p.out16)[row * p.out16_ld + qc + c] = cvt_bf16_sat(0.f);
          else reinterpret_cast<__half*>(p.out16)[row * p.out16_ld + qc + c] = cvt_f16_sat(0.f);
        }
      }
    }
    return;
  }

  for (int i = tid; i < kAQ * kAD; i += 256) {
    const int r = i / kAD, c = i - r * kAD, t = q0 + r;
    s.Qs[r][c] = (t < p.T) ? base[static_cast<size_t>(t) * p.qkv_ld + qc + c] / inv : 0.f;
  }
  for (int i = tid; i < kAR * kAD; i += 256) {
    (&s.relk[0][0])[i] = p.rel_k[i];
    (&s.relv[0][0])[i] = p.rel_v[i];
  }
  __syncthreads();
  for (int i = tid; i < kAQ * kAR; i += 256) {
    const int r = i / kAR, w = i - r * kAR;
    float a = 0.f;
    for (int c = 0; c < kAD; ++c) a = fmaf(s.Qs[r][c], s.relk[w][c], a);
    s.qrel[r][w] = a;
    s.band[r][w] = -INFINITY;
  }

  float m_run[4], l_run[4], o[4][6];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    m_run[i] = -INFINITY; l_run[i] = 0.f;
#pragma unroll
    for (int j = 0; j < 6; ++j) o[i][j] = 0.f;
  }

  const int n_ktiles = (len + kAK - 1) / kAK;
  for (int kt = 0; kt < n_ktiles; ++kt) {
    const int k0 = kt * kAK;
    __syncthreads();   // previous tile fully consumed (also orders qrel/band init on the first pass)
    for (int i = tid; i < kAK * kAD; i += 256) {
      const int r = i / kAD, c = i - r * kAD, t = k0 + r;
      const bool ok = t < len;
      const float* rowp = base + static_cast<size_t>(ok ? t : 0) * p.qkv_ld;
      s.Kt[c][r] = ok ? rowp[kc + c] : 0.f;
      s.Vs[r][c] = ok ? rowp[vc + c] : 0.f;
    }
    __syncthreads();

    float sc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) sc[i][j] = 0.f;
#pragma unroll 8
    for (int c = 0; c < kAD; ++c) {
      const float4 kq = *reinterpret_cast<const float4*>(&s.Kt[c][tx * 4]);
      float a[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = s.Qs[ty * 4 + i][c];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        sc[i][0] = fmaf(a[i], kq.x, sc[i][0]); sc[i][1] = fmaf(a[i], kq.y, sc[i][1]);
        sc[i][2] = fmaf(a[i], kq.z, sc[i][2]); sc[i][3] = fmaf(a[i], kq.w, sc[i][3]);
      }
    }
    // relative-key band, padding mask, band bookkeeping, online softmax
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int r = ty * 4 + i, gi = q0 + r;
      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int gj = k0 + tx * 4 + j;
        const int d = gj - gi;
        float v = sc[i][j];
        if (d >= -kAW && d <= kAW) v += s.qrel[r][d + kAW];
        if (gj >= len) v = -1e4f;          // masked_fill(mask == 0, -1e4)
        if (gj >= p.T) v = -INFINITY;       // beyond the tensor: not a key at all
        if (d >= -kAW && d <= kAW && gj < p.T) s.band[r][d + kAW] = v;
        sc[i][j] = v;
        mx = fmaxf(mx, v);
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
      const float m_new = fmaxf(m_run[i], mx);
      const float alpha = (m_run[i] == -INFINITY) ? 0.f : expf(m_run[i] - m_new);
      float sum = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float pv = (sc[i][j] == -INFINITY) ? 0.f : expf(sc[i][j] - m_new);
        s.Ps[r][tx * 4 + j] = pv;
        sum += pv;
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
      l_run[i] = l_run[i] * alpha + sum;
      m_run[i] = m_new;
#pragma unroll
      for (int j = 0; j < 6; ++j) o[i][j] *= alpha;
    }
    __syncthreads();
#pragma unroll 4
    for (int k = 0; k < kAK; ++k) {
      float vv[6];
      const float2 v0 = *reinterpret_cast<const float2*>(&s.Vs[k][tx * 6]);
      const float2 v1 = *reinterpret_cast<const float2*>(&s.Vs[k][tx * 6 + 2]);
      const float2 v2 = *reinterpret_cast<const float2*>(&s.Vs[k][tx * 6 + 4]);
      vv[0] = v0.x; vv[1] = v0.y; vv[2] = v1.x; vv[3] = v1.y; vv[4] = v2.x; vv[5] = v2.y;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float pr = s.Ps[ty * 4 + i][k];
#pragma unroll
        for (int j = 0; j < 6; ++j) o[i][j] = fmaf(pr, vv[j], o[i][j]);
      }
    }
  }
  __syncthreads();   // band[] complete

#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = ty * 4 + i, t = q0 + r;
    if (t >= p.T) continue;
    const bool live = t < len;
    float res[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) res[j] = o[i][j];
    if (live) {
#pragma unroll
      for (int w = 0; w < kAR; ++w) {
        const float sv = s.band[r][w];
        if (sv == -INFINITY) continue;
        const float pw = expf(sv - m_run[i]);
#pragma unroll
        for (int j = 0; j < 6; ++j) res[j] = fmaf(pw, s.relv[w][tx * 6 + j], res[j]);
      }
    }
    const float invl = live ? 1.0f / l_run[i] : 0.f;
    const size_t row = static_cast<size_t>(b) * p.T + t;
#pragma unroll
    for (int j = 0; j < 6; ++j) {
      const float v = live ? res[j] * invl : 0.f;
      const int c = qc + tx * 6 + j;
      p.out[row * p.out_ld + c] = v;
      if (p.out16) {
        if (p.out16_dtype == B200SVS_DT_BF16) reinterpret_cast<__nv_bfloat16*>(p.out16)[row * p.out16_ld + c] = cvt_bf16_sat(v);
        else reinterpret_cast<__half*>(p.out16)[row * p.out16_ld + c] = cvt_f16_sat(v);
      }
    }
  }
}

}  // namespace b200svs

using namespace b200svs;

extern "C" int b200svs_attention_f32(const float* qkv, int32_t qkv_ld, float* out, int32_t out_ld, void* out16,
                                     int32_t out16_ld, int32_t out16_dtype, const float* rel_k,
                                     const float* rel_v, const int32_t* lens, int32_t B, int32_t T, int32_t C,
                                     int32_t H, int32_t window, void* stream) {
  B200SVS_REQUIRE(qkv && out && rel_k && rel_v, "attention_f32: null argument");
  B200SVS_REQUIRE(B > 0 && T > 0, "attention_f32: empty shape B=%d T=%d", B, T);
  B200SVS_REQUIRE(C == H * kAD && window == kAW, "attention_f32: built for head dim 96, window 4 (C=%d H=%d window=%d)", C, H, window);
  AttnParams p;
  p.qkv = qkv; p.qkv_ld = qkv_ld; p.out = out; p.out_ld = out_ld; p.out16 = out16; p.out16_ld = out16_ld;
  p.out16_dtype = out16_dtype; p.rel_k = rel_k; p.rel_v = rel_v; p.lens = lens; p.B = B; p.T = T; p.C = C; p.H = H;
  const size_t smem = sizeof(AttnSmem);
  B200SVS_SMEM_OPTIN(attention_f32_kernel, "cudaFuncSetAttribute(attention_f32)");
  dim3 grid((T + kAQ - 1) / kAQ, H, B);
  attention_f32_kernel<<<grid, 256, smem, as_stream(stream)>>>(p);
  return check_launch("attention_f32_kernel");
}
