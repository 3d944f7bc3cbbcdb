// Bandwidth-bound and integer kernels of the VISinger forward: LayerNorm, embeddings, duration
// quantisation, the length regulator (bit-exact gather), f0 quantisation, prior sampling, conv_post.
// All are coalesced / 128-bit vectorised where the layout allows; none reshapes work into a GEMM.
#include "common.cuh"
#include "api_internal.h"

namespace b200svs {

__device__ __forceinline__ void store16(void* base, size_t idx, float v, int dtype) {
  if (dtype == B200SVS_DT_BF16) reinterpret_cast<__nv_bfloat16*>(base)[idx] = cvt_bf16_sat(v);
  else reinterpret_cast<__half*>(base)[idx] = cvt_f16_sat(v);
}
// 16-bit store of column c of a C-column row at `row_off`; B200SVS_DT_F16_SPLIT writes [hi | lo | hi]
__device__ __forceinline__ void store16c(void* base, size_t row_off, int c, int C, float v, int dtype) {
  if (dtype != B200SVS_DT_F16_SPLIT) { store16(base, row_off + c, v, dtype); return; }
  __half* o = reinterpret_cast<__half*>(base) + row_off;
  const __half hi = cvt_f16_sat(v);
  o[c] = hi;
  o[C + c] = cvt_f16_sat(v - __half2float(hi));
  o[2 * C + c] = hi;
}
__device__ __forceinline__ float load16(const void* base, size_t idx, int dtype) {
  if (dtype == B200SVS_DT_BF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[idx]);
  return __half2float(reinterpret_cast<const __half*>(base)[idx]);
}

// ------------------------------------------------------------------ LayerNorm over channels
// One warp per (b,t) row; C <= 32*8.  Two-pass (mean, then centred variance) like ATen's CPU kernel.
struct LnParams {
  const float* x; int x_ld; const float* res; int res_ld; const float* gamma; const float* beta; float eps;
  const int32_t* lens; float* out32; int out32_ld; void* out16; int out16_ld; int out16_dtype; int B, T, C;
};
__global__ void __launch_bounds__(256) layernorm_kernel(const LnParams p) {
  const int warps_per_block = blockDim.x >> 5;
  const long long row = static_cast<long long>(blockIdx.x) * warps_per_block + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= static_cast<long long>(p.B) * p.T) return;
  const int b = static_cast<int>(row / p.T), t = static_cast<int>(row - static_cast<long long>(b) * p.T);
  const bool live = p.lens ? (t < p.lens[b]) : true;
  float v[8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = lane + i * 32;
    float a = 0.f;
    if (c < p.C) {
      a = p.x[row * p.x_ld + c];
      if (p.res) a += p.res[row * p.res_ld + c];
    }
    v[i] = a;
    sum += a;
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
  const float mean = sum / p.C;
  float var = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = lane + i * 32;
    if (c < p.C) { const float d = v[i] - mean; var += d * d; }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) var += __shfl_xor_sync(0xffffffffu, var, off);
  const float rstd = rsqrtf(var / p.C + p.eps);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = lane + i * 32;
    if (c < p.C) {
      float y = (v[i] - mean) * rstd * p.gamma[c] + p.beta[c];
      if (!live) y = 0.f;
      if (p.out32) p.out32[row * p.out32_ld + c] = y;
      if (p.out16) store16c(p.out16, row * p.out16_ld, c, p.C, y, p.out16_dtype);
    }
  }
}

// Vectorised variant for C % 4 == 0 (C <= 256) with 16-byte aligned rows: lane l owns channels [4l, 4l+4) and [128 + 4l, ..),
// float4 loads / stores, 8-byte 16-bit stores.  The scalar kernel above spent 423 warp instructions per row and ran with its
// issue slots 82 % busy at 4.8 TB/s (ncu); same two-pass arithmetic, the partial sums are grouped per lane differently.
__global__ void __launch_bounds__(256) layernorm_v4_kernel(const LnParams p) {
  const long long row = static_cast<long long>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= static_cast<long long>(p.B) * p.T) return;
  const int b = static_cast<int>(row / p.T), t = static_cast<int>(row - static_cast<long long>(b) * p.T);
  const bool live = p.lens ? (t < __ldg(p.lens + b)) : true;
  float4 v[2];
  bool on[2];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int c = 4 * lane + 128 * i;
    on[i] = c < p.C;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    if (on[i]) {
      a = __ldg(reinterpret_cast<const float4*>(p.x + row * p.x_ld + c));
      if (p.res) {
        const float4 r = __ldg(reinterpret_cast<const float4*>(p.res + row * p.res_ld + c));
        a.x += r.x; a.y += r.y; a.z += r.z; a.w += r.w;
      }
    }
    v[i] = a;
    sum += (a.x + a.y) + (a.z + a.w);
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, off);
  const float mean = sum / p.C;
  float var = 0.f;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    if (on[i]) {
      const float d0 = v[i].x - mean, d1 = v[i].y - mean, d2 = v[i].z - mean, d3 = v[i].w - mean;
      var += (d0 * d0 + d1 * d1) + (d2 * d2 + d3 * d3);
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) var += __shfl_xor_sync(0xffffffffu, var, off);
  const float rstd = rsqrtf(var / p.C + p.eps);
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    if (!on[i]) continue;
    const int c = 4 * lane + 128 * i;
    const float4 g = __ldg(reinterpret_cast<const float4*>(p.gamma + c)), be = __ldg(reinterpret_cast<const float4*>(p.beta + c));
    float4 y;
    y.x = live ? (v[i].x - mean) * rstd * g.x + be.x : 0.f;
    y.y = live ? (v[i].y - mean) * rstd * g.y + be.y : 0.f;
    y.z = live ? (v[i].z - mean) * rstd * g.z + be.z : 0.f;
    y.w = live ? (v[i].w - mean) * rstd * g.w + be.w : 0.f;
    if (p.out32) *reinterpret_cast<float4*>(p.out32 + row * p.out32_ld + c) = y;
    if (p.out16) {
      uint16_t* o = reinterpret_cast<uint16_t*>(p.out16) + row * p.out16_ld + c;
      if (p.out16_dtype == B200SVS_DT_BF16) {
        *reinterpret_cast<uint2*>(o) = make_uint2(cvt_bf16x2_sat(y.x, y.y), cvt_bf16x2_sat(y.z, y.w));
      } else {
        const uint2 hi = make_uint2(cvt_f16x2_sat(y.x, y.y), cvt_f16x2_sat(y.z, y.w));
        *reinterpret_cast<uint2*>(o) = hi;
        if (p.out16_dtype == B200SVS_DT_F16_SPLIT) {
          const float2 h0 = __half22float2(*reinterpret_cast<const __half2*>(&hi.x));
          const float2 h1 = __half22float2(*reinterpret_cast<const __half2*>(&hi.y));
          *reinterpret_cast<uint2*>(o + p.C) = make_uint2(cvt_f16x2_sat(y.x - h0.x, y.y - h0.y), cvt_f16x2_sat(y.z - h1.x, y.w - h1.y));
          *reinterpret_cast<uint2*>(o + 2 * p.C) = hi;
        }
      }
    }
  }
}

// ------------------------------------------------------------------ TextEncoder input embedding
__global__ void text_embed_kernel(const int64_t* phon, const int64_t* pitch, const int64_t* tone, const float* dur,
                                  const float* e_ph, const float* e_pitch, const float* e_tone, const float* w_dur,
                                  const float* b_dur, const int64_t* x_lengths, int32_t* lens32, float* out, int B,
                                  int P, int C, float scale, int n_ph, int n_pitch, int n_tone, int32_t* status) {
  const int bp = blockIdx.x;
  const int b = bp / P, i = bp - b * P;
  const int len = static_cast<int>(x_lengths[b]);
  if (i == 0 && threadIdx.x == 0 && lens32) lens32[b] = len;
  const bool live = i < len;
  long long ip = phon[bp], it = pitch[bp], is = tone[bp];
  // nn.Embedding raises on an id outside its table; here a live out-of-range id sets status[1] bit 0 (the host raises
  // at its next sync) and is clamped so that the read stays inside the table
  const bool bad = live && (ip < 0 || ip >= n_ph || it < 0 || it >= n_pitch || is < 0 || is >= n_tone);
  if (bad && threadIdx.x == 0 && status) atomicOr(status + 1, 1);
  ip = ip < 0 ? 0 : (ip >= n_ph ? n_ph - 1 : ip);
  it = it < 0 ? 0 : (it >= n_pitch ? n_pitch - 1 : it);
  is = is < 0 ? 0 : (is >= n_tone ? n_tone - 1 : is);
  const float d = dur[bp];
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    // (emb_phoneme + emb_note_duration + emb_note_pitch + emb_phone_shengdiao) * sqrt(C), same add order
    // as models_ae.py:213;  Linear(1->C): d * w[c] + b[c]
    float v = e_ph[ip * C + c] + (d * w_dur[c] + b_dur[c]);
    v = v + e_pitch[it * C + c];
    v = v + e_tone[is * C + c];
    v = v * scale;
    out[static_cast<size_t>(bp) * C + c] = live ? v : 0.f;
  }
}

// ------------------------------------------------------------------ duration quantisation
__global__ void duration_quantize_kernel(const float* pred, const float* note_dur, int64_t* out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  // torch: floor((pred * note_dur) / 0.0116) in fp32 (0.0116 rounded to fp32), no FMA contraction
  const float prod = __fmul_rn(pred[i], note_dur[i]);
  const float q = __fdiv_rn(prod, 0.0116f);
  out[i] = static_cast<int64_t>(floorf(q));
}

// ------------------------------------------------------------------ length regulator
// scan: one warp per utterance, inclusive prefix sum of clamped durations.
__global__ void lr_scan_kernel(const int64_t* dur, int B, int P, int32_t* ends, int64_t* mel64, int32_t* mel32,
                               int32_t* status) {
  const int b = blockIdx.x;
  const int lane = threadIdx.x;
  int carry = 0;
  for (int base = 0; base < P; base += 32) {
    const int i = base + lane;
    long long dl = (i < P) ? dur[static_cast<size_t>(b) * P + i] : 0;
    int d = dl > 0 ? static_cast<int>(dl) : 0;
    int x = d;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, off);
      if (lane >= off) x += y;
    }
    if (i < P) ends[static_cast<size_t>(b) * P + i] = carry + x;
    carry += __shfl_sync(0xffffffffu, x, 31);
  }
  if (lane == 0) {
    if (mel64) mel64[b] = carry;
    if (mel32) mel32[b] = carry;
    if (status) atomicMax(status, carry);      // status[0] = max_b mel_len[b]: the padded frame count T
  }
}
// gather: one warp per output frame; binary search of t in ends[b,:], then a 128-bit row copy.
__global__ void __launch_bounds__(256) lr_gather_kernel(const float* h, int h_ld, const int32_t* ends,
                                                        const int32_t* mel32, int B, int P, int T, int C, float* out,
                                                        int out_ld, int32_t* index_out) {
  const long long row = static_cast<long long>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= static_cast<long long>(B) * T) return;
  const int b = static_cast<int>(row / T), t = static_cast<int>(row - static_cast<long long>(b) * T);
  const int32_t* e = ends + static_cast<size_t>(b) * P;
  int idx = -1;
  if (t < mel32[b]) {
    int lo = 0, hi = P - 1;   // first i with ends[i] > t
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (e[mid] > t) hi = mid; else lo = mid + 1;
    }
    idx = lo;
  }
  if (lane == 0 && index_out) index_out[row] = idx;
  float* o = out + row * out_ld;
  if (idx < 0) {
    for (int c = lane; c < C; c += 32) o[c] = 0.f;
  } else {
    const float* src = h + (static_cast<size_t>(b) * P + idx) * h_ld;
    for (int c = lane; c < C; c += 32) o[c] = src[c];
  }
}

// ------------------------------------------------------------------ (x + pos[t]) * mask
__global__ void add_pos_mask_kernel(const float* x, int x_ld, const float* pos, const int32_t* lens, float* out32,
                                    int out32_ld, void* out16, int out16_ld, int dt, int B, int T, int C) {
  const long long n = static_cast<long long>(B) * T * C;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long row = i / C;
    const int c = static_cast<int>(i - row * C);
    const int b = static_cast<int>(row / T), t = static_cast<int>(row - static_cast<long long>(b) * T);
    float v = x[row * x_ld + c] + pos[static_cast<size_t>(t) * C + c];
    if (lens && t >= lens[b]) v = 0.f;
    if (out32) out32[row * out32_ld + c] = v;
    if (out16) store16c(out16, row * out16_ld, c, C, v, dt);
  }
}

// ------------------------------------------------------------------ f0_to_coarse + embedding gather
__device__ __forceinline__ long long f0_coarse_of(float lf0) {
  // utils/pitch_utils.py:22-31 on f0 = exp(log_f0) - 1, fp32 op for op
  const float f0 = expf(lf0) - 1.0f;
  float mel = 1127.0f * logf(1.0f + f0 / 700.0f);
  const float mel_min = 1127.0 * 0.06899287148695143;   // ln(1 + 50/700)
  const float denom = static_cast<float>(1127.0 * 0.9444616088408514 - 1127.0 * 0.06899287148695143);
  if (mel > 0.f) mel = (mel - mel_min) * 254.0f / denom + 1.0f;
  if (mel <= 1.f) mel = 1.f;
  if (mel > 255.f) mel = 255.f;
  return static_cast<long long>(mel + 0.5f);
}
__global__ void __launch_bounds__(256) f0_embed_kernel(const float* log_f0, const int64_t* coarse_in, const float* f0_emb,
                                                       const int32_t* lens, int64_t* coarse_out, float* out32,
                                                       int out32_ld, void* out16, int out16_ld, int dt, int B, int T,
                                                       int C, int n_rows) {
  const long long row = static_cast<long long>(blockIdx.x) * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= static_cast<long long>(B) * T) return;
  const int b = static_cast<int>(row / T), t = static_cast<int>(row - static_cast<long long>(b) * T);
  const bool live = lens ? t < lens[b] : true;
  long long k = coarse_in ? coarse_in[row] : f0_coarse_of(log_f0[row]);
  if (lane == 0 && coarse_out) coarse_out[row] = k;
  k = k < 0 ? 0 : (k >= n_rows ? n_rows - 1 : k);      // a user f0_score outside the table (validated on the host) never reads out of bounds
  const float* src = f0_emb + static_cast<size_t>(k) * C;
  for (int c = lane; c < C; c += 32) {
    const float v = live ? src[c] : 0.f;
    if (out32) out32[row * out32_ld + c] = v;
    if (out16) store16(out16, row * out16_ld + c, v, dt);
  }
}

// ------------------------------------------------------------------ z_p = m + noise*exp(logs)*scale
// stats is channels-last [B,T,2C]; noise is the reference layout [B,C,T] (tile transpose via smem).
__global__ void __launch_bounds__(256) sample_prior_kernel(const float* stats, int stats_ld, const float* noise,
                                                           float noise_scale, float* zp32, int zp_ld, void* zp16,
                                                           int zp16_ld, int dt, int B, int T, int C,
                                                           const int32_t* lens) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z, t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
  for (int j = ty; j < 32; j += 8) {
    const int c = c0 + j, t = t0 + tx;
    tile[j][tx] = (c < C && t < T) ? noise[(static_cast<size_t>(b) * C + c) * T + t] : 0.f;
  }
  __syncthreads();
  for (int j = ty; j < 32; j += 8) {
    const int t = t0 + j, c = c0 + tx;
    if (t < T && c < C) {
      const size_t row = static_cast<size_t>(b) * T + t;
      const float m = stats[row * stats_ld + c], logs = stats[row * stats_ld + C + c];
      // torch: m_p + randn * exp(logs_p) * noise_scale  (left to right)
      float v = m + (tile[tx][j] * expf(logs)) * noise_scale;
      if (lens && t >= lens[b]) v = 0.f;      // PosteriorEncoder: (m + randn * exp(logs)) * x_mask (models_ae.py:290)
      if (zp32) zp32[row * zp_ld + c] = v;
      if (zp16) store16(zp16, row * zp16_ld + c, v, dt);
    }
  }
}

// ------------------------------------------------------------------ conv_post (C -> 1, k taps) + tanh
// A block owns kPostTile consecutive input rows of one utterance (kPostTile - (k - 1) outputs).  Every input row (C contiguous 16-bit
// channels, 128-bit loads) is read ONCE: its thread forms the k partial dot products row . w[tap] and leaves
// them in shared memory; after a barrier each output sample sums its k partials (tap j of sample t is the
// partial j of row t + j - pad) and applies tanh.  Weights are broadcast reads from shared memory.
constexpr int kPostThreads = 128;
constexpr int kPostRows = 4;                          // input rows per thread: every weight read from shared memory feeds 4 rows
constexpr int kPostTile = kPostThreads * kPostRows;   // input rows per block
constexpr int kPostMaxTaps = 16;
constexpr int kPostMaxC = 64;
// KT: compile-time tap count (7 for the shipped conv_post), 0 = runtime `ktaps`.  The first version ran one row per thread:
// 823 instructions per warp-row for 224 FMAs, the L1 / shared pipe 83 % busy with the broadcast weight reads (ncu).
template <bool BF, int KT>
__global__ void __launch_bounds__(kPostThreads) conv_post_tanh_kernel(const void* x, int x_ld, const float* w, int B, int L,
                                                                       int C, int ktaps_rt, int tiles_per_row, float* out) {
  extern __shared__ float smem_post[];
  const int ktaps = KT ? KT : ktaps_rt;
  constexpr int kAcc = KT ? KT : kPostMaxTaps;
  float* ws = smem_post;                               // [ktaps][C]
  float* part = ws + ktaps * C;                        // [ktaps][kPostTile + kPostMaxTaps]
  constexpr int kStride = kPostTile + kPostMaxTaps;
  for (int i = threadIdx.x; i < ktaps * C; i += blockDim.x) ws[i] = w[i];
  __syncthreads();
  // a block covers kPostTile input rows = kPostTile - (ktaps - 1) outputs; thread r owns rows r, r + 128, r + 256, r + 384
  const int pad = ktaps / 2;
  const int outs = kPostTile - (ktaps - 1);
  const int b = blockIdx.x / tiles_per_row;
  const int t0 = (blockIdx.x - b * tiles_per_row) * outs;
  const uint16_t* xb = reinterpret_cast<const uint16_t*>(x) + static_cast<size_t>(b) * L * x_ld;
  {
    const int r = threadIdx.x;
    float acc[kPostRows][kAcc];
    const uint4* src[kPostRows];
    bool ok[kPostRows];
#pragma unroll
    for (int i = 0; i < kPostRows; ++i) {
      const int t = t0 - pad + r + i * kPostThreads;
      ok[i] = t >= 0 && t < L;
      src[i] = reinterpret_cast<const uint4*>(xb + static_cast<size_t>(ok[i] ? t : 0) * x_ld);
#pragma unroll
      for (int k = 0; k < kAcc; ++k) acc[i][k] = 0.f;
    }
    const int n8 = C / 8;
    for (int c8 = 0; c8 < n8; ++c8) {
      float xv[kPostRows][8];
#pragma unroll
      for (int i = 0; i < kPostRows; ++i) {
        const uint4 q = ok[i] ? __ldg(src[i] + c8) : make_uint4(0, 0, 0, 0);      // rows outside the sequence: zeros (conv padding)
        const uint32_t u[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (BF) {
            xv[i][2 * j] = __uint_as_float(u[j] << 16); xv[i][2 * j + 1] = __uint_as_float(u[j] & 0xFFFF0000u);
          } else {
            const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&u[j]));
            xv[i][2 * j] = f.x; xv[i][2 * j + 1] = f.y;
          }
        }
      }
#pragma unroll
      for (int k = 0; k < kAcc; ++k) {
        if (KT || k < ktaps) {
          const float4 w0 = *reinterpret_cast<const float4*>(ws + k * C + c8 * 8);
          const float4 w1 = *reinterpret_cast<const float4*>(ws + k * C + c8 * 8 + 4);
#pragma unroll
          for (int i = 0; i < kPostRows; ++i) {
            float a = acc[i][k];
            a = fmaf(xv[i][0], w0.x, a); a = fmaf(xv[i][1], w0.y, a); a = fmaf(xv[i][2], w0.z, a); a = fmaf(xv[i][3], w0.w, a);
            a = fmaf(xv[i][4], w1.x, a); a = fmaf(xv[i][5], w1.y, a); a = fmaf(xv[i][6], w1.z, a); a = fmaf(xv[i][7], w1.w, a);
            acc[i][k] = a;
          }
        }
      }
    }
#pragma unroll
    for (int i = 0; i < kPostRows; ++i)
#pragma unroll
      for (int k = 0; k < kAcc; ++k)
        if (KT || k < ktaps) part[k * kStride + r + i * kPostThreads] = acc[i][k];
  }
  __syncthreads();
#pragma unroll
  for (int i = 0; i < kPostRows; ++i) {
    const int o = threadIdx.x + i * kPostThreads;
    const int t = t0 + o;
    if (o < outs && t < L) {
      float v = 0.f;
      for (int k = 0; k < ktaps; ++k) v += part[k * kStride + o + k];     // same tap order as a direct conv
      out[static_cast<size_t>(b) * L + t] = tanhf(v);
    }
  }
}

// ------------------------------------------------------------------ layout helpers (32x32 smem transpose)
__global__ void __launch_bounds__(256) bct_to_btc_kernel(const float* in, int B, int C, int T, float* out32, int out32_ld,
                                                         void* out16, int out16_ld, int dt, const int32_t* lens) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z, t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int j = ty; j < 32; j += 8) {
    const int c = c0 + j, t = t0 + tx;
    tile[j][tx] = (c < C && t < T) ? in[(static_cast<size_t>(b) * C + c) * T + t] : 0.f;
  }
  __syncthreads();
  for (int j = ty; j < 32; j += 8) {
    const int t = t0 + j, c = c0 + tx;
    if (t < T && c < C) {
      float v = tile[tx][j];
      if (lens && t >= lens[b]) v = 0.f;
      const size_t row = static_cast<size_t>(b) * T + t;
      if (out32) out32[row * out32_ld + c] = v;
      if (out16) store16(out16, row * out16_ld + c, v, dt);
    }
  }
}
__global__ void __launch_bounds__(256) btc_to_bct_kernel(const float* in, int in_ld, int B, int C, int T, float* out) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z, t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int j = ty; j < 32; j += 8) {
    const int t = t0 + j, c = c0 + tx;
    tile[j][tx] = (c < C && t < T) ? in[(static_cast<size_t>(b) * T + t) * in_ld + c] : 0.f;
  }
  __syncthreads();
  for (int j = ty; j < 32; j += 8) {
    const int c = c0 + j, t = t0 + tx;
    if (t < T && c < C) out[(static_cast<size_t>(b) * C + c) * T + t] = tile[tx][j];
  }
}

static inline int grid_1d(long long n, int block, int cap) {
  long long g = (n + block - 1) / block;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return static_cast<int>(g);
}

}  // namespace b200svs

using namespace b200svs;

extern "C" {

int b200svs_layernorm(const float* x, int32_t x_ld, const float* res, int32_t res_ld, const float* gamma,
                      const float* beta, float eps, const int32_t* lens, float* out32, int32_t out32_ld, void* out16,
                      int32_t out16_ld, int32_t out16_dtype, int32_t B, int32_t T, int32_t C, void* stream) {
  B200SVS_REQUIRE(x && gamma && beta && (out32 || out16), "layernorm: null argument");
  B200SVS_REQUIRE(B > 0 && T > 0 && C > 0 && C <= 256, "layernorm: B=%d T=%d C=%d (C <= 256)", B, T, C);
  LnParams p{x, x_ld, res, res_ld, gamma, beta, eps, lens, out32, out32_ld, out16, out16_ld, out16_dtype, B, T, C};
  const long long rows = static_cast<long long>(B) * T;
  auto al = [](const void* q, int a) { return (reinterpret_cast<uintptr_t>(q) & (a - 1)) == 0; };
  const bool v4 = C % 4 == 0 && x_ld % 4 == 0 && al(x, 16) && (!res || (res_ld % 4 == 0 && al(res, 16))) && al(gamma, 16) && al(beta, 16) &&
                  (!out32 || (out32_ld % 4 == 0 && al(out32, 16))) && (!out16 || (out16_ld % 4 == 0 && al(out16, 8)));
  if (v4) layernorm_v4_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, as_stream(stream)>>>(p);
  else layernorm_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, as_stream(stream)>>>(p);
  return check_launch("layernorm_kernel");
}

int b200svs_text_embed(const int64_t* phon, const int64_t* pitch, const int64_t* tone, const float* dur,
                       const float* e_ph, const float* e_pitch, const float* e_tone, const float* w_dur,
                       const float* b_dur, const int64_t* x_lengths, int32_t* lens32_out, float* out, int32_t B,
                       int32_t P, int32_t C, int32_t n_ph, int32_t n_pitch, int32_t n_tone, int32_t* status, void* stream) {
  B200SVS_REQUIRE(phon && pitch && tone && dur && e_ph && e_pitch && e_tone && w_dur && b_dur && x_lengths && out,
                  "text_embed: null argument");
  B200SVS_REQUIRE(B > 0 && P > 0 && C > 0, "text_embed: B=%d P=%d C=%d", B, P, C);
  B200SVS_REQUIRE(n_ph > 0 && n_pitch > 0 && n_tone > 0, "text_embed: table sizes %d %d %d", n_ph, n_pitch, n_tone);
  text_embed_kernel<<<B * P, 64, 0, as_stream(stream)>>>(phon, pitch, tone, dur, e_ph, e_pitch, e_tone, w_dur, b_dur,
                                                        x_lengths, lens32_out, out, B, P, C, sqrtf((float)C), n_ph, n_pitch,
                                                        n_tone, status);
  return check_launch("text_embed_kernel");
}

int b200svs_duration_quantize(const float* pred, const float* note_dur, int64_t* dur_out, int32_t n, void* stream) {
  B200SVS_REQUIRE(pred && note_dur && dur_out && n > 0, "duration_quantize: bad argument");
  duration_quantize_kernel<<<(n + 255) / 256, 256, 0, as_stream(stream)>>>(pred, note_dur, dur_out, n);
  return check_launch("duration_quantize_kernel");
}

int b200svs_lr_scan(const int64_t* dur, int32_t B, int32_t P, int32_t* ends, int64_t* mel_len64, int32_t* mel_len32,
                    int32_t* status, void* stream) {
  B200SVS_REQUIRE(dur && ends && B > 0 && P > 0, "lr_scan: bad argument");
  lr_scan_kernel<<<B, 32, 0, as_stream(stream)>>>(dur, B, P, ends, mel_len64, mel_len32, status);
  return check_launch("lr_scan_kernel");
}

int b200svs_lr_gather(const float* h, int32_t h_ld, const int32_t* ends, const int32_t* mel_len32, int32_t B,
                      int32_t P, int32_t T, int32_t C, float* out, int32_t out_ld, int32_t* index_out, void* stream) {
  B200SVS_REQUIRE(h && ends && mel_len32 && out, "lr_gather: null argument");
  B200SVS_REQUIRE(B > 0 && P > 0 && C > 0, "lr_gather: B=%d P=%d C=%d", B, P, C);
  if (T <= 0) return B200SVS_OK;   // every utterance has zero frames
  const long long rows = static_cast<long long>(B) * T;
  lr_gather_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, as_stream(stream)>>>(h, h_ld, ends, mel_len32, B, P,
                                                                                       T, C, out, out_ld, index_out);
  return check_launch("lr_gather_kernel");
}

int b200svs_add_pos_mask(const float* x, int32_t x_ld, const float* pos, const int32_t* lens, float* out32,
                         int32_t out32_ld, void* out16, int32_t out16_ld, int32_t out16_dtype, int32_t B, int32_t T,
                         int32_t C, void* stream) {
  B200SVS_REQUIRE(x && pos && (out32 || out16) && B > 0 && T > 0 && C > 0, "add_pos_mask: bad argument");
  const long long n = static_cast<long long>(B) * T * C;
  add_pos_mask_kernel<<<grid_1d(n, 256, num_sms() * 16), 256, 0, as_stream(stream)>>>(x, x_ld, pos, lens, out32, out32_ld,
                                                                                     out16, out16_ld, out16_dtype, B, T, C);
  return check_launch("add_pos_mask_kernel");
}

int b200svs_f0_embed(const float* log_f0, const int64_t* coarse_in, const float* f0_emb, const int32_t* lens,
                     int64_t* coarse_out, float* out32, int32_t out32_ld, void* out16, int32_t out16_ld,
                     int32_t out16_dtype, int32_t B, int32_t T, int32_t C, int32_t n_rows, void* stream) {
  B200SVS_REQUIRE((log_f0 || coarse_in) && f0_emb && (out32 || out16) && B > 0 && T > 0 && n_rows > 0, "f0_embed: bad argument");
  const long long rows = static_cast<long long>(B) * T;
  f0_embed_kernel<<<static_cast<unsigned>((rows + 7) / 8), 256, 0, as_stream(stream)>>>(
      log_f0, coarse_in, f0_emb, lens, coarse_out, out32, out32_ld, out16, out16_ld, out16_dtype, B, T, C, n_rows);
  return check_launch("f0_embed_kernel");
}

int b200svs_sample_prior(const float* stats, int32_t stats_ld, const float* noise_bct, float noise_scale, float* zp32,
                         int32_t zp_ld, void* zp16, int32_t zp16_ld, int32_t out16_dtype, int32_t B, int32_t T,
                         int32_t C, const int32_t* lens, void* stream) {
  B200SVS_REQUIRE(stats && noise_bct && (zp32 || zp16) && B > 0 && T > 0 && C > 0, "sample_prior: bad argument");
  dim3 grid((T + 31) / 32, (C + 31) / 32, B);
  sample_prior_kernel<<<grid, 256, 0, as_stream(stream)>>>(stats, stats_ld, noise_bct, noise_scale, zp32, zp_ld, zp16,
                                                          zp16_ld, out16_dtype, B, T, C, lens);
  return check_launch("sample_prior_kernel");
}

int b200svs_conv_post_tanh(const void* x, int32_t x_dtype, int32_t x_ld, const float* w, int32_t B, int32_t L,
                           int32_t C, int32_t ktaps, float* out, void* stream) {
  B200SVS_REQUIRE(x && w && out && B > 0 && L > 0, "conv_post_tanh: bad argument");
  B200SVS_REQUIRE(C % 8 == 0 && x_ld % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0,
                  "conv_post_tanh: C=%d ld=%d must be multiples of 8, x 16-B aligned", C, x_ld);
  B200SVS_REQUIRE(ktaps >= 1 && ktaps <= kPostMaxTaps && C <= kPostMaxC, "conv_post_tanh: ktaps=%d (<= %d), C=%d (<= %d)",
                  ktaps, kPostMaxTaps, C, kPostMaxC);
  const int outs = kPostTile - (ktaps - 1);
  const int tiles_per_row = (L + outs - 1) / outs;
  const size_t smem = (static_cast<size_t>(ktaps) * C + static_cast<size_t>(ktaps) * (kPostTile + kPostMaxTaps)) * sizeof(float);
  const unsigned grid = static_cast<unsigned>(B) * tiles_per_row;
  if (x_dtype == B200SVS_DT_BF16) {
    if (ktaps == 7) conv_post_tanh_kernel<true, 7><<<grid, kPostThreads, smem, as_stream(stream)>>>(x, x_ld, w, B, L, C, ktaps, tiles_per_row, out);
    else conv_post_tanh_kernel<true, 0><<<grid, kPostThreads, smem, as_stream(stream)>>>(x, x_ld, w, B, L, C, ktaps, tiles_per_row, out);
  } else {
    if (ktaps == 7) conv_post_tanh_kernel<false, 7><<<grid, kPostThreads, smem, as_stream(stream)>>>(x, x_ld, w, B, L, C, ktaps, tiles_per_row, out);
    else conv_post_tanh_kernel<false, 0><<<grid, kPostThreads, smem, as_stream(stream)>>>(x, x_ld, w, B, L, C, ktaps, tiles_per_row, out);
  }
  return check_launch("conv_post_tanh_kernel");
}

int b200svs_bct_to_btc(const float* in_bct, int32_t B, int32_t C, int32_t T, float* out32, int32_t out32_ld,
                       void* out16, int32_t out16_ld, int32_t out16_dtype, const int32_t* lens, void* stream) {
  B200SVS_REQUIRE(in_bct && (out32 || out16) && B > 0 && C > 0 && T > 0, "bct_to_btc: bad argument");
  dim3 grid((T + 31) / 32, (C + 31) / 32, B);
  bct_to_btc_kernel<<<grid, 256, 0, as_stream(stream)>>>(in_bct, B, C, T, out32, out32_ld, out16, out16_ld, out16_dtype, lens);
  return check_launch("bct_to_btc_kernel");
}

int b200svs_btc_to_bct(const float* in_btc, int32_t in_ld, int32_t B, int32_t C, int32_t T, float* out_bct,
                       void* stream) {
  B200SVS_REQUIRE(in_btc && out_bct && B > 0 && C > 0 && T > 0, "btc_to_bct: bad argument");
  dim3 grid((T + 31) / 32, (C + 31) / 32, B);
  btc_to_bct_kernel<<<grid, 256, 0, as_stream(stream)>>>(in_btc, in_ld, B, C, T, out_bct);
  return check_launch("btc_to_bct_kernel");
}

}  // extern "C"
