// conv1d_f32: fp32 SIMT implicit-GEMM 1-D convolution (channels-last), same fused epilogue contract as
// conv1d_tc (GATE excluded).  Used where the reference's float->int boundaries demand fp32 arithmetic:
// TextEncoder / DurationPredictor (floor(), VISinger/models_ae.py:624-625) and the F0 predictor
// (f0_to_coarse, VISinger/utils/pitch_utils.py:22-31) - SURVEY.md 7.3-B/C.
//
// Tile: 64 time rows x 64 output channels per CTA (256 threads, 4x4 outputs per thread), K step =
// 16 input channels of one tap staged in shared memory.  w is packed [ntaps][Cin][Cout_pad].
#include "common.cuh"
#include "api_internal.h"

namespace b200svs {

constexpr int kFM = 64, kFN = 64, kFK = 16;

struct ConvF32Params {
  const float* x;
  const float* w;
  int B, T, Cin, x_ld, ntaps, cout_pad, dil, pad_left, tiles_t;
  b200svs_epilogue ep;
};

__device__ __forceinline__ float act_f32(float v, int act) {
  switch (act) {
    case B200SVS_ACT_RELU: return fmaxf(v, 0.f);
    case B200SVS_ACT_GELU: return gelu_erf(v);
    case B200SVS_ACT_SIGMOID: return 1.0f / (1.0f + expf(-v));
    case B200SVS_ACT_TANH: return tanhf(v);
    default: return v;
  }
}

__global__ void __launch_bounds__(256) conv1d_f32_kernel(const ConvF32Params p) {
  __shared__ float Xs[kFM][kFK + 1];
  __shared__ __align__(16) float Ws[kFK][kFN];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int b = blockIdx.x / p.tiles_t;
  const int t0 = (blockIdx.x - b * p.tiles_t) * kFM;
  const int n0 = blockIdx.y * kFN;
  const float* xb = p.x + static_cast<size_t>(b) * p.T * p.x_ld;

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  const int lr = tid >> 2;          // X loader: row in tile
  const int lc = (tid & 3) * 4;     // X loader: first channel of 4
  const int wr = tid >> 4;          // W loader: k row
  const int wc = (tid & 15) * 4;    // W loader: first column of 4

  for (int tap = 0; tap < p.ntaps; ++tap) {
    const int shift = tap * p.dil - p.pad_left;
    const int tr = t0 + lr + shift;
    const bool rok = tr >= 0 && tr < p.T;
    const float* wt = p.w + static_cast<size_t>(tap) * p.Cin * p.cout_pad;
    for (int c0 = 0; c0 < p.Cin; c0 += kFK) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int c = c0 + lc + i;
        Xs[lr][lc + i] = (rok && c < p.Cin) ? xb[static_cast<size_t>(tr) * p.x_ld + c] : 0.f;
      }
      {
        const int c = c0 + wr;
        float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c < p.Cin) q = *reinterpret_cast<const float4*>(wt + static_cast<size_t>(c) * p.cout_pad + n0 + wc);
        *reinterpret_cast<float4*>(&Ws[wr][wc]) = q;
      }
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < kFK; ++kk) {
        const float4 bq = *reinterpret_cast<const float4*>(&Ws[kk][tx * 4]);
        float a[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = Xs[ty * 4 + i][kk];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          acc[i][0] = fmaf(a[i], bq.x, acc[i][0]);
          acc[i][1] = fmaf(a[i], bq.y, acc[i][1]);
          acc[i][2] = fmaf(a[i], bq.z, acc[i][2]);
          acc[i][3] = fmaf(a[i], bq.w, acc[i][3]);
        }
      }
      __syncthreads();
    }
  }

  const b200svs_epilogue& ep = p.ep;
  const int col0 = n0 + tx * 4;
  const int len_b = ep.lens ? ep.lens[b] : p.T;
  const float* gb = ep.gbias ? ep.gbias + static_cast<size_t>(b) * ep.gbias_ld : nullptr;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int t = t0 + ty * 4 + i;
    if (t >= p.T) continue;
    const size_t row = static_cast<size_t>(b) * p.T + t;
    const float s = (t < len_b) ? ep.scale : 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = col0 + j;
      if (c >= ep.n_valid) continue;
      float v = acc[i][j];
      if (ep.bias) v += ep.bias[c];
      if (gb) v += gb[c];
      v = act_f32(v, ep.act);
      v *= ep.alpha;
      if (ep.res1) v += ep.res1[row * ep.res1_ld + c];
      if (ep.res2) v += ep.res2[row * ep.res2_ld + c];
      v *= s;
      if (ep.out32) ep.out32[row * ep.out32_ld + c] = v;
      if (ep.out16) {
        const float o = ep.post_lrelu ? lrelu(v, ep.post_slope) : v;
        if (ep.out16_dtype == B200SVS_DT_BF16)
          reinterpret_cast<__nv_bfloat16*>(ep.out16)[row * ep.out16_ld + c] = cvt_bf16_sat(o);
        else
          reinterpret_cast<__half*>(ep.out16)[row * ep.out16_ld + c] = cvt_f16_sat(o);
      }
    }
  }
}

}  // namespace b200svs

using namespace b200svs;

extern "C" int b200svs_conv1d_f32(const b200svs_conv_args* a, void* stream) {
  B200SVS_REQUIRE(a && a->x && a->w, "conv1d_f32: null argument");
  B200SVS_REQUIRE(a->B > 0 && a->T > 0 && a->Cin > 0, "conv1d_f32: empty shape B=%d T=%d Cin=%d", a->B, a->T, a->Cin);
  B200SVS_REQUIRE(a->Cout_pad % kFN == 0, "conv1d_f32: Cout_pad=%d must be a multiple of 64", a->Cout_pad);
  B200SVS_REQUIRE((reinterpret_cast<uintptr_t>(a->w) & 15) == 0, "conv1d_f32: w must be 16-B aligned");
  B200SVS_REQUIRE(a->ep.act != B200SVS_ACT_GATE, "conv1d_f32: GATE epilogue is tensor-core only");
  B200SVS_REQUIRE(a->ep.out32 || a->ep.out16, "conv1d_f32: no output");
  B200SVS_REQUIRE(!a->ep.out16 || a->ep.out16_dtype != B200SVS_DT_F16_SPLIT, "conv1d_f32: split 16-bit output is tensor-core only");
  B200SVS_REQUIRE(a->ep.n_valid > 0 && a->ep.n_valid <= a->Cout_pad, "conv1d_f32: n_valid=%d", a->ep.n_valid);
  ConvF32Params p;
  p.x = static_cast<const float*>(a->x);
  p.w = static_cast<const float*>(a->w);
  p.B = a->B; p.T = a->T; p.Cin = a->Cin; p.x_ld = a->x_ld; p.ntaps = a->ntaps; p.cout_pad = a->Cout_pad;
  p.dil = a->dil; p.pad_left = a->pad_left;
  p.tiles_t = (a->T + kFM - 1) / kFM;
  p.ep = a->ep;
  dim3 grid(p.tiles_t * a->B, (a->ep.n_valid + kFN - 1) / kFN);
  conv1d_f32_kernel<<<grid, 256, 0, as_stream(stream)>>>(p);
  return check_launch("conv1d_f32_kernel");
}
