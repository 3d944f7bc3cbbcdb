// conv1d_tc: implicit-GEMM 1-D convolution on the 5th-gen tensor cores (tcgen05 + TMEM), operands
// staged by TMA (SWIZZLE_128B, K-major), fp16/bf16 in, fp32 accumulate, fused epilogue.
//
//   D[t, co] = sum_tap sum_ci  X[b, t + tap*dil - pad, ci] * W[tap][co][ci]
//
//   GEMM view per CTA tile:  M = 128 time rows (TMEM lanes), N = block_n output channels (TMEM
//   columns), K = 64 input channels per pipeline step, ntaps * ceil(Cin/64) steps.  The zero padding
//   of the convolution and the batch boundaries come for free from TMA out-of-bounds zero fill on a
//   rank-3 tensor map {C, T, B}.
//
//   Warp roles (256 threads, 1 CTA/SM, persistent over tiles):
//     warp 0 lane 0 : TMA producer           (smem ring: full/empty mbarriers)
//     warp 1 lane 0 : tcgen05.mma issuer     (accumulator ring in TMEM: 2 x block_n columns)
//     warp 2        : TMEM alloc / dealloc
//     warps 4..7    : epilogue, one TMEM lane (= one time row) per thread: tcgen05.ld -> fused
//                     bias/act/residual/mask -> vectorised global stores (fp32 and/or 16-bit)
//
// Replaces: nn.Conv1d/ConvTranspose1d + surrounding pointwise ops of the reference
// (VISinger/models_ae.py:327-346, modules.py:148-176,210-223,329-349, attentions.py:138-146,277-285).
#include "common.cuh"
#include "api_internal.h"

namespace b200svs {

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;
constexpr int kATileBytes = kBlockM * kBlockK * 2;  // 16 KB
constexpr int kMaxStages = 8;
constexpr int kTcThreads = 256;

struct ConvTcParams {
  int B, T, tiles_t, m_tiles, n_tiles, total_tiles;
  int ntaps, nchunks, dil, pad_left, cout_pad, block_n;
  int num_stages, stage_bytes, tmem_cols;
  uint32_t idesc;
  b200svs_epilogue ep;
};

struct __align__(8) TcBarriers {
  uint64_t full[kMaxStages];
  uint64_t empty[kMaxStages];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
};

__device__ __forceinline__ float apply_act(float v, int act) {
  switch (act) {
    case B200SVS_ACT_RELU: return fmaxf(v, 0.f);
    case B200SVS_ACT_GELU: return gelu_erf(v);
    case B200SVS_ACT_SIGMOID: return 1.0f / (1.0f + expf(-v));
    case B200SVS_ACT_TANH: return tanhf(v);
    default: return v;
  }
}

__device__ __forceinline__ uint32_t pack16(float a, float b, int dtype) {
  if (dtype == B200SVS_DT_BF16) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// Finishes 16 consecutive output columns [col0, col0+16) of row (b,t): v[] already holds
// act(acc + bias).  Residual add, scale, mask and the (vectorised) stores.
__device__ __forceinline__ void finish_store16(float* v, const b200svs_epilogue& ep, int b, int T, int t,
                                               int col0, bool row_ok, bool row_live) {
  if (!row_ok) return;
  const size_t row = static_cast<size_t>(b) * T + t;
  const int nv = ep.n_valid - col0;  // valid columns in this chunk
  if (nv <= 0) return;
  const bool full = nv >= 16;
  if (ep.res1) {
    const float* r = ep.res1 + row * ep.res1_ld + col0;
    if (full) {
#pragma unroll
      for (int i = 0; i < 16; i += 4) {
        float4 q = *reinterpret_cast<const float4*>(r + i);
        v[i] = ep.alpha * v[i] + q.x; v[i + 1] = ep.alpha * v[i + 1] + q.y;
        v[i + 2] = ep.alpha * v[i + 2] + q.z; v[i + 3] = ep.alpha * v[i + 3] + q.w;
      }
    } else {
      for (int i = 0; i < nv; ++i) v[i] = ep.alpha * v[i] + r[i];
    }
  } else if (ep.alpha != 1.0f) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] *= ep.alpha;
  }
  if (ep.res2) {
    const float* r = ep.res2 + row * ep.res2_ld + col0;
    if (full) {
#pragma unroll
      for (int i = 0; i < 16; i += 4) {
        float4 q = *reinterpret_cast<const float4*>(r + i);
        v[i] += q.x; v[i + 1] += q.y; v[i + 2] += q.z; v[i + 3] += q.w;
      }
    } else {
      for (int i = 0; i < nv; ++i) v[i] += r[i];
    }
  }
  const float s = row_live ? ep.scale : 0.0f;
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] *= s;
  if (ep.out32) {
    float* o = ep.out32 + row * ep.out32_ld + col0;
    if (full) {
#pragma unroll
      for (int i = 0; i < 16; i += 4) *reinterpret_cast<float4*>(o + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
    } else {
      for (int i = 0; i < nv; ++i) o[i] = v[i];
    }
  }
  if (ep.out16) {
    if (ep.post_lrelu) {
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = lrelu(v[i], ep.post_slope);
    }
    uint16_t* o = reinterpret_cast<uint16_t*>(ep.out16) + row * ep.out16_ld + col0;
    if (full) {
      uint4 q0, q1;
      q0.x = pack16(v[0], v[1], ep.out16_dtype); q0.y = pack16(v[2], v[3], ep.out16_dtype);
      q0.z = pack16(v[4], v[5], ep.out16_dtype); q0.w = pack16(v[6], v[7], ep.out16_dtype);
      q1.x = pack16(v[8], v[9], ep.out16_dtype); q1.y = pack16(v[10], v[11], ep.out16_dtype);
      q1.z = pack16(v[12], v[13], ep.out16_dtype); q1.w = pack16(v[14], v[15], ep.out16_dtype);
      *reinterpret_cast<uint4*>(o) = q0;
      *reinterpret_cast<uint4*>(o + 8) = q1;
    } else {
      for (int i = 0; i < nv; ++i) {
        uint32_t pk = pack16(v[i], 0.f, ep.out16_dtype);
        o[i] = static_cast<uint16_t>(pk & 0xFFFF);
      }
    }
  }
}

__global__ void __launch_bounds__(kTcThreads, 1)
conv1d_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                 const ConvTcParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  TcBarriers* bars = reinterpret_cast<TcBarriers*>(smem + static_cast<size_t>(p.num_stages) * p.stage_bytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmW);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(&bars->full[s], 1);
      mbar_init(&bars->empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&bars->tmem_full[s], 1);
      mbar_init(&bars->tmem_empty[s], 4);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(&bars->tmem_base, p.tmem_cols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;
  const int ksteps = p.ntaps * p.nchunks;

  if (warp == 0) {
    if (lane == 0) {
      // ------------------------------------------------------------ TMA producer
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x) {
        const int n_idx = tile / p.m_tiles;
        const int m = tile - n_idx * p.m_tiles;
        const int b = m / p.tiles_t;
        const int t0 = (m - b * p.tiles_t) * kBlockM;
        for (int tap = 0; tap < p.ntaps; ++tap) {
          const int trow = t0 + tap * p.dil - p.pad_left;
          const int wrow = tap * p.cout_pad + n_idx * p.block_n;
          for (int ch = 0; ch < p.nchunks; ++ch) {
            mbar_wait(&bars->empty[stage], phase ^ 1);
            uint8_t* sa = smem + static_cast<size_t>(stage) * p.stage_bytes;
            mbar_arrive_expect_tx(&bars->full[stage], p.stage_bytes);
            tma_load_3d(sa, &tmA, &bars->full[stage], ch * kBlockK, trow, b);
            tma_load_2d(sa + kATileBytes, &tmW, &bars->full[stage], ch * kBlockK, wrow);
            if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ------------------------------------------------------------ MMA issuer
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        mbar_wait(&bars->tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * p.block_n;
        for (int ks = 0; ks < ksteps; ++ks) {
          mbar_wait(&bars->full[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + static_cast<size_t>(stage) * p.stage_bytes);
          const uint64_t da = make_desc_k128(sa);
          const uint64_t db = make_desc_k128(sa + kATileBytes);
#pragma unroll
          for (int k = 0; k < kBlockK / 16; ++k) {
            // advance 16 elements (32 B) along K inside the 128-B swizzle row: +2 in 16-B units
            umma_f16(d_tmem, da + 2 * k, db + 2 * k, p.idesc, (ks | k) != 0);
          }
          umma_commit(&bars->empty[stage]);
          if (++stage == p.num_stages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&bars->tmem_full[acc]);
      }
    }
  } else if (warp >= 4) {
    // -------------------------------------------------------------- epilogue
    const int q = warp & 3;  // TMEM lane quarter this warp may access
    const int row_in_tile = q * 32 + lane;
    const b200svs_epilogue& ep = p.ep;
    const bool gate = ep.act == B200SVS_ACT_GATE;
    const int half_n = p.block_n >> 1;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      const int n_idx = tile / p.m_tiles;
      const int m = tile - n_idx * p.m_tiles;
      const int b = m / p.tiles_t;
      const int t = (m - b * p.tiles_t) * kBlockM + row_in_tile;
      const bool row_ok = t < p.T;
      bool row_live = true;
      if (ep.lens) row_live = t < ep.lens[b];
      mbar_wait(&bars->tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * p.block_n;
      const float* gb = ep.gbias ? ep.gbias + static_cast<size_t>(b) * ep.gbias_ld : nullptr;
      if (!gate) {
        for (int c0 = 0; c0 < p.block_n; c0 += 16) {
          float v[16];
          tmem_ld16(taddr + c0, v);
          const int col0 = n_idx * p.block_n + c0;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float x = v[i];
            if (ep.bias) x += __ldg(ep.bias + col0 + i);
            if (gb) x += __ldg(gb + col0 + i);
            v[i] = apply_act(x, ep.act);
          }
          finish_store16(v, ep, b, p.T, t, col0, row_ok, row_live);
        }
      } else {
        for (int c0 = 0; c0 < half_n; c0 += 16) {
          float va[16], vb[16];
          tmem_ld16(taddr + c0, va);
          tmem_ld16(taddr + half_n + c0, vb);
          const int pa = n_idx * p.block_n + c0;  // packed column of the tanh half
          const int pb = pa + half_n;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float xa = va[i], xb = vb[i];
            if (ep.bias) { xa += __ldg(ep.bias + pa + i); xb += __ldg(ep.bias + pb + i); }
            if (gb) { xa += __ldg(gb + pa + i); xb += __ldg(gb + pb + i); }
            va[i] = tanhf(xa) * (1.0f / (1.0f + expf(-xb)));
          }
          finish_store16(va, ep, b, p.T, t, n_idx * half_n + c0, row_ok, row_live);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->tmem_empty[acc]);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

static int next_pow2_cols(int c) {
  int v = 32;
  while (v < c) v <<= 1;
  return v;
}

}  // namespace b200svs

using namespace b200svs;

extern "C" int b200svs_conv1d_tc(const b200svs_conv_args* a, void* stream) {
  B200SVS_REQUIRE(a && a->x && a->w, "conv1d_tc: null argument");
  B200SVS_REQUIRE(a->B > 0 && a->T > 0 && a->Cin > 0, "conv1d_tc: empty shape B=%d T=%d Cin=%d", a->B, a->T, a->Cin);
  B200SVS_REQUIRE(a->Cin_pad % kBlockK == 0 && a->Cin_pad >= a->Cin, "conv1d_tc: Cin_pad=%d must be a multiple of 64 >= Cin=%d", a->Cin_pad, a->Cin);
  B200SVS_REQUIRE(a->block_n >= 16 && a->block_n <= 256 && a->block_n % 16 == 0, "conv1d_tc: block_n=%d", a->block_n);
  B200SVS_REQUIRE(a->Cout_pad % a->block_n == 0, "conv1d_tc: Cout_pad=%d not a multiple of block_n=%d", a->Cout_pad, a->block_n);
  B200SVS_REQUIRE(a->x_ld % 8 == 0 && (reinterpret_cast<uintptr_t>(a->x) & 15) == 0, "conv1d_tc: x must be 16-B aligned with ld %% 8 == 0 (ld=%d)", a->x_ld);
  B200SVS_REQUIRE((reinterpret_cast<uintptr_t>(a->w) & 15) == 0, "conv1d_tc: w must be 16-B aligned");
  B200SVS_REQUIRE(a->ntaps >= 1 && a->dil >= 1, "conv1d_tc: ntaps=%d dil=%d", a->ntaps, a->dil);
  const b200svs_epilogue& ep = a->ep;
  const bool gate = ep.act == B200SVS_ACT_GATE;
  B200SVS_REQUIRE(!gate || a->block_n % 32 == 0, "conv1d_tc: GATE needs block_n %% 32 == 0");
  B200SVS_REQUIRE(ep.out32 || ep.out16, "conv1d_tc: no output");
  B200SVS_REQUIRE(!ep.out32 || (ep.out32_ld % 4 == 0 && (reinterpret_cast<uintptr_t>(ep.out32) & 15) == 0), "conv1d_tc: out32 alignment");
  B200SVS_REQUIRE(!ep.out16 || (ep.out16_ld % 8 == 0 && (reinterpret_cast<uintptr_t>(ep.out16) & 15) == 0), "conv1d_tc: out16 alignment");
  B200SVS_REQUIRE(!ep.res1 || (ep.res1_ld % 4 == 0 && (reinterpret_cast<uintptr_t>(ep.res1) & 15) == 0), "conv1d_tc: res1 alignment");
  B200SVS_REQUIRE(!ep.res2 || (ep.res2_ld % 4 == 0 && (reinterpret_cast<uintptr_t>(ep.res2) & 15) == 0), "conv1d_tc: res2 alignment");
  B200SVS_REQUIRE(ep.n_valid > 0, "conv1d_tc: n_valid=%d", ep.n_valid);

  CUtensorMap tmA, tmW;
  {
    uint64_t dims[3] = {(uint64_t)a->Cin, (uint64_t)a->T, (uint64_t)a->B};
    uint64_t strides[2] = {(uint64_t)a->x_ld * 2, (uint64_t)a->T * a->x_ld * 2};
    uint32_t box[3] = {(uint32_t)kBlockK, (uint32_t)kBlockM, 1};
    int r = make_tmap_16bit(&tmA, a->x_dtype, a->x, 3, dims, strides, box);
    if (r) return r;
  }
  {
    uint64_t dims[2] = {(uint64_t)a->Cin_pad, (uint64_t)a->ntaps * a->Cout_pad};
    uint64_t strides[1] = {(uint64_t)a->Cin_pad * 2};
    uint32_t box[2] = {(uint32_t)kBlockK, (uint32_t)a->block_n};
    int r = make_tmap_16bit(&tmW, a->x_dtype, a->w, 2, dims, strides, box);
    if (r) return r;
  }

  ConvTcParams p;
  p.B = a->B; p.T = a->T;
  p.tiles_t = (a->T + kBlockM - 1) / kBlockM;
  p.m_tiles = p.tiles_t * a->B;
  p.n_tiles = a->Cout_pad / a->block_n;
  p.total_tiles = p.m_tiles * p.n_tiles;
  p.ntaps = a->ntaps;
  p.nchunks = (a->Cin + kBlockK - 1) / kBlockK;
  p.dil = a->dil; p.pad_left = a->pad_left; p.cout_pad = a->Cout_pad; p.block_n = a->block_n;
  p.stage_bytes = kATileBytes + a->block_n * kBlockK * 2;
  const int budget = max_smem_optin() - 2048;   // 1 KB alignment slack + barriers
  int stages = budget / p.stage_bytes;
  if (stages > kMaxStages) stages = kMaxStages;
  B200SVS_REQUIRE(stages >= 2, "conv1d_tc: not enough shared memory for 2 stages");
  p.num_stages = stages;
  p.tmem_cols = next_pow2_cols(2 * a->block_n);
  p.idesc = make_idesc_f16(kBlockM, a->block_n, a->x_dtype == B200SVS_DT_BF16);
  p.ep = ep;
  if (p.ep.lens_div <= 0) p.ep.lens_div = 1;

  const size_t smem = (size_t)stages * p.stage_bytes + sizeof(TcBarriers) + 1024;
  static size_t configured = 0;
  if (smem > configured) {
    int r = check_cuda(cudaFuncSetAttribute(conv1d_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                       "cudaFuncSetAttribute(conv1d_tc)");
    if (r) return r;
    configured = smem;
  }
  int grid = p.total_tiles < num_sms() ? p.total_tiles : num_sms();
  conv1d_tc_kernel<<<grid, kTcThreads, smem, as_stream(stream)>>>(tmA, tmW, p);
  return check_launch("conv1d_tc_kernel");
}
