// Shared between attention_tc.cu (operands from shared memory) and attention_ts.cu (Q and P operands from TMEM):
// tile constants, the kernel parameter block and the 16-bit helpers.
#pragma once
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
constexpr int kSoftWarps = 8;              // two per scheduler: warp w and w + 4 share query rows, split the columns
constexpr int kAttnThreads = kSoftWarps * 32 + 64;    // softmax warps + MMA warp + TMA warp
constexpr int kMmaWarp = kSoftWarps, kTmaWarp = kSoftWarps + 1;
constexpr float kLog2e = 1.4426950408889634f;
#ifndef B200SVS_ATTN_POLL
#define B200SVS_ATTN_POLL 0
#endif
__device__ __forceinline__ void attn_wait(uint64_t* bar, uint32_t parity) {
  if (B200SVS_ATTN_POLL) mbar_wait_poll(bar, parity); else mbar_wait(bar, parity);
}

struct AttnTcParams {
  const void* qk_hi; const void* qk_lo; int qk_ld;   // [B,T,2C] 16-bit (q | k)
  float* out32; int out32_ld;
  void* out16; int out16_ld; int out16_dtype; int out16_split;
  const float* rel_k; const float* rel_v;
  const int32_t* lens;
  int B, T, C, H, dtype;
  long long* trace; int trace_cap;    // diagnostic (attention_ts): [trace_cap tiles][16] clock64 stamps of CTA (1,0,0); NULL: off
};

struct __align__(8) AttnBars {
  uint64_t qbar;
  uint64_t k_full[2], k_empty[2];
  uint64_t v_full, v_empty;
  uint64_t s_full[2], s_free[2];     // S_j in TMEM ready / read out by the softmax warps
  uint64_t p_full, o_full;           // P_j published in shared memory / PV_j complete in TMEM
  uint32_t tmem_base;
};

__device__ __forceinline__ float ld16f(const void* base, size_t idx, int dtype) {
  if (dtype == B200SVS_DT_BF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(base)[idx]);
  return __half2float(reinterpret_cast<const __half*>(base)[idx]);
}
__device__ __forceinline__ uint32_t pk16(float a, float b, int dtype) {
  return dtype == B200SVS_DT_BF16 ? cvt_bf16x2_sat(a, b) : cvt_f16x2_sat(a, b);      // saturating (common.cuh)
}
__device__ __forceinline__ float rt16(float a, int dtype) {   // round trip through the 16-bit type
  if (dtype == B200SVS_DT_BF16) return __bfloat162float(cvt_bf16_sat(a));
  return __half2float(cvt_f16_sat(a));
}
__device__ __forceinline__ float ex2f(float x) {              // 2^x, 2 ulp; ex2(-inf) = 0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}


// attention_ts.cu: same contract as attention_tc_kernel, launched by b200svs_attention_tc when B200SVS_ATTN_TS != 0
int launch_attention_ts(const CUtensorMap* tmQ, const CUtensorMap* tmK, const CUtensorMap* tmV, const AttnTcParams& p,
                        bool split, bool wide, int parts, cudaStream_t stream);

}  // namespace b200svs
