// Hardware probe (debug only, not on the product path): does a UMMA shared-memory descriptor whose start
// address is shifted by `shift` rows (shift*128 B, not 1024-B aligned) inside a SWIZZLE_128B K-major
// tile address rows [shift, shift+128) correctly, and does it want base_offset = (addr>>7)&7 or 0?
// The answer decides whether the conv kernel can reuse ONE activation tile (rows t0-pad .. t0+127+pad)
// for every tap instead of re-loading it per tap ("A-tile reuse").
//
// out[128][64] = A[shift : shift+128, 0:64] . B[0:64, 0:64]^T   (fp16 operands, fp32 accumulate)
#include "common.cuh"
#include "api_internal.h"
#include "../../include/b200svs_debug.h"

namespace b200svs {

constexpr int kProbeRowsMax = 160;

__global__ void __launch_bounds__(128) probe_rowshift_kernel(const __half* a, const __half* bm, float* out, int rows_a,
                                                             int shift, int base_offset_mode) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;                           // kProbeRowsMax x 128 B
  uint8_t* sB = smem + kProbeRowsMax * 128;     // 64 x 128 B
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;

  // manual SWIZZLE_128B: 16-byte chunk j of row r lives at r*128 + ((j ^ (r & 7)) << 4)
  for (int i = tid; i < kProbeRowsMax * 8; i += 128) {
    const int r = i >> 3, j = i & 7;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (r < rows_a) v = *reinterpret_cast<const uint4*>(a + static_cast<size_t>(r) * 64 + j * 8);
    *reinterpret_cast<uint4*>(sA + r * 128 + ((j ^ (r & 7)) << 4)) = v;
  }
  for (int i = tid; i < 64 * 8; i += 128) {
    const int r = i >> 3, j = i & 7;
    const uint4 v = *reinterpret_cast<const uint4*>(bm + static_cast<size_t>(r) * 64 + j * 8);
    *reinterpret_cast<uint4*>(sB + r * 128 + ((j ^ (r & 7)) << 4)) = v;
  }
  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(&tmem_base_s, 64);
    tmem_relinquish();
  }
  fence_proxy_async();   // generic-proxy smem writes -> visible to the tensor-core (async) proxy
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;

  if (tid == 0) {
    const uint32_t a_addr = smem_u32(sA) + shift * 128;
    const uint32_t bo = base_offset_mode ? ((a_addr >> 7) & 7) : 0;
    const uint64_t da = make_desc_k128(a_addr, bo);
    const uint64_t db = make_desc_k128(smem_u32(sB));
    const uint32_t idesc = make_idesc_f16(128, 64, false);
#pragma unroll
    for (int k = 0; k < 4; ++k) umma_f16(tmem_base, da + 2 * k, db + 2 * k, idesc, k != 0);
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  const int row = warp * 32 + (tid & 31);
  for (int c0 = 0; c0 < 64; c0 += 16) {
    float v[16];
    tmem_ld16(tmem_base + (static_cast<uint32_t>(warp * 32) << 16) + c0, v);
#pragma unroll
    for (int i = 0; i < 16; ++i) out[row * 64 + c0 + i] = v[i];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 64);
  }
}

}  // namespace b200svs

using namespace b200svs;

extern "C" int b200svs_probe_rowshift(const void* a_f16, const void* b_f16, float* out, int32_t rows_a, int32_t shift,
                                      int32_t base_offset_mode, void* stream) {
  B200SVS_REQUIRE(a_f16 && b_f16 && out, "probe_rowshift: null argument");
  B200SVS_REQUIRE(rows_a > 0 && rows_a <= kProbeRowsMax && shift >= 0 && shift + 128 <= kProbeRowsMax,
                  "probe_rowshift: rows_a=%d shift=%d", rows_a, shift);
  const size_t smem = (kProbeRowsMax + 64) * 128 + 1024;
  int r = check_cuda(cudaFuncSetAttribute(probe_rowshift_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                     "cudaFuncSetAttribute(probe)");
  if (r) return r;
  probe_rowshift_kernel<<<1, 128, smem, as_stream(stream)>>>(static_cast<const __half*>(a_f16),
                                                            static_cast<const __half*>(b_f16), out, rows_a, shift,
                                                            base_offset_mode);
  return check_launch("probe_rowshift_kernel");
}
