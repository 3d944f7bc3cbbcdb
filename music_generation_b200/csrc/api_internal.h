// Internal host-side helpers shared by the .cu translation units (error reporting, driver entry).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include "../../include/b200svs.h"

namespace b200svs {

int set_error(int code, const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);
// Checks the launch that was just issued (cudaGetLastError; no sync).
int check_launch(const char* what);

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode_tiled();

// 16-bit tensor map with 128-B swizzle; dims/strides innermost first; strides in BYTES for dims 1..rank-1.
int make_tmap_16bit(CUtensorMap* out, int dtype, const void* base, int rank, const uint64_t* dims,
                    const uint64_t* strides_bytes, const uint32_t* box);

// General tiled tensor map: elem 0 = fp16, 1 = bf16, 2 = fp32; swizzle_bytes in {0, 32, 64, 128}.
int make_tmap(CUtensorMap* out, int elem, const void* base, int rank, const uint64_t* dims,
              const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes);

int num_sms();
int max_smem_optin();

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

}  // namespace b200svs

#define B200SVS_REQUIRE(cond, ...)                                  \
  do {                                                              \
    if (!(cond)) return ::b200svs::set_error(B200SVS_EINVAL, __VA_ARGS__); \
  } while (0)
