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
// Adds one to the process-wide count of kernels launched through the library (check_launch does it for <<<>>> launches).
void count_launch();

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

// Per-DEVICE cached attributes of the current device (a process may drive several GPUs: SingerBank(device=...)).
constexpr int kMaxDevices = 64;
int device_ordinal();            // cudaGetDevice(), clamped to [0, kMaxDevices)
int num_sms();
int max_smem_optin();
// Overflow canary of the current device: one library-owned uint32 in device memory (allocated on first use - the only
// allocation this library makes; NULL if that fails, the kernels then skip the count).
unsigned int* sat_counter();

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

}  // namespace b200svs

// One-time (per device and kernel) opt-in to the device's full dynamic shared memory; returns from the caller on error.
#define B200SVS_SMEM_OPTIN(kernel, what)                                                                    \
  do {                                                                                                      \
    static bool done_[::b200svs::kMaxDevices] = {};                                                         \
    const int d_ = ::b200svs::device_ordinal();                                                             \
    if (!done_[d_]) {                                                                                       \
      int r_ = ::b200svs::check_cuda(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                                          ::b200svs::max_smem_optin()), what);              \
      if (r_) return r_;                                                                                    \
      done_[d_] = true;                                                                                     \
    }                                                                                                       \
  } while (0)

#define B200SVS_REQUIRE(cond, ...)                                  \
  do {                                                              \
    if (!(cond)) return ::b200svs::set_error(B200SVS_EINVAL, __VA_ARGS__); \
  } while (0)
