// C-ABI plumbing: error strings, device checks, tensor-map encoding through the driver entry point
// (resolved at run time so the library loads - and exports every symbol - on a machine without a GPU).
#include "api_internal.h"
#include <string.h>

namespace b200svs {

static thread_local char g_err[512] = "";

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return B200SVS_OK;
  return set_error(B200SVS_ECUDA, "%s: %s", what, cudaGetErrorString(e));
}

static long long g_launches = 0;     // kernels launched by this process through the library (bench.py's gpu_launches)
void count_launch() { __atomic_fetch_add(&g_launches, 1ll, __ATOMIC_RELAXED); }
int check_launch(const char* what) {
  const int r = check_cuda(cudaGetLastError(), what);
  if (r == B200SVS_OK) count_launch();
  return r;
}

EncodeTiledFn get_encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres);
    if (e == cudaSuccess && qres == cudaDriverEntryPointSuccess) fn = reinterpret_cast<EncodeTiledFn>(p);
    (void)cudaGetLastError();
  }
  return fn;
}

int make_tmap_16bit(CUtensorMap* out, int dtype, const void* base, int rank, const uint64_t* dims,
                    const uint64_t* strides_bytes, const uint32_t* box) {
  return make_tmap(out, dtype == B200SVS_DT_BF16 ? 1 : 0, base, rank, dims, strides_bytes, box, 128);
}

int make_tmap(CUtensorMap* out, int elem, const void* base, int rank, const uint64_t* dims,
              const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes) {
  EncodeTiledFn enc = get_encode_tiled();
  if (!enc) return set_error(B200SVS_ENODEV, "cuTensorMapEncodeTiled driver entry point not available");
  cuuint64_t gdim[5];
  cuuint64_t gstr[5];
  cuuint32_t bx[5];
  cuuint32_t es[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bx[i] = box[i];
    es[i] = 1;
    if (i > 0) gstr[i - 1] = strides_bytes[i - 1];
  }
  CUtensorMapDataType dt = elem == 2 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                           : elem == 1 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  CUtensorMapSwizzle sw = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                          : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B
                          : swizzle_bytes == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = enc(out, dt, rank, const_cast<void*>(base), gdim, gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return set_error(B200SVS_ECUDA, "cuTensorMapEncodeTiled failed (CUresult %d) rank %d dims %llu,%llu,%llu", (int)r, rank,
                     (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
                     (unsigned long long)(rank > 2 ? dims[2] : 0));
  return B200SVS_OK;
}

struct DevInfo { int sms, smem, cc; };
static DevInfo g_dev[kMaxDevices] = {};
int device_ordinal() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) { (void)cudaGetLastError(); return 0; }
  return dev < 0 ? 0 : (dev >= kMaxDevices ? kMaxDevices - 1 : dev);
}
static const DevInfo& query_device() {
  const int dev = device_ordinal();
  DevInfo& d = g_dev[dev];
  if (d.sms) return d;
  int sms = 0, smem = 0, major = 0, minor = 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) { (void)cudaGetLastError(); return d; }
  cudaDeviceGetAttribute(&smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
  cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
  d.smem = smem; d.cc = major * 10 + minor; d.sms = sms;
  return d;
}
int num_sms() { const int v = query_device().sms; return v > 0 ? v : 148; }
int max_smem_optin() { const int v = query_device().smem; return v > 0 ? v : 227 * 1024; }
static int device_cc() { return query_device().cc; }

static unsigned int* g_sat[kMaxDevices] = {};
unsigned int* sat_counter() {
  const int dev = device_ordinal();
  if (!g_sat[dev]) {
    unsigned int* p = nullptr;
    if (cudaMalloc(&p, sizeof(unsigned int)) != cudaSuccess || cudaMemset(p, 0, sizeof(unsigned int)) != cudaSuccess) {
      (void)cudaGetLastError();
      return nullptr;
    }
    g_sat[dev] = p;
  }
  return g_sat[dev];
}

}  // namespace b200svs

extern "C" {

const char* b200svs_last_error(void) { return b200svs::g_err; }
int b200svs_abi_version(void) { return B200SVS_ABI_VERSION; }
int64_t b200svs_launch_count(void) { return __atomic_load_n(&b200svs::g_launches, __ATOMIC_RELAXED); }
int b200svs_abi_struct_size(int32_t which) {
  switch (which) {
    case 0: return (int)sizeof(b200svs_epilogue);
    case 1: return (int)sizeof(b200svs_conv_args);
    case 2: return (int)sizeof(b200svs_conv_w);
    case 3: return (int)sizeof(b200svs_generator_w);
    case 4: return (int)sizeof(b200svs_wn_w);
    case 5: return (int)sizeof(b200svs_encoder_layer_w);
    default: return -1;
  }
}

int b200svs_saturation_events(uint32_t* host_out, int32_t reset, void* stream) {
  unsigned int* c = b200svs::sat_counter();
  if (!c) return b200svs::set_error(B200SVS_ECUDA, "saturation_events: counter allocation failed");
  cudaStream_t s = b200svs::as_stream(stream);
  if (host_out) {
    int r = b200svs::check_cuda(cudaMemcpyAsync(host_out, c, sizeof(uint32_t), cudaMemcpyDeviceToHost, s), "cudaMemcpyAsync(sat)");
    if (r) return r;
    r = b200svs::check_cuda(cudaStreamSynchronize(s), "cudaStreamSynchronize(sat)");
    if (r) return r;
  }
  if (reset) return b200svs::check_cuda(cudaMemsetAsync(c, 0, sizeof(uint32_t), s), "cudaMemsetAsync(sat)");
  return B200SVS_OK;
}

int b200svs_check_device(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    (void)cudaGetLastError();
    return b200svs::set_error(B200SVS_ENODEV, "no CUDA device visible");
  }
  if (b200svs::device_cc() / 10 != 10)
    return b200svs::set_error(B200SVS_ENODEV, "device is sm_%d, this library is sm_100a only", b200svs::device_cc());
  if (!b200svs::get_encode_tiled())
    return b200svs::set_error(B200SVS_ENODEV, "driver lacks cuTensorMapEncodeTiled");
  return B200SVS_OK;
}

}  // extern "C"
