// song_assemble: the post-processing the reference does on the host, per lyric line, after every infer()
// (pipline_inference_svs_v5.py:134-139 and :225-243, :241-258; same in the v6-multispeaker driver
// pipline_inference_svs_v6_multispeaker.py:240-258), as two kernels over a ragged batch that is already in HBM:
//
//   audio  = o[b, 0, 0 : mel_len[b]*hop]                                   valid samples of the line
//   audio *= 32767 / max(0.01, max|audio|)        Synthesiser.syn (the per-line .wav path is always given)
//   audio *= 32767 / max(0.01, max|audio|)        again in the song loop (a factor ~1.0, kept for bit parity)
//   song   = concat([0.0] * int(0.2 * 22050), audio_0, audio_1, ...).astype(int16)     (C truncation)
//
// max|.| is order-independent and x -> fl(x*s), s > 0, is monotonic, so the second peak is fl(peak*s1): one
// reduction pass (atomicMax on the bit pattern of |x|, which orders like the float for x >= 0) and one
// scale/convert/scatter pass.  The scalar 32767/peak is formed in fp32 (NumPy >= 2 scalar semantics, what the
// reference's statements do in this image) or in fp64 then rounded (legacy NumPy value-based casting).
#include "api_internal.h"

namespace b200svs {

__global__ void __launch_bounds__(256) song_peak_kernel(const float* o, long long o_ld, const int64_t* mel_len,
                                                        const int32_t* order, int hop, unsigned int* peak_bits) {
  const int line = blockIdx.y;
  const int b = order ? order[line] : line;
  const long long n = mel_len[b] * static_cast<long long>(hop);
  const float* src = o + static_cast<long long>(b) * o_ld;
  float m = 0.f;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x * 4;
  for (long long i = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * 4; i < n; i += stride) {
    if (i + 3 < n && ((reinterpret_cast<uintptr_t>(src + i) & 15) == 0)) {
      const float4 v = *reinterpret_cast<const float4*>(src + i);
      m = fmaxf(fmaxf(m, fabsf(v.x)), fmaxf(fabsf(v.y), fmaxf(fabsf(v.z), fabsf(v.w))));
    } else {
      for (long long j = i; j < n && j < i + 4; ++j) m = fmaxf(m, fabsf(src[j]));
    }
  }
  for (int s = 16; s > 0; s >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, s));
  if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(peak_bits + line, __float_as_uint(m));
}

__device__ __forceinline__ float peak_scale(float peak, int legacy) {
  // 32767 / max(0.01, peak): Python's max() returns the double 0.01 when it wins, and 32767 / 0.01 = 3276700.0 is
  // exact in fp32 either way.
  if (legacy) {     // NumPy < 2: python-scalar arithmetic in fp64, rounded to fp32 by the in-place multiply
    if (!(static_cast<double>(peak) > 0.01)) return 3276700.0f;
    return static_cast<float>(32767.0 / static_cast<double>(peak));
  }
  if (!(peak > 0.01f)) return 3276700.0f;     // NumPy >= 2 (NEP 50): the comparison and the division are fp32
  return __fdiv_rn(32767.0f, peak);
}

__global__ void __launch_bounds__(256) song_write_kernel(const float* o, long long o_ld, const int64_t* mel_len,
                                                         const int32_t* order, const int64_t* offsets, int hop,
                                                         int lead_in, const unsigned int* peak_bits, int legacy,
                                                         int16_t* song, float* scales_out) {
  const int line = blockIdx.y;
  const int b = order ? order[line] : line;
  const long long n = mel_len[b] * static_cast<long long>(hop);
  const float peak = __uint_as_float(peak_bits[line]);
  const float s1 = peak_scale(peak, legacy);
  const float peak2 = __fmul_rn(peak, s1);
  const float s2 = peak_scale(peak2, legacy);
  if (blockIdx.x == 0 && threadIdx.x == 0 && scales_out) { scales_out[2 * line] = s1; scales_out[2 * line + 1] = s2; }
  const float* src = o + static_cast<long long>(b) * o_ld;
  int16_t* dst = song + lead_in + offsets[line];
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float v = __fmul_rn(__fmul_rn(src[i], s1), s2);
    dst[i] = static_cast<int16_t>(static_cast<int>(v));     // astype(np.int16): truncation toward zero
  }
  if (line == 0)                                            // the 0.2 s lead-in
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < lead_in; i += stride) song[i] = 0;
}

}  // namespace b200svs

using namespace b200svs;

extern "C" int b200svs_song_assemble(const float* o, int64_t o_ld, const int64_t* mel_len, const int32_t* order,
                                     const int64_t* offsets, int32_t n_lines, int32_t hop, int32_t lead_in,
                                     int32_t legacy_scalar, uint32_t* peak_ws, int16_t* song, float* scales_out,
                                     void* stream) {
  B200SVS_REQUIRE(o && mel_len && offsets && peak_ws && song, "song_assemble: null argument");
  B200SVS_REQUIRE(n_lines > 0 && hop > 0 && lead_in >= 0 && o_ld > 0, "song_assemble: n_lines=%d hop=%d lead_in=%d", n_lines,
                  hop, lead_in);
  int r = check_cuda(cudaMemsetAsync(peak_ws, 0, sizeof(uint32_t) * n_lines, as_stream(stream)), "cudaMemsetAsync(peak_ws)");
  if (r) return r;
  const int bx = num_sms() * 4 / n_lines > 0 ? (num_sms() * 4 + n_lines - 1) / n_lines : 1;
  dim3 grid(bx, n_lines);
  song_peak_kernel<<<grid, 256, 0, as_stream(stream)>>>(o, o_ld, mel_len, order, hop, peak_ws);
  r = check_launch("song_peak_kernel");
  if (r) return r;
  song_write_kernel<<<grid, 256, 0, as_stream(stream)>>>(o, o_ld, mel_len, order, offsets, hop, lead_in, peak_ws,
                                                        legacy_scalar, song, scales_out);
  return check_launch("song_write_kernel");
}

// ---------------------------------------------------------------------------------------------------------------------
// b200svs_song_assemble_noalloc: the same two kernels with the per-line peak scratch in a file-scope device array instead
// of a caller workspace, for callers that cannot allocate (CUDA graph capture, tight serving loops).  No counterpart in
// the reference.  The array lives in every GPU that loads the library
// until exit (16 KB); all calls share it, so two calls that overlap on one GPU (different streams) race.
namespace b200svs {
constexpr int kSongNoallocMaxLines = 4096;
static __device__ __align__(256) unsigned int g_song_peak_scratch[kSongNoallocMaxLines];

// In host code `g_song_peak_scratch` is a host-side stand-in, not a device address: resolve it once per device, with
// that device current, and hand the kernels the resolved pointer.
static int song_scratch_ptr(unsigned int** out) {
  static unsigned int* ptrs[kMaxDevices] = {};
  const int dev = device_ordinal();
  if (!ptrs[dev]) {
    unsigned int* p = nullptr;
    int r = check_cuda(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_song_peak_scratch),
                       "cudaGetSymbolAddress(g_song_peak_scratch)");
    if (r) return r;
    if (!p) return set_error(B200SVS_ECUDA, "song_assemble_noalloc: cudaGetSymbolAddress returned NULL");
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return B200SVS_OK;
}
}  // namespace b200svs

extern "C" int b200svs_song_assemble_noalloc(const float* o, int64_t o_ld, const int64_t* mel_len, const int32_t* order,
                                             const int64_t* offsets, int32_t n_lines, int32_t hop, int32_t lead_in,
                                             int32_t legacy_scalar, int16_t* song, float* scales_out, void* stream) {
  B200SVS_REQUIRE(o && mel_len && offsets && song, "song_assemble_noalloc: null argument");
  B200SVS_REQUIRE(n_lines > 0 && hop > 0 && lead_in >= 0 && o_ld > 0, "song_assemble_noalloc: n_lines=%d hop=%d lead_in=%d",
                  n_lines, hop, lead_in);
  B200SVS_REQUIRE(n_lines <= kSongNoallocMaxLines,
                  "song_assemble_noalloc: n_lines=%d exceeds the static scratch of %d lines (use b200svs_song_assemble with a "
                  "caller workspace)", n_lines, kSongNoallocMaxLines);
  unsigned int* peak_ws = nullptr;
  int r = song_scratch_ptr(&peak_ws);
  if (r) return r;
  // the array keeps the previous call's peaks: the atomicMax reduction must start from zero
  r = check_cuda(cudaMemsetAsync(peak_ws, 0, sizeof(unsigned int) * n_lines, as_stream(stream)), "cudaMemsetAsync(song scratch)");
  if (r) return r;
  const int bx = (num_sms() * 4 + n_lines - 1) / n_lines;
  dim3 grid(bx > 0 ? bx : 1, n_lines);
  song_peak_kernel<<<grid, 256, 0, as_stream(stream)>>>(o, o_ld, mel_len, order, hop, peak_ws);
  r = check_launch("song_peak_kernel (noalloc)");
  if (r) return r;
  song_write_kernel<<<grid, 256, 0, as_stream(stream)>>>(o, o_ld, mel_len, order, offsets, hop, lead_in, peak_ws,
                                                        legacy_scalar, song, scales_out);
  return check_launch("song_write_kernel (noalloc)");
}
