/* b200svs - C ABI of the B200-native VISinger synthesis forward (sm_100a only).
 *
 * The reference (GuluOfficial/music_generation) has no FFI/plugin registry: its boundary for this
 * path is the Python method SynthesizerTrn.infer (VISinger/models_ae.py:608, VISinger/models.py:622).
 * music_generation_b200.SynthesizerTrn mirrors that method and drives the entry points below through
 * ctypes.  Each entry point replaces the PyTorch/cuDNN/cuBLAS call sequence cited beside it.
 *
 * Conventions
 *  - Every pointer is a DEVICE pointer owned by the caller; nothing here allocates or frees.
 *  - Activations are channels-last: x[b][t][c] at  x + (b*T + t)*ld + c   ("[B,T,C] with row stride ld").
 *    The reference's [B,C,T] tensors are the transposes of these.
 *  - All work is enqueued on `stream` (a cudaStream_t passed as void*); calls are re-entrant and
 *    keep no state between calls apart from a per-thread last-error string.
 *  - Return value: 0 on success, a negative B200SVS_E* code otherwise (see b200svs_last_error()).
 *  - There is no CPU fallback: on a machine without an sm_100 device every compute call returns
 *    B200SVS_ENODEV / a CUDA error.
 */
#ifndef B200SVS_H_
#define B200SVS_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200SVS_ABI_VERSION 2

enum {
  B200SVS_OK = 0,
  B200SVS_EINVAL = -1,  /* bad shape / alignment / argument */
  B200SVS_ECUDA = -2,   /* CUDA runtime or driver error (text in b200svs_last_error) */
  B200SVS_ENODEV = -3,  /* no sm_100 device / driver entry point missing */
  B200SVS_EUNSUPPORTED = -4
};

/* pre-activation applied to (acc + bias) */
enum {
  B200SVS_ACT_NONE = 0,
  B200SVS_ACT_RELU = 1,
  B200SVS_ACT_GELU = 2,    /* exact erf GELU  (ext_modules/frame_prior_network.py:38) */
  B200SVS_ACT_SIGMOID = 3, /* DurationPredictor (models_ae.py:147) */
  B200SVS_ACT_TANH = 4,
  B200SVS_ACT_GATE = 5     /* tanh(a)*sigmoid(b): commons.fused_add_tanh_sigmoid_multiply (commons.py:100-107);
                              packed weight tile columns [0,N/2) = tanh half, [N/2,N) = sigmoid half */
};
/* B200SVS_DT_F16_SPLIT (16-bit OUTPUTS only): the destination has 3x the columns, laid out [hi | lo | hi]
 * with hi = fp16(v), lo = fp16(v - hi).  A conv whose weights are packed [w_hi | w_hi | w_lo] along Cin then
 * computes x_hi.w_hi + x_lo.w_hi + x_hi.w_lo: tensor-core speed at fp32-class accuracy (F0 predictor). */
enum { B200SVS_DT_F16 = 0, B200SVS_DT_BF16 = 1, B200SVS_DT_F16_SPLIT = 2 };

/* Fused epilogue shared by the conv1d kernels.  For output row (b,t) and packed column c:
 *   v = acc + bias[c] + gbias[b*gbias_ld + c]
 *   v = act(v)                                   (GATE halves the column count)
 *   v = alpha*v + res1[b,t,c'] + res2[b,t,c'] + res16[b,t,c']    (c' = output column)
 *   v = v*scale;  v *= (t < lens[b])             (lens == NULL: no mask)
 *   out32[b,t,c'] = v;  out16r[b,t,c'] = cvt16(v);  out16[b,t,c'] = cvt16(post_lrelu ? lrelu(v, post_slope) : v)
 * Columns c' >= n_valid are not stored by conv1d_f32; conv1d_tc stores through TMA, which clips at
 * 16-byte granularity: columns up to the next multiple of 4 (fp32) / 8 (16-bit) may be written with the
 * zero-padded channels' values, so out32_ld / out16_ld must cover that rounding.  out32 may alias res1
 * (in-place update of the same tile). */
typedef struct b200svs_epilogue {
  const float* bias;
  const float* gbias;
  int32_t gbias_ld;
  int32_t act;
  float alpha;
  float scale;
  const float* res1;
  int32_t res1_ld;
  const float* res2;
  int32_t res2_ld;
  const int32_t* lens;
  int32_t lens_div;   /* row t is valid iff t / lens_div < lens[b]  (1 normally) */
  float* out32;
  int32_t out32_ld;
  void* out16;
  int32_t out16_ld;
  int32_t out16_dtype; /* B200SVS_DT_* */
  int32_t post_lrelu;
  float post_slope;
  int32_t n_valid;
  /* conv1d_tc only (act NONE, non-split): 16-bit residual stream.  res16 is added like res1 (same 16-bit type
   * as the operands); out16r receives cvt16(v) WITHOUT the post-activation and may alias res16. */
  const void* res16;
  int32_t res16_ld;
  void* out16r;
  int32_t out16r_ld;
  /* conv1d_tc only: res2 points to a 16-bit tensor of the operand type instead of fp32 (the Generator's MRF sum
   * `xs += resblock_j(x)`, models_ae.py:335-341, kept as a 16-bit stream; same restrictions as res16). */
  int32_t res2_is16;
  /* conv1d_tc only: res16 holds leaky_relu(x, 1 / res16_unlrelu) instead of x (the 16-bit copy the previous conv
   * wrote for the NEXT conv's operand); the epilogue adds x = min(r, r * res16_unlrelu).  <= 1: res16 is x itself.
   * Lets a ResBlock pair `x + c2(lrelu(c1(lrelu(x))))` (modules.py:210-223) take its residual from the tensor c1
   * consumed, so no un-activated 16-bit stream is written: 10 instead of 12 bytes per element and pair. */
  float res16_unlrelu;
} b200svs_epilogue;

/* 1-D convolution over time as an implicit GEMM (replaces nn.Conv1d / ConvTranspose1d + the pointwise
 * ops around it: models_ae.py:327-346, modules.py:148-176,210-223,329-349, attentions.py:138-146,277-285).
 *   out[b,t,:] = sum_{tap<ntaps} W[tap] . x[b, t + tap*dil - pad_left, 0:Cin]      (zero outside [0,T))
 * w is packed [ntaps][Cout_pad][Cin_pad] (Cin contiguous), Cin_pad % 64 == 0, Cout_pad % block_n == 0.
 * A ConvTranspose1d of stride s is issued as a 3-tap conv over input time with Cout = s*Cout'
 * (polyphase) writing a [B,T,s*Cout'] view of the [B,s*T,Cout'] output. */
typedef struct b200svs_conv_args {
  const void* x; /* tc: 16-bit (x_dtype); f32 variant: float */
  int32_t x_dtype;
  int32_t B, T, Cin, x_ld;
  const void* w;
  int32_t ntaps, Cout_pad, Cin_pad, dil, pad_left;
  int32_t block_n; /* tc only: N tile (16..256, multiple of 16; GATE needs multiple of 32) */
  b200svs_epilogue ep;
} b200svs_conv_args;

const char* b200svs_last_error(void);
int b200svs_abi_version(void);
/* Kernels launched through this library by the calling process so far (every entry point counts its own launches). */
int64_t b200svs_launch_count(void);
/* sizeof of the ABI structs as the library was compiled (0 epilogue, 1 conv_args, 2 conv_w, 3 generator_w, 4 wn_w,
 * 5 encoder_layer_w): lets a binding verify its mirror of the layouts. */
int b200svs_abi_struct_size(int32_t which);
/* 0 if an sm_100 device is current and the driver exports cuTensorMapEncodeTiled */
int b200svs_check_device(void);

/* Numeric-range canary: every fp32 -> 16-bit conversion in the library saturates (+-65504 for fp16)
 * instead of overflowing to inf; conv1d_tc additionally counts, in a library-owned device word of the current device,
 * the epilogue threads that stored a saturated or NaN fp16 value.  Copies the count to *host_out (nullable;
 * synchronises `stream`) and optionally resets it.  0 after a forward = no 16-bit tensor clipped. */
int b200svs_saturation_events(uint32_t* host_out, int32_t reset, void* stream);

/* tcgen05/TMEM/TMA implicit-GEMM conv, fp16|bf16 operands, fp32 accumulate */
int b200svs_conv1d_tc(const b200svs_conv_args* a, void* stream);
/* Diagnostic (tuning aid, not part of the data path): when dev_buf != NULL every following conv1d_tc launch
 * writes per-tile clock64() stamps of its pipeline roles into dev_buf[grid][tiles_per_cta][16] (int64);
 * NULL switches it off.  Slot meanings: tools/trace_conv.py. */
int b200svs_conv1d_tc_set_trace(void* dev_buf, int32_t tiles_per_cta);
/* fp32 SIMT conv (phoneme-level + F0 path where the float->int boundaries need fp32);
 * w packed [ntaps][Cin][Cout_pad] fp32 (Cout contiguous) */
int b200svs_conv1d_f32(const b200svs_conv_args* a, void* stream);

/* Windowed relative-position multi-head self-attention (attentions.py:148-179,199-243), fp32, flash style
 * (never materialises [B,H,T,T]).  qkv: [B,T,3*C] fp32 (q|k|v), out: [B,T,C]; rel_k/rel_v: [2*window+1, C/H];
 * keys >= lens[b] are masked with -1e4 exactly like masked_fill; rows >= lens[b] are written as 0. */
int b200svs_attention_f32(const float* qkv, int32_t qkv_ld, float* out, int32_t out_ld, void* out16,
                          int32_t out16_ld, int32_t out16_dtype, const float* rel_k, const float* rel_v,
                          const int32_t* lens, int32_t B, int32_t T, int32_t C, int32_t H, int32_t window,
                          void* stream);

/* Tensor-core version of the same attention (tcgen05 QK^T and PV, softmax in registers).
 * b200svs_attn_prep splits the fp32 qkv [B,T,3C] into (q|k) 16-bit [B,T,2C] and V^T 16-bit [B,C,Tp]
 * (Tp >= T, multiple of 8).  With qk_lo / vt_lo != NULL every operand is an fp16 (hi, lo) pair and every
 * product three MMAs (hi.hi + hi.lo + lo.hi): fp32-class accuracy for the F0 predictor, whose output is
 * quantised by f0_to_coarse (utils/pitch_utils.py:22-31).  out16_split: out16 is [B,T,3C] = hi | lo | hi
 * (fp16), the K-concatenated operand of a split-precision conv. */
int b200svs_attn_prep(const float* qkv, int32_t qkv_ld, void* qk_hi, void* qk_lo, void* vt_hi, void* vt_lo,
                      int32_t dtype, int32_t B, int32_t T, int32_t Tp, int32_t C, void* stream);
int b200svs_attention_tc(const void* qk_hi, const void* qk_lo, const void* vt_hi, const void* vt_lo,
                         int32_t dtype, int32_t Tp, float* out32, int32_t out32_ld, void* out16, int32_t out16_ld,
                         int32_t out16_dtype, int32_t out16_split, const float* rel_k, const float* rel_v,
                         const int32_t* lens, int32_t B, int32_t T, int32_t C, int32_t H, int32_t window,
                         void* stream);

/* y = LayerNorm_C(x (+ res)) * mask  (modules.py:29-32 + the residual adds of attentions.py:42,46) */
int b200svs_layernorm(const float* x, int32_t x_ld, const float* res, int32_t res_ld, const float* gamma,
                      const float* beta, float eps, const int32_t* lens, float* out32, int32_t out32_ld,
                      void* out16, int32_t out16_ld, int32_t out16_dtype, int32_t B, int32_t T, int32_t C,
                      void* stream);

/* TextEncoder input (models_ae.py:204-215): (E_ph[x] + W_dur*dur + b_dur + E_pitch[p] + E_tone[s])*sqrt(C)*mask.
 * n_ph / n_pitch / n_tone: rows of the three tables.  nn.Embedding raises on an id outside its table; here a live
 * out-of-range id is clamped and sets bit 0 of status[1] (status: nullable int32[2], zeroed by the caller, shared with
 * b200svs_lr_scan) so that the host can raise at its next synchronisation instead of reading out of bounds. */
int b200svs_text_embed(const int64_t* phon, const int64_t* pitch, const int64_t* tone, const float* dur,
                       const float* e_ph, const float* e_pitch, const float* e_tone, const float* w_dur,
                       const float* b_dur, const int64_t* x_lengths, int32_t* lens32_out, float* out,
                       int32_t B, int32_t P, int32_t C, int32_t n_ph, int32_t n_pitch, int32_t n_tone,
                       int32_t* status, void* stream);

/* floor(pred * note_dur / 0.0116) -> int64 (models_ae.py:624-625), bit-exact fp32 op order */
int b200svs_duration_quantize(const float* pred, const float* note_dur, int64_t* dur_out, int32_t n,
                              void* stream);

/* Length regulator (models_ae.py:689-721 + utils/utils.py:344-361) as a gather.
 * pass 1: mel_len[b] = sum_i max(dur[b,i],0), ends = inclusive cumsum (int32 workspace [B,P]).
 * pass 2: out[b,t,:] = h[b, idx(b,t), :] for t < mel_len[b] else 0; index[b,t] = idx or -1. */
/* status (nullable int32[2], zeroed by the caller): status[0] = max_b mel_len[b] (atomicMax) - the one number the
 * host has to read back (the padded frame count T). */
int b200svs_lr_scan(const int64_t* dur, int32_t B, int32_t P, int32_t* ends, int64_t* mel_len64,
                    int32_t* mel_len32, int32_t* status, void* stream);
int b200svs_lr_gather(const float* h, int32_t h_ld, const int32_t* ends, const int32_t* mel_len32,
                      int32_t B, int32_t P, int32_t T, int32_t C, float* out, int32_t out_ld,
                      int32_t* index_out, void* stream);

/* (x + pos[t]) * mask  (ext_modules/text_post_net.py:67-75) -> fp32 + optional 16-bit copy */
int b200svs_add_pos_mask(const float* x, int32_t x_ld, const float* pos, const int32_t* lens, float* out32,
                         int32_t out32_ld, void* out16, int32_t out16_ld, int32_t out16_dtype, int32_t B,
                         int32_t T, int32_t C, void* stream);

/* f0_to_coarse(exp(log_f0)-1) (utils/pitch_utils.py:22-31) + f0_emb gather * mask
 * (text_post_net.py:80-85).  coarse_in != NULL overrides the prediction (f0_score argument); n_rows = rows of the
 * f0_emb table (indices are clamped to it). */
int b200svs_f0_embed(const float* log_f0, const int64_t* coarse_in, const float* f0_emb,
                     const int32_t* lens, int64_t* coarse_out, float* out32, int32_t out32_ld, void* out16,
                     int32_t out16_ld, int32_t out16_dtype, int32_t B, int32_t T, int32_t C, int32_t n_rows,
                     void* stream);

/* z_p = m_p + noise * exp(logs_p) * noise_scale (models_ae.py:643); stats: [B,T,2C] (m|logs);
 * noise is the reference layout [B,C,T]; writes z_p fp32 [B,T,C] (+16-bit copy).  lens (nullable): rows t >= lens[b]
 * are written as 0 - PosteriorEncoder's z = (m + randn * exp(logs)) * x_mask (models_ae.py:290). */
int b200svs_sample_prior(const float* stats, int32_t stats_ld, const float* noise_bct, float noise_scale,
                         float* zp32, int32_t zp_ld, void* zp16, int32_t zp16_ld, int32_t out16_dtype,
                         int32_t B, int32_t T, int32_t C, const int32_t* lens, void* stream);

/* Generator tail (models_ae.py:342-344): tanh(conv_post_k7(x)) with x = leaky_relu(.,0.01) already
 * applied by the producer; x: 16-bit [B,L,C], w: fp32 [7][C]; out: fp32 [B,L] */
int b200svs_conv_post_tanh(const void* x, int32_t x_dtype, int32_t x_ld, const float* w, int32_t B,
                           int32_t L, int32_t C, int32_t ktaps, float* out, void* stream);

/* layout helpers: [B,C,T] fp32 <-> channels-last fp32/16-bit */
int b200svs_bct_to_btc(const float* in_bct, int32_t B, int32_t C, int32_t T, float* out32, int32_t out32_ld,
                       void* out16, int32_t out16_ld, int32_t out16_dtype, const int32_t* lens, void* stream);
int b200svs_btc_to_bct(const float* in_btc, int32_t in_ld, int32_t B, int32_t C, int32_t T, float* out_bct,
                       void* stream);

/* Song assembly: what the reference does on the host after every per-line infer()
 * (pipline_inference_svs_v5.py:134-139,225-243,241-258; pipline_inference_svs_v6_multispeaker.py:240-258), on a
 * ragged batch in HBM: for line i (utterance order[i], or i when order == NULL) the valid samples
 * o[b, 0 : mel_len[b]*hop] are peak-normalised twice (audio *= 32767 / max(0.01, max|audio|), fp32), truncated to
 * int16 and written at song[lead_in + offsets[i] ...]; song[0 : lead_in] = 0 (the 0.2 s = 4410-sample lead-in).
 * legacy_scalar: 0 = NumPy >= 2 scalar semantics (32767/peak in fp32), 1 = NumPy < 2 (fp64 then rounded).
 * peak_ws: n_lines uint32 of scratch; scales_out (nullable): [n_lines][2] the two factors, for parity checks. */
int b200svs_song_assemble(const float* o, int64_t o_ld, const int64_t* mel_len, const int32_t* order,
                          const int64_t* offsets, int32_t n_lines, int32_t hop, int32_t lead_in, int32_t legacy_scalar,
                          uint32_t* peak_ws, int16_t* song, float* scales_out, void* stream);

/* ------------------------------------------------------------------------------------------------------------------
 * Stage-level entry points: one call enqueues every kernel of a stage from a table of packed weights and
 * ONE caller-owned workspace - no host language in the loop, so a C / C++ / Go / Rust host can drive the path, and a
 * small (B = 1) call does not pay an interpreter round trip per kernel.  Nothing is allocated; the workspace size comes
 * from the matching *_workspace_bytes function and must be 256-byte aligned.  All calls enqueue on `stream` and return
 * without synchronising. */

/* Packed weights of one convolution (music_generation_b200/packer.py produces them; layout as in b200svs_conv_args). */
typedef struct b200svs_conv_w {
  const void* w;      /* tc: 16-bit [ntaps][cout_pad][cin_pad]; f32: float [ntaps][cin][cout_pad] */
  const float* bias;  /* [cout_pad] or NULL */
  int32_t ntaps, dil, pad_left, cin, cin_pad, cout, cout_pad, block_n;
  int32_t gate;       /* GATE epilogue (WN in_layers): n_out = cout / 2 */
  int32_t f32;        /* 1: fp32 SIMT conv */
  int32_t fold;       /* 2: time-pair folded (dense [B,T,C] tensors are passed as [B,T/2,2C]) */
  int32_t n_out;      /* columns the epilogue stores */
  float out_scale;    /* split-precision convs: folded into the epilogue's alpha */
} b200svs_conv_w;

#define B200SVS_MAX_STAGES 6
#define B200SVS_MAX_KERNELS 4
#define B200SVS_MAX_PAIRS 4

/* Generator.forward (VISinger/models_ae.py:327-346) with ResBlock1 (modules.py:210-223), the default 16-bit streams
 * (residual recovered from the activated tensor, 16-bit MRF sum): conv_pre -> n_stages x { ConvTranspose (polyphase),
 * n_kernels ResBlocks of n_pairs pairs, MRF average } -> leaky-ReLU(0.01) -> conv_post -> tanh. */
typedef struct b200svs_generator_w {
  int32_t n_stages, n_kernels, n_pairs;
  int32_t dtype;                       /* B200SVS_DT_F16 (the 16-bit residual recovery needs fp16 operands) */
  int32_t up_rate[B200SVS_MAX_STAGES];
  int32_t channels[B200SVS_MAX_STAGES]; /* output channels of stage i */
  int32_t pre_channels;                /* upsample_initial_channel */
  int32_t chunk_frames, wide_chunk_frames, wide_stages;   /* utterance chunking */
  int32_t fuse_pairs;                  /* 1: the 32- and 64-channel stages run b200svs_resblock_pair (one kernel per pair)
                                          where both weight sets fit in shared memory */
  float lrelu_slope;                   /* 0.1 (modules.py:17) */
  b200svs_conv_w pre;
  b200svs_conv_w ups[B200SVS_MAX_STAGES];
  b200svs_conv_w rb[B200SVS_MAX_STAGES][B200SVS_MAX_KERNELS][B200SVS_MAX_PAIRS][2];   /* [stage][kernel][pair][conv1, conv2] */
  const float* post_w;                 /* conv_post [taps][C] fp32 */
  int32_t post_taps;
} b200svs_generator_w;

size_t b200svs_generator_workspace_bytes(const b200svs_generator_w* w, int32_t B, int32_t T);
/* x16: [B,T,inter] 16-bit masked Generator input (row stride x_ld); gbias: nullable [B, pre_channels] speaker term added
 * to conv_pre (Generator.cond, models_ae.py:329-330); out: [B, T * prod(up_rate)] fp32. */
int b200svs_generator_fwd(const b200svs_generator_w* w, const void* x16, int32_t x_ld, int32_t B, int32_t T,
                          const float* gbias, int32_t gbias_ld, float* out, void* workspace, size_t workspace_bytes,
                          void* stream);

/* modules.WN.forward (VISinger/modules.py:148-176): n_layers x { gate conv (tanh * sigmoid fused), res 1x1 conv,
 * x = (x + res) * mask }, skip sum as ONE 1x1 conv over the K-concatenation of every layer's gate output. */
#define B200SVS_MAX_WN_LAYERS 16
typedef struct b200svs_wn_w {
  int32_t n_layers, hidden, dtype;
  b200svs_conv_w in[B200SVS_MAX_WN_LAYERS];
  b200svs_conv_w rs[B200SVS_MAX_WN_LAYERS];   /* rows [0, hidden) = residual half (unused for the last layer) */
  b200svs_conv_w skip_cat;                    /* [hidden, n_layers * hidden] */
} b200svs_wn_w;
size_t b200svs_wn_workspace_bytes(const b200svs_wn_w* w, int32_t B, int32_t T);
/* x32 / x16: [B,T,hidden] masked input, fp32 + 16-bit copy, both updated in place; lens: [B] valid rows; gcond: nullable
 * [B, n_layers * 2 * hidden] speaker term (cond_layer output in gate order); out16: [B,T,hidden] masked skip sum. */
int b200svs_wn_stack_fwd(const b200svs_wn_w* w, float* x32, void* x16, const int32_t* lens, int32_t B, int32_t T,
                         const float* gcond, int32_t gcond_ld, void* out16, void* workspace, size_t workspace_bytes,
                         void* stream);

/* One attentions.Encoder layer (VISinger/attentions.py:35-47,138-179,277-285), post-LN:
 *   x = LN1(x + conv_o(MHA(x)));  x = LN2(x + FFN(x)).
 * mode 0: fp32 SIMT convs + fp32 attention (x16 unused); 1: tensor cores, 16-bit operands (x16 = 16-bit copy of x32);
 * 2: split precision (x16 = fp16 [hi | lo | hi], 3 * hidden columns). */
typedef struct b200svs_encoder_layer_w {
  int32_t hidden, filter, heads, window, mode, dtype;
  b200svs_conv_w qkv, o, ffn1, ffn2;
  const float *rel_k, *rel_v;          /* [2 * window + 1, hidden / heads] */
  const float *ln1_g, *ln1_b, *ln2_g, *ln2_b;
} b200svs_encoder_layer_w;
size_t b200svs_encoder_workspace_bytes(const b200svs_encoder_layer_w* w, int32_t B, int32_t T);
/* Runs n_layers consecutive layers (w[0..n_layers)) in place on x32 (and x16); lens: [B] int32. */
int b200svs_encoder_fwd(const b200svs_encoder_layer_w* w, int32_t n_layers, float* x32, void* x16, const int32_t* lens,
                        int32_t B, int32_t T, void* workspace, size_t workspace_bytes, void* stream);

/* Weight-stationary (transposed) conv for the 64-channel, dilation-1 ResBlock convs (VISinger/modules.py:210-223): the
 * weights are the tensor core's A operand, resident in tensor memory as the banded [128 x (k+1)*64] matrix `w_img`
 * (music_generation_b200/packer.py pack_conv_ws: row (s, co), column (j, ci) = W[co, ci, j - s] for 0 <= j - s < k), the
 * activations the B operand straight from their channels-last layout.
 *   res == NULL:  out = cvt16(out_lrelu ? lrelu(conv(x) + bias, post_slope) : conv(x) + bias)
 *   res != NULL:  v = (conv(x) + bias + min(r, r * res_unlrelu) [+ mrf]) * scale;  out = cvt16(out_lrelu ? lrelu(v) : v)
 * x / res / mrf / out: dense [B, T, 64] fp16, T even; ntaps 3, 7 or 11; out must not alias x or res (it may alias mrf).
 * Same result as b200svs_conv1d_tc up to the fp32 summation order inside the tensor core (not bit-identical). */
int b200svs_conv1d_ws(const void* w_img, const float* bias, int32_t ntaps, const void* x, int32_t B, int32_t T,
                      const void* res, float res_unlrelu, const void* mrf, void* out, int32_t out_lrelu,
                      float post_slope, float scale, void* stream);

/* One ResBlock1 pair of the Generator's 32- or 64-channel stage, fused (VISinger/modules.py:210-223):
 *   v   = x + conv2(lrelu(conv1(lrelu(x)), 0.1))  [+ mrf]      x = min(r, r * res_unlrelu) recovered from the input tensor,
 *   out = cvt16(out_lrelu ? lrelu(v * scale, post_slope) : v * scale)
 * with the intermediate tile in shared memory (4 B of HBM traffic per element instead of the 10 of two conv1d_tc
 * launches; bit-identical to them).  x / mrf / out: dense [B, L, C] fp16; x holds lrelu(x, 1 / res_unlrelu).
 * C = 32 (L even): c2 must be the time-pair-folded conv (fold == 2), c1 folded (dilation 1) or plain (any dilation).
 * C = 64: two plain 64 -> 64 convs whose weights fit in shared memory together (taps1 + taps2 <= 14).
 * out must not alias x; it may alias mrf. */
int b200svs_resblock_pair(const b200svs_conv_w* c1, const b200svs_conv_w* c2, const void* x, int32_t B, int32_t L,
                          const void* mrf, void* out, int32_t out_lrelu, float post_slope, float scale,
                          float res_unlrelu, void* stream);

/* Allocation-free variant of b200svs_song_assemble.  It has no counterpart in the reference and was added
 * for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc): same arguments minus `peak_ws` - the per-line peak scratch is a file-scope
 * `static __device__ __align__(256)` array of B200SVS_SONG_NOALLOC_MAX_LINES uint32 (16 KB, resident in every GPU that
 * loads the library until exit), resolved with cudaGetSymbolAddress once per device.  n_lines beyond the maximum is
 * B200SVS_EINVAL.  The call makes no cudaMalloc / cudaFree / allocating library call; it clears the n_lines words it uses
 * (the array keeps the previous call's contents).  Every call on a GPU shares the array: calls that overlap on the GPU
 * (different streams) race - serialise them, or use b200svs_song_assemble with one workspace per stream. */
#define B200SVS_SONG_NOALLOC_MAX_LINES 4096
int b200svs_song_assemble_noalloc(const float* o, int64_t o_ld, const int64_t* mel_len, const int32_t* order,
                                  const int64_t* offsets, int32_t n_lines, int32_t hop, int32_t lead_in,
                                  int32_t legacy_scalar, int16_t* song, float* scales_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200SVS_H_ */
