// Stage-level entry points (include/b200svs.h): the per-stage orchestration of SynthesizerTrn.infer in
// C++, on top of the library's own per-op entry points.  One call = every launch of a Generator / WN stack / encoder
// forward, from a table of packed weights and one caller-owned workspace: no interpreter between the launches (a B = 1
// call is ~300 launches; a Python round trip per launch costs more than the kernels), and a non-Python host can drive
// the path.  Host code only - the kernels are in conv1d_tc.cu, conv1d_f32.cu, attention*.cu and pointwise.cu.
//
// Replaces the Python loops of music_generation_b200/synthesizer.py (_generator, _wn, _encoder), which remain as the
// reference implementation of the non-default precision knobs; tests/test_gpu_stages.py checks bit-identity.
#include <string.h>
#include "api_internal.h"

namespace b200svs {

struct Arena {                      // bump allocator over the caller's workspace, 256-byte granules
  uint8_t* base;
  size_t cap, off;
  bool ok;
  Arena(void* p, size_t n) : base(static_cast<uint8_t*>(p)), cap(n), off(0), ok(true) {}
  void* take(size_t bytes) {
    const size_t b = (bytes + 255) & ~size_t(255);
    if (off + b > cap) { ok = false; return base; }       // base != NULL: the error is reported once, after planning
    void* r = base ? base + off : nullptr;
    off += b;
    return r;
  }
};
static size_t align256(size_t b) { return (b + 255) & ~size_t(255); }

static void init_ep(b200svs_epilogue* e) {
  memset(e, 0, sizeof(*e));
  e->alpha = 1.0f; e->scale = 1.0f; e->lens_div = 1;
}

// One conv launch.  x: [B, T, cw.cin (x fold)] with row stride x_ld; every tensor of `ep` has the same [B, T] geometry.
// fold == 2 (time-pair folded conv): dense tensors are re-viewed as [B, T/2, 2C] (ld doubles) - the caller passes the
// UNFOLDED geometry.
static int run_conv(const b200svs_conv_w& cw, const void* x, int x_dtype, int x_ld, int B, int T, b200svs_epilogue ep,
                    void* stream) {
  b200svs_conv_args a;
  memset(&a, 0, sizeof(a));
  const int f = cw.fold > 1 ? cw.fold : 1;
  a.x = x; a.x_dtype = x_dtype; a.B = B; a.T = T / f; a.Cin = cw.cin; a.x_ld = x_ld * f;
  a.w = cw.w; a.ntaps = cw.ntaps; a.Cout_pad = cw.cout_pad; a.Cin_pad = cw.cin_pad; a.dil = cw.dil; a.pad_left = cw.pad_left;
  a.block_n = cw.block_n;
  ep.bias = cw.bias;
  if (cw.gate) ep.act = B200SVS_ACT_GATE;
  ep.alpha *= cw.out_scale;
  if (ep.n_valid == 0) ep.n_valid = cw.n_out;
  if (f > 1) {
    if (T % f) return set_error(B200SVS_EINVAL, "folded conv needs T %% %d == 0 (T=%d)", f, T);
    ep.res1_ld *= f; ep.res2_ld *= f; ep.out32_ld *= f; ep.out16_ld *= f; ep.res16_ld *= f; ep.out16r_ld *= f;
  }
  a.ep = ep;
  return cw.f32 ? b200svs_conv1d_f32(&a, stream) : b200svs_conv1d_tc(&a, stream);
}

static int check_conv_w(const b200svs_conv_w& c, const char* what) {
  if (!c.w || c.ntaps < 1 || c.cin < 1 || c.cout_pad < 1)
    return set_error(B200SVS_EINVAL, "%s: weight table entry is empty", what);
  return B200SVS_OK;
}

// ------------------------------------------------------------------------------------------------ Generator
struct GenPlan {
  int per_w, per, n_wide, hop;
  size_t pre_bytes, wide_arena, nar_arena, total;
};

static int plan_generator(const b200svs_generator_w* w, int B, int T, GenPlan* g) {
  B200SVS_REQUIRE(w && B > 0 && T > 0, "generator: bad argument (B=%d T=%d)", B, T);
  B200SVS_REQUIRE(w->n_stages >= 1 && w->n_stages <= B200SVS_MAX_STAGES && w->n_kernels >= 1 && w->n_kernels <= B200SVS_MAX_KERNELS &&
                  w->n_pairs >= 1 && w->n_pairs <= B200SVS_MAX_PAIRS, "generator: n_stages=%d n_kernels=%d n_pairs=%d",
                  w->n_stages, w->n_kernels, w->n_pairs);
  B200SVS_REQUIRE(w->dtype == B200SVS_DT_F16, "generator_fwd: the 16-bit residual recovery needs fp16 operands");
  g->n_wide = w->wide_stages < 0 ? 0 : (w->wide_stages > w->n_stages ? w->n_stages : w->wide_stages);
  const int chunk = w->chunk_frames > 0 ? w->chunk_frames : 32768;
  const int wide = g->n_wide ? (w->wide_chunk_frames > chunk ? w->wide_chunk_frames : chunk) : chunk;
  g->per_w = wide / T > 0 ? wide / T : 1;
  if (g->per_w > B) g->per_w = B;
  g->per = chunk / T > 0 ? chunk / T : 1;
  if (g->per > g->per_w) g->per = g->per_w;
  g->hop = 1;
  size_t el_wide = 0, el_nar = 0;          // largest per-utterance stage tensor (elements) in the wide / narrow phase
  long long L = T;
  for (int i = 0; i < w->n_stages; ++i) {
    B200SVS_REQUIRE(w->up_rate[i] >= 1 && w->channels[i] >= 8, "generator: stage %d rate %d channels %d", i, w->up_rate[i], w->channels[i]);
    L *= w->up_rate[i];
    g->hop *= w->up_rate[i];
    const size_t el = static_cast<size_t>(L) * w->channels[i];
    if (i < g->n_wide) { if (el > el_wide) el_wide = el; } else { if (el > el_nar) el_nar = el; }
  }
  // per stage five 16-bit tensors: xl (activated input of the ResBlocks), cl, t (pair intermediate), mrf, nxt
  g->pre_bytes = align256(static_cast<size_t>(g->per_w) * T * w->pre_channels * 2);
  g->wide_arena = 5 * align256(static_cast<size_t>(g->per_w) * el_wide * 2);
  g->nar_arena = 5 * align256(static_cast<size_t>(g->per) * el_nar * 2);
  g->total = g->pre_bytes + 2 * g->wide_arena + 2 * g->nar_arena;
  return B200SVS_OK;
}

// Upsampling stages st0 .. st1-1 on the activated 16-bit tensor cur [nb, L, cin_ch]; arenas ping-pong per stage.
static int generator_stages(const b200svs_generator_w* w, const void* cur, int cur_ch, int nb, long long L, int st0, int st1,
                            uint8_t* arena0, uint8_t* arena1, size_t arena_bytes, const void** out_cur, long long* out_L,
                            void* stream) {
  const int dt = w->dtype;
  const float slope = w->lrelu_slope;
  for (int i = st0; i < st1; ++i) {
    const int u = w->up_rate[i], ch = w->channels[i];
    const long long Lo = L * u;
    B200SVS_REQUIRE(Lo < (1ll << 31), "generator: stage %d length %lld exceeds int32", i, Lo);
    Arena ar(((i - st0) & 1) ? arena1 : arena0, arena_bytes);
    const size_t tb = static_cast<size_t>(nb) * Lo * ch * 2;
    void* xl = ar.take(tb); void* cl = ar.take(tb); void* t16 = ar.take(tb); void* mrf = ar.take(tb); void* nxt = ar.take(tb);
    if (!ar.ok) return set_error(B200SVS_EINVAL, "generator: workspace too small (stage %d)", i);
    {   // ConvTranspose1d as a 3-tap conv writing the [nb, L, u*ch] view of [nb, L*u, ch]; next conv's leaky-ReLU applied
      b200svs_epilogue e; init_ep(&e);
      e.out16 = xl; e.out16_ld = u * ch; e.out16_dtype = dt; e.post_lrelu = 1; e.post_slope = slope;
      int r = run_conv(w->ups[i], cur, dt, cur_ch, nb, static_cast<int>(L), e, stream);
      if (r) return r;
    }
    const float next_slope = i < w->n_stages - 1 ? slope : 0.01f;       // F.leaky_relu default before conv_post (models_ae.py:342)
    const int T2 = static_cast<int>(Lo);
    for (int j = 0; j < w->n_kernels; ++j) {
      const void* src = xl;
      for (int m = 0; m < w->n_pairs; ++m) {
        const bool last = m == w->n_pairs - 1;
        const b200svs_conv_w& c1 = w->rb[i][j][m][0];
        const b200svs_conv_w& c2 = w->rb[i][j][m][1];
        const bool fusable = (ch == 32 && c2.fold == 2 && T2 % 2 == 0) ||
                             (ch == 64 && c1.fold <= 1 && c2.fold <= 1 && c1.ntaps + c2.ntaps <= 14);      // both weight sets resident
        if (w->fuse_pairs && fusable) {
          // one kernel per pair (resblock_pair.cu); the output must not alias the input: cl and t16 ping-pong
          void* dst = (src == cl) ? t16 : cl;
          int r;
          if (!last) r = b200svs_resblock_pair(&c1, &c2, src, nb, T2, nullptr, dst, 1, slope, 1.0f, 1.0f / slope, stream);
          else if (j < w->n_kernels - 1)
            r = b200svs_resblock_pair(&c1, &c2, src, nb, T2, j > 0 ? mrf : nullptr, mrf, 0, 0.0f, 1.0f, 1.0f / slope, stream);
          else
            r = b200svs_resblock_pair(&c1, &c2, src, nb, T2, j > 0 ? mrf : nullptr, nxt, 1, next_slope, 1.0f / w->n_kernels,
                                      1.0f / slope, stream);
          if (r) return r;
          src = dst;
          continue;
        }
        {   // xt = lrelu(conv1(lrelu(x)))
          b200svs_epilogue e; init_ep(&e);
          e.out16 = t16; e.out16_ld = ch; e.out16_dtype = dt; e.post_lrelu = 1; e.post_slope = slope;
          int r = run_conv(c1, src, dt, ch, nb, T2, e, stream);
          if (r) return r;
        }
        b200svs_epilogue e; init_ep(&e);
        e.res16 = src; e.res16_ld = ch; e.res16_unlrelu = 1.0f / slope;     // x recovered from lrelu(x): min(r, r / slope)
        if (!last) {
          e.out16 = cl; e.out16_ld = ch; e.out16_dtype = dt; e.post_lrelu = 1; e.post_slope = slope;
        } else {                       // MRF: xs = sum_j resblock_j(x); x = xs / n_kernels (models_ae.py:335-341)
          if (j > 0) { e.res2 = static_cast<const float*>(mrf); e.res2_ld = ch; e.res2_is16 = 1; }
          if (j < w->n_kernels - 1) {
            e.out16r = mrf; e.out16r_ld = ch;
          } else {
            e.scale = 1.0f / w->n_kernels;
            e.out16 = nxt; e.out16_ld = ch; e.out16_dtype = dt; e.post_lrelu = 1; e.post_slope = next_slope;
          }
        }
        int r = run_conv(c2, t16, dt, ch, nb, T2, e, stream);
        if (r) return r;
        src = cl;
      }
    }
    cur = nxt; cur_ch = ch; L = Lo;
  }
  *out_cur = cur; *out_L = L;
  return B200SVS_OK;
}

}  // namespace b200svs

using namespace b200svs;

extern "C" size_t b200svs_generator_workspace_bytes(const b200svs_generator_w* w, int32_t B, int32_t T) {
  GenPlan g = {};
  if (plan_generator(w, B, T, &g)) return 0;
  return g.total;
}

extern "C" int b200svs_generator_fwd(const b200svs_generator_w* w, const void* x16, int32_t x_ld, int32_t B, int32_t T,
                                     const float* gbias, int32_t gbias_ld, float* out, void* workspace,
                                     size_t workspace_bytes, void* stream) {
  GenPlan g = {};
  int r = plan_generator(w, B, T, &g);
  if (r) return r;
  B200SVS_REQUIRE(x16 && out && workspace, "generator_fwd: null argument");
  B200SVS_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0 && workspace_bytes >= g.total,
                  "generator_fwd: workspace must be 256-byte aligned and >= %zu bytes (got %zu)", g.total, workspace_bytes);
  r = check_conv_w(w->pre, "generator.pre");
  if (r) return r;
  uint8_t* ws = static_cast<uint8_t*>(workspace);
  uint8_t* pre_buf = ws;
  uint8_t* wide0 = ws + g.pre_bytes;
  uint8_t* wide1 = wide0 + g.wide_arena;
  uint8_t* nar0 = wide1 + g.wide_arena;
  uint8_t* nar1 = nar0 + g.nar_arena;
  const int dt = w->dtype, C0 = w->pre_channels;
  const size_t x_row = static_cast<size_t>(x_ld) * 2;
  const long long S = static_cast<long long>(T) * g.hop;       // samples per utterance
  for (int w0 = 0; w0 < B; w0 += g.per_w) {
    const int nw = w0 + g.per_w <= B ? g.per_w : B - w0;
    {   // conv_pre (+ speaker term), leaky-ReLU for stage 0 applied
      b200svs_epilogue e; init_ep(&e);
      e.out16 = pre_buf; e.out16_ld = C0; e.out16_dtype = dt; e.post_lrelu = 1; e.post_slope = w->lrelu_slope;
      if (gbias) { e.gbias = gbias + static_cast<size_t>(w0) * gbias_ld; e.gbias_ld = gbias_ld; }
      r = run_conv(w->pre, static_cast<const uint8_t*>(x16) + static_cast<size_t>(w0) * T * x_row, dt, x_ld, nw, T, e, stream);
      if (r) return r;
    }
    const void* cur = pre_buf;
    long long L = T;
    int cur_ch = C0;
    r = generator_stages(w, cur, cur_ch, nw, L, 0, g.n_wide, wide0, wide1, g.wide_arena, &cur, &L, stream);
    if (r) return r;
    if (g.n_wide) cur_ch = w->channels[g.n_wide - 1];
    for (int b0 = 0; b0 < nw; b0 += g.per) {
      const int nb = b0 + g.per <= nw ? g.per : nw - b0;
      const void* c = static_cast<const uint8_t*>(cur) + static_cast<size_t>(b0) * L * cur_ch * 2;
      const void* oc = c;
      long long Lc = L;
      r = generator_stages(w, c, cur_ch, nb, L, g.n_wide, w->n_stages, nar0, nar1, g.nar_arena, &oc, &Lc, stream);
      if (r) return r;
      const int ch = w->channels[w->n_stages - 1];
      r = b200svs_conv_post_tanh(oc, dt, ch, w->post_w, nb, static_cast<int>(Lc), ch, w->post_taps,
                                 out + static_cast<size_t>(w0 + b0) * S, stream);
      if (r) return r;
    }
  }
  return B200SVS_OK;
}

// ------------------------------------------------------------------------------------------------ WN stack
extern "C" size_t b200svs_wn_workspace_bytes(const b200svs_wn_w* w, int32_t B, int32_t T) {
  if (!w || B <= 0 || T <= 0 || w->n_layers < 1 || w->n_layers > B200SVS_MAX_WN_LAYERS) return 0;
  return align256(static_cast<size_t>(B) * T * w->n_layers * w->hidden * 2);      // every layer's gate output, 16-bit
}

extern "C" int b200svs_wn_stack_fwd(const b200svs_wn_w* w, float* x32, void* x16, const int32_t* lens, int32_t B,
                                    int32_t T, const float* gcond, int32_t gcond_ld, void* out16, void* workspace,
                                    size_t workspace_bytes, void* stream) {
  B200SVS_REQUIRE(w && x32 && x16 && out16 && workspace && B > 0 && T > 0, "wn_stack_fwd: bad argument");
  B200SVS_REQUIRE(w->n_layers >= 1 && w->n_layers <= B200SVS_MAX_WN_LAYERS && w->hidden > 0, "wn_stack_fwd: n_layers=%d hidden=%d",
                  w->n_layers, w->hidden);
  const size_t need = b200svs_wn_workspace_bytes(w, B, T);
  B200SVS_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0 && workspace_bytes >= need,
                  "wn_stack_fwd: workspace must be 256-byte aligned and >= %zu bytes (got %zu)", need, workspace_bytes);
  const int n = w->n_layers, H = w->hidden, dt = w->dtype;
  uint8_t* acts = static_cast<uint8_t*>(workspace);              // [B, T, n*H]: layer i in columns i*H ..
  for (int i = 0; i < n; ++i) {
    void* a_i = acts + static_cast<size_t>(i) * H * 2;
    {   // acts_i = tanh(.) * sigmoid(.) of the k-tap gate conv (+ speaker term), fused in the epilogue
      b200svs_epilogue e; init_ep(&e);
      if (gcond) { e.gbias = gcond + static_cast<size_t>(i) * 2 * H; e.gbias_ld = gcond_ld; }
      e.out16 = a_i; e.out16_ld = n * H; e.out16_dtype = dt;
      int r = run_conv(w->in[i], x16, dt, H, B, T, e, stream);
      if (r) return r;
    }
    if (i < n - 1) {   // x = (x + res_i) * mask: the residual half (rows [0, H)) of res_skip_layers[i]
      b200svs_conv_w c = w->rs[i];
      c.cout_pad = H; c.n_out = H; c.cout = H;
      if (c.block_n > H) c.block_n = H;
      b200svs_epilogue e; init_ep(&e);
      e.res1 = x32; e.res1_ld = H; e.lens = lens; e.out32 = x32; e.out32_ld = H; e.out16 = x16; e.out16_ld = H; e.out16_dtype = dt;
      int r = run_conv(c, a_i, dt, n * H, B, T, e, stream);
      if (r) return r;
    }
  }
  b200svs_epilogue e; init_ep(&e);      // skip sum of all layers as one 1x1 conv over K = n*H
  e.lens = lens; e.out16 = out16; e.out16_ld = H; e.out16_dtype = dt;
  return run_conv(w->skip_cat, acts, dt, n * H, B, T, e, stream);
}

// ------------------------------------------------------------------------------------------------ encoder layers
namespace b200svs {
struct EncPlan {
  size_t qkv, tmp, att, hid, qk, vt, total;
  int Tp;
};
static int plan_encoder(const b200svs_encoder_layer_w* w, int B, int T, EncPlan* p) {
  B200SVS_REQUIRE(w && B > 0 && T > 0 && w->hidden > 0 && w->filter > 0 && w->mode >= 0 && w->mode <= 2,
                  "encoder: bad argument (B=%d T=%d)", B, T);
  const size_t C = w->hidden, F = w->filter, rows = static_cast<size_t>(B) * T;
  p->Tp = (T + 7) / 8 * 8;
  p->qkv = align256(rows * 3 * C * 4);
  p->tmp = align256(rows * C * 4);
  const bool f32 = w->mode == 0, split = w->mode == 2;
  p->att = align256(f32 ? rows * C * 4 : rows * C * 2 * (split ? 3 : 1));
  p->hid = align256(f32 ? rows * F * 4 : rows * F * 2 * (split ? 3 : 1));
  p->qk = f32 ? 0 : align256(rows * 2 * C * 2);
  p->vt = f32 ? 0 : align256(static_cast<size_t>(B) * C * p->Tp * 2);
  p->total = p->qkv + p->tmp + p->att + p->hid + (split ? 2 : 1) * (p->qk + p->vt);
  return B200SVS_OK;
}
}  // namespace b200svs

extern "C" size_t b200svs_encoder_workspace_bytes(const b200svs_encoder_layer_w* w, int32_t B, int32_t T) {
  EncPlan p = {};
  if (plan_encoder(w, B, T, &p)) return 0;
  return p.total;
}

extern "C" int b200svs_encoder_fwd(const b200svs_encoder_layer_w* w, int32_t n_layers, float* x32, void* x16,
                                   const int32_t* lens, int32_t B, int32_t T, void* workspace, size_t workspace_bytes,
                                   void* stream) {
  EncPlan p = {};
  int r = plan_encoder(w, B, T, &p);
  if (r) return r;
  B200SVS_REQUIRE(n_layers >= 1 && x32 && workspace && (w->mode == 0 || x16), "encoder_fwd: bad argument");
  B200SVS_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0 && workspace_bytes >= p.total,
                  "encoder_fwd: workspace must be 256-byte aligned and >= %zu bytes (got %zu)", p.total, workspace_bytes);
  Arena ar(workspace, workspace_bytes);
  const int C = w->hidden, F = w->filter, mode = w->mode;
  const bool f32 = mode == 0, split = mode == 2;
  float* qkv = static_cast<float*>(ar.take(p.qkv));
  float* tmp = static_cast<float*>(ar.take(p.tmp));
  void* att = ar.take(p.att);
  void* hid = ar.take(p.hid);
  void *qk_hi = nullptr, *qk_lo = nullptr, *vt_hi = nullptr, *vt_lo = nullptr;
  if (!f32) {
    qk_hi = ar.take(p.qk); vt_hi = ar.take(p.vt);
    if (split) { qk_lo = ar.take(p.qk); vt_lo = ar.take(p.vt); }
  }
  for (int l = 0; l < n_layers; ++l) {
    const b200svs_encoder_layer_w& L = w[l];
    B200SVS_REQUIRE(L.hidden == C && L.filter == F && L.mode == mode, "encoder_fwd: layer %d differs in shape / mode", l);
    const int dt = split ? B200SVS_DT_F16 : L.dtype;
    const void* xin = f32 ? static_cast<const void*>(x32) : x16;
    const int xin_ld = split ? 3 * C : C;
    const int o16 = split ? B200SVS_DT_F16_SPLIT : dt;
    {   // q | k | v = 1x1 convs (attentions.py:138-146)
      b200svs_epilogue e; init_ep(&e);
      e.out32 = qkv; e.out32_ld = 3 * C;
      r = run_conv(L.qkv, xin, dt, xin_ld, B, T, e, stream);
      if (r) return r;
    }
    if (f32) {
      r = b200svs_attention_f32(qkv, 3 * C, static_cast<float*>(att), C, nullptr, 0, 0, L.rel_k, L.rel_v, lens, B, T, C, L.heads,
                                L.window, stream);
      if (r) return r;
    } else {
      r = b200svs_attn_prep(qkv, 3 * C, qk_hi, qk_lo, vt_hi, vt_lo, dt, B, T, p.Tp, C, stream);
      if (r) return r;
      r = b200svs_attention_tc(qk_hi, qk_lo, vt_hi, vt_lo, dt, p.Tp, nullptr, 0, att, split ? 3 * C : C, dt, split ? 1 : 0, L.rel_k,
                               L.rel_v, lens, B, T, C, L.heads, L.window, stream);
      if (r) return r;
    }
    {   // x + conv_o(attention)
      b200svs_epilogue e; init_ep(&e);
      e.res1 = x32; e.res1_ld = C; e.out32 = tmp; e.out32_ld = C;
      r = run_conv(L.o, att, dt, split ? 3 * C : C, B, T, e, stream);
      if (r) return r;
    }
    r = b200svs_layernorm(tmp, C, nullptr, 0, L.ln1_g, L.ln1_b, 1e-5f, lens, x32, C, f32 ? nullptr : x16, f32 ? 0 : xin_ld,
                          f32 ? 0 : o16, B, T, C, stream);
    if (r) return r;
    {   // FFN (attentions.py:277-285): conv k -> ReLU -> mask -> conv k -> mask
      b200svs_epilogue e; init_ep(&e);
      e.act = B200SVS_ACT_RELU; e.lens = lens;
      if (f32) { e.out32 = static_cast<float*>(hid); e.out32_ld = F; }
      else { e.out16 = hid; e.out16_ld = split ? 3 * F : F; e.out16_dtype = o16; }
      r = run_conv(L.ffn1, xin, dt, xin_ld, B, T, e, stream);
      if (r) return r;
    }
    {
      b200svs_epilogue e; init_ep(&e);
      e.res1 = x32; e.res1_ld = C; e.lens = lens; e.out32 = tmp; e.out32_ld = C;
      r = run_conv(L.ffn2, hid, dt, split ? 3 * F : F, B, T, e, stream);
      if (r) return r;
    }
    r = b200svs_layernorm(tmp, C, nullptr, 0, L.ln2_g, L.ln2_b, 1e-5f, lens, x32, C, f32 ? nullptr : x16, f32 ? 0 : xin_ld,
                          f32 ? 0 : o16, B, T, C, stream);
    if (r) return r;
  }
  return B200SVS_OK;
}
