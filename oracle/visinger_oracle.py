"""TEST INFRASTRUCTURE ONLY - CPU/fp32 restatement of the VISinger synthesis forward.

This is the *oracle* for the B200 kernels: a functional, state-dict driven PyTorch fp32
restatement of `SynthesizerTrn.infer` (reference VISinger/models_ae.py:608-651 and
VISinger/models.py:622-673).  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / `--impl reference` legs may import it; the product package
(music_generation_b200/) never does.

Parity status: PINNED against the real reference *run in the build container*
(oracle/ref_harness.py imports /root/reference unmodified; oracle/make_golden.py writes
tests/golden/*.npz; tests/test_oracle.py re-checks this file against those fixtures).
The reference itself ships no golden vectors / known-answer tests.

Every function cites the reference lines it restates.  Tensors are [B, C, T] fp32 like the
reference.  `taps` (optional dict) collects named intermediates for per-stage parity.
"""
import math

import torch
import torch.nn.functional as F

LRELU_SLOPE = 0.1          # VISinger/modules.py:17
FRAME_SECONDS = 0.0116     # VISinger/models_ae.py:625
WINDOW = 4                 # VISinger/attentions.py:14 (window_size default)

F0_BIN = 256               # VISinger/utils/pitch_utils.py:15-19
F0_MAX = 1100.0
F0_MIN = 50.0
F0_MEL_MIN = 1127 * math.log(1 + F0_MIN / 700)
F0_MEL_MAX = 1127 * math.log(1 + F0_MAX / 700)


# --------------------------------------------------------------------------------------
# parameter helpers
# --------------------------------------------------------------------------------------
def wn_weight(sd, prefix):
    """Effective conv weight.  Old-style torch weight_norm (dim=0): w = g * v / ||v||,
    norm over every dim but 0 (VISinger/models_ae.py:311, modules.py:135,145,191-206)."""
    if prefix + ".weight" in sd:
        return sd[prefix + ".weight"]
    v = sd[prefix + ".weight_v"]
    g = sd[prefix + ".weight_g"]
    return torch._weight_norm(v, g, 0)


def _bias(sd, prefix):
    return sd.get(prefix + ".bias")


def conv(sd, prefix, x, dilation=1, padding=0):
    return F.conv1d(x, wn_weight(sd, prefix), _bias(sd, prefix), dilation=dilation, padding=padding)


def sequence_mask(length, max_length):
    """VISinger/commons.py:121-125"""
    r = torch.arange(max_length, dtype=length.dtype, device=length.device)
    return r.unsqueeze(0) < length.unsqueeze(1)


def layer_norm_c(sd, prefix, x, eps=1e-5):
    """Channel LayerNorm, VISinger/modules.py:29-32"""
    c = x.shape[1]
    y = F.layer_norm(x.transpose(1, -1), (c,), sd[prefix + ".gamma"], sd[prefix + ".beta"], eps)
    return y.transpose(1, -1)


# --------------------------------------------------------------------------------------
# attentions.Encoder (VISinger/attentions.py:13-47, 101-254, 257-303)
# --------------------------------------------------------------------------------------
def _rel_embeddings(emb, length):
    """attentions.py:199-212: pad/slice the [1, 9, d] table to [1, 2L-1, d]."""
    pad_length = max(length - (WINDOW + 1), 0)
    start = max((WINDOW + 1) - length, 0)
    end = start + 2 * length - 1
    if pad_length > 0:
        emb = F.pad(emb, [0, 0, pad_length, pad_length, 0, 0])
    return emb[:, start:end]


def _rel_to_abs(x):
    """attentions.py:214-229"""
    b, h, l, _ = x.size()
    x = F.pad(x, [0, 1, 0, 0, 0, 0, 0, 0])
    x_flat = x.view([b, h, l * 2 * l])
    x_flat = F.pad(x_flat, [0, l - 1, 0, 0, 0, 0])
    return x_flat.view([b, h, l + 1, 2 * l - 1])[:, :, :l, l - 1:]


def _abs_to_rel(x):
    """attentions.py:231-243"""
    b, h, l, _ = x.size()
    x = F.pad(x, [0, l - 1, 0, 0, 0, 0, 0, 0])
    x_flat = x.view([b, h, l ** 2 + l * (l - 1)])
    x_flat = F.pad(x_flat, [l, 0, 0, 0, 0, 0])
    return x_flat.view([b, h, l, 2 * l])[:, :, :, 1:]


def multi_head_attention(sd, prefix, x, attn_mask, n_heads=2):
    """attentions.py:138-179 (self-attention, window_size=4, heads_share=True)."""
    q = conv(sd, prefix + ".conv_q", x)
    k = conv(sd, prefix + ".conv_k", x)
    v = conv(sd, prefix + ".conv_v", x)
    b, d, t = k.size()
    dk = d // n_heads
    q = q.view(b, n_heads, dk, t).transpose(2, 3)
    k = k.view(b, n_heads, dk, t).transpose(2, 3)
    v = v.view(b, n_heads, dk, t).transpose(2, 3)
    qs = q / math.sqrt(dk)
    scores = torch.matmul(qs, k.transpose(-2, -1))
    rel_k = _rel_embeddings(sd[prefix + ".emb_rel_k"], t)
    rel_logits = torch.matmul(qs, rel_k.unsqueeze(0).transpose(-2, -1))
    scores = scores + _rel_to_abs(rel_logits)
    scores = scores.masked_fill(attn_mask == 0, -1e4)
    p = F.softmax(scores, dim=-1)
    out = torch.matmul(p, v)
    rel_w = _abs_to_rel(p)
    rel_v = _rel_embeddings(sd[prefix + ".emb_rel_v"], t)
    out = out + torch.matmul(rel_w, rel_v.unsqueeze(0))
    out = out.transpose(2, 3).contiguous().view(b, d, t)
    return conv(sd, prefix + ".conv_o", out)


def ffn(sd, prefix, x, x_mask, kernel_size=3):
    """attentions.py:277-303 (same padding, ReLU)."""
    pl, pr = (kernel_size - 1) // 2, kernel_size // 2
    y = conv(sd, prefix + ".conv_1", F.pad(x * x_mask, [pl, pr]))
    y = torch.relu(y)
    y = conv(sd, prefix + ".conv_2", F.pad(y * x_mask, [pl, pr]))
    return y * x_mask


def encoder(sd, prefix, x, x_mask, n_layers):
    """attentions.py:35-47 (post-LN transformer, dropout = identity in eval)."""
    attn_mask = x_mask.unsqueeze(2) * x_mask.unsqueeze(-1)
    x = x * x_mask
    for i in range(n_layers):
        y = multi_head_attention(sd, f"{prefix}.attn_layers.{i}", x, attn_mask)
        x = layer_norm_c(sd, f"{prefix}.norm_layers_1.{i}", x + y)
        y = ffn(sd, f"{prefix}.ffn_layers.{i}", x, x_mask)
        x = layer_norm_c(sd, f"{prefix}.norm_layers_2.{i}", x + y)
    return x * x_mask


def n_encoder_layers(sd, prefix):
    n = 0
    while f"{prefix}.attn_layers.{n}.conv_q.weight" in sd:
        n += 1
    return n


# --------------------------------------------------------------------------------------
# phoneme level: TextEncoder, DurationPredictor, quantisation, LengthRegulator
# --------------------------------------------------------------------------------------
def text_encoder(sd, x, note_pitch_id, note_duration, x_lengths, phone_shengdiao, p="text_encoder"):
    """VISinger/models_ae.py:203-220"""
    h = sd[p + ".emb_phoneme.weight"].shape[1]
    e_ph = F.embedding(x, sd[p + ".emb_phoneme.weight"])
    e_pitch = F.embedding(note_pitch_id, sd[p + ".emb_note_pitch.weight"])
    e_dur = F.linear(note_duration.unsqueeze(-1), sd[p + ".emb_note_duration.weight"],
                     sd[p + ".emb_note_duration.bias"])
    e_sd = F.embedding(phone_shengdiao, sd[p + ".emb_phone_shengdiao.weight"])
    emb = (e_ph + e_dur + e_pitch + e_sd) * math.sqrt(h)
    xx = emb.transpose(1, -1)
    x_mask = sequence_mask(x_lengths, xx.size(2)).unsqueeze(1).to(xx.dtype)
    y = encoder(sd, p + ".encoder", xx * x_mask, x_mask, n_encoder_layers(sd, p + ".encoder"))
    return y * x_mask, x_mask


def duration_predictor(sd, x, x_mask, p="duration_predictor"):
    """VISinger/models_ae.py:126-150"""
    for i in (1, 2, 3):
        x = conv(sd, f"{p}.conv_{i}", x * x_mask, padding=1)
        x = torch.relu(x)
        x = layer_norm_c(sd, f"{p}.norm_{i}", x)
    x = conv(sd, p + ".proj", x * x_mask)
    x = torch.sigmoid(x) * x_mask
    return x.squeeze(1)


def quantize_durations(pred, note_duration):
    """VISinger/models_ae.py:624-625 - the float->int boundary."""
    return torch.floor(pred * note_duration / FRAME_SECONDS).long()


def length_regulate(h_text, durations):
    """VISinger/models_ae.py:689-721 + utils/utils.py:344-361, restated as an index gather:
    frame t of utterance b copies phoneme column i where cumsum(dur)[i-1] <= t < cumsum(dur)[i];
    frames >= mel_len[b] are zero.  Returns ([B,C,T], mel_len[B] int64, index[B,T] int64 (-1 = pad))."""
    b, c, p = h_text.shape
    dur = durations.clamp_min(0).to(torch.int64)
    mel_len = dur.sum(1)
    t_max = int(mel_len.max().item()) if b > 0 else 0
    ends = dur.cumsum(1)                                  # [B,P]
    frames = torch.arange(t_max, device=h_text.device).unsqueeze(0).expand(b, t_max)
    idx = torch.searchsorted(ends, frames.contiguous(), right=True)     # first i with ends[i] > t
    valid = frames < mel_len.unsqueeze(1)
    idx = torch.where(valid, idx, torch.full_like(idx, -1))
    gathered = torch.gather(h_text, 2, idx.clamp_min(0).unsqueeze(1).expand(b, c, t_max))
    out = gathered * valid.unsqueeze(1).to(h_text.dtype)
    return out, mel_len, idx


# --------------------------------------------------------------------------------------
# frame level: TextPostNet (ext_modules/text_post_net.py:64-94)
# --------------------------------------------------------------------------------------
def f0_to_coarse(f0):
    """VISinger/utils/pitch_utils.py:22-31 (torch branch), without boolean-mask indexing."""
    f0_mel = 1127 * (1 + f0 / 700).log()
    scaled = (f0_mel - F0_MEL_MIN) * (F0_BIN - 2) / (F0_MEL_MAX - F0_MEL_MIN) + 1
    f0_mel = torch.where(f0_mel > 0, scaled, f0_mel)
    f0_mel = torch.where(f0_mel <= 1, torch.ones_like(f0_mel), f0_mel)
    f0_mel = torch.where(f0_mel > F0_BIN - 1, torch.full_like(f0_mel, F0_BIN - 1), f0_mel)
    return (f0_mel + 0.5).long()


def f0_predictor(sd, x, x_mask, p="text_post_net.f0_predictor"):
    """ext_modules/f0_predictor.py:35-41"""
    h = encoder(sd, p + ".encoder", x * x_mask, x_mask, n_encoder_layers(sd, p + ".encoder"))
    lf0 = conv(sd, p + ".output_layer", h * x_mask) * x_mask
    return lf0.squeeze(1)


def mel_predict_network(sd, x, cond, x_mask, p="text_post_net.mel_predict_net"):
    """ext_modules/mel_predict_network.py:37-45"""
    x = x + conv(sd, p + ".condition_proj", cond)
    x = x * x_mask
    x = encoder(sd, p + ".encoder", x, x_mask, n_encoder_layers(sd, p + ".encoder")) * x_mask
    return conv(sd, p + ".output_layer", x) * x_mask


def frame_prior_network(sd, x, cond, x_mask, p="text_post_net.frame_prior_net"):
    """ext_modules/frame_prior_network.py:31-41 (exact erf GELU)"""
    x = x + conv(sd, p + ".condition_proj", cond)
    x = x * x_mask
    i = 0
    while f"{p}.conv_layers.{i}.weight" in sd:
        x = F.gelu(conv(sd, f"{p}.conv_layers.{i}", x)) * x_mask
        i += 1
    return x


def text_post_net(sd, x, x_mask, f0_score=None, taps=None, p="text_post_net"):
    """ext_modules/text_post_net.py:64-94.  x: [B,192,T] LR output; x_mask [B,1,T] (int or float)."""
    t = x.shape[-1]
    pos = sd[p + ".position_embeddings.weight"][:t]           # raises if T > table (reference: 2000)
    if pos.shape[0] != t:
        raise RuntimeError(f"T={t} exceeds position table {pos.shape[0]} (text_post_net.py:26)")
    x_emb = (x + pos.t().unsqueeze(0)) * x_mask
    log_f0 = f0_predictor(sd, x_emb, x_mask)
    if taps is not None:
        taps["f0_mel_pre"] = log_f0
    if f0_score is None:
        f0_score = f0_to_coarse(torch.exp(log_f0) - 1.0)
    f0_emb = F.embedding(f0_score, sd[p + ".f0_emb.weight"]).transpose(1, -1) * x_mask
    mel = mel_predict_network(sd, x_emb, f0_emb, x_mask)
    hid = frame_prior_network(sd, mel, f0_emb, x_mask)
    stats = conv(sd, p + ".proj", hid * x_mask) * x_mask
    c = stats.shape[1] // 2
    if taps is not None:
        taps.update(x_emb=x_emb, f0_score=f0_score, f0_emb=f0_emb, frame_prior=hid)
    return stats[:, :c], stats[:, c:], log_f0, mel, f0_score


# --------------------------------------------------------------------------------------
# WN / reverse flow / SpecDecoder (modules.py:148-176, 329-349; models_ae.py:246-253;
# ext_modules/spec_decoder.py:41-46)
# --------------------------------------------------------------------------------------
def wn(sd, prefix, x, x_mask, g=None):
    """modules.py:148-176; fused_add_tanh_sigmoid_multiply = commons.py:100-107."""
    hc = x.shape[1]
    n_layers = 0
    while f"{prefix}.in_layers.{n_layers}.bias" in sd:
        n_layers += 1
    out = torch.zeros_like(x)
    if g is not None:
        g = conv(sd, prefix + ".cond_layer", g)
    for i in range(n_layers):
        k = wn_weight(sd, f"{prefix}.in_layers.{i}").shape[-1]
        x_in = conv(sd, f"{prefix}.in_layers.{i}", x, padding=(k - 1) // 2)   # dilation_rate 1
        if g is not None:
            x_in = x_in + g[:, i * 2 * hc:(i + 1) * 2 * hc]
        acts = torch.tanh(x_in[:, :hc]) * torch.sigmoid(x_in[:, hc:])
        rs = conv(sd, f"{prefix}.res_skip_layers.{i}", acts)
        if i < n_layers - 1:
            x = (x + rs[:, :hc]) * x_mask
            out = out + rs[:, hc:]
        else:
            out = out + rs
    return out * x_mask


def coupling_reverse(sd, prefix, x, x_mask, g=None):
    """modules.py:329-349 with reverse=True, mean_only=True (logs == 0)."""
    half = x.shape[1] // 2
    x0, x1 = x[:, :half], x[:, half:]
    h = conv(sd, prefix + ".pre", x0) * x_mask
    h = wn(sd, prefix + ".enc", h, x_mask, g=g)
    m = conv(sd, prefix + ".post", h) * x_mask
    x1 = (x1 - m) * x_mask
    return torch.cat([x0, x1], 1)


def flow_reverse(sd, z_p, x_mask, g=None, p="spec_flow"):
    """models_ae.py:246-253, reverse: Flip, RCL3, Flip, RCL2, Flip, RCL1, Flip, RCL0."""
    n = 0
    while f"{p}.spec_flows.{2 * n}.pre.weight" in sd:
        n += 1
    x = z_p
    for i in reversed(range(n)):
        x = torch.flip(x, [1])
        x = coupling_reverse(sd, f"{p}.spec_flows.{2 * i}", x, x_mask, g=g)
    return x


def coupling_forward(sd, prefix, x, x_mask, g=None):
    """modules.py:329-349 with reverse=False, mean_only=True (logs == 0, logdet unused here)."""
    half = x.shape[1] // 2
    x0, x1 = x[:, :half], x[:, half:]
    h = conv(sd, prefix + ".pre", x0) * x_mask
    h = wn(sd, prefix + ".enc", h, x_mask, g=g)
    m = conv(sd, prefix + ".post", h) * x_mask
    x1 = m + x1 * x_mask
    return torch.cat([x0, x1], 1)


def flow_forward(sd, z, x_mask, g=None, p="spec_flow"):
    """models_ae.py:246-253, forward (training side: z_p = flow(z)): RCL0, Flip, RCL1, Flip, ..."""
    n = 0
    while f"{p}.spec_flows.{2 * n}.pre.weight" in sd:
        n += 1
    x = z
    for i in range(n):
        x = coupling_forward(sd, f"{p}.spec_flows.{2 * i}", x, x_mask, g=g)
        x = torch.flip(x, [1])
    return x


def spec_decoder(sd, z, mel_len, g=None, p="spec_decoder"):
    x_mask = sequence_mask(mel_len, z.size(2)).unsqueeze(1).to(z.dtype)
    x = conv(sd, p + ".pre", z) * x_mask
    x = wn(sd, p + ".enc", x, x_mask, g=g)
    return conv(sd, p + ".proj", x) * x_mask


# --------------------------------------------------------------------------------------
# Generator (models_ae.py:327-346; ResBlock1/2 modules.py:210-223, 243-252)
# --------------------------------------------------------------------------------------
def posterior_encoder(sd, y, y_lengths, noise, g=None, p="spec_encoder"):
    """PosteriorEncoder.forward (models_ae.py:284-291); `noise` replaces torch.randn_like(m).
    Returns (z, m, logs, x_mask, x) like the reference."""
    x_mask = sequence_mask(y_lengths, y.size(2)).unsqueeze(1).to(y.dtype)
    x = conv(sd, p + ".pre", y) * x_mask
    x = wn(sd, p + ".enc", x, x_mask, g=g)
    stats = conv(sd, p + ".proj", x) * x_mask
    m, logs = torch.split(stats, stats.size(1) // 2, dim=1)
    z = (m + noise * torch.exp(logs)) * x_mask
    return z, m, logs, x_mask, x


def resblock1(sd, prefix, x, dilations=(1, 3, 5)):
    for j, d in enumerate(dilations):
        k = wn_weight(sd, f"{prefix}.convs1.{j}").shape[-1]
        xt = F.leaky_relu(x, LRELU_SLOPE)
        xt = conv(sd, f"{prefix}.convs1.{j}", xt, dilation=d, padding=(k * d - d) // 2)
        xt = F.leaky_relu(xt, LRELU_SLOPE)
        xt = conv(sd, f"{prefix}.convs2.{j}", xt, padding=(k - 1) // 2)
        x = xt + x
    return x


def resblock2(sd, prefix, x, dilations=(1, 3)):
    for j, d in enumerate(dilations):
        k = wn_weight(sd, f"{prefix}.convs.{j}").shape[-1]
        xt = F.leaky_relu(x, LRELU_SLOPE)
        xt = conv(sd, f"{prefix}.convs.{j}", xt, dilation=d, padding=(k * d - d) // 2)
        x = xt + x
    return x


def generator(sd, x, hp, g=None, taps=None, p="decoder"):
    rates, ksz = hp["upsample_rates"], hp["upsample_kernel_sizes"]
    rks, rds = hp["resblock_kernel_sizes"], hp["resblock_dilation_sizes"]
    x = conv(sd, p + ".conv_pre", x, padding=3)
    if g is not None:
        x = x + conv(sd, p + ".cond", g)
    nk = len(rks)
    for i, (u, k) in enumerate(zip(rates, ksz)):
        x = F.leaky_relu(x, LRELU_SLOPE)
        x = F.conv_transpose1d(x, wn_weight(sd, f"{p}.ups.{i}"), _bias(sd, f"{p}.ups.{i}"),
                               stride=u, padding=(k - u) // 2)
        if taps is not None:
            taps[f"ups{i}"] = x
        xs = None
        for j in range(nk):
            pre = f"{p}.resblocks.{i * nk + j}"
            y = resblock1(sd, pre, x, rds[j]) if str(hp["resblock"]) == "1" else resblock2(sd, pre, x, rds[j])
            xs = y if xs is None else xs + y
        x = xs / nk
        if taps is not None:
            taps[f"stage{i}"] = x
    x = F.leaky_relu(x)                      # default slope 0.01 (models_ae.py:342)
    x = conv(sd, p + ".conv_post", x, padding=3)
    return torch.tanh(x)


# --------------------------------------------------------------------------------------
# SynthesizerTrn.infer
# --------------------------------------------------------------------------------------
DEFAULT_HP = dict(resblock="1", resblock_kernel_sizes=[3, 7, 11],
                  resblock_dilation_sizes=[[1, 3, 5]] * 3, upsample_rates=[8, 8, 2, 2],
                  upsample_kernel_sizes=[16, 16, 4, 4])


def infer(sd, x, x_lengths, note_pitch_id, note_duration, phone_shengdiao, noise, ph_dur=None,
          noise_scale=1.0, max_len=None, variant="ae", hp=None, g=None, f0_score=None, taps=None):
    """models_ae.py:608-651 (variant 'ae', 6-tuple) / models.py:622-673 (variant 'v4', 5-tuple).
    `noise` replaces torch.randn_like(m_p); `g`, `f0_score` are oracle-only injections
    (the reference hard-codes g=None and predicts f0)."""
    hp = hp or DEFAULT_HP
    h_text, x_mask = text_encoder(sd, x, note_pitch_id, note_duration, x_lengths, phone_shengdiao)
    dur_pred = None
    if ph_dur is None:
        dur_pred = duration_predictor(sd, h_text, x_mask)
        durations = quantize_durations(dur_pred, note_duration)
    else:
        durations = ph_dur
    h_lr, mel_len, lr_index = length_regulate(h_text, durations)
    y_mask = sequence_mask(mel_len, h_lr.size(2)).unsqueeze(1).to(x.dtype)
    m_p, logs_p, log_f0, mel, f0_used = text_post_net(sd, h_lr, y_mask, f0_score=f0_score, taps=taps)
    z_p = m_p + noise * torch.exp(logs_p) * noise_scale
    z = flow_reverse(sd, z_p, y_mask, g=g)
    if taps is not None:
        taps.update(h_text=h_text, dur_pred=dur_pred, durations=durations, mel_len=mel_len,
                    lr_index=lr_index, h_lr=h_lr, f0_coarse=f0_used)
    if variant == "ae":
        spec = spec_decoder(sd, z, mel_len, g=g)
        hid = conv(sd, "spec_proj", spec)
        if taps is not None:
            taps["spec_hidden"] = hid
        o = generator(sd, (hid * y_mask)[:, :, :max_len], hp, g=g, taps=taps)
        return o, y_mask, (z, z_p, m_p, logs_p), log_f0, spec, mel
    o = generator(sd, (z * y_mask)[:, :, :max_len], hp, g=g, taps=taps)
    return o, y_mask, (z, z_p, m_p, logs_p), log_f0, mel


def infer_long(sd, x, x_lengths, note_pitch_id, note_duration, phone_shengdiao, noise, plan, ph_dur=None, noise_scale=1.0,
               variant="ae", hp=None, g=None, taps=None):
    """Composition checked by tests/test_gpu_long.py for utterances beyond the reference's 2000-frame position table:
    the reference's modules exactly as `infer` calls them, except that TextPostNet (the only module with the limit,
    text_post_net.py:26,69-72) is evaluated per segment of `plan` = [(b, core0, core1, ctx0, ctx1)] on the
    length-regulated frames [ctx0, ctx1) - positions restart at ctx0 - and the core frames' m_p / logs_p / log_f0 / mel
    are concatenated; flow, SpecDecoder and Generator then run over the full length (they are convolutional)."""
    hp = hp or DEFAULT_HP
    h_text, x_mask = text_encoder(sd, x, note_pitch_id, note_duration, x_lengths, phone_shengdiao)
    if ph_dur is None:
        durations = quantize_durations(duration_predictor(sd, h_text, x_mask), note_duration)
    else:
        durations = ph_dur
    h_lr, mel_len, lr_index = length_regulate(h_text, durations)
    B, _, T = h_lr.shape
    y_mask = sequence_mask(mel_len, T).unsqueeze(1).to(x.dtype)
    m_p, logs_p = torch.zeros(B, 192, T), torch.zeros(B, 192, T)
    log_f0, mel, f0c = torch.zeros(B, T), torch.zeros(B, 80, T), torch.zeros(B, T, dtype=torch.long)
    for b, c0, c1, x0, x1 in plan:
        seg = h_lr[b:b + 1, :, x0:x1].clone()
        ones = torch.ones(1, 1, x1 - x0, dtype=x.dtype)
        sm, sl, sf, smel, sc = text_post_net(sd, seg, ones)
        a, e = c0 - x0, c1 - x0
        m_p[b, :, c0:c1], logs_p[b, :, c0:c1], log_f0[b, c0:c1] = sm[0, :, a:e], sl[0, :, a:e], sf[0, a:e]
        mel[b, :, c0:c1], f0c[b, c0:c1] = smel[0, :, a:e], sc[0, a:e]
    z_p = m_p + noise * torch.exp(logs_p) * noise_scale
    z = flow_reverse(sd, z_p, y_mask, g=g)
    if taps is not None:
        taps.update(durations=durations, mel_len=mel_len, lr_index=lr_index, f0_coarse=f0c)
    if variant == "ae":
        spec = spec_decoder(sd, z, mel_len, g=g)
        hid = conv(sd, "spec_proj", spec)
        o = generator(sd, hid * y_mask, hp, g=g)
        return o, y_mask, (z, z_p, m_p, logs_p), log_f0, spec, mel
    o = generator(sd, z * y_mask, hp, g=g)
    return o, y_mask, (z, z_p, m_p, logs_p), log_f0, mel
