"""Synthetic inputs and the seeded weight recipe used by tests and bench.

No trained checkpoint ships with the reference (.MISSING_LARGE_BLOBS), so parity and
throughput are measured on random-init weights of the shipped config.  Plain random init
makes parts of the graph vacuous: `ResidualCouplingLayer.post` is
zero-initialised (VISinger/modules.py:326-327), Generator convs have std 0.01
(VISinger/commons.py:8-11) and log_f0 ~ 0 collapses every f0 bucket to 1.  `apply_recipe`
de-zeroes / rescales those tensors *in the state-dict that both sides load*.
"""
import math

import torch

# integer input domains: VISinger/phone_set.json (61 phones + 3 reserved),
# VISinger/phone_shengdiao_set.json (247 + 3), cpop_text_preprocessor.py:79-87 (midi, 0 = rest)
PHONE_LO, PHONE_HI = 3, 63
TONE_LO, TONE_HI = 3, 249
PITCH_LO, PITCH_HI = 48, 71


def make_inputs(batch, n_phones, seed=1234, ragged=False, device="cpu"):
    """Phoneme-level inputs with the value domains of SURVEY.md 8(d)."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randint(PHONE_LO, PHONE_HI + 1, (batch, n_phones), generator=g)
    sd = torch.randint(TONE_LO, TONE_HI + 1, (batch, n_phones), generator=g)
    pitch = torch.randint(PITCH_LO, PITCH_HI + 1, (batch, n_phones), generator=g)
    rest = torch.rand(batch, n_phones, generator=g) < 0.10
    pitch = torch.where(rest, torch.zeros_like(pitch), pitch)
    dur = (0.15 + 0.65 * torch.rand(batch, n_phones, generator=g)).float()
    if ragged:
        lo = max(1, n_phones // 2)
        xl = torch.randint(lo, n_phones + 1, (batch,), generator=g)
        xl[0] = n_phones
    else:
        xl = torch.full((batch,), n_phones, dtype=torch.int64)
    out = dict(x=x, x_lengths=xl, note_pitch_id=pitch, note_duration=dur, phone_shengdiao=sd)
    return {k: v.to(device) for k, v in out.items()}


def make_ph_dur(x_lengths, n_phones, frames, seed=1234):
    """Random positive integer durations per phoneme with sum == frames for the full-length
    rows (shorter rows get proportionally fewer frames); padded phoneme slots get 0."""
    g = torch.Generator().manual_seed(seed + 1)
    b = x_lengths.shape[0]
    out = torch.zeros(b, n_phones, dtype=torch.int64)
    for i in range(b):
        p = int(x_lengths[i])
        total = max(p, int(round(frames * p / n_phones)))
        w = torch.rand(p, generator=g) + 0.2
        d = torch.floor(w / w.sum() * (total - p)).long() + 1
        d[: total - int(d.sum())] += 1
        out[i, :p] = d
    return out


def make_noise(batch, channels, frames, seed=1234, device="cpu"):
    g = torch.Generator().manual_seed(seed + 2)
    return torch.randn(batch, channels, frames, generator=g).to(device)


def apply_recipe(sd, seed=1234, generator_gain=0.7, f0_bias=math.log(250.0), positions=None, pre_gain=None, post_gain=1.0):
    """Seeded, in-place edit of a state-dict (SURVEY.md 7.3-A / 8c "weight recipe").

    * spec_flow.*.post.{weight,bias} ~ N(0, 0.05)   (zero-init in the reference => flow vacuous)
    * decoder weight_v / weight_g rescaled so activations stay O(1) (He-style gain on g)
    * f0_predictor.output_layer.bias = log(250)      (spread f0 buckets)
    * optional: enlarge the 2000-row position table for T > 2000
    * pre_gain / post_gain scale conv_pre / conv_post: pre_gain * s with post_gain / s keeps the waveform O(1) while the
      Generator's activations grow by s (numeric-range tests: fp16 tops out at 65504)
    """
    g = torch.Generator().manual_seed(seed + 7)
    for k in list(sd.keys()):
        if k.startswith("spec_flow.") and (k.endswith(".post.weight") or k.endswith(".post.bias")):
            sd[k] = torch.randn(sd[k].shape, generator=g) * 0.05
    if generator_gain:
        for k in list(sd.keys()):
            if not k.startswith("decoder."):
                continue
            if k.endswith(".weight_g"):
                v = sd[k[:-1] + "v"]
                fan = v[0].numel() if "ups" not in k else v.shape[0] * v.shape[2] / max(1, _stride_of(k))
                # weight_norm: |w_row| = g.  Target per-weight std = gain * sqrt(2 / fan_in)
                n_per_g = v.numel() / sd[k].numel()
                sd[k] = torch.full_like(sd[k], generator_gain * math.sqrt(2.0 / fan) * math.sqrt(n_per_g))
            elif k in ("decoder.conv_pre.weight", "decoder.conv_post.weight"):
                w = sd[k]
                fan = w[0].numel()
                # generator input std at random init: spec_hidden ~0.14 (ae), z ~1.0 (v4, no spec_proj)
                default_pre = 4.0 if "spec_proj.weight" in sd else 0.6
                gain = (pre_gain or default_pre) if "pre" in k else post_gain
                sd[k] = torch.randn(w.shape, generator=g) * gain * math.sqrt(1.0 / fan)
            elif k.endswith(".bias"):
                sd[k] = torch.randn(sd[k].shape, generator=g) * 0.05
    k = "text_post_net.f0_predictor.output_layer.bias"
    if k in sd and f0_bias is not None:
        sd[k] = torch.full_like(sd[k], f0_bias)
        kw = "text_post_net.f0_predictor.output_layer.weight"
        sd[kw] = torch.randn(sd[kw].shape, generator=g) * 0.04
    if positions is not None:
        kp = "text_post_net.position_embeddings.weight"
        if sd[kp].shape[0] < positions:
            extra = torch.randn(positions - sd[kp].shape[0], sd[kp].shape[1], generator=g)
            sd[kp] = torch.cat([sd[kp], extra], 0)
            sd["text_post_net.position_ids"] = torch.arange(positions).unsqueeze(0)
    return sd


def _stride_of(key):
    # decoder.ups.{i}: upsample_rates [8, 8, 2, 2]
    i = int(key.split(".")[2])
    return [8, 8, 2, 2][i]
