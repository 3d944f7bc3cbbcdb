"""TEST DOUBLE of music_generation_b200.ops: a plain-torch CPU emulation of every C-ABI call's
*documented* semantics (include/b200svs.h).  It lets the CPU test-suite run the whole host-side
orchestration of SynthesizerTrn.infer (weight packing, Flip folding, polyphase ConvTranspose, buffer
aliasing, epilogue wiring) against the oracle without a GPU.  It is never importable from the product
package and is not a fallback: tests install it explicitly with `install(monkeypatch)`.
"""
import math

import torch
import torch.nn.functional as F

from music_generation_b200 import ops as real_ops
from music_generation_b200._lib import ACT_GATE, ACT_GELU, ACT_NONE, ACT_RELU, ACT_SIGMOID, ACT_TANH


def _mask(lens, B, T):
    if lens is None:
        return None
    return (torch.arange(T).unsqueeze(0) < lens.to(torch.int64).unsqueeze(1)).to(torch.float32).unsqueeze(-1)


def _store(dst, val, split16=False):
    if dst is None:
        return
    if split16:      # fp16 [hi | lo | hi]
        hi = val.half()
        lo = (val - hi.float()).half()
        dst.copy_(torch.cat([hi, lo, hi], -1))
    else:
        dst.copy_(val.to(dst.dtype))


def conv1d(pc, x, *, act=ACT_NONE, gbias=None, alpha=1.0, scale=1.0, res1=None, res2=None, lens=None,
           out32=None, out16=None, post_lrelu=None, w_rows=None, n_valid=None, split16=False, res16=None,
           out16r=None, res16_unlrelu=None):
    B, T, cin = x.shape
    assert cin == pc.cin
    xf = x.float()
    if pc.f32:
        w = pc.w.float()[:, :, :]                    # [k, Cin, Cout_pad]
        wk = w.permute(2, 1, 0)                      # [Cout_pad, Cin, k]
    else:
        assert x.dtype == pc.w.dtype
        wk = pc.w.float()[:, :, :cin].permute(1, 2, 0)   # [Cout_pad, Cin, k]
    bias = pc.bias
    if w_rows is not None:
        r0, nr = w_rows
        wk = wk[r0:r0 + nr]
        bias = None if bias is None else bias[r0:r0 + nr]
    k, d, pl = pc.ntaps, pc.dil, pc.pad_left
    pr = (k - 1) * d - pl
    xp = F.pad(xf.transpose(1, 2), [pl, max(pr, 0)])
    y = F.conv1d(xp, wk, None, dilation=d)[:, :, :T].transpose(1, 2)      # [B,T,Cout_pad]
    if bias is not None:
        y = y + bias.float()
    if gbias is not None:
        y = y + gbias.float().unsqueeze(1)
    gate = pc.gate and w_rows is None
    if gate:
        bn = pc.block_n
        yb = y.reshape(B, T, -1, bn)
        y = (torch.tanh(yb[..., :bn // 2]) * torch.sigmoid(yb[..., bn // 2:])).reshape(B, T, -1)
    elif act == ACT_RELU:
        y = torch.relu(y)
    elif act == ACT_GELU:
        y = F.gelu(y)
    elif act == ACT_SIGMOID:
        y = torch.sigmoid(y)
    elif act == ACT_TANH:
        y = torch.tanh(y)
    nv = n_valid if n_valid is not None else (pc.n_out if w_rows is None else w_rows[1])
    y = y[:, :, :nv] * (alpha * pc.out_scale)
    if res1 is not None:
        y = y + res1.float()[:, :, :nv]
    if res2 is not None:
        y = y + res2.float()[:, :, :nv]
    if res16 is not None:
        r = res16.float()[:, :, :nv]
        if res16_unlrelu is not None and res16_unlrelu > 1:
            r = torch.minimum(r, r * res16_unlrelu)
        y = y + r
    y = y * scale
    m = _mask(lens, B, T)
    if m is not None:
        y = y * m
    if out32 is not None:
        out32[:, :, :nv] = y
    if out16r is not None:
        out16r[:, :, :nv] = y.to(out16r.dtype)
    if out16 is not None:
        o = F.leaky_relu(y, post_lrelu) if post_lrelu is not None else y
        if split16:
            _store(out16, o, True)
        else:
            out16[:, :, :nv] = o.to(out16.dtype)


def attention(qkv, out32, out16, rel_k, rel_v, lens, heads=2, window=4):
    B, T, c3 = qkv.shape
    c = c3 // 3
    dk = c // heads
    q, k, v = (qkv[:, :, i * c:(i + 1) * c].reshape(B, T, heads, dk).permute(0, 2, 1, 3) for i in range(3))
    qs = q / math.sqrt(dk)
    scores = qs @ k.transpose(-1, -2)
    idx = torch.arange(T)
    rel = idx.unsqueeze(0) - idx.unsqueeze(1)                       # j - i
    band = rel.abs() <= window
    rk = rel_k[(rel.clamp(-window, window) + window)]               # [T,T,dk]
    bias = torch.einsum("bhid,ijd->bhij", qs, rk) * band
    scores = scores + bias
    ln = lens.to(torch.int64) if lens is not None else torch.full((B,), T)
    key_ok = (idx.unsqueeze(0) < ln.unsqueeze(1))                   # [B,T]
    scores = scores.masked_fill(~key_ok[:, None, None, :], -1e4)
    p = torch.softmax(scores, -1)
    out = p @ v
    rv = rel_v[(rel.clamp(-window, window) + window)]
    out = out + torch.einsum("bhij,ijd->bhid", p * band, rv)
    out = out.permute(0, 2, 1, 3).reshape(B, T, c) * key_ok.unsqueeze(-1)
    _store(out32, out)
    _store(out16, out)


def attention_tc(qkv, out32, out16, rel_k, rel_v, lens, *, dtype, split=False, out16_split=False, ws=None, heads=2,
                 window=4):
    """Semantics of ops.attention_tc: 16-bit rounding of q,k,v in plain mode; split mode keeps ~fp32."""
    q = qkv if split else qkv.to(dtype).float()
    B, T, c3 = qkv.shape
    tmp = torch.empty(B, T, c3 // 3)
    attention(q, tmp, None, rel_k, rel_v, lens, heads, window)
    _store(out32, tmp)
    if out16 is not None:
        if out16_split:
            hi = tmp.half()
            lo = (tmp - hi.float()).half()
            out16.copy_(torch.cat([hi, lo, hi], -1))
        else:
            _store(out16, tmp)


def layernorm(x, gamma, beta, lens, out32=None, out16=None, res=None, eps=1e-5, split16=False):
    B, T, c = x.shape
    y = x if res is None else x + res
    y = F.layer_norm(y, (c,), gamma, beta, eps)
    m = _mask(lens, B, T)
    if m is not None:
        y = y * m
    _store(out32, y)
    _store(out16, y, split16)


def text_embed(phon, pitch, tone, dur, e_ph, e_pitch, e_tone, w_dur, b_dur, x_lengths, lens32, out, status=None):
    B, P = phon.shape
    live = _mask(x_lengths, B, P).squeeze(-1) > 0
    bad = live & ((phon < 0) | (phon >= e_ph.shape[0]) | (pitch < 0) | (pitch >= e_pitch.shape[0]) | (tone < 0) |
                  (tone >= e_tone.shape[0]))
    if status is not None and bool(bad.any()):
        status[1] |= 1
    phon, pitch, tone = (t.clamp(0, n - 1) for t, n in ((phon, e_ph.shape[0]), (pitch, e_pitch.shape[0]), (tone, e_tone.shape[0])))
    v = (e_ph[phon] + (dur.unsqueeze(-1) * w_dur + b_dur) + e_pitch[pitch] + e_tone[tone]) * math.sqrt(out.shape[-1])
    lens32.copy_(x_lengths.to(torch.int32))
    out.copy_(v * _mask(x_lengths, B, P))


def duration_quantize(pred, note_dur, out):
    out.copy_(torch.floor(pred.reshape(note_dur.shape) * note_dur / 0.0116).long())


def lr_scan(dur, ends, mel64, mel32, status=None):
    e = dur.clamp_min(0).cumsum(1)
    if status is not None:
        status[0] = torch.maximum(status[0], e[:, -1].max().to(torch.int32))
    ends.copy_(e.to(torch.int32))
    mel64.copy_(e[:, -1])
    mel32.copy_(e[:, -1].to(torch.int32))


def lr_gather(h, ends, mel32, out, index_out=None):
    B, P, c = h.shape
    T = out.shape[1]
    t = torch.arange(T).unsqueeze(0).expand(B, T).contiguous()
    idx = torch.searchsorted(ends.to(torch.int64), t, right=True)
    valid = t < mel32.to(torch.int64).unsqueeze(1)
    idx = torch.where(valid, idx, torch.full_like(idx, -1))
    g = torch.gather(h, 1, idx.clamp_min(0).unsqueeze(-1).expand(B, T, c)) * valid.unsqueeze(-1)
    out.copy_(g)
    if index_out is not None:
        index_out.copy_(idx.to(torch.int32))


def add_pos_mask(x, pos, lens, out32=None, out16=None, split16=False):
    B, T, c = x.shape
    y = (x + pos[:T].unsqueeze(0)) * _mask(lens, B, T)
    _store(out32, y)
    _store(out16, y, split16)


def f0_embed(log_f0, coarse_in, table, lens, coarse_out, out32=None, out16=None):
    o = out32 if out32 is not None else out16
    B, T, c = o.shape
    if coarse_in is None:
        f0 = torch.exp(log_f0) - 1.0
        mel = 1127 * (1 + f0 / 700).log()
        mn, mx = 1127 * math.log(1 + 50 / 700), 1127 * math.log(1 + 1100 / 700)
        mel = torch.where(mel > 0, (mel - mn) * 254 / (mx - mn) + 1, mel)
        mel = mel.clamp(1, 255)
        coarse_in = (mel + 0.5).long()
    coarse_out.copy_(coarse_in)
    y = table[coarse_in] * _mask(lens, B, T)
    _store(out32, y)
    _store(out16, y)


def sample_prior(stats, noise_bct, noise_scale, zp32=None, zp16=None, lens=None):
    c = stats.shape[-1] // 2
    y = stats[:, :, :c] + noise_bct.transpose(1, 2) * torch.exp(stats[:, :, c:]) * noise_scale
    if lens is not None:
        y = y * _mask(lens, y.shape[0], y.shape[1])
    _store(zp32, y)
    _store(zp16, y)


def bct_to_btc(x_bct, out32=None, out16=None, lens=None):
    y = x_bct.transpose(1, 2)
    if lens is not None:
        y = y * _mask(lens, y.shape[0], y.shape[1])
    _store(out32, y)
    _store(out16, y)


def resblock_pair(c1, c2, x, out, *, mrf=None, out_lrelu=True, post_slope=0.1, scale=1.0, res_unlrelu=10.0):
    """b200svs_resblock_pair = the two conv launches it fuses (documented as bit-identical to them)."""
    f = lambda pc, t: t.view(t.shape[0], t.shape[1] // pc.fold, t.shape[2] * pc.fold) if pc.fold > 1 else t
    t16 = torch.empty_like(x)
    conv1d(c1, f(c1, x), out16=f(c1, t16), post_lrelu=0.1)
    kw = dict(res16=f(c2, x), res16_unlrelu=res_unlrelu, scale=scale)
    if mrf is not None:
        kw["res2"] = f(c2, mrf.clone())
    tmp = torch.empty_like(x)
    if out_lrelu:
        kw.update(out16=f(c2, tmp), post_lrelu=post_slope)
    else:
        kw.update(out16r=f(c2, tmp))
    conv1d(c2, f(c2, t16), **kw)
    out.copy_(tmp)


def conv_post_tanh(x16, w, out):
    B, L, c = x16.shape
    k = w.shape[0]
    y = F.conv1d(x16.float().transpose(1, 2), w.t().unsqueeze(0), padding=k // 2)
    out.copy_(torch.tanh(y[:, 0]))


def install(monkeypatch):
    """Route music_generation_b200.ops.* to the emulation above (CPU tests only)."""
    import music_generation_b200.synthesizer as syn
    for name in ("conv1d", "attention", "attention_tc", "layernorm", "text_embed", "duration_quantize", "lr_scan", "lr_gather",
                 "add_pos_mask", "f0_embed", "sample_prior", "conv_post_tanh", "bct_to_btc", "resblock_pair"):
        monkeypatch.setattr(real_ops, name, globals()[name])
    monkeypatch.setattr(syn, "_require_cuda", lambda t: None)
    monkeypatch.setattr(syn, "_STAGE_API", False)          # the emulation covers the per-op entry points
    import contextlib
    monkeypatch.setattr(syn, "_device_guard", lambda dev: contextlib.nullcontext())
