"""Thin torch-tensor wrappers over the C ABI (include/b200svs.h).

Tensors are channels-last [B, T, C] views (last dim contiguous, rows `ld` elements apart, batch
stride = T*ld).  PyTorch is used for device memory and streams only; every numeric op below is a
hand-written sm_100a kernel in libb200svs.so.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import (ACT_GATE, ACT_GELU, ACT_NONE, ACT_RELU, ACT_SIGMOID, ACT_TANH, DT_BF16, DT_F16,
                   ConvArgs, Epilogue)

__all__ = ["conv1d", "attention", "layernorm", "ACT_NONE", "ACT_RELU", "ACT_GELU", "ACT_SIGMOID",
           "ACT_TANH", "ACT_GATE"]


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def out16_code(t, split16):
    """dtype code of a 16-bit OUTPUT tensor; split16: fp16 [hi | lo | hi] with 3x the logical columns."""
    if t is None:
        return 0
    if split16:
        if t.dtype != torch.float16:
            raise TypeError("split 16-bit outputs are fp16")
        return _lib.DT_F16_SPLIT
    return dt_code(t.dtype)


def dt_code(dtype):
    if dtype == torch.float16:
        return DT_F16
    if dtype == torch.bfloat16:
        return DT_BF16
    raise TypeError(f"16-bit operand dtype expected, got {dtype}")


def _btc(t, name):
    """Checks a [B,T,C] channels-last view and returns its row stride."""
    if t is None:
        return 0
    if t.dim() != 3 or t.stride(2) != 1 or (t.shape[0] > 1 and t.stride(0) != t.shape[1] * t.stride(1)):
        raise ValueError(f"{name}: expected a [B,T,C] view with unit channel stride and dense batch stride, "
                         f"got shape {tuple(t.shape)} strides {t.stride()}")
    if not t.is_cuda:
        raise ValueError(f"{name}: CUDA tensor required (there is no CPU path)")
    return t.stride(1)


class PackedConv:
    """Device-resident packed weights of one convolution (see packer.py)."""
    __slots__ = ("w", "bias", "ntaps", "dil", "pad_left", "cin", "cin_pad", "cout", "cout_pad", "block_n",
                 "gate", "f32", "n_out", "alg_macs_per_row", "tag", "out_scale", "fold")

    def __init__(self, w, bias, ntaps, dil, pad_left, cin, cin_pad, cout, cout_pad, block_n, gate, f32):
        self.w, self.bias, self.ntaps, self.dil, self.pad_left = w, bias, ntaps, dil, pad_left
        self.cin, self.cin_pad, self.cout, self.cout_pad = cin, cin_pad, cout, cout_pad
        self.block_n, self.gate, self.f32 = block_n, gate, f32
        self.n_out = cout // 2 if gate else cout      # columns the epilogue stores
        # algorithmic MACs per input row (the reference conv's own count; the polyphase ConvTranspose
        # form executes 1.5x of that, packer.pack_conv_transpose overrides it) - used by bench.py only
        self.alg_macs_per_row = cin * cout * ntaps
        self.tag = ""
        self.fold = 1               # 2: time-pair folded (packer.fold_time_pairs); tensors are passed as [B, T/2, 2C] views
        self.out_scale = 1.0        # split-precision convs: power-of-two factor folded into the epilogue's alpha

    def to(self, device):
        self.w = self.w.to(device)
        self.bias = None if self.bias is None else self.bias.to(device)
        return self


def conv1d(pc, x, *, act=ACT_NONE, gbias=None, alpha=1.0, scale=1.0, res1=None, res2=None, lens=None,
           out32=None, out16=None, post_lrelu=None, w_rows=None, n_valid=None, split16=False, res16=None,
           out16r=None, res16_unlrelu=None):
    """out = epilogue(conv(x, pc)).  See b200svs_epilogue in include/b200svs.h for the formula.
    `w_rows=(row0, nrows)` restricts the conv to a block of packed output channels (WN res/skip split)."""
    B, T, cin = x.shape
    x_ld = _btc(x, "x")
    if cin != pc.cin:
        raise ValueError(f"conv1d: input has {cin} channels, weights expect {pc.cin}")
    a = ConvArgs()
    a.x = x.data_ptr()
    a.B, a.T, a.Cin, a.x_ld = B, T, cin, x_ld
    a.ntaps, a.Cin_pad, a.dil, a.pad_left = pc.ntaps, pc.cin_pad, pc.dil, pc.pad_left
    bias = pc.bias
    cout_pad, n_out, w = pc.cout_pad, pc.n_out, pc.w
    if w_rows is not None:
        if pc.f32 or pc.gate:
            raise ValueError("w_rows is only supported for plain tensor-core convs")
        r0, nr = w_rows
        # packed [ntaps, Cout_pad, Cin_pad]: a row block is a strided view -> only valid for 1 tap
        if pc.ntaps != 1:
            raise ValueError("w_rows needs a 1-tap conv")
        w = pc.w[0, r0:r0 + nr]
        bias = None if bias is None else bias[r0:r0 + nr]
        cout_pad, n_out = nr, nr
    a.w = w.data_ptr()
    a.Cout_pad = cout_pad
    a.block_n = pc.block_n if w_rows is None else min(pc.block_n, cout_pad)
    e = a.ep
    e.bias = None if bias is None else bias.data_ptr()
    if gbias is not None:
        e.gbias, e.gbias_ld = gbias.data_ptr(), gbias.stride(0)
    e.act = ACT_GATE if pc.gate else act
    e.alpha, e.scale = alpha * pc.out_scale, scale
    if res1 is not None:
        e.res1, e.res1_ld = res1.data_ptr(), _btc(res1, "res1")
    if res2 is not None:
        e.res2, e.res2_ld = res2.data_ptr(), _btc(res2, "res2")
        e.res2_is16 = 0 if res2.dtype == torch.float32 else 1      # 16-bit MRF stream (operand dtype)
    if lens is not None:
        e.lens = lens.data_ptr()
    e.lens_div = 1
    if out32 is not None:
        e.out32, e.out32_ld = out32.data_ptr(), _btc(out32, "out32")
    if out16 is not None:
        e.out16, e.out16_ld, e.out16_dtype = out16.data_ptr(), _btc(out16, "out16"), out16_code(out16, split16)
    if post_lrelu is not None:
        e.post_lrelu, e.post_slope = 1, float(post_lrelu)
    if res16 is not None:      # 16-bit residual input (same dtype as the operands), added like res1
        e.res16, e.res16_ld = res16.data_ptr(), _btc(res16, "res16")
        if res16_unlrelu is not None:      # res16 holds leaky_relu(x, 1 / res16_unlrelu): the epilogue inverts it
            e.res16_unlrelu = float(res16_unlrelu)
    if out16r is not None:     # 16-bit copy of the result WITHOUT the post-activation (may alias res16)
        e.out16r, e.out16r_ld = out16r.data_ptr(), _btc(out16r, "out16r")
    e.n_valid = n_out if n_valid is None else n_valid
    if _lib.PROFILE is not None:
        frac = 1.0 if w_rows is None else w_rows[1] / pc.cout_pad
        n = float(B) * T * e.n_valid
        xb = x.element_size()
        nbytes = (float(B) * T * cin * xb + w.numel() * w.element_size()
                  + n * (4 * (res1 is not None) + (0 if res2 is None else res2.element_size()) + 2 * (res16 is not None) + 4 * (out32 is not None)
                         + 2 * (out16r is not None) + (6 if split16 else 2) * (out16 is not None)))
        _lib.NEXT_META = (pc.tag, 2.0 * B * T * pc.alg_macs_per_row * frac, nbytes)
    if pc.f32:
        if x.dtype != torch.float32:
            raise TypeError("fp32 conv needs fp32 activations")
        _lib.call("b200svs_conv1d_f32", C.byref(a), _stream())
    else:
        a.x_dtype = dt_code(x.dtype)
        if pc.w.dtype != x.dtype:
            raise TypeError(f"operand dtype mismatch: x {x.dtype}, w {pc.w.dtype}")
        _lib.call("b200svs_conv1d_tc", C.byref(a), _stream())


def attention(qkv, out32, out16, rel_k, rel_v, lens, heads=2, window=4):
    B, T, c3 = qkv.shape
    c = c3 // 3
    _lib.call("b200svs_attention_f32", _ptr(qkv), _btc(qkv, "qkv"), _ptr(out32), _btc(out32, "out32"),
              _ptr(out16), _btc(out16, "out16"), 0 if out16 is None else dt_code(out16.dtype),
              _ptr(rel_k), _ptr(rel_v), _ptr(lens), B, T, c, heads, window, _stream())


def attention_tc(qkv, out32, out16, rel_k, rel_v, lens, *, dtype, split=False, out16_split=False, ws=None, heads=2,
                 window=4):
    """Tensor-core attention: qkv fp32 [B,T,3C] -> out32 fp32 [B,T,C] and/or out16.  `split`: fp16 (hi, lo)
    operands, three MMAs per product.  `ws`: optional dict reused across layers for the 16-bit operand copies."""
    B, T, c3 = qkv.shape
    c = c3 // 3
    Tp = (T + 7) // 8 * 8
    odt = torch.float16 if split else dtype
    ws = ws if ws is not None else {}
    key = (B, T, c, split, odt)
    if ws.get("key") != key:
        ws["key"] = key
        ws["qk_hi"] = torch.empty(B, T, 2 * c, device=qkv.device, dtype=odt)
        ws["vt_hi"] = torch.empty(B, c, Tp, device=qkv.device, dtype=odt)
        ws["qk_lo"] = torch.empty_like(ws["qk_hi"]) if split else None
        ws["vt_lo"] = torch.empty_like(ws["vt_hi"]) if split else None
    code = dt_code(odt)
    _lib.call("b200svs_attn_prep", _ptr(qkv), _btc(qkv, "qkv"), _ptr(ws["qk_hi"]), _ptr(ws["qk_lo"]), _ptr(ws["vt_hi"]),
              _ptr(ws["vt_lo"]), code, B, T, Tp, c, _stream())
    _lib.call("b200svs_attention_tc", _ptr(ws["qk_hi"]), _ptr(ws["qk_lo"]), _ptr(ws["vt_hi"]), _ptr(ws["vt_lo"]), code, Tp,
              _ptr(out32), _btc(out32, "out32"), _ptr(out16), _btc(out16, "out16"),
              0 if out16 is None else dt_code(out16.dtype), 1 if out16_split else 0, _ptr(rel_k), _ptr(rel_v),
              _ptr(lens), B, T, c, heads, window, _stream())


def layernorm(x, gamma, beta, lens, out32=None, out16=None, res=None, eps=1e-5, split16=False):
    B, T, c = x.shape
    _lib.call("b200svs_layernorm", _ptr(x), _btc(x, "x"), _ptr(res), _btc(res, "res"), _ptr(gamma), _ptr(beta),
              eps, _ptr(lens), _ptr(out32), _btc(out32, "out32"), _ptr(out16), _btc(out16, "out16"),
              out16_code(out16, split16), B, T, c, _stream())


def text_embed(phon, pitch, tone, dur, e_ph, e_pitch, e_tone, w_dur, b_dur, x_lengths, lens32, out, status=None):
    B, P = phon.shape
    _lib.call("b200svs_text_embed", _ptr(phon), _ptr(pitch), _ptr(tone), _ptr(dur), _ptr(e_ph), _ptr(e_pitch),
              _ptr(e_tone), _ptr(w_dur), _ptr(b_dur), _ptr(x_lengths), _ptr(lens32), _ptr(out), B, P,
              out.shape[-1], e_ph.shape[0], e_pitch.shape[0], e_tone.shape[0], _ptr(status), _stream())


def duration_quantize(pred, note_dur, out):
    _lib.call("b200svs_duration_quantize", _ptr(pred), _ptr(note_dur), _ptr(out), pred.numel(), _stream())


def lr_scan(dur, ends, mel64, mel32, status=None):
    B, P = dur.shape
    _lib.call("b200svs_lr_scan", _ptr(dur), B, P, _ptr(ends), _ptr(mel64), _ptr(mel32), _ptr(status), _stream())


def lr_gather(h, ends, mel32, out, index_out=None):
    B, P, c = h.shape
    T = out.shape[1]
    _lib.call("b200svs_lr_gather", _ptr(h), _btc(h, "h"), _ptr(ends), _ptr(mel32), B, P, T, c, _ptr(out),
              _btc(out, "out"), _ptr(index_out), _stream())


def add_pos_mask(x, pos, lens, out32=None, out16=None, split16=False):
    B, T, c = x.shape
    _lib.call("b200svs_add_pos_mask", _ptr(x), _btc(x, "x"), _ptr(pos), _ptr(lens), _ptr(out32),
              _btc(out32, "out32"), _ptr(out16), _btc(out16, "out16"),
              out16_code(out16, split16), B, T, c, _stream())


def f0_embed(log_f0, coarse_in, table, lens, coarse_out, out32=None, out16=None):
    o = out32 if out32 is not None else out16
    B, T, c = o.shape
    _lib.call("b200svs_f0_embed", _ptr(log_f0), _ptr(coarse_in), _ptr(table), _ptr(lens), _ptr(coarse_out),
              _ptr(out32), _btc(out32, "out32"), _ptr(out16), _btc(out16, "out16"),
              0 if out16 is None else dt_code(out16.dtype), B, T, c, table.shape[0], _stream())


def sample_prior(stats, noise_bct, noise_scale, zp32=None, zp16=None, lens=None):
    B, T, c2 = stats.shape
    _lib.call("b200svs_sample_prior", _ptr(stats), _btc(stats, "stats"), _ptr(noise_bct), float(noise_scale),
              _ptr(zp32), _btc(zp32, "zp32"), _ptr(zp16), _btc(zp16, "zp16"),
              0 if zp16 is None else dt_code(zp16.dtype), B, T, c2 // 2, _ptr(lens), _stream())


def conv_post_tanh(x16, w, out):
    B, L, c = x16.shape
    _lib.call("b200svs_conv_post_tanh", _ptr(x16), dt_code(x16.dtype), _btc(x16, "x"), _ptr(w), B, L, c,
              w.shape[0], _ptr(out), _stream())


def bct_to_btc(x_bct, out32=None, out16=None, lens=None):
    B, c, T = x_bct.shape
    _lib.call("b200svs_bct_to_btc", _ptr(x_bct), B, c, T, _ptr(out32), _btc(out32, "out32"), _ptr(out16),
              _btc(out16, "out16"), 0 if out16 is None else dt_code(out16.dtype), _ptr(lens), _stream())


def btc_to_bct(x_btc, out_bct):
    B, T, c = x_btc.shape
    _lib.call("b200svs_btc_to_bct", _ptr(x_btc), _btc(x_btc, "x"), B, c, T, _ptr(out_bct), _stream())


def song_assemble(o, mel_len, offsets, song, *, order=None, hop=256, lead_in=4410, legacy_scalar=False, scales_out=None):
    """Peak-normalise (twice, as the reference's driver does), truncate to int16 and concatenate the valid samples of
    every line of a ragged batch (b200svs_song_assemble).  o: [B,1,S] fp32; mel_len: [B] int64; offsets: [n] int64
    start sample of each line in the song (without the lead-in); order: [n] int32 utterance index per line."""
    n = offsets.numel()
    ws = torch.empty(n, device=o.device, dtype=torch.int32)
    _lib.call("b200svs_song_assemble", _ptr(o), o.stride(0), _ptr(mel_len), _ptr(order), _ptr(offsets), n, hop, lead_in,
              1 if legacy_scalar else 0, _ptr(ws), _ptr(song), _ptr(scales_out), _stream())


# ---------------------------------------------------------------------------------------------- stage-level API
def conv_w(pc):
    """ctypes mirror (b200svs_conv_w) of a PackedConv; the PackedConv must stay alive as long as the table is used."""
    c = _lib.ConvW()
    c.w, c.bias = pc.w.data_ptr(), (0 if pc.bias is None else pc.bias.data_ptr())
    c.ntaps, c.dil, c.pad_left, c.cin, c.cin_pad = pc.ntaps, pc.dil, pc.pad_left, pc.cin, pc.cin_pad
    c.cout, c.cout_pad, c.block_n = pc.cout, pc.cout_pad, pc.block_n
    c.gate, c.f32, c.fold, c.n_out, c.out_scale = int(pc.gate), int(pc.f32), pc.fold, pc.n_out, pc.out_scale
    return c


_WS = {}


def workspace(nbytes, device):
    """Caller-owned workspace of the stage-level entry points: one cached uint8 tensor per (device, stream), grown on
    demand (stream-ordered reuse: consecutive stage calls on a stream may share it)."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), torch.cuda.current_stream(device).cuda_stream)
    t = _WS.get(key)
    if t is None or t.numel() < nbytes:
        t = _WS[key] = torch.empty(max(int(nbytes), 1 << 20), device=device, dtype=torch.uint8)
    return t


def generator_fwd(table, x16, gbias, out):
    B, T, _ = x16.shape
    n = _lib.lib().b200svs_generator_workspace_bytes(C.byref(table), B, T)
    if n == 0:
        _lib.check(-1, "generator_workspace_bytes")
    ws = workspace(n, x16.device)
    _lib.call("b200svs_generator_fwd", C.byref(table), _ptr(x16), _btc(x16, "x16"), B, T, _ptr(gbias),
              0 if gbias is None else gbias.stride(0), _ptr(out), _ptr(ws), ws.numel(), _stream())


def wn_stack_fwd(table, x32, x16, lens, gcond, out16):
    B, T, _ = x32.shape
    n = _lib.lib().b200svs_wn_workspace_bytes(C.byref(table), B, T)
    ws = workspace(n, x32.device)
    _lib.call("b200svs_wn_stack_fwd", C.byref(table), _ptr(x32), _ptr(x16), _ptr(lens), B, T, _ptr(gcond),
              0 if gcond is None else gcond.stride(0), _ptr(out16), _ptr(ws), ws.numel(), _stream())


def encoder_fwd(tables, n_layers, x32, x16, lens):
    B, T, _ = x32.shape
    n = _lib.lib().b200svs_encoder_workspace_bytes(tables, B, T)
    ws = workspace(n, x32.device)
    _lib.call("b200svs_encoder_fwd", tables, n_layers, _ptr(x32), _ptr(x16), _ptr(lens), B, T, _ptr(ws), ws.numel(), _stream())


def pair_fusable(c1, c2, ch):
    """Shapes b200svs_resblock_pair takes: the 32-channel stage (conv2 time-pair folded) and the 64-channel stage while both
    weight sets stay resident in shared memory (kernel sizes up to 7)."""
    if c2 is None or c1.f32 or c2.f32:
        return False
    if ch == 32:
        return c2.fold == 2
    return ch == 64 and c1.fold == 1 and c2.fold == 1 and c1.ntaps + c2.ntaps <= 14


def resblock_pair(c1, c2, x, out, *, mrf=None, out_lrelu=True, post_slope=0.1, scale=1.0, res_unlrelu=10.0):
    """Fused ResBlock1 pair of the 32-channel Generator stage (b200svs_resblock_pair): out = x + c2(lrelu(c1(lrelu(x))))
    [+ mrf], x recovered from the activated 16-bit tensor.  x / out / mrf: dense [B, L, 32] fp16."""
    B, L, ch = x.shape
    if ch not in (32, 64) or not x.is_contiguous() or not out.is_contiguous() or x.dtype != torch.float16:
        raise ValueError("resblock_pair: dense [B, L, 32 | 64] fp16 tensors expected")
    if _lib.PROFILE is not None:
        n = float(B) * L * ch
        _lib.NEXT_META = (c2.tag, 2.0 * B * L * (c1.alg_macs_per_row / c1.fold + c2.alg_macs_per_row / c2.fold),
                          2.0 * n * (2 + (mrf is not None)) + c1.w.numel() * 2 + c2.w.numel() * 2)
    _lib.call("b200svs_resblock_pair", C.byref(conv_w(c1)), C.byref(conv_w(c2)), _ptr(x), B, L, _ptr(mrf), _ptr(out),
              1 if out_lrelu else 0, float(post_slope), float(scale), float(res_unlrelu), _stream())


def conv1d_ws(cw, x, out, *, res=None, mrf=None, out_lrelu=True, post_slope=0.1, scale=1.0, res_unlrelu=10.0):
    """Weight-stationary 64 -> 64 dilation-1 conv (b200svs_conv1d_ws; packer.pack_conv_ws).  res is None: out =
    [lrelu](conv(x) + b); else out = [lrelu]((conv(x) + b + unlrelu(res) [+ mrf]) * scale).  Dense [B, T, 64] fp16, T even."""
    B, T, ch = x.shape
    if ch != 64 or x.dtype != torch.float16 or not x.is_contiguous() or not out.is_contiguous():
        raise ValueError("conv1d_ws: dense [B, T, 64] fp16 tensors expected")
    if _lib.PROFILE is not None:
        n = float(B) * T * ch
        _lib.NEXT_META = (cw.tag, 2.0 * B * T * 64 * 64 * cw.ntaps, 2.0 * n * (2 + (res is not None) + (mrf is not None)))
    _lib.call("b200svs_conv1d_ws", _ptr(cw.img), _ptr(cw.bias), cw.ntaps, _ptr(x), B, T, _ptr(res), float(res_unlrelu),
              _ptr(mrf), _ptr(out), 1 if out_lrelu else 0, float(post_slope), float(scale), _stream())
    return out
