"""Load-time weight packing: fold weight-norm once, lay weights out for the kernels.

The reference never calls remove_weight_norm at inference, so every forward recomputes
w = g * v / ||v|| for ~100 convs (VISinger/models_ae.py:311, modules.py:135,145,
191-206).  Here that happens once per load_state_dict.

Layouts (see include/b200svs.h):
  tensor-core convs : [ntaps][Cout_pad][Cin_pad] 16-bit, Cin contiguous (K-major B operand for TMA)
  fp32 SIMT convs   : [ntaps][Cin][Cout_pad] fp32, Cout contiguous
ConvTranspose1d(stride s, kernel 2s, padding s/2) becomes a 3-tap conv over input time with s*Cout
output columns (polyphase).  WN in_layers are packed so each N tile holds a tanh
half and the matching sigmoid half (GATE epilogue).  The flow's channel Flips are folded into the
pre/post 1x1 weights.
"""
import torch

from .ops import PackedConv


def fold_weight_norm(sd, prefix):
    """Old-style torch.nn.utils.weight_norm, dim=0: w = g * v / ||v||_2 over all dims but 0."""
    if prefix + ".weight" in sd:
        return sd[prefix + ".weight"].detach().float()
    v = sd[prefix + ".weight_v"].detach().float()
    g = sd[prefix + ".weight_g"].detach().float()
    norm = v.reshape(v.shape[0], -1).norm(dim=1).reshape(-1, *([1] * (v.dim() - 1)))
    return v * (g / norm)


def _round_up(x, m):
    return (x + m - 1) // m * m


def choose_block_n(cout):
    """N tile of the tensor-core conv: a multiple of 32 (epilogue chunk), <= 256.
    256 when it divides cout: an N=256 MMA reads 12 KB of shared memory per 128 cycles and runs at ~94 % of the
    tensor peak, an N=128 one needs 8 KB per 64 cycles - all of the SM's shared-memory bandwidth - and measures
    79-126 cycles next to the TMA / epilogue traffic (tools/trace_conv.py bn).  Otherwise 128 is preferred for
    wide layers: TMEM then holds 2 x 2 accumulators and the kernel runs 256-row tiles in which every weight
    tile feeds two MMAs (conv1d_tc.cu)."""
    if cout <= 128:
        return _round_up(cout, 32)
    if cout % 256 == 0:
        return 256
    best = None
    for bn in (128, 192, 160, 224, 256):
        pad = _round_up(cout, bn)
        if best is None or pad < best[0]:
            best = (pad, bn)
    return best[1]


def pack_conv(weight, bias, *, dil=1, pad_left=None, dtype=torch.float16, f32=False, gate=False,
              block_n=None, split=False):
    """weight: [Cout, Cin, k] effective (weight-norm folded) Conv1d weight.

    split=True: split-precision tensor-core conv (F0 predictor).  The activation operand is the fp16
    K-concatenation [x_hi | x_lo | x_hi] (3*Cin channels) and the weights are packed [w_hi | w_hi | w_lo],
    so one conv computes x_hi.w_hi + x_lo.w_hi + x_hi.w_lo ~ x.w to 2^-22 with fp32 accumulation."""
    weight = weight.detach().float()
    if split:
        if f32 or gate:
            raise ValueError("split packing is for plain tensor-core convs")
        # The tensor cores flush fp16 subnormals, and w_lo ~ w * 2^-12 is subnormal for |w| < 0.25: scale the
        # weights (and the bias) by a power of two so that max|w| sits just below 2^15; the epilogue's alpha
        # undoes it exactly (act in {NONE, RELU} commutes with a positive scale).
        import math
        out_scale = 2.0 ** -math.floor(math.log2(32768.0 / max(float(weight.abs().max()), 1e-30)))
        weight = weight / out_scale
        if bias is not None:
            bias = bias.detach().float() / out_scale
        w_hi = weight.half().float()
        w_lo = (weight - w_hi).half().float()
        weight = torch.cat([w_hi, w_hi, w_lo], 1)
        dtype = torch.float16
    cout, cin, k = weight.shape
    if pad_left is None:
        pad_left = (k * dil - dil) // 2          # commons.get_padding (VISinger/commons.py:14-15)
    b = None if bias is None else bias.detach().float()
    if f32:
        cout_pad = _round_up(cout, 64)
        w = torch.zeros(k, cin, cout_pad)
        w[:, :, :cout] = weight.permute(2, 1, 0)
        bp = None
        if b is not None:
            bp = torch.zeros(cout_pad)
            bp[:cout] = b
        return PackedConv(w.contiguous(), bp, k, dil, pad_left, cin, cin, cout, cout_pad, 64, False, True)
    cin_pad = _round_up(cin, 64)
    if gate:
        half = cout // 2
        bn = block_n or gate_block_n(half)
        hb = bn // 2
        if half % hb:
            raise ValueError(f"gate packing needs hidden {half} divisible by {hb}")
        order = []
        for n in range(half // hb):
            order += list(range(n * hb, (n + 1) * hb)) + list(range(half + n * hb, half + (n + 1) * hb))
        idx = torch.tensor(order)
        weight = weight[idx]
        if b is not None:
            b = b[idx]
        cout_pad = cout
    else:
        bn = block_n or choose_block_n(cout)
        cout_pad = _round_up(cout, bn)
    w = torch.zeros(k, cout_pad, cin_pad)
    w[:, :cout, :cin] = weight.permute(2, 0, 1)
    bp = None
    if b is not None:
        bp = torch.zeros(cout_pad)
        bp[:cout] = b
    if dtype == torch.float16 and float(w.abs().max()) > 65504.0:
        # an fp16 weight would become inf and every zero-padded input row NaN (0 * inf): refuse instead
        raise ValueError(f"conv weight magnitude {float(w.abs().max()):.3g} exceeds the fp16 range (65504): "
                         "build the model with operand_dtype='bf16'")
    pc = PackedConv(w.to(dtype).contiguous(), bp, k, dil, pad_left, cin, cin_pad, cout, cout_pad, bn, gate, False)
    if split:
        pc.alg_macs_per_row = (cin // 3) * cout * k      # algorithmic (fp32-equivalent) MACs, not the 3x executed
        pc.out_scale = out_scale
    return pc


def fold_time_pairs(weight, bias, pad_left=None):
    """Time-pair folding of a dilation-1 conv: a channels-last [B, T, C] tensor with even T IS a [B, T/2, 2C] tensor
    whose row u holds time steps (2u, 2u+1), so
        y[2u+e, co] = sum_j sum_ci x[2u+e+j-p, ci] w[co, ci, j]
    is a conv over u with 2*Cin input and 2*Cout output channels and ceil((k+1)/2) taps (block-Toeplitz weights):
        W'[e*Cout+co, f*Cin+ci, v-vmin] = w[co, ci, 2v+f-e+p]   (0 elsewhere),  pad_left' = -vmin.
    For the 32- and 64-channel Generator stages (modules.py:210-223) this doubles the MMA's N (32 -> 64, 64 -> 128: an
    M=128 MMA costs >= 64 cycles whatever its N), trims the K steps of a k-tap conv from 2k to k+1(+1) per time pair
    and fills the 128-byte TMA rows a 32-channel tensor leaves half empty.  Returns (weight', bias', pad_left')."""
    weight = weight.detach().float()
    cout, cin, k = weight.shape
    p = (k - 1) // 2 if pad_left is None else pad_left
    vmin, vmax = (0 - p) // 2, (1 + k - 1 - p) // 2
    w2 = torch.zeros(2 * cout, 2 * cin, vmax - vmin + 1)
    for e in range(2):
        for v in range(vmin, vmax + 1):
            for f in range(2):
                j = 2 * v + f - e + p
                if 0 <= j < k:
                    w2[e * cout:(e + 1) * cout, f * cin:(f + 1) * cin, v - vmin] = weight[:, :, j]
    b2 = None if bias is None else bias.detach().float().repeat(2)
    return w2, b2, -vmin


def pack_conv_folded(weight, bias, *, dtype=torch.float16, block_n=None):
    """pack_conv of the time-pair-folded form of a dilation-1 conv; the caller passes [B, T/2, 2C] views (PackedConv.fold == 2)."""
    cout, cin, k = weight.shape
    w2, b2, pl = fold_time_pairs(weight, bias)
    pc = pack_conv(w2, b2, dil=1, pad_left=pl, dtype=dtype, block_n=block_n)
    pc.alg_macs_per_row = 2 * cin * cout * k      # the unfolded conv's own MACs for the two time steps of a row
    pc.fold = 2
    return pc


class PackedConvWs:
    """Weights of a 64 -> 64 dilation-1 conv for b200svs_conv1d_ws: `img` int32 [128, (k+1)*32] is the tensor-memory image
    of the banded A operand (row s*64 + co, fp16 pair w of chunk j = W[co, 2w / 2w+1, j - s], zero outside 0 <= j - s < k)."""

    def __init__(self, img, bias, ntaps, tag=""):
        self.img, self.bias, self.ntaps, self.tag = img, bias, ntaps, tag

    def to(self, device):
        return PackedConvWs(self.img.to(device), self.bias.to(device), self.ntaps, self.tag)


def pack_conv_ws(weight, bias, tag=""):
    """weight [64, 64, k] (k in 3, 7, 11), effective (weight-norm folded), dilation 1, 'same' padding."""
    weight = weight.detach().float()
    cout, cin, k = weight.shape
    if cout != 64 or cin != 64 or k not in (3, 7, 11):
        raise ValueError("pack_conv_ws: a 64 -> 64 conv with kernel size 3, 7 or 11 is expected")
    if float(weight.abs().max()) > 65504.0:
        raise ValueError("pack_conv_ws: |w| exceeds the fp16 range")
    nj = k + 1
    a = torch.zeros(2, cout, nj, cin)
    for s_out in range(2):
        for j in range(nj):
            tap = j - s_out
            if 0 <= tap < k:
                a[s_out, :, j, :] = weight[:, :, tap]
    img = a.reshape(128, nj * cin).to(torch.float16).contiguous().view(torch.int32)
    return PackedConvWs(img.contiguous(), bias.detach().float().contiguous(), k, tag)


def gate_block_n(hidden):
    """N tile of a gated conv: 192 (two tiles for hidden 192, N=192 MMAs) when hidden splits into 96-wide halves, else 128."""
    return 192 if hidden % 96 == 0 else 128


def gate_order(hidden, block_n=None):
    """Packed column order of a gated [tanh | sigmoid] 2*hidden conv (also used for cond biases)."""
    block_n = block_n or gate_block_n(hidden)
    hb = block_n // 2
    order = []
    for n in range(hidden // hb):
        order += list(range(n * hb, (n + 1) * hb)) + list(range(hidden + n * hb, hidden + (n + 1) * hb))
    return torch.tensor(order)


def pack_conv_transpose(weight, bias, stride, padding, *, dtype=torch.float16, block_n=None):
    """weight: [Cin, Cout, K] ConvTranspose1d weight with K == 2*stride, padding == stride//2... any
    K <= 2*stride + padding works as long as every output phase needs only input offsets -1, 0, +1.

    y[q*s + r, co] = sum_{j in 0..2} sum_ci x[q + j - 1, ci] * w[ci, co, (1 - j)*s + r + p]
    -> a 3-tap conv (pad_left 1) with output column r*Cout + co; the [B, Tin, s*Cout] result is the
    [B, s*Tin, Cout] tensor."""
    weight = weight.detach().float()
    cin, cout, K = weight.shape
    s, p = stride, padding
    w3 = torch.zeros(s * cout, cin, 3)
    for r in range(s):
        for j in range(3):
            k = (1 - j) * s + r + p
            if 0 <= k < K:
                w3[r * cout:(r + 1) * cout, :, j] = weight[:, :, k].t()
    # every kernel tap must be reachable with offsets -1..+1
    reach = {(1 - j) * s + r + p for r in range(s) for j in range(3)}
    if not set(range(K)) <= reach:
        raise ValueError(f"ConvTranspose1d(k={K}, s={s}, p={p}) does not fit the 3-tap polyphase form")
    b3 = None if bias is None else bias.detach().float().repeat(s)
    pc = pack_conv(w3, b3, dil=1, pad_left=1, dtype=dtype, block_n=block_n)
    pc.alg_macs_per_row = cin * cout * K      # the ConvTranspose1d's own MAC count per input frame
    return pc
