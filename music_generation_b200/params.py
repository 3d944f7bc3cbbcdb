"""Parameter tree of SynthesizerTrn, generated from a flat spec.

The drop-in contract is the reference's state-dict: 873 keys for the shipped v5 config
(`...weight_g` / `...weight_v` for weight-normed convs, VISinger/utils/utils.py:33-58 loads them by
name).  Instead of re-declaring the reference's nn.Module classes, the tree is described as
(dotted key, shape, initialiser) rows and materialised into generic containers, so
`state_dict()` / `load_state_dict()` / `.cuda()` behave exactly like the reference object's.
Initialisers follow the reference's distributions (not its RNG stream): PyTorch Conv/Linear default
init, N(0, 0.01) for Generator convs (VISinger/commons.py:8-11), N(0, h^-0.5) embeddings
(VISinger/models_ae.py:189-190), xavier q/k/v (attentions.py:130-132), zero flow `post`
(modules.py:326-327).
"""
import math

import torch
from torch import nn


class Node(nn.Module):
    """Generic container: children and parameters are attached by name from the spec."""


def _conv_default(shape, gen):
    fan_in = 1
    for s in shape[1:]:
        fan_in *= s
    bound = 1.0 / math.sqrt(fan_in)
    return (torch.rand(shape, generator=gen) * 2 - 1) * bound


def _normal(std):
    return lambda shape, gen: torch.randn(shape, generator=gen) * std


def _xavier(shape, gen):
    fan_out = shape[0] * (shape[2] if len(shape) > 2 else 1)
    fan_in = shape[1] * (shape[2] if len(shape) > 2 else 1)
    a = math.sqrt(6.0 / (fan_in + fan_out))
    return (torch.rand(shape, generator=gen) * 2 - 1) * a


def _const(v):
    return lambda shape, gen: torch.full(shape, float(v))


class SpecBuilder:
    def __init__(self):
        self.rows = []          # (key, shape, init, is_buffer)
        self.wn = []            # (prefix, dim0 size) weight-normed convs: g initialised to ||v||

    def add(self, key, shape, init, buffer=False):
        self.rows.append((key, tuple(shape), init, buffer))

    def conv(self, p, cout, cin, k=1, bias=True, init=None, bias_fan=None):
        self.add(p + ".weight", (cout, cin, k), init or _conv_default)
        if bias:
            fan = bias_fan or cin * k
            self.add(p + ".bias", (cout,), lambda s, g, f=fan: (torch.rand(s, generator=g) * 2 - 1) / math.sqrt(f))

    def wn_conv(self, p, vshape, bias_n, init=None):
        """weight_norm(conv): keys bias [bias_n], weight_g [d0,1,1], weight_v vshape (in that order)."""
        fan = vshape[1] * vshape[2]
        self.add(p + ".bias", (bias_n,), lambda s, g, f=fan: (torch.rand(s, generator=g) * 2 - 1) / math.sqrt(f))
        self.add(p + ".weight_g", (vshape[0], 1, 1), None)
        self.add(p + ".weight_v", tuple(vshape), init or _conv_default)
        self.wn.append(p)

    def layernorm(self, p, c):
        self.add(p + ".gamma", (c,), _const(1))
        self.add(p + ".beta", (c,), _const(0))

    def encoder(self, p, hidden, filt, heads, layers, k, window=4):
        dk = hidden // heads
        for i in range(layers):
            a = f"{p}.attn_layers.{i}"
            self.add(a + ".emb_rel_k", (1, 2 * window + 1, dk), _normal(dk ** -0.5))
            self.add(a + ".emb_rel_v", (1, 2 * window + 1, dk), _normal(dk ** -0.5))
            for n in ("q", "k", "v"):
                self.conv(f"{a}.conv_{n}", hidden, hidden, 1, init=_xavier)
            self.conv(a + ".conv_o", hidden, hidden, 1)
        for i in range(layers):
            self.layernorm(f"{p}.norm_layers_1.{i}", hidden)
        for i in range(layers):
            self.conv(f"{p}.ffn_layers.{i}.conv_1", filt, hidden, k)
            self.conv(f"{p}.ffn_layers.{i}.conv_2", hidden, filt, k)
        for i in range(layers):
            self.layernorm(f"{p}.norm_layers_2.{i}", hidden)

    def wn_stack(self, p, hidden, k, layers, gin):
        if gin:
            self.wn_conv(p + ".cond_layer", (2 * hidden * layers, gin, 1), 2 * hidden * layers)
        for i in range(layers):
            self.wn_conv(f"{p}.in_layers.{i}", (2 * hidden, hidden, k), 2 * hidden)
        for i in range(layers):
            n = 2 * hidden if i < layers - 1 else hidden
            self.wn_conv(f"{p}.res_skip_layers.{i}", (n, hidden, 1), n)


def build_spec(*, variant, n_vocab, spec_channels, inter, hidden, filt, heads, n_layers, ksize, resblock,
               rb_kernels, rb_dils, up_rates, up_init, up_kernels, gin, n_vocab_note, max_pos):
    s = SpecBuilder()
    # text_encoder (models_ae.py:153-201)
    s.add("text_encoder.emb_phoneme.weight", (n_vocab, hidden), _normal(hidden ** -0.5))
    s.add("text_encoder.emb_note_pitch.weight", (n_vocab_note, hidden), _normal(hidden ** -0.5))
    s.add("text_encoder.emb_phone_shengdiao.weight", (300, hidden), _normal(1.0))
    s.add("text_encoder.emb_note_duration.weight", (hidden, 1), _conv_default)
    s.add("text_encoder.emb_note_duration.bias", (hidden,), lambda sh, g: torch.rand(sh, generator=g) * 2 - 1)
    s.encoder("text_encoder.encoder", hidden, filt, heads, n_layers, ksize)
    # decoder = Generator (models_ae.py:294-325)
    s.conv("decoder.conv_pre", up_init, inter, 7)
    ch = up_init
    for i, (u, k) in enumerate(zip(up_rates, up_kernels)):
        cin, cout = up_init // (2 ** i), up_init // (2 ** (i + 1))
        # ConvTranspose1d weight is [Cin, Cout, k]; weight_norm dim 0 => g is [Cin,1,1]; bias [Cout]
        s.wn_conv(f"decoder.ups.{i}", (cin, cout, k), cout, init=_normal(0.01))
    for i in range(len(up_rates)):
        ch = up_init // (2 ** (i + 1))
        for j, (k, d) in enumerate(zip(rb_kernels, rb_dils)):
            p = f"decoder.resblocks.{i * len(rb_kernels) + j}"
            if str(resblock) == "1":
                for m in range(len(d)):
                    s.wn_conv(f"{p}.convs1.{m}", (ch, ch, k), ch, init=_normal(0.01))
                for m in range(len(d)):
                    s.wn_conv(f"{p}.convs2.{m}", (ch, ch, k), ch, init=_normal(0.01))
            else:
                for m in range(len(d)):
                    s.wn_conv(f"{p}.convs.{m}", (ch, ch, k), ch, init=_normal(0.01))
    s.conv("decoder.conv_post", 1, ch, 7, bias=False)
    if gin:
        s.conv("decoder.cond", up_init, gin, 1)
    # duration_predictor (models_ae.py:102-124)
    for i, cin in ((1, hidden), (2, 256), (3, 256)):
        s.conv(f"duration_predictor.conv_{i}", 256, cin, 3)
        s.layernorm(f"duration_predictor.norm_{i}", 256)
    s.conv("duration_predictor.proj", 1, 256, 1)
    if gin:
        s.conv("duration_predictor.cond", hidden, gin, 1)
    # text_post_net (ext_modules/text_post_net.py:17-62)
    s.add("text_post_net.position_ids", (1, max_pos), lambda sh, g: torch.arange(sh[1]).unsqueeze(0), buffer=True)
    s.add("text_post_net.position_embeddings.weight", (max_pos, hidden), _normal(1.0))
    s.encoder("text_post_net.f0_predictor.encoder", hidden, filt, heads, 6, ksize)
    s.conv("text_post_net.f0_predictor.output_layer", 1, hidden, 1)
    s.encoder("text_post_net.mel_predict_net.encoder", hidden, filt, heads, 4, ksize)
    s.conv("text_post_net.mel_predict_net.condition_proj", hidden, hidden, 1)
    s.conv("text_post_net.mel_predict_net.output_layer", 80, hidden, 1)
    for i in range(4):
        s.conv(f"text_post_net.frame_prior_net.conv_layers.{i}", hidden, 80 if i == 0 else hidden, 1)
    s.conv("text_post_net.frame_prior_net.condition_proj", 80, hidden, 1)
    s.conv("text_post_net.proj", 2 * inter, hidden, 1)
    s.add("text_post_net.f0_emb.weight", (300, hidden), _normal(1.0))
    # spec_encoder = PosteriorEncoder: training only, kept for state-dict compatibility (models_ae.py:256-291)
    s.conv("spec_encoder.pre", hidden, spec_channels, 1)
    s.wn_stack("spec_encoder.enc", hidden, 5, 16, gin)
    s.conv("spec_encoder.proj", 2 * inter, hidden, 1)
    # spec_flow (models_ae.py:223-244; modules.py:299-327)
    for i in range(4):
        p = f"spec_flow.spec_flows.{2 * i}"
        s.conv(p + ".pre", hidden, inter // 2, 1)
        s.wn_stack(p + ".enc", hidden, 5, 4, gin)
        s.add(p + ".post.weight", (inter // 2, hidden, 1), _const(0))
        s.add(p + ".post.bias", (inter // 2,), _const(0))
    if variant == "ae":
        s.conv("spec_decoder.pre", hidden, inter, 1)
        s.wn_stack("spec_decoder.enc", hidden, 5, 16, gin)
        s.conv("spec_decoder.proj", spec_channels, hidden, 1)
        s.conv("spec_proj", inter, spec_channels, 1)
    return s


def materialise(root, spec, seed=1234):
    """Creates parameters/buffers on `root` following the dotted keys of `spec`."""
    gen = torch.Generator().manual_seed(seed)
    values = {}
    for key, shape, init, is_buf in spec.rows:
        if init is not None:
            values[key] = init(shape, gen)
    for p in spec.wn:   # weight_norm initialises g to the per-dim-0 norm of v
        v = values[p + ".weight_v"]
        values[p + ".weight_g"] = v.reshape(v.shape[0], -1).norm(dim=1).reshape(-1, 1, 1)
    for key, shape, init, is_buf in spec.rows:
        parts = key.split(".")
        node = root
        for name in parts[:-1]:
            if not hasattr(node, name):
                node.add_module(name, Node())
            node = getattr(node, name)
        t = values[key]
        assert tuple(t.shape) == shape, (key, t.shape, shape)
        if is_buf:
            node.register_buffer(parts[-1], t)
        else:
            node.register_parameter(parts[-1], nn.Parameter(t, requires_grad=False))
    return root
