"""SynthesizerTrn - drop-in for the reference's VISinger synthesis model on B200.

Mirrors the constructor and `infer(...)` of VISinger/models_ae.py:446-651 (variant "ae", SVS v5/v6,
6-tuple result) and VISinger/models.py:622-673 (variant "v4", 5-tuple result).  Same state-dict keys,
same `.cuda()/.eval()/.load_state_dict()` behaviour, so pipline_inference_svs_*.Synthesiser works
unchanged.  Everything numeric runs in the hand-written sm_100a kernels of
libb200svs.so; there is no PyTorch/CPU fallback - without the library or a B200 `infer` raises.

Precision plan (DESIGN.md "Numerics"): phoneme-level path and F0 predictor in fp32 (their outputs
cross float->int boundaries), every other conv on the tcgen05 tensor cores with 16-bit operands and
fp32 accumulation, residual streams kept in fp32.
"""
import math

import torch
from torch import nn

from . import _lib, ops, packer, params
from .ops import ACT_GELU, ACT_NONE, ACT_RELU, ACT_SIGMOID

LRELU_SLOPE = 0.1   # VISinger/modules.py:17


def _require_cuda(t):
    if not t.is_cuda:
        raise RuntimeError("SynthesizerTrn.infer: inputs must be CUDA tensors - this build has no CPU path")


_STAGE_API = True       # module switch: the CPU test-suite's emulated ops (tests/emu_ops.py) have no stage-level entry points


def _device_guard(dev):
    """Makes `dev` the current CUDA device for the duration of a call: the kernels of libb200svs launch on the current
    device's stream (ops._stream) and must match the device of the tensors they are handed."""
    if dev.type != "cuda":
        raise RuntimeError("SynthesizerTrn.infer: the model is not on a CUDA device (call .cuda() / .to(device))")
    return torch.cuda.device(dev)


class _Enc:
    """Packed weights of one attentions.Encoder layer."""
    __slots__ = ("qkv", "o", "ffn1", "ffn2", "rel_k", "rel_v", "ln1", "ln2")


class SynthesizerTrn(nn.Module):
    variant = "ae"
    GENERATOR_HALO_FRAMES = 32          # > receptive field of the Generator in frames (conv_pre 3 + stage 0: 7.5 + ...)

    def __init__(self, n_vocab, spec_channels, segment_size, inter_channels, hidden_channels, filter_channels,
                 n_heads, n_layers, kernel_size, p_dropout, resblock, resblock_kernel_sizes,
                 resblock_dilation_sizes, upsample_rates, upsample_initial_channel, upsample_kernel_sizes,
                 n_speakers=0, gin_channels=0, use_sdp=True, n_vocab_note=100, *, operand_dtype="fp16",
                 max_position_embeddings=2000, generator_chunk_frames=32768, generator_wide_chunk_frames=196608,
                 generator_wide_stages=1, f0_precision="split",
                 text_precision="split", generator_residual="act16", generator_mrf="fp16", generator_fold=32, wn_skip="gemm", cuda_graphs=False,
                 fuse_spec_proj=False, stage_api=True, generator_bucket=False, generator_pair_fused=True,
                 seed=1234, **kwargs):
        super().__init__()
        if hidden_channels != 192 or n_heads != 2:
            raise ValueError("kernels are specialised for hidden_channels=192, n_heads=2 (ours_ae.json)")
        self.n_vocab, self.spec_channels, self.segment_size = n_vocab, spec_channels, segment_size
        self.inter_channels, self.hidden_channels, self.filter_channels = inter_channels, hidden_channels, filter_channels
        self.n_heads, self.n_layers, self.kernel_size, self.p_dropout = n_heads, n_layers, kernel_size, p_dropout
        self.resblock = str(resblock)
        self.resblock_kernel_sizes = list(resblock_kernel_sizes)
        self.resblock_dilation_sizes = [list(d) for d in resblock_dilation_sizes]
        self.upsample_rates, self.upsample_kernel_sizes = list(upsample_rates), list(upsample_kernel_sizes)
        self.upsample_initial_channel = upsample_initial_channel
        self.n_speakers, self.gin_channels, self.use_sdp = n_speakers, gin_channels, use_sdp
        self.operand_dtype = {"fp16": torch.float16, "bf16": torch.bfloat16}[operand_dtype]
        self.generator_chunk_frames = generator_chunk_frames
        self.generator_time_chunk_frames = 65536       # per utterance; longer inputs are cut in time with exact halos
        self.generator_wide_chunk_frames, self.generator_wide_stages = generator_wide_chunk_frames, generator_wide_stages
        if f0_precision not in ("split", "fp32"):
            raise ValueError("f0_precision must be 'split' (tensor cores, fp16 hi/lo operands) or 'fp32' (SIMT)")
        self.f0_precision = f0_precision
        if text_precision not in ("split", "fp32"):
            raise ValueError("text_precision must be 'split' or 'fp32'")
        self.text_precision = text_precision
        if generator_residual not in ("act16", "fp16", "fp32"):
            raise ValueError("generator_residual must be 'act16' (the residual of a ResBlock pair is recovered from the "
                             "leaky-ReLU'd 16-bit tensor its first conv consumed; fp16 operand mode only), 'fp16' "
                             "(separate un-activated 16-bit residual stream) or 'fp32'")
        self.generator_residual = generator_residual
        if generator_mrf not in ("fp16", "fp32"):
            raise ValueError("generator_mrf must be 'fp16' (the MRF sum xs += resblock_j(x) as a 16-bit stream; only "
                             "with generator_residual='fp16') or 'fp32'")
        self.generator_mrf = generator_mrf
        if wn_skip not in ("gemm", "accumulate"):
            raise ValueError("wn_skip must be 'gemm' (WN skip sum as one 1x1 conv over all layers' gate outputs) or 'accumulate'")
        self.wn_skip = wn_skip
        # ResBlock convs with dilation 1 and <= generator_fold channels run on [B, T/2, 2C] views (0: never)
        # or {channels: largest kernel size to fold}, e.g. {32: 11, 64: 3}
        # dict values: largest kernel size to fold, or an explicit list of kernel sizes (e.g. {32: 11, 64: [7, 11]})
        self.generator_fold = ({int(c): (sorted(int(v) for v in k) if isinstance(k, (list, tuple)) else int(k))
                                for c, k in generator_fold.items()} if isinstance(generator_fold, dict)
                               else int(generator_fold))
        # every call site of the reference discards spec_predict; with fuse_spec_proj=True the two
        # 513-wide 1x1 convs SpecDecoder.proj -> spec_proj collapse into one 192x192 conv (see _pack) and infer
        # returns None in spec_predict's place.  Off by default: the result tuple then matches the reference's.
        self.fuse_spec_proj = bool(fuse_spec_proj)
        # stage-level C entry points (b200svs_generator_fwd / _wn_stack_fwd / _encoder_fwd): one library call per stage
        # instead of one per kernel; False keeps the per-op Python loops (bit-identical; also used by bench.py's
        # per-launch instrumentation and by the non-default precision knobs)
        self.stage_api = bool(stage_api)
        # ragged batches: run the Generator on length-sorted utterance groups, each over its own longest length + halo
        # instead of the batch's padded length.  Valid samples are bit-identical; samples of the PADDED region (which the
        # reference fills with bias-driven values and every consumer masks) are written as 0.  Off by default so that
        # `infer` matches the reference over the whole padded tensor; pipeline.Synthesiser turns it on.
        self.generator_bucket = bool(generator_bucket)
        # 32- and 64-channel stages (kernel sizes whose two weight sets fit in shared memory): one fused kernel per ResBlock pair (b200svs_resblock_pair, 4 instead of 10 B of HBM traffic per
        # element, bit-identical to the two conv launches; -1 % on the whole step).
        self.generator_pair_fused = bool(generator_pair_fused)
        # launch-bound (small) calls: replay the frame-level part of infer from a CUDA graph keyed by its shapes
        self.cuda_graphs = bool(cuda_graphs)
        self.cuda_graph_max_frames = 32768
        self.cuda_graph_cache = 8
        self._graphs = {}
        self.hop = 1
        for u in self.upsample_rates:
            self.hop *= u
        spec = params.build_spec(
            variant=self.variant, n_vocab=n_vocab, spec_channels=spec_channels, inter=inter_channels,
            hidden=hidden_channels, filt=filter_channels, heads=n_heads, n_layers=n_layers, ksize=kernel_size,
            resblock=self.resblock, rb_kernels=self.resblock_kernel_sizes, rb_dils=self.resblock_dilation_sizes,
            up_rates=self.upsample_rates, up_init=upsample_initial_channel, up_kernels=self.upsample_kernel_sizes,
            gin=gin_channels, n_vocab_note=n_vocab_note, max_pos=max_position_embeddings)
        params.materialise(self, spec, seed=seed)
        self._packed = None
        self.last_aux = {}

    # ------------------------------------------------------------------ nn.Module plumbing
    def _apply(self, fn, *a, **k):
        self._packed = None
        self._graphs = {}
        return super()._apply(fn, *a, **k)

    def load_state_dict(self, state_dict, strict=True, **k):
        self._packed = None
        self._graphs = {}
        # tolerate a position table enlarged for T > 2000
        kp = "text_post_net.position_embeddings.weight"
        if kp in state_dict and state_dict[kp].shape != self.text_post_net.position_embeddings.weight.shape:
            n = state_dict[kp].shape[0]
            dev = self.text_post_net.position_embeddings.weight.device
            self.text_post_net.position_embeddings.weight = nn.Parameter(
                torch.empty(n, self.hidden_channels, device=dev), requires_grad=False)
            self.text_post_net.position_ids = torch.arange(n, device=dev).unsqueeze(0)
        return super().load_state_dict(state_dict, strict=strict, **k)

    @property
    def device(self):
        return self.text_post_net.f0_emb.weight.device

    def _pack_wn_stack(self, sd, prefix, dev, dt):
        """Packed weights of one modules.WN stack (modules.py:110-176): gate-interleaved in_layers, res/skip 1x1 convs, the
        skip sum as one K-concatenated 1x1 conv, the speaker cond_layer in gate order."""
        W = lambda p: packer.fold_weight_norm(sd, p)
        Bv = lambda p: sd.get(p + ".bias")
        n = 0
        while f"{prefix}.in_layers.{n}.bias" in sd:
            n += 1
        st = {"in": [], "rs": [], "n": n, "cond": None}
        for i in range(n):
            st["in"].append(packer.pack_conv(W(f"{prefix}.in_layers.{i}"), Bv(f"{prefix}.in_layers.{i}"),
                                             dtype=dt, gate=True).to(dev))
            st["rs"].append(packer.pack_conv(W(f"{prefix}.res_skip_layers.{i}"), Bv(f"{prefix}.res_skip_layers.{i}"),
                                             dtype=dt, block_n=192).to(dev))
        # skip path as ONE GEMM: output = sum_i W_skip_i . acts_i + sum_i b_skip_i (modules.py:165-176 accumulates it
        # layer by layer) = a 1x1 conv over the K-concatenation of all layers' gate outputs
        H = self.hidden_channels
        ws, bs = [], 0.0
        for i in range(n):
            w_i, b_i = W(f"{prefix}.res_skip_layers.{i}"), Bv(f"{prefix}.res_skip_layers.{i}")
            lo = H if i < n - 1 else 0          # the last layer's 192 outputs are all skip
            ws.append(w_i[lo:lo + H])
            bs = bs + b_i[lo:lo + H].double()
        st["skip_cat"] = packer.pack_conv(torch.cat(ws, 1), bs.float(), dtype=dt).to(dev)
        if self.gin_channels:
            wc, bc = W(prefix + ".cond_layer"), Bv(prefix + ".cond_layer")
            h2 = 2 * self.hidden_channels
            order = packer.gate_order(self.hidden_channels)
            idx = torch.cat([order + i * h2 for i in range(n)])
            st["cond"] = packer.pack_conv(wc[idx], bc[idx], f32=True).to(dev)
        return st

    # ------------------------------------------------------------------ weight packing (once per load)
    def _pack(self):
        sd = {k: v.detach().float().cpu() for k, v in self.state_dict().items()}
        dev, dt = self.device, self.operand_dtype
        W = lambda p: packer.fold_weight_norm(sd, p)
        Bv = lambda p: sd.get(p + ".bias")
        P = {}

        def enc(prefix, n_layers, f32, split=False):
            layers = []
            for i in range(n_layers):
                a, e = f"{prefix}.attn_layers.{i}", _Enc()
                wqkv = torch.cat([W(f"{a}.conv_{n}") for n in "qkv"], 0)
                bqkv = torch.cat([Bv(f"{a}.conv_{n}") for n in "qkv"], 0)
                e.qkv = packer.pack_conv(wqkv, bqkv, dtype=dt, f32=f32, split=split).to(dev)
                e.o = packer.pack_conv(W(a + ".conv_o"), Bv(a + ".conv_o"), dtype=dt, f32=f32, split=split).to(dev)
                f = f"{prefix}.ffn_layers.{i}"
                e.ffn1 = packer.pack_conv(W(f + ".conv_1"), Bv(f + ".conv_1"), dtype=dt, f32=f32, split=split).to(dev)
                e.ffn2 = packer.pack_conv(W(f + ".conv_2"), Bv(f + ".conv_2"), dtype=dt, f32=f32, split=split).to(dev)
                e.rel_k = sd[a + ".emb_rel_k"][0].contiguous().to(dev)
                e.rel_v = sd[a + ".emb_rel_v"][0].contiguous().to(dev)
                e.ln1 = (sd[f"{prefix}.norm_layers_1.{i}.gamma"].to(dev), sd[f"{prefix}.norm_layers_1.{i}.beta"].to(dev))
                e.ln2 = (sd[f"{prefix}.norm_layers_2.{i}.gamma"].to(dev), sd[f"{prefix}.norm_layers_2.{i}.beta"].to(dev))
                layers.append(e)
            return layers

        def conv(p, f32=False, **kw):
            return packer.pack_conv(W(p), Bv(p), dtype=dt, f32=f32, **kw).to(dev)

        def wn_stack(prefix):
            return self._pack_wn_stack(sd, prefix, dev, dt)

        te = "text_encoder"
        P["emb"] = tuple(sd[f"{te}.{n}"].contiguous().to(dev) for n in
                         ("emb_phoneme.weight", "emb_note_pitch.weight", "emb_phone_shengdiao.weight"))
        P["emb_dur"] = (sd[te + ".emb_note_duration.weight"].reshape(-1).contiguous().to(dev),
                        sd[te + ".emb_note_duration.bias"].contiguous().to(dev))
        t_split = self.text_precision == "split"
        P["text_enc"] = enc(te + ".encoder", self.n_layers, not t_split, split=t_split)
        P["text_mode"] = "split" if t_split else "f32"
        dp = "duration_predictor"
        P["dp"] = [(conv(f"{dp}.conv_{i}", f32=True), sd[f"{dp}.norm_{i}.gamma"].to(dev), sd[f"{dp}.norm_{i}.beta"].to(dev))
                   for i in (1, 2, 3)]
        P["dp_proj"] = conv(dp + ".proj", f32=True)
        tp = "text_post_net"
        P["pos"] = sd[tp + ".position_embeddings.weight"].contiguous().to(dev)
        P["zero_pos"] = torch.zeros(512, self.hidden_channels, device=dev)      # "no position table" operand of add_pos_mask
        f0_split = self.f0_precision == "split"
        P["f0_enc"] = enc(tp + ".f0_predictor.encoder", 6, not f0_split, split=f0_split)
        P["f0_mode"] = "split" if f0_split else "f32"
        P["f0_out"] = conv(tp + ".f0_predictor.output_layer", f32=True)
        P["f0_emb"] = sd[tp + ".f0_emb.weight"].contiguous().to(dev)
        P["mel_cond"] = conv(tp + ".mel_predict_net.condition_proj")
        P["mel_enc"] = enc(tp + ".mel_predict_net.encoder", 4, False)
        P["mel_out"] = conv(tp + ".mel_predict_net.output_layer")
        P["fp_cond"] = conv(tp + ".frame_prior_net.condition_proj")
        P["fp"] = []
        i = 0
        while f"{tp}.frame_prior_net.conv_layers.{i}.weight" in sd:
            P["fp"].append(conv(f"{tp}.frame_prior_net.conv_layers.{i}"))
            i += 1
        P["proj"] = conv(tp + ".proj")
        # flow: Flip folded into pre (input order) / post (output order); reverse order of application
        half = self.inter_channels // 2
        n_flows = 0
        while f"spec_flow.spec_flows.{2 * n_flows}.pre.weight" in sd:
            n_flows += 1
        P["flow"] = []
        for i in reversed(range(n_flows)):
            p = f"spec_flow.spec_flows.{2 * i}"
            flipped = (n_flows - 1 - i) % 2 == 0
            wpre, wpost, bpost = W(p + ".pre"), W(p + ".post"), Bv(p + ".post")
            if flipped:
                wpre, wpost, bpost = wpre.flip(1), wpost.flip(0), bpost.flip(0)
            P["flow"].append(dict(
                pre=packer.pack_conv(wpre, Bv(p + ".pre"), dtype=dt).to(dev),
                post=packer.pack_conv(wpost, bpost, dtype=dt).to(dev),
                wn=wn_stack(p + ".enc"),
                in_off=half if flipped else 0, out_off=0 if flipped else half))
        if self.variant == "ae":
            P["sd_pre"] = conv("spec_decoder.pre")
            P["sd_wn"] = wn_stack("spec_decoder.enc")
            P["sd_proj"] = conv("spec_decoder.proj")
            P["spec_proj"] = conv("spec_proj")
            # spec_proj(proj(x) * m) * m == m * (W2 W1 x + W2 b1 + b2) for a 0/1 frame mask m
            # (ext_modules/spec_decoder.py:41-46, models_ae.py:647-650): one masked 1x1 conv
            w1, b1 = W("spec_decoder.proj")[:, :, 0], Bv("spec_decoder.proj")
            w2, b2 = W("spec_proj")[:, :, 0], Bv("spec_proj")
            P["sd_fused"] = packer.pack_conv((w2.double() @ w1.double()).float().unsqueeze(-1),
                                             (w2.double() @ b1.double() + b2.double()).float(), dtype=dt).to(dev)
        # generator
        d = "decoder"
        P["g_pre"] = conv(d + ".conv_pre")
        P["g_cond"] = conv(d + ".cond", f32=True) if self.gin_channels else None
        P["g_ups"] = []
        for i, (u, k) in enumerate(zip(self.upsample_rates, self.upsample_kernel_sizes)):
            P["g_ups"].append(packer.pack_conv_transpose(W(f"{d}.ups.{i}"), Bv(f"{d}.ups.{i}"), u, (k - u) // 2,
                                                         dtype=dt).to(dev))
        P["g_rb"] = []
        nk = len(self.resblock_kernel_sizes)
        for i in range(len(self.upsample_rates)):
            stage = []
            for j, (k, dils) in enumerate(zip(self.resblock_kernel_sizes, self.resblock_dilation_sizes)):
                p = f"{d}.resblocks.{i * nk + j}"
                ch_i = self.upsample_initial_channel // (2 ** (i + 1))

                def rconv(name, dl, ch_i=ch_i, k=k):
                    # dilation-1 convs of the narrow stages run time-pair folded (packer.fold_time_pairs)
                    gf = self.generator_fold
                    sel = gf.get(ch_i, 0) if isinstance(gf, dict) else None
                    on = ((k in sel) if isinstance(sel, list) else sel >= k) if isinstance(gf, dict) else ch_i <= gf
                    if dl == 1 and on and math.prod(self.upsample_rates[:i + 1]) % 2 == 0:
                        return packer.pack_conv_folded(W(name), Bv(name), dtype=dt).to(dev)
                    return conv(name, dil=dl)
                if self.resblock == "1":
                    pairs = [(rconv(f"{p}.convs1.{m}", dl), rconv(f"{p}.convs2.{m}", 1)) for m, dl in enumerate(dils)]
                else:
                    pairs = [(rconv(f"{p}.convs.{m}", dl), None) for m, dl in enumerate(dils)]
                stage.append(pairs)
            P["g_rb"].append(stage)
        wpost = W(d + ".conv_post")                     # [1, C, 7] -> [7][C] fp32
        P["g_post"] = wpost[0].t().contiguous().to(dev)
        # stage tags for bench.py's per-stage breakdown
        def tag(obj, name):
            if isinstance(obj, ops.PackedConv):
                obj.tag = name
            elif isinstance(obj, _Enc):
                for a in ("qkv", "o", "ffn1", "ffn2"):
                    getattr(obj, a).tag = name
            elif isinstance(obj, dict):
                for v in obj.values():
                    tag(v, name)
            elif isinstance(obj, (list, tuple)):
                for v in obj:
                    tag(v, name)
        for key, name in (("text_enc", "text"), ("dp", "text"), ("dp_proj", "text"), ("f0_enc", "f0"), ("f0_out", "f0"),
                          ("mel_cond", "mel"), ("mel_enc", "mel"), ("mel_out", "mel"), ("fp_cond", "prior"),
                          ("fp", "prior"), ("proj", "prior"), ("flow", "flow"), ("sd_pre", "specdec"),
                          ("sd_wn", "specdec"), ("sd_proj", "specdec"), ("spec_proj", "specdec"), ("sd_fused", "specdec"),
                          ("g_pre", "gen_pre"),
                          ("g_cond", "gen_pre")):
            if key in P and P[key] is not None:
                tag(P[key], name)
        for i, pc in enumerate(P["g_ups"]):
            tag(pc, f"gen_up{i}")
        for i, st in enumerate(P["g_rb"]):
            tag(st, f"gen_rb{i}")
        self._build_stage_tables(P)
        self._packed = P
        return P

    def _use_stage_api(self):
        # bench.py's per-launch instrumentation (_lib.PROFILE) needs one library call per kernel
        return self.stage_api and _STAGE_API and _lib.PROFILE is None

    def _build_stage_tables(self, P):
        """b200svs_generator_w / _wn_w / _encoder_layer_w tables over the packed weights (include/b200svs.h)."""
        dtc = ops.dt_code(self.operand_dtype)
        H = self.hidden_channels

        def wn_table(st):
            t = _lib.WnW()
            t.n_layers, t.hidden, t.dtype = st["n"], H, dtc
            for i in range(st["n"]):
                t.in_[i], t.rs[i] = ops.conv_w(st["in"][i]), ops.conv_w(st["rs"][i])
            t.skip_cat = ops.conv_w(st["skip_cat"])
            st["c"] = t if st["n"] <= _lib.MAX_WN_LAYERS else None
        for fl in P["flow"]:
            wn_table(fl["wn"])
        if "sd_wn" in P:
            wn_table(P["sd_wn"])
        if "pe_wn" in P:
            wn_table(P["pe_wn"])

        def enc_table(layers, mode):
            arr = (_lib.EncoderLayerW * len(layers))()
            for t, e in zip(arr, layers):
                t.hidden, t.filter, t.heads, t.window = H, self.filter_channels, self.n_heads, 4
                t.mode, t.dtype = {"f32": 0, "tc": 1, "split": 2}[mode], dtc
                t.qkv, t.o, t.ffn1, t.ffn2 = (ops.conv_w(getattr(e, a)) for a in ("qkv", "o", "ffn1", "ffn2"))
                t.rel_k, t.rel_v = e.rel_k.data_ptr(), e.rel_v.data_ptr()
                t.ln1_g, t.ln1_b, t.ln2_g, t.ln2_b = (x.data_ptr() for x in (*e.ln1, *e.ln2))
            return arr
        P["c_enc"] = {id(P[k]): enc_table(P[k], m) for k, m in (("text_enc", P["text_mode"]), ("f0_enc", P["f0_mode"]),
                                                                  ("mel_enc", "tc"))}
        # Generator: the default 16-bit streams only (act16 residual + fp16 MRF need fp16 operands), ResBlock1
        P["c_gen"] = None
        n_st, nk = len(self.upsample_rates), len(self.resblock_kernel_sizes)
        npairs = len(self.resblock_dilation_sizes[0]) if self.resblock_dilation_sizes else 0
        if (self.resblock == "1" and self.operand_dtype == torch.float16 and self.generator_residual == "act16"
                and self.generator_mrf == "fp16" and n_st <= _lib.MAX_STAGES and nk <= _lib.MAX_KERNELS
                and npairs <= _lib.MAX_PAIRS and all(len(d) == npairs for d in self.resblock_dilation_sizes)):
            g = _lib.GeneratorW()
            g.n_stages, g.n_kernels, g.n_pairs, g.dtype = n_st, nk, npairs, dtc
            g.pre_channels, g.lrelu_slope = self.upsample_initial_channel, LRELU_SLOPE
            g.pre = ops.conv_w(P["g_pre"])
            for i in range(n_st):
                g.up_rate[i] = self.upsample_rates[i]
                g.channels[i] = self.upsample_initial_channel // (2 ** (i + 1))
                g.ups[i] = ops.conv_w(P["g_ups"][i])
                for j in range(nk):
                    for m in range(npairs):
                        g.rb[i][j][m][0], g.rb[i][j][m][1] = (ops.conv_w(c) for c in P["g_rb"][i][j][m])
            g.post_w, g.post_taps = P["g_post"].data_ptr(), P["g_post"].shape[0]
            P["c_gen"] = g

    # ------------------------------------------------------------------ building blocks
    def _encoder(self, layers, x32, x16, lens, mode):
        """attentions.Encoder (VISinger/attentions.py:35-47), channels-last, in place on x32 (/x16).
        x is masked on entry and after every LayerNorm; valid rows equal the reference's.
        mode "f32"  : fp32 SIMT convs + fp32 attention (phoneme level: durations are floor()ed bit-exactly)
        mode "tc"   : tensor-core convs / attention with 16-bit operands (mel predictor); x16 = 16-bit copy of x32
        mode "split": tensor cores with fp16 (hi, lo) split operands, fp32-class accuracy (F0 predictor, whose
                      output is quantised by f0_to_coarse); x16 = [B,T,3C] fp16 hi|lo|hi copy of x32"""
        B, T, Cc = x32.shape
        dev = x32.device
        dt = self.operand_dtype
        f32, split = mode == "f32", mode == "split"
        tab = self._packed["c_enc"].get(id(layers)) if self._packed else None
        if self._use_stage_api() and tab is not None:
            ops.encoder_fwd(tab, len(layers), x32, x16, lens)
            return
        qkv = torch.empty(B, T, 3 * Cc, device=dev, dtype=torch.float32)
        tmp = torch.empty(B, T, Cc, device=dev, dtype=torch.float32)
        F_ = self.filter_channels
        if f32:
            att = torch.empty(B, T, Cc, device=dev, dtype=torch.float32)
            hid = torch.empty(B, T, F_, device=dev, dtype=torch.float32)
        elif split:
            att = torch.empty(B, T, 3 * Cc, device=dev, dtype=torch.float16)
            hid = torch.empty(B, T, 3 * F_, device=dev, dtype=torch.float16)
        else:
            att = torch.empty(B, T, Cc, device=dev, dtype=dt)
            hid = torch.empty(B, T, F_, device=dev, dtype=dt)
        ws = {}
        for e in layers:
            xin = x32 if f32 else x16
            ops.conv1d(e.qkv, xin, out32=qkv)
            if f32:
                ops.attention(qkv, att, None, e.rel_k, e.rel_v, lens)
            else:
                ops.attention_tc(qkv, None, att, e.rel_k, e.rel_v, lens, dtype=dt, split=split, out16_split=split, ws=ws)
            ops.conv1d(e.o, att, res1=x32, out32=tmp)
            ops.layernorm(tmp, e.ln1[0], e.ln1[1], lens, out32=x32, out16=x16, split16=split)
            xin = x32 if f32 else x16
            if f32:
                ops.conv1d(e.ffn1, xin, act=ACT_RELU, lens=lens, out32=hid)
            else:
                ops.conv1d(e.ffn1, xin, act=ACT_RELU, lens=lens, out16=hid, split16=split)
            ops.conv1d(e.ffn2, hid, res1=x32, lens=lens, out32=tmp)
            ops.layernorm(tmp, e.ln2[0], e.ln2[1], lens, out32=x32, out16=x16, split16=split)

    def _wn(self, st, x32, x16, lens, out16, gcond=None):
        """modules.WN (VISinger/modules.py:148-176): x32/x16 hold the masked input and are updated in
        place; the masked skip sum is written to out16 (16-bit, consumer is always a 1x1 conv)."""
        B, T, H = x32.shape
        dev = x32.device
        n = st["n"]
        if self._use_stage_api() and self.wn_skip == "gemm" and st.get("c") is not None:
            ops.wn_stack_fwd(st["c"], x32, x16, lens, gcond, out16)
            return
        if self.wn_skip == "gemm":
            # every layer's gate output is kept ([B, T, n*H], layer i in columns i*H..) and the skip sum is one 1x1
            # conv over all of it: no fp32 skip accumulator read and written per layer, one launch less per layer
            acts_all = torch.empty(B, T, n * H, device=dev, dtype=self.operand_dtype)
            for i in range(n):
                gb = None if gcond is None else gcond[:, i * 2 * H:(i + 1) * 2 * H]
                a_i = acts_all[:, :, i * H:(i + 1) * H]
                ops.conv1d(st["in"][i], x16, gbias=gb, out16=a_i)
                if i < n - 1:
                    ops.conv1d(st["rs"][i], a_i, w_rows=(0, H), res1=x32, lens=lens, out32=x32, out16=x16)
            ops.conv1d(st["skip_cat"], acts_all, lens=lens, out16=out16)
            return
        acts = torch.empty(B, T, H, device=dev, dtype=self.operand_dtype)
        skip = torch.empty(B, T, H, device=dev, dtype=torch.float32)
        for i in range(n):
            gb = None if gcond is None else gcond[:, i * 2 * H:(i + 1) * 2 * H]
            ops.conv1d(st["in"][i], x16, gbias=gb, out16=acts)
            rs = st["rs"][i]
            if i < n - 1:
                ops.conv1d(rs, acts, w_rows=(0, H), res1=x32, lens=lens, out32=x32, out16=x16)
                ops.conv1d(rs, acts, w_rows=(H, H), res1=skip if i > 0 else None, out32=skip)
            else:
                ops.conv1d(rs, acts, res1=skip if i > 0 else None, lens=lens, out16=out16)

    def _generator_bucketed(self, x16, gbias, frame_lens):
        """Length-sorted utterance groups, each synthesised over min(T, its longest utterance + halo) frames."""
        B, T, _ = x16.shape
        hl = self.GENERATOR_HALO_FRAMES
        order = sorted(range(B), key=lambda b: -int(frame_lens[b]))
        groups, i = [], 0
        while i < B:
            tc = min(T, int(frame_lens[order[i]]) + hl)
            tc += tc & 1                                    # even: time-pair folded convs
            tc = min(tc, T)
            n = max(1, self.generator_chunk_frames // max(tc, 1))
            groups.append((order[i:i + n], tc))
            i += n
        if sum(len(g) * tc for g, tc in groups) > 0.93 * B * T:
            return None                                     # not ragged enough to pay for the gathers
        out = torch.zeros(B, 1, T * self.hop, device=x16.device, dtype=torch.float32)
        for idx, tc in groups:
            ix = torch.tensor(idx, device=x16.device)
            part = self._generator(x16.index_select(0, ix)[:, :tc].contiguous(), None if gbias is None else gbias.index_select(0, ix))
            out[ix, :, :tc * self.hop] = part
        return out

    def _generator(self, x16, gbias=None, frame_lens=None):
        """Generator.forward (VISinger/models_ae.py:327-346).  x16: [B,T,C] masked input; returns [B,1,T*hop].

        Utterance chunks bound the activation working set (stage tensors are hop*C*T elements).  Two chunk sizes:
        conv_pre and the first `generator_wide_stages` upsampling stages (few, tensor-bound launches whose fixed
        costs want long launches) run over `generator_wide_chunk_frames`, the later HBM-bound stages over
        `generator_chunk_frames` (measured: stage 0 19.0 -> 16.5 ms on the whole 64 x 30 s batch, while stages 2-3
        lose 10 % beyond ~32 k frames per chunk)."""
        P = self._packed
        B, T, _ = x16.shape
        if frame_lens is not None and self.generator_bucket and B > 1:
            out = self._generator_bucketed(x16, gbias, frame_lens)
            if out is not None:
                return out
        tc = self.generator_time_chunk_frames
        if tc and T > tc + 2 * self.GENERATOR_HALO_FRAMES:
            # one very long utterance: cut in time with exact halos (the Generator's receptive field is < 16 frames per
            # side; conv zero padding is only ever seen at the true ends) - bit-identical to the uncut call
            out = torch.empty(B, 1, T * self.hop, device=x16.device, dtype=torch.float32)
            hl = self.GENERATOR_HALO_FRAMES
            for t0 in range(0, T, tc):
                t1 = min(T, t0 + tc)
                s0, s1 = max(0, t0 - hl), min(T, t1 + hl)
                part = self._generator(x16[:, s0:s1].contiguous(), gbias)
                out[:, :, t0 * self.hop:t1 * self.hop] = part[:, :, (t0 - s0) * self.hop:(t1 - s0) * self.hop]
            return out
        out = torch.empty(B, 1, T * self.hop, device=x16.device, dtype=torch.float32)
        if self._use_stage_api() and P.get("c_gen") is not None:
            g = P["c_gen"]
            g.chunk_frames, g.wide_chunk_frames = int(self.generator_chunk_frames), int(self.generator_wide_chunk_frames)
            g.wide_stages = int(self.generator_wide_stages)
            g.fuse_pairs = int(self.generator_pair_fused)
            ops.generator_fwd(g, x16, gbias, out)
            return out
        n_st = len(self.upsample_rates)
        n_wide = min(max(int(self.generator_wide_stages), 0), n_st)
        wide = max(self.generator_wide_chunk_frames, self.generator_chunk_frames) if n_wide else self.generator_chunk_frames
        per_w = max(1, wide // max(T, 1))
        per = max(1, self.generator_chunk_frames // max(T, 1))
        for w0 in range(0, B, per_w):
            w1 = min(B, w0 + per_w)
            cur16 = torch.empty(w1 - w0, T, self.upsample_initial_channel, device=x16.device, dtype=self.operand_dtype)
            ops.conv1d(P["g_pre"], x16[w0:w1], gbias=None if gbias is None else gbias[w0:w1], out16=cur16,
                       post_lrelu=LRELU_SLOPE)
            cur16, L = self._generator_stages(P, cur16, T, 0, n_wide)
            for b0 in range(w0, w1, per):
                b1 = min(w1, b0 + per)
                c16, Lc = self._generator_stages(P, cur16[b0 - w0:b1 - w0], L, n_wide, n_st)
                ops.conv_post_tanh(c16, P["g_post"], out[b0:b1].view(b1 - b0, Lc))
        return out

    @staticmethod
    def _rconv(pc, x, **kw):
        """ops.conv1d for a ResBlock conv; a time-pair-folded conv (PackedConv.fold == 2) gets [B, T/2, 2C] views of
        every [B, T, C] tensor (the stage tensors are dense and T = hop-multiple is even)."""
        if pc.fold == 1:
            return ops.conv1d(pc, x, **kw)
        f = pc.fold
        B, T, C = x.shape

        def v(t):
            return t.view(B, T // f, f * t.shape[2]) if torch.is_tensor(t) and t.dim() == 3 else t
        return ops.conv1d(pc, v(x), **{k: v(a) for k, a in kw.items()})

    def _generator_stages(self, P, cur16, L, st0, st1):
        """Upsampling stages st0 .. st1-1 (ups[i] + MRF of ResBlocks) on the activated 16-bit tensor cur16 [B, L, C_in];
        returns (activated 16-bit stage output, its length)."""
        B = cur16.shape[0]
        dev, dt = cur16.device, self.operand_dtype
        up0 = self.upsample_initial_channel
        n_st = len(self.upsample_rates)
        r16 = self.generator_residual in ("fp16", "act16") and dt == torch.float16
        # act16: no un-activated stream at all - the pair's second conv reads its residual from the activated tensor
        # the first conv consumed and inverts the leaky-ReLU (x = min(r, r / slope)): 10 B per element and pair
        ract = r16 and self.generator_residual == "act16"
        for i in range(st0, st1):
            u = self.upsample_rates[i]
            ch = up0 // (2 ** (i + 1))
            Lo = L * u
            # stage residual stream: fp32 (x32) or, by default in fp16 mode, 16-bit (xr16: HBM traffic of a
            # ResBlock pair drops from 16 to 12 bytes per element); xl16 = leaky-ReLU'd copy the convs consume
            x32 = None if r16 else torch.empty(B, Lo, ch, device=dev, dtype=torch.float32)
            xr16 = torch.empty(B, Lo, ch, device=dev, dtype=dt) if r16 and not ract else None
            xl16 = torch.empty(B, Lo, ch, device=dev, dtype=dt)
            # ConvTranspose1d as a 3-tap conv writing the [B, L, u*ch] view of [B, L*u, ch]
            if ract:
                ops.conv1d(P["g_ups"][i], cur16, out16=xl16.view(B, L, u * ch), post_lrelu=LRELU_SLOPE)
            elif r16:
                ops.conv1d(P["g_ups"][i], cur16, out16r=xr16.view(B, L, u * ch), out16=xl16.view(B, L, u * ch),
                           post_lrelu=LRELU_SLOPE)
            else:
                ops.conv1d(P["g_ups"][i], cur16, out32=x32.view(B, L, u * ch), out16=xl16.view(B, L, u * ch),
                           post_lrelu=LRELU_SLOPE)
            c32 = None if r16 else torch.empty_like(x32)
            cr16 = torch.empty_like(xl16) if r16 and not ract else None
            cl16 = torch.empty_like(xl16)
            t16 = torch.empty_like(xl16)
            # MRF accumulator xs (models_ae.py:335-341): 16-bit in the 16-bit-residual mode (16 -> 8 B per element
            # of MRF traffic), else fp32
            mrf16 = r16 and self.generator_mrf == "fp16"
            mrf = torch.empty(B, Lo, ch, device=dev, dtype=dt if mrf16 else torch.float32)
            nxt16 = torch.empty_like(xl16)
            nk = len(P["g_rb"][i])
            next_slope = LRELU_SLOPE if i < n_st - 1 else 0.01      # F.leaky_relu default (models_ae.py:342)
            for j, pairs in enumerate(P["g_rb"][i]):
                src32, srcr, src16 = x32, xr16, xl16
                for m, (c1, c2) in enumerate(pairs):
                    last = m == len(pairs) - 1
                    kw = {}
                    if last:
                        # MRF: xs = sum_j resblock_j(x); x = xs / num_kernels (models_ae.py:335-341)
                        if j > 0:
                            kw["res2"] = mrf
                        if j < nk - 1:
                            kw.update(out16r=mrf) if mrf16 else kw.update(out32=mrf)
                        else:
                            kw.update(scale=1.0 / nk, out16=nxt16, post_lrelu=next_slope)
                    elif ract:
                        kw.update(out16=cl16, post_lrelu=LRELU_SLOPE)
                    elif r16:
                        kw.update(out16r=cr16, out16=cl16, post_lrelu=LRELU_SLOPE)
                    else:
                        kw.update(out32=c32, out16=cl16, post_lrelu=LRELU_SLOPE)
                    if ract:
                        kw.update(res16=src16, res16_unlrelu=1.0 / LRELU_SLOPE)
                    elif r16:
                        kw["res16"] = srcr
                    else:
                        kw["res1"] = src32
                    if c2 is not None and ract and mrf16 and self.generator_pair_fused and ops.pair_fusable(c1, c2, ch):
                        # one fused kernel per pair (resblock_pair.cu); output must not alias the input: cl16 / t16 ping-pong
                        dst16 = t16 if src16 is cl16 else cl16
                        if not last:
                            ops.resblock_pair(c1, c2, src16, dst16, post_slope=LRELU_SLOPE)
                        elif j < nk - 1:
                            ops.resblock_pair(c1, c2, src16, mrf, mrf=mrf if j > 0 else None, out_lrelu=False)
                        else:
                            ops.resblock_pair(c1, c2, src16, nxt16, mrf=mrf if j > 0 else None, post_slope=next_slope,
                                              scale=1.0 / nk)
                        src32, srcr, src16 = c32, cr16, dst16
                    elif c2 is not None:      # ResBlock1 (modules.py:210-223)
                        self._rconv(c1, src16, out16=t16, post_lrelu=LRELU_SLOPE)
                        self._rconv(c2, t16, **kw)
                        src32, srcr, src16 = c32, cr16, cl16
                    else:                   # ResBlock2 (modules.py:243-252): ping-pong the 16-bit stream,
                        dst16 = t16 if src16 is cl16 else cl16      # the conv reads halo rows of its input
                        if "out16" in kw and not last:
                            kw["out16"] = dst16
                        self._rconv(c1, src16, **kw)
                        src32, srcr, src16 = c32, cr16, dst16
            cur16, L = nxt16, Lo
        return cur16, L

    # ------------------------------------------------------------------ training-side piece that shares the kernels (8f-4)
    def _pack_posterior(self):
        """PosteriorEncoder = pre (1x1) + WN x 16 + proj (models_ae.py:256-291), packed on first use only: infer never
        touches spec_encoder, so its 7.2 M parameters are not packed by `_pack`."""
        P = self._packed or self._pack()
        if "pe_pre" in P:
            return P
        sd = {k: v.detach().float().cpu() for k, v in self.state_dict().items() if k.startswith("spec_encoder.")}
        dev, dt = self.device, self.operand_dtype
        W = lambda p: packer.fold_weight_norm(sd, p)
        P["pe_pre"] = packer.pack_conv(W("spec_encoder.pre"), sd["spec_encoder.pre.bias"], dtype=dt).to(dev)
        P["pe_proj"] = packer.pack_conv(W("spec_encoder.proj"), sd["spec_encoder.proj.bias"], dtype=dt).to(dev)
        P["pe_wn"] = self._pack_wn_stack(sd, "spec_encoder.enc", dev, dt)
        for pc in [P["pe_pre"], P["pe_proj"], *P["pe_wn"]["in"], *P["pe_wn"]["rs"], P["pe_wn"]["skip_cat"]]:
            pc.tag = "posterior"
        self._build_stage_tables(P)
        return P

    @torch.no_grad()
    def posterior_encode(self, y, y_lengths, g=None, *, noise=None):
        """`self.spec_encoder(y, y_lengths, g=g)` of the reference (PosteriorEncoder.forward, models_ae.py:284-291):
        y [B, spec_channels, T] fp32 linear spectrogram -> (z, m, logs, y_mask, x) in the reference's [B,C,T] layout.
        `noise` [B, inter, T] replaces torch.randn_like (parity runs).  Inference-only restatement of the training-side
        encoder on the same kernels as the SpecDecoder."""
        _require_cuda(y)
        dev, dt, f32 = self.device, self.operand_dtype, torch.float32
        with _device_guard(dev):
            P = self._pack_posterior()
            B, Cs, T = y.shape
            H, Ci = self.hidden_channels, self.inter_channels
            lens = y_lengths.to(device=dev, dtype=torch.int32).contiguous()
            y_mask = (torch.arange(T, device=dev).unsqueeze(0) < lens.unsqueeze(1)).unsqueeze(1).to(y.dtype)
            ld = (Cs + 63) // 64 * 64
            y16 = torch.empty(B, T, ld, device=dev, dtype=dt)
            ops.bct_to_btc(y.contiguous().float(), out16=y16[:, :, :Cs])
            hx32 = torch.empty(B, T, H, device=dev, dtype=f32)
            hx16 = torch.empty(B, T, H, device=dev, dtype=dt)
            wo16 = torch.empty(B, T, H, device=dev, dtype=dt)
            ops.conv1d(P["pe_pre"], y16[:, :, :Cs], lens=lens, out32=hx32, out16=hx16)
            gcond = None
            if g is not None:
                if not self.gin_channels:
                    raise ValueError("model was built with gin_channels=0")
                g_btc = g.to(device=dev, dtype=f32).reshape(B, 1, self.gin_channels).contiguous()
                gc = torch.empty(B, 1, P["pe_wn"]["cond"].cout, device=dev, dtype=f32)
                ops.conv1d(P["pe_wn"]["cond"], g_btc, out32=gc)
                gcond = gc.view(B, -1)
            self._wn(P["pe_wn"], hx32, hx16, lens, wo16, gcond)
            stats = torch.empty(B, T, 2 * Ci, device=dev, dtype=f32)
            ops.conv1d(P["pe_proj"], wo16, lens=lens, out32=stats)
            if noise is None:
                noise = torch.randn(B, Ci, T, device=dev, dtype=f32)
            z32 = torch.empty(B, T, Ci, device=dev, dtype=f32)
            ops.sample_prior(stats, noise.to(device=dev, dtype=f32).contiguous(), 1.0, zp32=z32, lens=lens)
        bct = lambda t: t.transpose(1, 2)
        return bct(z32), bct(stats[:, :, :Ci]), bct(stats[:, :, Ci:]), y_mask, bct(wo16.float())

    @torch.no_grad()
    def flow_forward(self, z, z_lengths, g=None):
        """`self.spec_flow(z, y_mask, g=g)` of the reference in the FORWARD direction (ResidualCouplingBlock.forward with
        reverse=False, models_ae.py:246-253; mean-only coupling layers, modules.py:329-349): the training-side map
        z -> z_p.  z [B, inter, T] fp32 -> z_p [B, inter, T].  It is the exact inverse of the reverse flow of `infer`, so it
        runs the same packed layers (Flips folded into the weights) in the opposite order with the coupling mean ADDED
        (SURVEY.md 8f-4)."""
        _require_cuda(z)
        dev, dt, f32 = self.device, self.operand_dtype, torch.float32
        with _device_guard(dev):
            P = self._packed or self._pack()
            B, Ci, T = z.shape
            H, half = self.hidden_channels, Ci // 2
            lens = z_lengths.to(device=dev, dtype=torch.int32).contiguous()
            z32 = torch.empty(B, T, Ci, device=dev, dtype=f32)
            z16 = torch.empty(B, T, Ci, device=dev, dtype=dt)
            ops.bct_to_btc(z.contiguous().float(), out32=z32, out16=z16, lens=lens)
            g_btc = None
            if g is not None:
                if not self.gin_channels:
                    raise ValueError("model was built with gin_channels=0")
                g_btc = g.to(device=dev, dtype=f32).reshape(B, 1, self.gin_channels).contiguous()
            hx32 = torch.empty(B, T, H, device=dev, dtype=f32)
            hx16 = torch.empty(B, T, H, device=dev, dtype=dt)
            wo16 = torch.empty(B, T, H, device=dev, dtype=dt)
            for fl in reversed(P["flow"]):
                io, oo = fl["in_off"], fl["out_off"]
                gcond = None
                if g_btc is not None:
                    gc = torch.empty(B, 1, fl["wn"]["cond"].cout, device=dev, dtype=f32)
                    ops.conv1d(fl["wn"]["cond"], g_btc, out32=gc)
                    gcond = gc.view(B, -1)
                ops.conv1d(fl["pre"], z16[:, :, io:io + half], lens=lens, out32=hx32, out16=hx16)
                self._wn(fl["wn"], hx32, hx16, lens, wo16, gcond)
                ops.conv1d(fl["post"], wo16, alpha=1.0, res1=z32[:, :, oo:oo + half], lens=lens,
                           out32=z32[:, :, oo:oo + half], out16=z16[:, :, oo:oo + half])
        return z32.transpose(1, 2)

    def _graphed(self, fn, args, shape_key):
        """Run fn(*args) through a cached CUDA graph.  args: tensors or None; the graph owns static copies of them and
        of its outputs, results are cloned so that a later replay cannot overwrite what the caller holds."""
        key = shape_key + tuple(None if a is None else (tuple(a.shape), str(a.dtype)) for a in args)
        ent = self._graphs.get(key)
        if ent is None:
            static = tuple(None if a is None else a.clone() for a in args)
            fn(*static)                               # eager warm-up: lazy one-time host setup must not be captured
            torch.cuda.synchronize()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                outs, faux = fn(*static)
            if len(self._graphs) >= self.cuda_graph_cache:
                self._graphs.pop(next(iter(self._graphs)))
            ent = self._graphs[key] = (gr, static, outs, faux)
        gr, static, outs, faux = ent
        for s_, a in zip(static, args):
            if a is not None:
                s_.copy_(a)
        gr.replay()
        return tuple(None if o_ is None else o_.clone() for o_ in outs), {k: v.clone() for k, v in faux.items()}

    def _text_post_net(self, h_lr, mlen, f0_score):
        """TextPostNet (ext_modules/text_post_net.py:64-94) on length-regulated frames h_lr [B,T,H] (fp32, rows >= mlen[b]
        are ignored).  Returns (stats [B,T,2*inter] = m_p | logs_p, log_f0 [B,T], mel [B,T,80], coarse f0 [B,T]).
        Pure device work (CUDA-graph capturable).  Row t of every utterance takes position embedding t."""
        P = self._packed
        B, T, H = h_lr.shape
        dev, dt, f32 = h_lr.device, self.operand_dtype, torch.float32
        faux = {}
        # 4. TextPostNet (ext_modules/text_post_net.py:64-94)
        xe32 = torch.empty(B, T, H, device=dev, dtype=f32)
        xe16 = torch.empty(B, T, H, device=dev, dtype=dt)
        ops.add_pos_mask(h_lr, P["pos"], mlen, out32=xe32, out16=xe16)
        # F0 predictor: tensor cores with fp16 (hi, lo) split operands (default), or fp32 SIMT (f0_precision="fp32")
        f0h = torch.empty_like(xe32)
        if P["f0_mode"] == "split":
            f0s = torch.empty(B, T, 3 * H, device=dev, dtype=torch.float16)
            ops.add_pos_mask(h_lr, P["pos"], mlen, out32=f0h, out16=f0s, split16=True)
        else:
            f0s = None
            ops.add_pos_mask(h_lr, P["pos"], mlen, out32=f0h)
        self._encoder(P["f0_enc"], f0h, f0s, mlen, P["f0_mode"])
        log_f0 = torch.empty(B, T, 1, device=dev, dtype=f32)
        ops.conv1d(P["f0_out"], f0h, lens=mlen, out32=log_f0)
        log_f0 = log_f0.view(B, T)
        coarse = torch.empty(B, T, device=dev, dtype=torch.int64)
        f0e16 = torch.empty(B, T, H, device=dev, dtype=dt)
        cin = None if f0_score is None else f0_score.to(device=dev, dtype=torch.int64).contiguous()
        ops.f0_embed(log_f0, cin, P["f0_emb"], mlen, coarse, out16=f0e16)
        # MelPredictNetwork (mel_predict_network.py:37-45)
        m32 = torch.empty(B, T, H, device=dev, dtype=f32)
        m16 = torch.empty(B, T, H, device=dev, dtype=dt)
        ops.conv1d(P["mel_cond"], f0e16, res1=xe32, lens=mlen, out32=m32, out16=m16)
        self._encoder(P["mel_enc"], m32, m16, mlen, "tc")
        mel32 = torch.empty(B, T, 80, device=dev, dtype=f32)
        ops.conv1d(P["mel_out"], m16, lens=mlen, out32=mel32)
        # FramePriorNetwork (frame_prior_network.py:31-41)
        fp_a = torch.empty(B, T, 80, device=dev, dtype=dt)
        ops.conv1d(P["fp_cond"], f0e16, res1=mel32, lens=mlen, out16=fp_a)
        src = fp_a
        for pc in P["fp"]:
            dst = torch.empty(B, T, H, device=dev, dtype=dt)
            ops.conv1d(pc, src, act=ACT_GELU, lens=mlen, out16=dst)
            src = dst
        stats = torch.empty(B, T, 2 * self.inter_channels, device=dev, dtype=f32)
        ops.conv1d(P["proj"], src, lens=mlen, out32=stats)
        Ci = self.inter_channels

        faux["f0_coarse"] = coarse
        return stats, log_f0, mel32, faux

    def _decode(self, stats, mlen, noise, g, noise_scale, max_len):
        """Prior sample -> reverse flow -> [SpecDecoder + spec_proj] -> Generator (models_ae.py:643-650) on frame-level
        statistics stats [B,T,2*inter].  Convolutional only: no limit on T.  Returns (o, z32, zp32, spec32 or None)."""
        P = self._packed
        B, T, _ = stats.shape
        H, Ci = self.hidden_channels, self.inter_channels
        dev, dt, f32 = stats.device, self.operand_dtype, torch.float32
        # 5. z_p = m_p + noise * exp(logs_p) * noise_scale (models_ae.py:643)
        if noise is None:
            noise = torch.randn(B, Ci, T, device=dev, dtype=f32)
        noise = noise.to(device=dev, dtype=f32).contiguous()
        zp32 = torch.empty(B, T, Ci, device=dev, dtype=f32)
        z16 = torch.empty(B, T, Ci, device=dev, dtype=dt)
        ops.sample_prior(stats, noise, float(noise_scale), zp32=zp32, zp16=z16)
        z32 = zp32.clone()

        # speaker conditioning (optional; WN.cond_layer modules.py:152-159, Generator.cond models_ae.py:329-330)
        g_btc = None
        if g is not None:
            if not self.gin_channels:
                raise ValueError("model was built with gin_channels=0")
            g_btc = g.to(device=dev, dtype=f32).reshape(B, 1, self.gin_channels).contiguous()

        def cond_of(st):
            if g_btc is None:
                return None
            out = torch.empty(B, 1, st["cond"].cout, device=dev, dtype=f32)
            ops.conv1d(st["cond"], g_btc, out32=out)
            return out.view(B, -1)

        # 6. reverse flow (models_ae.py:246-253; modules.py:329-349), Flips folded into the weights
        half = Ci // 2
        hx32 = torch.empty(B, T, H, device=dev, dtype=f32)
        hx16 = torch.empty(B, T, H, device=dev, dtype=dt)
        wo16 = torch.empty(B, T, H, device=dev, dtype=dt)
        for fl in P["flow"]:
            io, oo = fl["in_off"], fl["out_off"]
            ops.conv1d(fl["pre"], z16[:, :, io:io + half], lens=mlen, out32=hx32, out16=hx16)
            self._wn(fl["wn"], hx32, hx16, mlen, wo16, cond_of(fl["wn"]))
            ops.conv1d(fl["post"], wo16, alpha=-1.0, res1=z32[:, :, oo:oo + half], lens=mlen,
                       out32=z32[:, :, oo:oo + half], out16=z16[:, :, oo:oo + half])

        # 7. SpecDecoder + spec_proj (ae only; spec_decoder.py:41-46, models_ae.py:647-648)
        gen_g = None
        if g_btc is not None:
            gg = torch.empty(B, 1, self.upsample_initial_channel, device=dev, dtype=f32)
            ops.conv1d(P["g_cond"], g_btc, out32=gg)
            gen_g = gg.view(B, -1)
        Tg = T if max_len is None else min(T, int(max_len))
        if self.variant == "ae":
            ops.conv1d(P["sd_pre"], z16, lens=mlen, out32=hx32, out16=hx16)
            self._wn(P["sd_wn"], hx32, hx16, mlen, wo16, cond_of(P["sd_wn"]))
            gin16 = torch.empty(B, T, Ci, device=dev, dtype=dt)
            spec32, sc = None, self.spec_channels
            if self.fuse_spec_proj:
                ops.conv1d(P["sd_fused"], wo16, lens=mlen, out16=gin16)
            else:
                sc_ld = 576
                spec32 = torch.empty(B, T, sc_ld, device=dev, dtype=f32)
                spec16 = torch.empty(B, T, sc_ld, device=dev, dtype=dt)
                ops.conv1d(P["sd_proj"], wo16, lens=mlen, out32=spec32[:, :, :sc], out16=spec16[:, :, :sc])
                ops.conv1d(P["spec_proj"], spec16[:, :, :sc], lens=mlen, out16=gin16)
            gen_in = gin16
        else:
            gen_in = z16          # z is already masked ((x1 - m) * mask; x0 masked upstream)
        if Tg != T:
            gen_in = gen_in[:, :Tg].contiguous()

        # 8. Generator
        lens_host = mlen.tolist() if (self.generator_bucket and B > 1 and not torch.cuda.is_current_stream_capturing()) else None
        if lens_host is not None and Tg != T:
            lens_host = [min(n, Tg) for n in lens_host]
        o = self._generator(gen_in, gen_g, lens_host)

        spec = spec32[:, :, :sc] if self.variant == "ae" and spec32 is not None else None
        return o, z32, zp32, spec

    # ------------------------------------------------------------------ long utterances
    @torch.no_grad()
    def infer_long(self, x, x_lengths, note_pitch_id, note_duration, phone_shengdiao, ph_dur=None, *, segment_frames=None,
                   overlap=0, noise_scale=1, max_len=None, noise=None, g=None):
        """`infer` for utterances longer than the position table (2000 frames = 23.2 s in the reference, which raises
        there: text_post_net.py:26,69-72).  Same inputs and result tuple as `infer`.

        Phoneme level (text encoder, durations, length regulator) runs over the whole utterance.  The frame-level
        transformers (TextPostNet) run per segment - `segment_frames` core frames plus `overlap` frames of context on
        each side (segments.plan_segments), all segments of all utterances as ONE ragged batch - and the core frames'
        statistics are written into full-length tensors.  Prior sample, reverse flow, SpecDecoder and Generator are
        convolutional and run over the full length, so there is no seam in the waveform; the Generator is cut in time
        (`generator_time_chunk_frames`, exact halos) only to bound memory.  With one segment this is `infer` bit for bit."""
        from . import segments as seg
        _require_cuda(x)
        dev = self.device
        with _device_guard(dev):
            y_dtype = x.dtype
            B = x.shape[0]
            h, ends, mel_len, mlen, T, _ = self._phoneme_level(x, x_lengths, note_pitch_id, note_duration, phone_shengdiao,
                                                               ph_dur)
            P = self._packed
            f32, H, Ci = torch.float32, self.hidden_channels, self.inter_channels
            n_pos = P["pos"].shape[0]
            if segment_frames is None:
                segment_frames = n_pos - 2 * overlap
            if segment_frames <= 0 or segment_frames + 2 * overlap > n_pos:
                raise ValueError(f"segment_frames={segment_frames} + 2 * overlap={overlap} exceeds the {n_pos}-row position table")
            h_lr = torch.empty(B, T, H, device=dev, dtype=f32)
            lr_index = torch.empty(B, T, device=dev, dtype=torch.int32)
            ops.lr_gather(h, ends, mlen, h_lr, lr_index)
            lens = mel_len.tolist()
            plan = [(b, *s_) for b in range(B) for s_ in seg.plan_segments(int(lens[b]), segment_frames, overlap)]
            Ts = max(c[4] - c[3] for c in plan)
            seg_h = torch.zeros(len(plan), Ts, H, device=dev, dtype=f32)
            for i, (b, c0, c1, x0, x1) in enumerate(plan):
                seg_h[i, :x1 - x0] = h_lr[b, x0:x1]
            seg_len = torch.tensor([c[4] - c[3] for c in plan], device=dev, dtype=torch.int32)
            s_stats, s_lf0, s_mel, faux = self._text_post_net(seg_h, seg_len, None)
            stats = torch.zeros(B, T, 2 * Ci, device=dev, dtype=f32)
            log_f0 = torch.zeros(B, T, device=dev, dtype=f32)
            mel32 = torch.zeros(B, T, 80, device=dev, dtype=f32)
            coarse = torch.zeros(B, T, device=dev, dtype=torch.int64)
            for i, (b, c0, c1, x0, x1) in enumerate(plan):
                a, e = c0 - x0, c1 - x0
                stats[b, c0:c1], log_f0[b, c0:c1], mel32[b, c0:c1] = s_stats[i, a:e], s_lf0[i, a:e], s_mel[i, a:e]
                coarse[b, c0:c1] = faux["f0_coarse"][i, a:e]
            o, z32, zp32, spec = self._decode(stats, mlen, noise, g, noise_scale, max_len)
            self.last_aux.update(lr_index=lr_index, f0_coarse=coarse, segments=plan)
        y_mask = (torch.arange(T, device=dev).unsqueeze(0) < mel_len.unsqueeze(1)).unsqueeze(1).to(y_dtype)
        bct = lambda t: t.transpose(1, 2)
        m_p, logs_p = bct(stats[:, :, :Ci]), bct(stats[:, :, Ci:])
        if self.variant == "ae":
            return o, y_mask, (bct(z32), bct(zp32), m_p, logs_p), log_f0, None if spec is None else bct(spec), bct(mel32)
        return o, y_mask, (bct(z32), bct(zp32), m_p, logs_p), log_f0, bct(mel32)

    # ------------------------------------------------------------------ the public entry point
    @torch.no_grad()
    def infer(self, x, x_lengths, note_pitch_id, note_duration, phone_shengdiao, ph_dur=None, lf0=None, sid=None,
              noise_scale=1, length_scale=1, noise_scale_w=1., max_len=None, *, noise=None, g=None, f0_score=None,
              check_range=False):
        """Same signature and result tuple as the reference (models_ae.py:608-651 / models.py:622-673).
        Extensions (keyword-only, not in the reference): `noise` [B,C,T] replaces torch.randn_like for
        parity runs; `g` [B,gin,1] speaker conditioning (the reference hard-codes g=None, :634);
        `f0_score` [B,T] int64 overrides the predicted coarse f0 (TextPostNet.forward's own argument);
        `check_range`: read the library's fp16 overflow canary after the call (one extra sync) into
        last_aux["saturation_events"] (0 = no 16-bit tensor of this forward clipped).

        The kernels launch on the device of the model's parameters (SURVEY 8b "device conventions"), whatever the
        current device is; inputs on another device are an error, id / length tensors are normalised to int64 like
        nn.Embedding / sequence_mask accept them."""
        _require_cuda(x)
        dev = self.device
        for name, t in (("x", x), ("x_lengths", x_lengths), ("note_pitch_id", note_pitch_id),
                        ("note_duration", note_duration), ("phone_shengdiao", phone_shengdiao)):
            if t.device != dev:
                raise RuntimeError(f"SynthesizerTrn.infer: {name} is on {t.device}, the model on {dev}")
        with _device_guard(dev):
            if check_range:
                _lib.check(_lib.lib().b200svs_saturation_events(None, 1, ops._stream()), "saturation_events")
            res = self._infer(x, x_lengths, note_pitch_id, note_duration, phone_shengdiao, ph_dur, lf0, noise_scale, max_len,
                              noise, g, f0_score)
            if check_range:
                import ctypes
                n = ctypes.c_uint32(0)
                _lib.check(_lib.lib().b200svs_saturation_events(ctypes.byref(n), 0, ops._stream()), "saturation_events")
                self.last_aux["saturation_events"] = int(n.value)
        return res

    def _phoneme_level(self, x, x_lengths, note_pitch_id, note_duration, phone_shengdiao, ph_dur, f0_score=None):
        """TextEncoder -> DurationPredictor -> floor -> LengthRegulator scan (models_ae.py:612-627).  Returns
        (h [B,P,H] fp32, ends, mel_len int64, mlen int32, T, f0_score) and fills last_aux; ONE host sync (T)."""
        P = self._packed or self._pack()
        dev, dt = self.device, self.operand_dtype
        f32 = torch.float32
        B, Pn = x.shape
        H = self.hidden_channels
        i64 = lambda t: t.to(device=dev, dtype=torch.int64).contiguous()
        x, note_pitch_id, phone_shengdiao, x_lengths = i64(x), i64(note_pitch_id), i64(phone_shengdiao), i64(x_lengths)
        note_duration = note_duration.contiguous().float()
        if f0_score is not None:              # user-supplied buckets: validated here (nn.Embedding would raise)
            f0_score = i64(f0_score)
            lo, hi = int(f0_score.min()), int(f0_score.max())
            if lo < 0 or hi >= P["f0_emb"].shape[0]:
                raise IndexError(f"f0_score out of range [0, {P['f0_emb'].shape[0]}): min {lo} max {hi}")
        aux = self.last_aux = {}
        status = torch.zeros(2, device=dev, dtype=torch.int32)     # [0] max frames T (lr_scan), [1] bad-id flag (text_embed)

        # 1. TextEncoder (models_ae.py:203-220), fp32
        plen = torch.empty(B, device=dev, dtype=torch.int32)
        h = torch.empty(B, Pn, H, device=dev, dtype=f32)
        ops.text_embed(x, note_pitch_id, phone_shengdiao, note_duration, *P["emb"], *P["emb_dur"], x_lengths, plen, h,
                       status=status)
        hs = None
        if P["text_mode"] == "split":     # fp16 (hi, lo) split copy for the tensor-core path: (h + 0) * mask == h
            hs = torch.empty(B, Pn, 3 * H, device=dev, dtype=torch.float16)
            if P["zero_pos"].shape[0] < Pn:
                P["zero_pos"] = torch.zeros(Pn, H, device=dev, dtype=f32)
            ops.add_pos_mask(h, P["zero_pos"], plen, out16=hs, split16=True)
        self._encoder(P["text_enc"], h, hs, plen, P["text_mode"])

        # 2. durations (models_ae.py:622-627)
        if ph_dur is None:
            d32 = torch.empty(B, Pn, 256, device=dev, dtype=f32)
            dn = torch.empty(B, Pn, 256, device=dev, dtype=f32)
            src = h
            for pc, gamma, beta in P["dp"]:
                ops.conv1d(pc, src, act=ACT_RELU, out32=d32)
                ops.layernorm(d32, gamma, beta, plen, out32=dn)
                src = dn
            pred = torch.empty(B, Pn, 1, device=dev, dtype=f32)
            ops.conv1d(P["dp_proj"], dn, act=ACT_SIGMOID, lens=plen, out32=pred)
            durations = torch.empty(B, Pn, device=dev, dtype=torch.int64)
            ops.duration_quantize(pred, note_duration, durations)
            aux["dur_pred"] = pred.view(B, Pn)
        else:
            durations = ph_dur.to(device=dev, dtype=torch.int64).contiguous()
        aux["durations"] = durations

        # 3. LengthRegulator (models_ae.py:689-721) - one host sync for T instead of B*P .item() calls
        ends = torch.empty(B, Pn, device=dev, dtype=torch.int32)
        mel_len = torch.empty(B, device=dev, dtype=torch.int64)
        mlen = torch.empty(B, device=dev, dtype=torch.int32)
        ops.lr_scan(durations, ends, mel_len, mlen, status=status)
        T, bad_ids = status.tolist()           # the one host sync of infer (the reference does B*P .item() calls)
        if bad_ids:
            raise IndexError("SynthesizerTrn.infer: a phoneme / note-pitch / tone id is outside its embedding table "
                             "(index out of range in self)")
        if T <= 0:
            raise RuntimeError("SynthesizerTrn.infer: every utterance has zero frames")
        return h, ends, mel_len, mlen, T, f0_score

    def _infer(self, x, x_lengths, note_pitch_id, note_duration, phone_shengdiao, ph_dur, lf0, noise_scale, max_len,
               noise, g, f0_score):
        y_dtype = x.dtype                     # y_mask takes x's dtype (commons.sequence_mask(...).to(x.dtype), models_ae.py:632)
        B, Pn = x.shape
        h, ends, mel_len, mlen, T, f0_score = self._phoneme_level(x, x_lengths, note_pitch_id, note_duration,
                                                                   phone_shengdiao, ph_dur, f0_score)
        P = self._packed
        dev, dt, f32 = self.device, self.operand_dtype, torch.float32
        H = self.hidden_channels
        aux = self.last_aux
        n_pos = P["pos"].shape[0]
        if T > n_pos:
            # the reference fails here too: position table has 2000 rows (text_post_net.py:26,69-72)
            raise RuntimeError(f"The size of tensor a ({T}) must match the size of tensor b ({n_pos}) at "
                               f"non-singleton dimension 1 (frames exceed the position table)")
        y_mask = (torch.arange(T, device=dev).unsqueeze(0) < mel_len.unsqueeze(1)).unsqueeze(1).to(y_dtype)
        Ci = self.inter_channels

        def frame_level(h, ends, mlen, noise, g, f0_score):
            """Everything after the host learns T: ~95 % of the launches.  Pure device work on its arguments (no host
            sync), so it can be captured in a CUDA graph."""
            h_lr = torch.empty(B, T, H, device=dev, dtype=f32)
            lr_index = torch.empty(B, T, device=dev, dtype=torch.int32)
            ops.lr_gather(h, ends, mlen, h_lr, lr_index)
            cin = None if f0_score is None else f0_score.to(device=dev, dtype=torch.int64).contiguous()
            stats, log_f0, mel32, faux = self._text_post_net(h_lr, mlen, cin)
            faux["lr_index"] = lr_index
            o, z32, zp32, spec = self._decode(stats, mlen, noise, g, noise_scale, max_len)
            return (o, z32, zp32, stats, log_f0, spec, mel32), faux

        args = (h, ends, mlen, noise, g, f0_score)
        if self.cuda_graphs and B * T <= self.cuda_graph_max_frames and _lib.PROFILE is None:
            outs, faux = self._graphed(frame_level, args, (B, Pn, T, float(noise_scale), max_len, str(x.dtype)))
        else:
            outs, faux = frame_level(*args)
        aux.update(faux)
        o, z32, zp32, stats, log_f0, spec, mel32 = outs
        bct = lambda t: t.transpose(1, 2)
        m_p, logs_p = bct(stats[:, :, :Ci]), bct(stats[:, :, Ci:])
        if lf0 is not None:
            log_f0 = lf0
        if self.variant == "ae":
            return o, y_mask, (bct(z32), bct(zp32), m_p, logs_p), log_f0, None if spec is None else bct(spec), bct(mel32)
        return o, y_mask, (bct(z32), bct(zp32), m_p, logs_p), log_f0, bct(mel32)
