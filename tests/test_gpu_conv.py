"""-m gpu: the conv kernels through the C ABI vs the documented semantics (tests/emu_ops.py, plain torch
fp32 on CPU) on identical seeded inputs.  Tolerances: operands are rounded to fp16 on both sides, so the
only difference is fp32 accumulation order -> 2e-3 relative to the output scale."""
import pytest
import torch

import emu_ops
from common import max_abs
from music_generation_b200 import ops, packer
from music_generation_b200._lib import ACT_GELU, ACT_NONE, ACT_RELU, ACT_SIGMOID

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rand(*shape, seed=0, scale=1.0):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale


def _to_dev(x, dev):
    """.to() makes non-dense views contiguous; keep the padded row stride instead."""
    if x.is_contiguous() or dev == "cpu":
        return x.to(dev)
    ld = x.stride(1)
    base = torch.zeros(x.shape[0], x.shape[1], ld, dtype=x.dtype, device=dev)
    base[:, :, :x.shape[2]] = x.to(dev)
    return base[:, :, :x.shape[2]]


def _run_both(pc_cpu, x_cpu, *, out_cols, want32=True, want16=False, res1=None, res2=None, lens=None, gbias=None,
              out_ld=None, **kw):
    """Runs the real kernel on the GPU and the emulation on the CPU; returns (real32, emu32, real16, emu16)."""
    import copy
    B, T, _ = x_cpu.shape
    ld = out_ld or out_cols
    pc_gpu = copy.copy(pc_cpu)
    pc_gpu = pc_gpu.to(DEV)
    outs = {}
    for side, dev, fn, pc in (("emu", "cpu", emu_ops.conv1d, pc_cpu), ("real", DEV, ops.conv1d, pc_gpu)):
        o32 = torch.full((B, T, ld), 7.0, device=dev) if want32 else None
        o16 = torch.full((B, T, ld), 7.0, device=dev, dtype=x_cpu.dtype if x_cpu.dtype != torch.float32 else torch.float16) if want16 else None
        args = dict(kw)
        if res1 is not None:
            args["res1"] = res1.to(dev)
        if res2 is not None:
            args["res2"] = res2.to(dev)
        if lens is not None:
            args["lens"] = lens.to(dev)
        if gbias is not None:
            args["gbias"] = gbias.to(dev)
        fn(pc, _to_dev(x_cpu, dev), out32=None if o32 is None else o32[:, :, :out_cols],
           out16=None if o16 is None else o16[:, :, :out_cols], **args)
        if dev != "cpu":
            torch.cuda.synchronize()
        outs[side] = (None if o32 is None else o32.cpu(), None if o16 is None else o16.float().cpu())
    return outs["real"][0], outs["emu"][0], outs["real"][1], outs["emu"][1]


TC_CASES = [
    # name, B, T, Cin, Cout, k, dil
    ("1x1_192", 2, 300, 192, 192, 1, 1),
    ("k3_ffn1", 1, 257, 192, 768, 3, 1),
    ("k3_ffn2", 1, 130, 768, 192, 3, 1),
    ("k7_pre", 2, 200, 192, 512, 7, 1),
    ("k11_d5_256", 1, 700, 256, 256, 11, 5),
    ("k7_d3_128", 2, 513, 128, 128, 7, 3),
    ("k3_d1_64", 1, 1000, 64, 64, 3, 1),
    ("k11_d1_32", 1, 900, 32, 32, 11, 1),
    ("cin80", 1, 140, 80, 192, 1, 1),
    ("cin513", 1, 140, 513, 192, 1, 1),
    ("cout513", 1, 140, 192, 513, 1, 1),
    ("cout96", 2, 140, 192, 96, 1, 1),
    ("tiny_T", 3, 5, 192, 192, 5, 1),
]


@pytest.mark.parametrize("name,B,T,cin,cout,k,dil", TC_CASES, ids=[c[0] for c in TC_CASES])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16], ids=["fp16", "bf16"])
def test_conv1d_tc_plain(name, B, T, cin, cout, k, dil, dtype):
    w = _rand(cout, cin, k, seed=1, scale=(cin * k) ** -0.5)
    b = _rand(cout, seed=2, scale=0.1)
    x = _rand(B, T, cin, seed=3).to(dtype)
    if cin % 8:     # TMA needs 16-B aligned rows: odd channel counts live in a padded buffer (spec16, ld 576)
        xp = torch.zeros(B, T, (cin + 7) // 8 * 8, dtype=dtype)
        xp[:, :, :cin] = x
        x = xp[:, :, :cin]
    pc = packer.pack_conv(w, b, dil=dil, dtype=dtype)
    ld = (cout + 7) // 8 * 8
    real, emu, _, _ = _run_both(pc, x, out_cols=cout, out_ld=ld)
    assert max_abs(emu[:, :, :cout], real[:, :, :cout]) < 3e-3
    if ld != cout:   # columns >= n_valid untouched
        # TMA stores clip at 16-byte granularity: up to 3 fp32 pad columns may be written (b200svs.h)
        assert float((real[:, :, (cout + 3) // 4 * 4:] - 7.0).abs().max()) == 0.0


def test_conv1d_tc_full_epilogue():
    """bias + residuals + alpha/scale + mask + both outputs + post leaky-relu on the 16-bit copy."""
    B, T, c = 3, 400, 128
    w, b = _rand(c, c, 7, seed=4, scale=(7 * c) ** -0.5), _rand(c, seed=5, scale=0.1)
    x = _rand(B, T, c, seed=6).half()
    pc = packer.pack_conv(w, b, dil=3)
    res1, res2 = _rand(B, T, c, seed=7), _rand(B, T, c, seed=8)
    lens = torch.tensor([400, 133, 0], dtype=torch.int32)
    r32, e32, r16, e16 = _run_both(pc, x, out_cols=c, want16=True, res1=res1, res2=res2, lens=lens, alpha=-1.0,
                                   scale=1.0 / 3, post_lrelu=0.1)
    assert max_abs(e32, r32) < 3e-3 and max_abs(e16, r16) < 6e-3
    assert float(r32[1, 133:].abs().max()) == 0.0 and float(r32[2].abs().max()) == 0.0


@pytest.mark.parametrize("act", [ACT_RELU, ACT_GELU, ACT_SIGMOID])
def test_conv1d_tc_activations(act):
    w, b = _rand(192, 192, 1, seed=9, scale=192 ** -0.5), _rand(192, seed=10, scale=0.1)
    x = _rand(2, 150, 192, seed=11).half()
    pc = packer.pack_conv(w, b)
    real, emu, _, _ = _run_both(pc, x, out_cols=192, act=act)
    assert max_abs(emu, real) < 3e-3


def test_conv1d_tc_gate_with_speaker_bias():
    H = 192
    w, b = _rand(2 * H, H, 5, seed=12, scale=(5 * H) ** -0.5), _rand(2 * H, seed=13, scale=0.1)
    x = _rand(2, 300, H, seed=14).half()
    pc = packer.pack_conv(w, b, gate=True)
    gb = _rand(2, 2 * H, seed=15, scale=0.3)[:, packer.gate_order(H)].contiguous()
    _, _, r16, e16 = _run_both(pc, x, out_cols=H, want32=False, want16=True, gbias=gb)
    assert max_abs(e16, r16) < 3e-3


def test_conv1d_tc_channel_views_and_in_place():
    """Flow coupling pattern: read a 96-channel view of a 192-wide tensor, write the other half in place."""
    B, T = 2, 260
    z32 = _rand(B, T, 192, seed=16)
    z16 = z32.half()
    wpre = _rand(192, 96, 1, seed=17, scale=96 ** -0.5)
    pc = packer.pack_conv(wpre, _rand(192, seed=18, scale=0.1))
    real, emu, _, _ = _run_both(pc, z16[:, :, 96:], out_cols=192)
    assert max_abs(emu, real) < 3e-3
    wpost = _rand(96, 192, 1, seed=19, scale=192 ** -0.5)
    pcp = packer.pack_conv(wpost, _rand(96, seed=20, scale=0.1)).to(DEV)
    h16 = _rand(B, T, 192, seed=21).half()
    zg32, zg16 = z32.to(DEV), z16.to(DEV)
    lens = torch.tensor([260, 100], dtype=torch.int32)
    ops.conv1d(pcp, h16.to(DEV), alpha=-1.0, res1=zg32[:, :, :96], lens=lens.to(DEV), out32=zg32[:, :, :96],
               out16=zg16[:, :, :96])
    torch.cuda.synchronize()
    zc32, zc16 = z32.clone(), z16.clone()
    pcc = packer.pack_conv(wpost, _rand(96, seed=20, scale=0.1))
    emu_ops.conv1d(pcc, h16, alpha=-1.0, res1=zc32[:, :, :96], lens=lens, out32=zc32[:, :, :96], out16=zc16[:, :, :96])
    assert max_abs(zc32, zg32.cpu()) < 3e-3 and max_abs(zc16.float(), zg16.float().cpu()) < 6e-3
    assert torch.equal(zg32[:, :, 96:].cpu(), z32[:, :, 96:])          # untouched half


@pytest.mark.parametrize("k,s,cin,cout", [(16, 8, 512, 256), (16, 8, 256, 128), (4, 2, 128, 64), (4, 2, 64, 32)])
def test_conv_transpose_polyphase_on_gpu(k, s, cin, cout):
    import torch.nn.functional as F
    w, b = _rand(cin, cout, k, seed=22, scale=(cin * k / s) ** -0.5), _rand(cout, seed=23, scale=0.1)
    B, T = 2, 150
    x = _rand(B, T, cin, seed=24).half()
    ref = F.conv_transpose1d(x.float().transpose(1, 2), w.half().float(), b, stride=s, padding=(k - s) // 2).transpose(1, 2)
    pc = packer.pack_conv_transpose(w, b, s, (k - s) // 2).to(DEV)
    out = torch.empty(B, T * s, cout, device=DEV)
    ops.conv1d(pc, x.to(DEV), out32=out.view(B, T, s * cout))
    torch.cuda.synchronize()
    assert max_abs(ref, out.cpu()) < 3e-3


def test_conv1d_tc_w_rows_split():
    H = 192
    w, b = _rand(2 * H, H, 1, seed=25, scale=H ** -0.5), _rand(2 * H, seed=26, scale=0.1)
    x = _rand(2, 200, H, seed=27).half()
    pc = packer.pack_conv(w, b, block_n=192)
    for r0 in (0, H):
        real, emu, _, _ = _run_both(pc, x, out_cols=H, w_rows=(r0, H))
        assert max_abs(emu, real) < 3e-3


F32_CASES = [("k3_192_256", 2, 70, 192, 256, 3), ("qkv", 1, 130, 192, 576, 1), ("k3_768_192", 2, 33, 768, 192, 3),
             ("to1", 3, 50, 256, 1, 1), ("gin", 4, 1, 256, 1536, 1)]


@pytest.mark.parametrize("name,B,T,cin,cout,k", F32_CASES, ids=[c[0] for c in F32_CASES])
def test_conv1d_f32(name, B, T, cin, cout, k):
    w, b = _rand(cout, cin, k, seed=28, scale=(cin * k) ** -0.5), _rand(cout, seed=29, scale=0.1)
    x = _rand(B, T, cin, seed=30)
    pc = packer.pack_conv(w, b, f32=True)
    res1 = _rand(B, T, cout, seed=31)
    lens = torch.tensor([T] + [max(T // 2, 0)] * (B - 1), dtype=torch.int32)
    real, emu, _, _ = _run_both(pc, x, out_cols=cout, res1=res1, lens=lens, act=ACT_RELU)
    assert max_abs(emu, real) < 2e-5


def test_bad_arguments_raise():
    pc = packer.pack_conv(_rand(192, 192, 1), None).to(DEV)
    x = torch.zeros(1, 10, 192, device=DEV, dtype=torch.float16)
    with pytest.raises(RuntimeError, match="no output"):
        ops.conv1d(pc, x)
    with pytest.raises(ValueError, match="channels"):
        ops.conv1d(pc, x[:, :, :96], out32=torch.zeros(1, 10, 192, device=DEV))
    with pytest.raises(TypeError):
        ops.conv1d(pc, x.bfloat16(), out32=torch.zeros(1, 10, 192, device=DEV))


@pytest.mark.parametrize("cin,cout,k,act", [(192, 576, 1, ACT_NONE), (192, 768, 3, ACT_RELU), (768, 192, 3, ACT_NONE)])
def test_conv1d_tc_split_precision(cin, cout, k, act):
    """fp16 (hi, lo) split operands: x_hi.w_hi + x_lo.w_hi + x_hi.w_lo on the tensor cores must agree with an
    fp32 convolution to fp32-class accuracy (F0 predictor path), and the [hi | lo | hi] output must re-compose."""
    import torch.nn.functional as F
    B, T = 2, 300
    w, b = _rand(cout, cin, k, seed=40, scale=(cin * k) ** -0.5), _rand(cout, seed=41, scale=0.1)
    x = _rand(B, T, cin, seed=42)
    res = _rand(B, T, cout, seed=43)
    lens = torch.tensor([300, 170], dtype=torch.int32)
    ref = F.conv1d(x.double().transpose(1, 2), w.double(), b.double(), padding=k // 2).transpose(1, 2)
    if act == ACT_RELU:
        ref = torch.relu(ref)
    ref = ((ref + res) * (torch.arange(T)[None, :, None] < lens[:, None, None])).float()
    hi = x.half()
    xs = torch.cat([hi, (x - hi.float()).half(), hi], -1).to(DEV)
    pc = packer.pack_conv(w, b, split=True).to(DEV)
    o32 = torch.empty(B, T, cout, device=DEV)
    o16 = torch.empty(B, T, 3 * cout, device=DEV, dtype=torch.float16)
    ops.conv1d(pc, xs, act=act, res1=res.to(DEV), lens=lens.to(DEV), out32=o32, out16=o16, split16=True)
    torch.cuda.synchronize()
    scale = float(ref.abs().max())
    # operand quantisation contributes ~2e-7; the rest is the tensor core's truncating fp32 accumulate, measured
    # (tools/diag_split.py) at <= 1.7e-7*|acc| per tcgen05.mma step and biased toward zero.  3*cin*k/16 steps.
    steps = 3 * cin * k // 16
    assert max_abs(ref, o32.cpu()) < (2e-6 + 2.5e-7 * steps) * max(scale, 1.0) / 4
    h, l, h2 = (o16[:, :, i * cout:(i + 1) * cout].float().cpu() for i in range(3))
    assert torch.equal(h, h2)
    assert max_abs(o32.cpu(), h + l) < 2e-6 * max(scale, 1.0)


@pytest.mark.parametrize("cout,k,dil,with_res2,scale", [(128, 3, 1, False, 1.0), (32, 7, 3, True, 1.0), (64, 11, 5, True, 1.0 / 3)])
def test_conv1d_tc_16bit_residual_stream(cout, k, dil, with_res2, scale):
    """res16 (16-bit residual in) + out16r (un-activated 16-bit copy out, in place over res16) + activated out16:
    the Generator's 16-bit residual stream."""
    import torch.nn.functional as F
    B, T, cin = 2, 700, cout
    w, b = _rand(cout, cin, k, seed=50, scale=(cin * k) ** -0.5), _rand(cout, seed=51, scale=0.1)
    x = _rand(B, T, cin, seed=52).half()
    r16 = _rand(B, T, cout, seed=53).half()
    r2 = _rand(B, T, cout, seed=54) if with_res2 else None
    ref = F.conv1d(x.float().transpose(1, 2), w.half().float(), b, padding=dil * (k // 2), dilation=dil).transpose(1, 2)
    ref = ref + r16.float() + (r2 if with_res2 else 0.0)
    ref = ref * scale
    pc = packer.pack_conv(w, b, dil=dil).to(DEV)
    raw = r16.clone().to(DEV)                      # updated in place
    act16 = torch.empty(B, T, cout, device=DEV, dtype=torch.float16)
    o32 = torch.empty(B, T, cout, device=DEV) if with_res2 else None
    ops.conv1d(pc, x.to(DEV), res16=raw, out16r=raw, out16=act16, post_lrelu=0.1, out32=o32,
               res2=None if r2 is None else r2.to(DEV), scale=scale)
    torch.cuda.synchronize()
    assert max_abs(ref.half().float(), raw.float().cpu()) < 4e-3
    assert max_abs(F.leaky_relu(ref, 0.1).half().float(), act16.float().cpu()) < 4e-3
    if with_res2:
        assert max_abs(ref, o32.cpu()) < 2e-3


@pytest.mark.gpu
@pytest.mark.parametrize("cout,k,with_mrf", [(32, 3, False), (64, 11, True), (128, 7, False)])
def test_conv1d_tc_residual_from_activated_stream(cout, k, with_mrf):
    """res16_unlrelu: res16 holds leaky_relu(x, 0.1) (the tensor the pair's first conv consumed) and the epilogue
    recovers x = min(r, 10 r); the activated result is written in place over it (generator_residual='act16')."""
    import torch.nn.functional as F
    B, T, cin = 2, 700, cout
    w, b = _rand(cout, cin, k, seed=60, scale=(cin * k) ** -0.5), _rand(cout, seed=61, scale=0.1)
    x = _rand(B, T, cin, seed=62).half()
    xres = _rand(B, T, cout, seed=63)
    act_in = F.leaky_relu(xres, 0.1).half()                       # what the previous conv stored
    xrec = torch.minimum(act_in.float(), act_in.float() * 10.0)    # what the epilogue must recover
    assert max_abs(xrec, xres) < 4e-3
    mrf = _rand(B, T, cout, seed=64).half() if with_mrf else None
    ref = F.conv1d(x.float().transpose(1, 2), w.half().float(), b, padding=k // 2).transpose(1, 2) + xrec
    if with_mrf:
        ref = ref + mrf.float()
    pc = packer.pack_conv(w, b).to(DEV)
    buf = act_in.clone().to(DEV)
    ops.conv1d(pc, x.to(DEV), res16=buf, res16_unlrelu=10.0, out16=buf, post_lrelu=0.1,
               res2=None if mrf is None else mrf.to(DEV))
    torch.cuda.synchronize()
    assert max_abs(F.leaky_relu(ref, 0.1).half().float(), buf.float().cpu()) < 4e-3


@pytest.mark.gpu
@pytest.mark.parametrize("c,k,T", [(32, 3, 700), (32, 7, 1026), (32, 11, 258), (64, 11, 700), (32, 11, 2)])
def test_conv1d_tc_time_pair_folded(c, k, T):
    """packer.pack_conv_folded: the dilation-1 conv run on [B, T/2, 2C] views (block-Toeplitz weights) is the same conv
    - checked against torch and against the unfolded kernel path, including the zero padding at both ends of every
    utterance (row -1 / row T/2 of the folded view) and the act16 residual epilogue on folded views."""
    import torch.nn.functional as F
    B = 3
    w, b = _rand(c, c, k, seed=70, scale=(c * k) ** -0.5), _rand(c, seed=71, scale=0.1)
    x = _rand(B, T, c, seed=72).half()
    res = F.leaky_relu(_rand(B, T, c, seed=73), 0.1).half()
    xrec = torch.minimum(res.float(), res.float() * 10.0)
    ref = F.conv1d(x.float().transpose(1, 2), w.half().float(), b, padding=k // 2).transpose(1, 2) + xrec
    ref16 = F.leaky_relu(ref, 0.1).half().float()
    pcf, pcu = packer.pack_conv_folded(w, b).to(DEV), packer.pack_conv(w, b).to(DEV)
    assert pcf.fold == 2 and pcf.cin == 2 * c
    outs = []
    for pc, f in ((pcf, 2), (pcu, 1)):
        buf = res.clone().to(DEV)
        v = lambda t: t.view(B, T // f, f * c)
        ops.conv1d(pc, v(x.to(DEV)), res16=v(buf), res16_unlrelu=10.0, out16=v(buf), post_lrelu=0.1)
        torch.cuda.synchronize()
        outs.append(buf.float().cpu())
        assert max_abs(ref16, outs[-1]) < 4e-3
    assert max_abs(outs[0], outs[1]) < 2e-3          # same products, different accumulation order + fp16 rounding


@pytest.mark.gpu
@pytest.mark.parametrize("cout,k,dil,scale", [(32, 3, 1, 1.0), (64, 7, 1, 1.0 / 3.0), (128, 11, 1, 1.0 / 3.0), (256, 3, 1, 1.0)])
def test_conv1d_tc_16bit_mrf_stream(cout, k, dil, scale):
    """res2 as a 16-bit tensor (b200svs_epilogue.res2_is16): the MRF sum xs += resblock_j(x) kept as a 16-bit
    stream - accumulate in place (res2 in, out16r out over the same buffer) and the final scaled, activated out16."""
    import torch.nn.functional as F
    B, T, cin = 2, 700, cout
    w, b = _rand(cout, cin, k, seed=60, scale=(cin * k) ** -0.5), _rand(cout, seed=61, scale=0.1)
    x = _rand(B, T, cin, seed=62).half()
    r16 = _rand(B, T, cout, seed=63).half()
    mrf = _rand(B, T, cout, seed=64).half()
    ref = F.conv1d(x.float().transpose(1, 2), w.half().float(), b, padding=dil * (k // 2), dilation=dil).transpose(1, 2)
    ref = (ref + r16.float() + mrf.float()) * scale
    pc = packer.pack_conv(w, b, dil=dil).to(DEV)
    # j = 1: xs += resblock_1(x), in place
    acc = mrf.clone().to(DEV)
    ops.conv1d(pc, x.to(DEV), res16=r16.to(DEV), res2=acc, out16r=acc, scale=scale)
    # j = 2: x = lrelu(xs / 3) for the next stage
    nxt = torch.empty(B, T, cout, device=DEV, dtype=torch.float16)
    ops.conv1d(pc, x.to(DEV), res16=r16.to(DEV), res2=mrf.to(DEV), out16=nxt, post_lrelu=0.1, scale=scale)
    torch.cuda.synchronize()
    assert max_abs(ref.half().float(), acc.float().cpu()) < 6e-3
    assert max_abs(F.leaky_relu(ref, 0.1).half().float(), nxt.float().cpu()) < 6e-3
