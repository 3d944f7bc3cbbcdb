"""-m gpu: attention, LayerNorm and the integer / bandwidth-bound kernels vs their CPU semantics."""
import ctypes as C
import json
import os

import pytest
import torch

import emu_ops
from common import max_abs
from music_generation_b200 import _lib, ops

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _rand(*shape, seed=0, scale=1.0):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale


@pytest.mark.parametrize("B,T,lens", [(2, 150, [150, 77]), (1, 64, [64]), (3, 200, [200, 1, 0]), (1, 3, [3]),
                                      (2, 513, [513, 300])])
def test_attention_matches_reference_semantics(B, T, lens):
    qkv = _rand(B, T, 576, seed=1)
    rel_k, rel_v = _rand(9, 96, seed=2, scale=96 ** -0.5), _rand(9, 96, seed=3, scale=96 ** -0.5)
    ln = torch.tensor(lens, dtype=torch.int32)
    e32 = torch.empty(B, T, 192)
    emu_ops.attention(qkv, e32, None, rel_k, rel_v, ln)
    r32 = torch.empty(B, T, 192, device=DEV)
    r16 = torch.empty(B, T, 192, device=DEV, dtype=torch.float16)
    ops.attention(qkv.to(DEV), r32, r16, rel_k.to(DEV), rel_v.to(DEV), ln.to(DEV))
    torch.cuda.synchronize()
    assert max_abs(e32, r32.cpu()) < 2e-5
    assert max_abs(e32, r16.float().cpu()) < 4e-3


@pytest.mark.parametrize("B,T,lens", [(2, 150, [150, 77]), (1, 64, [64]), (3, 200, [200, 1, 0]), (1, 3, [3]),
                                      (2, 513, [513, 300]), (1, 1100, [1100])])
@pytest.mark.parametrize("split", [False, True], ids=["plain", "split"])
def test_attention_tc_matches_reference_semantics(B, T, lens, split):
    """tcgen05 attention vs the fp32 semantics.  plain: fp16 operand rounding of q,k,v,p -> 4e-3;
    split (hi,lo) operands -> fp32-class agreement."""
    qkv = _rand(B, T, 576, seed=1)
    rel_k, rel_v = _rand(9, 96, seed=2, scale=96 ** -0.5), _rand(9, 96, seed=3, scale=96 ** -0.5)
    ln = torch.tensor(lens, dtype=torch.int32)
    e32 = torch.empty(B, T, 192)
    emu_ops.attention(qkv, e32, None, rel_k, rel_v, ln)
    r32 = torch.full((B, T, 192), 7.0, device=DEV)
    r16 = torch.full((B, T, 576 if split else 192), 7.0, device=DEV, dtype=torch.float16)
    ops.attention_tc(qkv.to(DEV), r32, r16, rel_k.to(DEV), rel_v.to(DEV), ln.to(DEV), dtype=torch.float16, split=split,
                     out16_split=split)
    torch.cuda.synchronize()
    tol = 2e-5 if split else 4e-3
    assert max_abs(e32, r32.cpu()) < tol
    if split:
        hi, lo, hi2 = (r16[:, :, i * 192:(i + 1) * 192].float().cpu() for i in range(3))
        assert torch.equal(hi, hi2)
        assert max_abs(e32, hi + lo) < 2e-5
    else:
        assert max_abs(e32, r16.float().cpu()) < 6e-3


@pytest.mark.parametrize("split", [False, True], ids=["plain", "split"])
def test_attention_tc_rising_scores_exercise_the_lazy_rescale(split):
    """attention_ts accumulates O in tensor memory and moves the running row maximum only when a key tile exceeds it by
    more than 2^8, rescaling O in place.  Random scores never do that, so: keys whose magnitude grows
    along the sequence (x1, x6, x40: the maximum jumps by far more than 8 in the log2 domain, twice, at different tiles
    for different queries), a stretch of tiny keys in between (maximum must NOT move), and a ragged second row."""
    B, T = 2, 700
    qkv = _rand(B, T, 576, seed=11)
    gain = torch.ones(T)
    gain[130:300] = 6.0
    gain[300:420] = 0.05
    gain[420:] = 40.0
    qkv[:, :, 192:384] *= gain[None, :, None]          # the k third of q|k|v
    qkv[1, :, 192:384] = qkv[1].flip(0)[:, 192:384]    # row 1: the large keys come FIRST (no rescale after tile 0..4)
    rel_k, rel_v = _rand(9, 96, seed=12, scale=96 ** -0.5), _rand(9, 96, seed=13, scale=96 ** -0.5)
    ln = torch.tensor([700, 555], dtype=torch.int32)
    # scores reach several hundred here, so rounding q / k to fp16 alone moves the fp32 result by 6e-2: the plain kernel is
    # compared with the fp32 semantics of the fp16-ROUNDED inputs (what it is given), the split kernel with the exact ones
    ref_in = qkv if split else qkv.half().float()
    e32 = torch.empty(B, T, 192)
    emu_ops.attention(ref_in, e32, None, rel_k, rel_v, ln)
    r32 = torch.full((B, T, 192), 7.0, device=DEV)
    ops.attention_tc(qkv.to(DEV), r32, None, rel_k.to(DEV), rel_v.to(DEV), ln.to(DEV), dtype=torch.float16, split=split)
    torch.cuda.synchronize()
    assert torch.isfinite(r32).all()
    err = max_abs(e32, r32.cpu())
    print(f"rising scores, {'split' if split else 'plain'}: max abs error {err:.3e}")
    assert err < (2e-4 if split else 8e-3)


@pytest.mark.parametrize("T", [256, 384, 512, 640, 768])
@pytest.mark.parametrize("split", [False, True], ids=["plain", "split"])
def test_attention_tc_quick_last_tile(T, split):
    """Regression (found by tools/fuzz_attn.py): with T % 128 == 0 the second-to-last query tile ends on a key tile that
    needs neither the band nor the length mask, the softmax warps reach the final PV wait early, and a parity wait on ONE
    PV barrier returned two phases early (O read before the last two PV products had landed; intermittent, the upper
    rows of that tile).  Repeated, because it was a race."""
    qkv = _rand(1, T, 576, seed=T, scale=0.5)
    rel_k, rel_v = _rand(9, 96, seed=2, scale=96 ** -0.5), _rand(9, 96, seed=3, scale=96 ** -0.5)
    ln = torch.tensor([T], dtype=torch.int32)
    e32 = torch.empty(1, T, 192)
    emu_ops.attention(qkv if split else qkv.half().float(), e32, None, rel_k, rel_v, ln)
    args = (rel_k.to(DEV), rel_v.to(DEV), ln.to(DEV))
    q = qkv.to(DEV)
    for _ in range(5):
        r32 = torch.full((1, T, 192), 7.0, device=DEV)
        ops.attention_tc(q, r32, None, *args, dtype=torch.float16, split=split)
        torch.cuda.synchronize()
        assert max_abs(e32, r32.cpu()) < (2e-5 if split else 4e-3)


def test_attention_tc_random_shapes():
    """Seeded fuzz over batch, length (around the 64-key / 128-query tile edges), ragged lengths and score scale."""
    import random
    rnd = random.Random(5)
    for it in range(24):
        B = rnd.randint(1, 3)
        T = rnd.choice([1, 2, 5, 63, 64, 65, 127, 128, 129, 191, 200, 257, 320, 449, 640])
        lens = [rnd.choice([T, rnd.randint(0, T), max(1, T // 2)]) for _ in range(B)]
        sc = rnd.choice([0.5, 1.0, 2.0])
        qkv = _rand(B, T, 576, seed=it, scale=sc)
        rel_k, rel_v = _rand(9, 96, seed=100 + it, scale=96 ** -0.5), _rand(9, 96, seed=200 + it, scale=96 ** -0.5)
        ln = torch.tensor(lens, dtype=torch.int32)
        for split in (False, True):
            e32 = torch.empty(B, T, 192)
            emu_ops.attention(qkv if split else qkv.half().float(), e32, None, rel_k, rel_v, ln)
            r32 = torch.full((B, T, 192), 7.0, device=DEV)
            ops.attention_tc(qkv.to(DEV), r32, None, rel_k.to(DEV), rel_v.to(DEV), ln.to(DEV), dtype=torch.float16, split=split)
            torch.cuda.synchronize()
            assert max_abs(e32, r32.cpu()) < (2e-4 if split else 1e-2), (B, T, lens, sc, split)


def test_attention_tc_bf16():
    B, T = 2, 300
    qkv = _rand(B, T, 576, seed=4)
    rel_k, rel_v = _rand(9, 96, seed=5, scale=96 ** -0.5), _rand(9, 96, seed=6, scale=96 ** -0.5)
    ln = torch.tensor([300, 190], dtype=torch.int32)
    e32 = torch.empty(B, T, 192)
    emu_ops.attention(qkv, e32, None, rel_k, rel_v, ln)
    r32 = torch.empty(B, T, 192, device=DEV)
    ops.attention_tc(qkv.to(DEV), r32, None, rel_k.to(DEV), rel_v.to(DEV), ln.to(DEV), dtype=torch.bfloat16)
    torch.cuda.synchronize()
    assert max_abs(e32, r32.cpu()) < 3e-2


def test_attention_agrees_with_oracle_mha():
    """Against the oracle's pad/view formulation (VISinger/attentions.py:148-243), not just the emulation."""
    from oracle import visinger_oracle as vo
    B, T = 2, 90
    sd = {f"a.conv_{n}.weight": _rand(192, 192, 1, seed=10 + i, scale=192 ** -0.5) for i, n in enumerate("qkvo")}
    sd.update({f"a.conv_{n}.bias": _rand(192, seed=20 + i, scale=0.1) for i, n in enumerate("qkvo")})
    sd["a.emb_rel_k"], sd["a.emb_rel_v"] = _rand(1, 9, 96, seed=30, scale=0.1), _rand(1, 9, 96, seed=31, scale=0.1)
    x = _rand(B, 192, T, seed=32)
    lens = torch.tensor([90, 41])
    m = vo.sequence_mask(lens, T).unsqueeze(1).float()
    attn_mask = m.unsqueeze(2) * m.unsqueeze(-1)
    ref = vo.multi_head_attention(sd, "a", x * m, attn_mask)              # includes conv_o
    xm = (x * m).transpose(1, 2)
    qkv = torch.cat([torch.nn.functional.conv1d(x * m, sd[f"a.conv_{n}.weight"], sd[f"a.conv_{n}.bias"]) for n in "qkv"], 1)
    out = torch.empty(B, T, 192, device=DEV)
    ops.attention(qkv.transpose(1, 2).contiguous().to(DEV), out, None, sd["a.emb_rel_k"][0].contiguous().to(DEV),
                  sd["a.emb_rel_v"][0].contiguous().to(DEV), lens.to(torch.int32).to(DEV))
    torch.cuda.synchronize()
    got = torch.nn.functional.conv1d(out.cpu().transpose(1, 2), sd["a.conv_o.weight"], sd["a.conv_o.bias"])
    valid = m.bool().expand_as(ref)
    assert max_abs(ref[valid], got[valid]) < 5e-5


@pytest.mark.parametrize("C_", [192, 256])
def test_layernorm(C_):
    B, T = 3, 77
    x, res = _rand(B, T, C_, seed=4), _rand(B, T, C_, seed=5)
    g, b = _rand(C_, seed=6) + 1, _rand(C_, seed=7)
    lens = torch.tensor([77, 30, 0], dtype=torch.int32)
    e = torch.empty(B, T, C_)
    emu_ops.layernorm(x, g, b, lens, out32=e, res=res)
    r = torch.empty(B, T, C_, device=DEV)
    r16 = torch.empty(B, T, C_, device=DEV, dtype=torch.float16)
    ops.layernorm(x.to(DEV), g.to(DEV), b.to(DEV), lens.to(DEV), out32=r, out16=r16, res=res.to(DEV))
    torch.cuda.synchronize()
    assert max_abs(e, r.cpu()) < 1e-5 and max_abs(e, r16.float().cpu()) < 4e-3


def test_length_regulator_bit_exact():
    B, P, Cc = 4, 37, 192
    g = torch.Generator().manual_seed(8)
    dur = torch.randint(0, 9, (B, P), generator=g)
    dur[1] = 0                       # utterance with no frames
    dur[2, 5:] = 0                   # padded phoneme slots
    dur[3, 0] = -2                   # negative -> treated as empty
    h = _rand(B, P, Cc, seed=9)
    ends_c, m64_c, m32_c = torch.empty(B, P, dtype=torch.int32), torch.empty(B, dtype=torch.int64), torch.empty(B, dtype=torch.int32)
    emu_ops.lr_scan(dur, ends_c, m64_c, m32_c)
    T = int(m64_c.max())
    out_c, idx_c = torch.empty(B, T, Cc), torch.empty(B, T, dtype=torch.int32)
    emu_ops.lr_gather(h, ends_c, m32_c, out_c, idx_c)
    ends = torch.empty(B, P, dtype=torch.int32, device=DEV)
    m64, m32 = torch.empty(B, dtype=torch.int64, device=DEV), torch.empty(B, dtype=torch.int32, device=DEV)
    ops.lr_scan(dur.to(DEV), ends, m64, m32)
    out, idx = torch.empty(B, T, Cc, device=DEV), torch.empty(B, T, dtype=torch.int32, device=DEV)
    ops.lr_gather(h.to(DEV), ends, m32, out, idx)
    torch.cuda.synchronize()
    assert torch.equal(ends.cpu(), ends_c) and torch.equal(m64.cpu(), m64_c)
    assert torch.equal(idx.cpu(), idx_c)
    assert torch.equal(out.cpu(), out_c)                                  # payload copied bit-exactly
    # against the oracle's formulation too
    from oracle import visinger_oracle as vo
    o2, ml2, i2 = vo.length_regulate(h.transpose(1, 2), dur)
    assert torch.equal(o2.transpose(1, 2), out.cpu()) and torch.equal(ml2, m64.cpu())


def test_duration_quantize_bit_exact():
    g = torch.Generator().manual_seed(10)
    pred, nd = torch.rand(4096, generator=g), 0.15 + 0.65 * torch.rand(4096, generator=g)
    ref = torch.floor(pred * nd / 0.0116).long()
    out = torch.empty(4096, dtype=torch.int64, device=DEV)
    ops.duration_quantize(pred.to(DEV), nd.to(DEV), out)
    torch.cuda.synchronize()
    assert torch.equal(out.cpu(), ref)


def test_text_embed_and_pos_and_f0_and_prior():
    B, P, Cc, T = 2, 11, 192, 50
    g = torch.Generator().manual_seed(11)
    ph, pi, to = (torch.randint(0, n, (B, P), generator=g) for n in (100, 100, 300))
    dur = torch.rand(B, P, generator=g)
    tabs = [_rand(n, Cc, seed=12 + i) for i, n in enumerate((100, 100, 300))]
    wd, bd = _rand(Cc, seed=15), _rand(Cc, seed=16)
    xl = torch.tensor([11, 6])
    l_c, o_c = torch.empty(B, dtype=torch.int32), torch.empty(B, P, Cc)
    emu_ops.text_embed(ph, pi, to, dur, *tabs, wd, bd, xl, l_c, o_c)
    l_g, o_g = torch.empty(B, dtype=torch.int32, device=DEV), torch.empty(B, P, Cc, device=DEV)
    ops.text_embed(ph.to(DEV), pi.to(DEV), to.to(DEV), dur.to(DEV), *[t.to(DEV) for t in tabs], wd.to(DEV), bd.to(DEV),
                   xl.to(DEV), l_g, o_g)
    assert torch.equal(l_g.cpu(), l_c) and max_abs(o_c, o_g.cpu()) < 1e-5
    # add_pos_mask
    x, pos, lens = _rand(B, T, Cc, seed=17), _rand(64, Cc, seed=18), torch.tensor([50, 20], dtype=torch.int32)
    e = torch.empty(B, T, Cc); emu_ops.add_pos_mask(x, pos, lens, out32=e)
    r = torch.empty(B, T, Cc, device=DEV); r16 = torch.empty(B, T, Cc, device=DEV, dtype=torch.float16)
    ops.add_pos_mask(x.to(DEV), pos.to(DEV), lens.to(DEV), out32=r, out16=r16)
    assert torch.equal(r.cpu(), e)
    # f0_embed: predicted buckets may differ only within an ulp of a bucket edge
    lf0 = 4.0 + 2.5 * torch.rand(B, T, generator=g)
    table = _rand(300, Cc, seed=19)
    c_c, e16 = torch.empty(B, T, dtype=torch.int64), torch.empty(B, T, Cc)
    emu_ops.f0_embed(lf0, None, table, lens, c_c, out32=e16)
    c_g, g16 = torch.empty(B, T, dtype=torch.int64, device=DEV), torch.empty(B, T, Cc, device=DEV)
    ops.f0_embed(lf0.to(DEV), None, table.to(DEV), lens.to(DEV), c_g, out32=g16)
    assert int((c_g.cpu() != c_c).sum()) <= 1 and c_g.min() >= 1 and c_g.max() <= 255
    inj = torch.randint(1, 256, (B, T), generator=g)
    ops.f0_embed(None, inj.to(DEV), table.to(DEV), lens.to(DEV), c_g, out32=g16)
    emu_ops.f0_embed(None, inj, table, lens, c_c, out32=e16)
    assert torch.equal(c_g.cpu(), inj) and torch.equal(g16.cpu(), e16)
    # sample_prior
    stats, noise = _rand(B, T, 2 * Cc, seed=20, scale=0.3), _rand(B, Cc, T, seed=21)
    e = torch.empty(B, T, Cc); emu_ops.sample_prior(stats, noise, 0.8, zp32=e)
    r = torch.empty(B, T, Cc, device=DEV); ops.sample_prior(stats.to(DEV), noise.to(DEV), 0.8, zp32=r)
    torch.cuda.synchronize()
    assert max_abs(e, r.cpu()) < 1e-5


def test_conv_post_tanh_and_layout_helpers():
    B, L, Cc = 2, 1000, 32
    x = _rand(B, L, Cc, seed=22).half()
    w = _rand(7, Cc, seed=23, scale=0.1)
    e = torch.empty(B, L); emu_ops.conv_post_tanh(x, w, e)
    r = torch.empty(B, L, device=DEV); ops.conv_post_tanh(x.to(DEV), w.to(DEV), r)
    torch.cuda.synchronize()
    assert max_abs(e, r.cpu()) < 1e-5
    a = _rand(2, 192, 77, seed=24)
    o = torch.empty(2, 77, 192, device=DEV); ops.bct_to_btc(a.to(DEV), out32=o)
    back = torch.empty(2, 192, 77, device=DEV); ops.btc_to_bct(o, back)
    torch.cuda.synchronize()
    assert torch.equal(o.cpu(), a.transpose(1, 2)) and torch.equal(back.cpu(), a)


def test_probe_rowshift_descriptor_semantics():
    """Pins how tcgen05 smem descriptors address a row-shifted SWIZZLE_128B tile; result recorded for DESIGN.md."""
    a = _rand(160, 64, seed=25).half()
    b = _rand(64, 64, seed=26).half()
    report = {}
    ag, bg = a.to(DEV), b.to(DEV)
    for mode in (0, 1):
        for shift in (0, 1, 2, 3, 4, 7, 8, 9, 16, 21):
            out = torch.zeros(128, 64, device=DEV)
            _lib.call("b200svs_probe_rowshift", C.c_void_p(ag.data_ptr()), C.c_void_p(bg.data_ptr()),
                      C.c_void_p(out.data_ptr()), 160, shift, mode, None)
            torch.cuda.synchronize()
            ref = a[shift:shift + 128].float() @ b.float().t()
            report[f"mode{mode}_shift{shift}"] = max_abs(ref, out.cpu())
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(report, open("gpurun_out/probe_rowshift.json", "w"), indent=1)
    assert report["mode0_shift0"] < 1e-3 and report["mode0_shift8"] < 1e-3 and report["mode0_shift16"] < 1e-3
