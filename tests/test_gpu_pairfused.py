"""-m gpu: b200svs_resblock_pair (one fused kernel per ResBlock1 pair of the 32-channel stage) against the two
conv1d_tc launches it replaces - bit-identical, for every kernel size / dilation of the shipped config, every epilogue
variant (plain pair, MRF first / middle / last), ragged tile counts and batch edges."""
import pytest
import torch

from music_generation_b200 import ops, packer

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _convs(k, d, seed):
    g = torch.Generator().manual_seed(seed)
    w1, b1 = torch.randn(32, 32, k, generator=g) * (32 * k) ** -0.5, torch.randn(32, generator=g) * 0.1
    w2, b2 = torch.randn(32, 32, k, generator=g) * (32 * k) ** -0.5, torch.randn(32, generator=g) * 0.1
    c1 = (packer.pack_conv_folded(w1, b1) if d == 1 else packer.pack_conv(w1, b1, dil=d)).to(DEV)
    c2 = packer.pack_conv_folded(w2, b2).to(DEV)
    return c1, c2


def _fold(pc, t):
    return t.view(t.shape[0], t.shape[1] // pc.fold, t.shape[2] * pc.fold) if pc.fold > 1 else t


def _two_launch(c1, c2, x, mode):
    """The per-op path of synthesizer._generator_stages (act16 residual, 16-bit MRF)."""
    t16 = torch.empty_like(x)
    ops.conv1d(c1, _fold(c1, x), out16=_fold(c1, t16), post_lrelu=0.1)
    out = torch.empty_like(x)
    kw = dict(res16=_fold(c2, x), res16_unlrelu=10.0)
    mrf = None
    if mode in ("mrf_mid", "mrf_last"):
        mrf = (torch.randn(x.shape, generator=torch.Generator().manual_seed(5)) * 0.5).half().to(DEV)
        kw["res2"] = _fold(c2, mrf)
    if mode == "pair":
        kw.update(out16=_fold(c2, out), post_lrelu=0.1)
    elif mode in ("mrf_first", "mrf_mid"):
        kw.update(out16r=_fold(c2, out))
    else:
        kw.update(scale=1.0 / 3, out16=_fold(c2, out), post_lrelu=0.01)
    ops.conv1d(c2, _fold(c2, t16), **kw)
    return out, mrf


@pytest.mark.parametrize("k,d", [(3, 1), (3, 3), (3, 5), (7, 1), (7, 3), (7, 5), (11, 1), (11, 3), (11, 5)])
@pytest.mark.parametrize("mode", ["pair", "mrf_first", "mrf_mid", "mrf_last"])
def test_fused_pair_is_bit_identical(k, d, mode):
    B, L = 3, 2 * 1237                                    # odd folded length: last tile partial, batch edges inside the grid
    c1, c2 = _convs(k, d, seed=100 * k + d)
    x = (torch.randn(B, L, 32, generator=torch.Generator().manual_seed(k * d)) * 0.8).half().to(DEV)
    want, mrf = _two_launch(c1, c2, x, mode)
    got = torch.full_like(x, 7.0)
    kw = dict(pair=dict(out_lrelu=True, post_slope=0.1), mrf_first=dict(out_lrelu=False), mrf_mid=dict(out_lrelu=False),
              mrf_last=dict(out_lrelu=True, post_slope=0.01, scale=1.0 / 3))[mode]
    ops.resblock_pair(c1, c2, x, got, mrf=mrf, **kw)
    torch.cuda.synchronize()
    assert torch.equal(got, want), float((got.float() - want.float()).abs().max())


@pytest.mark.parametrize("k,d", [(3, 1), (3, 5), (7, 1), (7, 3), (7, 5)])
@pytest.mark.parametrize("mode", ["pair", "mrf_mid", "mrf_last"])
def test_fused_pair_64_channels_is_bit_identical(k, d, mode):
    """The 64-channel stage: rows are time steps, two plain 64 -> 64 convs (kernel sizes up to 7 keep both weight sets
    resident)."""
    B, L = 2, 2500 + 7
    g = torch.Generator().manual_seed(10 * k + d)
    w1, b1 = torch.randn(64, 64, k, generator=g) * (64 * k) ** -0.5, torch.randn(64, generator=g) * 0.1
    w2, b2 = torch.randn(64, 64, k, generator=g) * (64 * k) ** -0.5, torch.randn(64, generator=g) * 0.1
    c1, c2 = packer.pack_conv(w1, b1, dil=d).to(DEV), packer.pack_conv(w2, b2).to(DEV)
    assert ops.pair_fusable(c1, c2, 64)
    x = (torch.randn(B, L, 64, generator=g) * 0.8).half().to(DEV)
    want, mrf = _two_launch(c1, c2, x, mode)
    got = torch.full_like(x, 7.0)
    kw = dict(pair=dict(out_lrelu=True, post_slope=0.1), mrf_mid=dict(out_lrelu=False),
              mrf_last=dict(out_lrelu=True, post_slope=0.01, scale=1.0 / 3))[mode]
    ops.resblock_pair(c1, c2, x, got, mrf=mrf, **kw)
    torch.cuda.synchronize()
    assert torch.equal(got, want), float((got.float() - want.float()).abs().max())
    w11 = torch.randn(64, 64, 11, generator=g)
    assert not ops.pair_fusable(packer.pack_conv(w11, None).to(DEV), packer.pack_conv(w11, None).to(DEV), 64)


def test_fused_pair_in_place_mrf_and_large_shape():
    """out aliasing mrf (the MRF sum updated in place) on a shape with several tiles per SM."""
    c1, c2 = _convs(7, 3, seed=9)
    B, L = 2, 2 * 40000
    x = (torch.randn(B, L, 32, generator=torch.Generator().manual_seed(1)) * 0.8).half().to(DEV)
    mrf = (torch.randn(B, L, 32, generator=torch.Generator().manual_seed(2)) * 0.5).half().to(DEV)
    t16 = torch.empty_like(x)
    ops.conv1d(c1, x, out16=t16, post_lrelu=0.1)
    want = mrf.clone()
    ops.conv1d(c2, _fold(c2, t16), res16=_fold(c2, x), res16_unlrelu=10.0, res2=_fold(c2, want), out16r=_fold(c2, want))
    got = mrf.clone()
    ops.resblock_pair(c1, c2, x, got, mrf=got, out_lrelu=False)
    torch.cuda.synchronize()
    assert torch.equal(got, want)
    with pytest.raises(RuntimeError, match="alias"):
        ops.resblock_pair(c1, c2, x, x)


@pytest.mark.parametrize("stage_api", [True, False])
def test_generator_with_fused_pairs_is_bit_identical(stage_api):
    """generator_pair_fused=True (32-channel stage: one kernel per ResBlock pair) through the whole Generator, via the
    stage-level C entry point and via the per-op Python loop."""
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from common import build_model
    a, _ = build_model("ae", device=DEV, stage_api=stage_api, generator_pair_fused=True)
    b, _ = build_model("ae", device=DEV, stage_api=stage_api, generator_pair_fused=False)
    a._pack(); b._pack()
    z = (torch.randn(3, 333, 192, generator=torch.Generator().manual_seed(3)) * 0.14).half().to(DEV)
    from music_generation_b200 import _lib
    l0 = _lib.launches(); oa = a._generator(z); l1 = _lib.launches(); ob = b._generator(z); l2 = _lib.launches()
    torch.cuda.synchronize()
    assert torch.equal(oa, ob)
    assert (l2 - l1) - (l1 - l0) == 9 + 6                 # nine pairs at 32 channels + the k = 3 / 7 pairs at 64: one launch each less
