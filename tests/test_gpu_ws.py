"""-m gpu: b200svs_conv1d_ws (weights as the tensor core's A operand in tensor memory, transposed product: DESIGN.md 4.1c)
against b200svs_conv1d_tc on the 64-channel dilation-1 ResBlock convs.  Same products, same fp32 accumulation, same
epilogue arithmetic: the results are compared bit for bit."""
import pytest
import torch

from music_generation_b200 import ops, packer

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _case(k, B, T, seed):
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(64, 64, k, generator=g) * (64 * k) ** -0.5
    b = torch.randn(64, generator=g) * 0.1
    x, src, mrf = ((torch.randn(B, T, 64, generator=g) * 0.8).half().to(DEV) for _ in range(3))
    return packer.pack_conv(w, b).to(DEV), packer.pack_conv_ws(w, b).to(DEV), x, src, mrf


@pytest.mark.parametrize("k", [3, 7, 11])
@pytest.mark.parametrize("B,T", [(1, 128), (2, 1000), (3, 2), (1, 20482)])
def test_conv1d_ws_is_bit_identical_to_conv1d_tc(k, B, T):
    pc, pw, x, src, mrf = _case(k, B, T, seed=100 * k + T)
    for kw_tc, kw_ws in (
            (dict(post_lrelu=0.1), dict()),                                                        # conv1 of a pair
            (dict(), dict(out_lrelu=False)),
            (dict(res16=src, res16_unlrelu=10.0, post_lrelu=0.1), dict(res=src)),                  # conv2
            (dict(res16=src, res16_unlrelu=10.0, res2=mrf, scale=1 / 3, post_lrelu=0.01),
             dict(res=src, mrf=mrf, scale=1 / 3, post_slope=0.01)),                                # last pair of the last ResBlock
            (dict(res16=src, res16_unlrelu=10.0, res2=mrf), dict(res=src, mrf=mrf, out_lrelu=False))):
        o_tc = torch.full_like(x, 7.0)
        o_ws = torch.full_like(x, 9.0)
        ops.conv1d(pc, x, out16=o_tc, **kw_tc)
        ops.conv1d_ws(pw, x, o_ws, **kw_ws)
        torch.cuda.synchronize()
        assert torch.equal(o_tc, o_ws), (k, B, T, sorted(kw_ws))


def test_conv1d_ws_rejects_what_it_cannot_do():
    pc, pw, x, src, mrf = _case(7, 1, 64, seed=1)
    with pytest.raises(ValueError):
        packer.pack_conv_ws(torch.zeros(32, 32, 7), torch.zeros(32))
    with pytest.raises(RuntimeError):
        ops.conv1d_ws(pw, x[:, :63].contiguous(), torch.empty(1, 63, 64, dtype=torch.float16, device=DEV))      # odd T
    with pytest.raises(RuntimeError):
        ops.conv1d_ws(pw, x, x)                                                                                  # aliasing
