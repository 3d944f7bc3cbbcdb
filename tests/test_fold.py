"""Time-pair folding of dilation-1 convs (packer.fold_time_pairs): the folded conv over [B, T/2, 2C] views must be the
same linear map as the reference Conv1d (VISinger/modules.py:191-206, padding = commons.get_padding)."""
import pytest
import torch
import torch.nn.functional as F

from music_generation_b200 import packer


@pytest.mark.parametrize("k", [1, 3, 5, 7, 11])
@pytest.mark.parametrize("cin,cout", [(32, 32), (64, 64), (8, 24)])
def test_fold_time_pairs_is_the_same_conv(k, cin, cout):
    g = torch.Generator().manual_seed(k * 100 + cin)
    w = torch.randn(cout, cin, k, generator=g, dtype=torch.float64)
    b = torch.randn(cout, generator=g, dtype=torch.float64)
    B, T = 3, 26
    x = torch.randn(B, T, cin, generator=g, dtype=torch.float64)                   # channels-last
    ref = F.conv1d(x.transpose(1, 2), w, b, padding=(k - 1) // 2).transpose(1, 2)  # [B, T, cout]
    w2, b2, pl = packer.fold_time_pairs(w, b)
    w2, b2 = w2.double(), b2.double()
    # fold_time_pairs returns fp32 copies; redo the fold in fp64 through the same index map for an exact check
    w2 = torch.zeros_like(w2)
    p = (k - 1) // 2
    for e in range(2):
        for v in range(-pl, w2.shape[2] - pl):
            for f in range(2):
                j = 2 * v + f - e + p
                if 0 <= j < k:
                    w2[e * cout:(e + 1) * cout, f * cin:(f + 1) * cin, v + pl] = w[:, :, j]
    assert torch.equal(w2.float(), packer.fold_time_pairs(w, b)[0])
    xf = x.reshape(B, T // 2, 2 * cin)                                             # the free reinterpretation
    nt = w2.shape[2]
    xp = F.pad(xf.transpose(1, 2), [pl, nt - 1 - pl])
    yf = F.conv1d(xp, w2, b.repeat(2)).transpose(1, 2)                              # [B, T/2, 2*cout]
    assert yf.shape == (B, T // 2, 2 * cout)
    assert torch.allclose(yf.reshape(B, T, cout), ref, atol=1e-12)
    assert nt == (k - p) // 2 - (-p) // 2 + 1 <= k // 2 + 2


def test_pack_conv_folded_metadata():
    w, b = torch.randn(32, 32, 11), torch.randn(32)
    pc = packer.pack_conv_folded(w, b)
    assert (pc.fold, pc.ntaps, pc.cin, pc.cout, pc.pad_left, pc.dil) == (2, 7, 64, 64, 3, 1)
    assert pc.alg_macs_per_row == 2 * 32 * 32 * 11
    assert pc.block_n == 64 and tuple(pc.w.shape) == (7, 64, 64)
