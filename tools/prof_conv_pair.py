"""Two Generator-shaped conv1d_tc launches for one ncu capture: a tensor-bound one (C=256, k=11, N tile 256) and an
HBM-bound one (C=64, k=3, 16-bit residual stream: res16 in, out16r + out16 out).  3 launches each (capture the last)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_generation_b200 import ops, packer

g = torch.Generator().manual_seed(0)
B = 4


def mk(C, k, dil, rows):
    T = rows // B
    pc = packer.pack_conv(torch.randn(C, C, k, generator=g) * (C * k) ** -0.5, torch.randn(C, generator=g) * 0.1, dil=dil).to("cuda")
    x = torch.randn(B, T, C, generator=g).half().cuda()
    return pc, x, T


pc1, x1, T1 = mk(256, 11, 1, 262144)
o1 = torch.empty(B, T1, 256, device="cuda", dtype=torch.float16)
pc2, x2, T2 = mk(64, 3, 1, 1048576)
raw = torch.randn(B, T2, 64, generator=g).half().cuda()
act = torch.empty_like(raw)
for _ in range(3):
    ops.conv1d(pc1, x1, out16=o1, post_lrelu=0.1)
for _ in range(3):
    ops.conv1d(pc2, x2, res16=raw, out16r=raw, out16=act, post_lrelu=0.1)
torch.cuda.synchronize()
print("ok")
