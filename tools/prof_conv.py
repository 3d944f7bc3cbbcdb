"""One Generator-shaped conv1d_tc launch for ncu: python tools/prof_conv.py C k dil res [rows] [fold]
   res: 0 none (conv1), 1 fp32 residual in place, 2 act16 conv2 (residual from the activated tensor, result in place)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_generation_b200 import ops, packer

C, k, dil, res = (int(v) for v in sys.argv[1:5])
rows = int(sys.argv[5]) if len(sys.argv) > 5 and int(sys.argv[5]) else 524288 * 128 // C
fold = len(sys.argv) > 6 and int(sys.argv[6])
B = 4
T = rows // B
g = torch.Generator().manual_seed(0)
w, b = torch.randn(C, C, k, generator=g) * (C * k) ** -0.5, torch.randn(C, generator=g) * 0.1
pc = (packer.pack_conv_folded(w, b) if fold else packer.pack_conv(w, b, dil=dil)).to("cuda")
if fold:
    T, C = T // 2, C * 2
x = torch.randn(B, T, C, generator=g).half().cuda()
o16 = torch.empty(B, T, C, device="cuda", dtype=torch.float16)
kw = dict(out16=o16, post_lrelu=0.1)
if res == 2:
    r = torch.randn(B, T, C, device="cuda").half()
    kw.update(res16=r, res16_unlrelu=10.0, out16=r)
elif res:
    r = torch.randn(B, T, C, device="cuda")
    kw.update(res1=r, out32=r)
for _ in range(3):
    ops.conv1d(pc, x, **kw)
torch.cuda.synchronize()
print("ok")
