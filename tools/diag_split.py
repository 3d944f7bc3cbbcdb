"""Diagnostic: where does the split-precision conv error come from (operand quantisation vs tensor-core
accumulation)?  Run on the GPU box."""
import sys, torch
sys.path.insert(0, "tests")
import torch.nn.functional as F
from music_generation_b200 import ops, packer

DEV = "cuda"
def rnd(*s, seed, scale=1.0):
    return torch.randn(*s, generator=torch.Generator().manual_seed(seed)) * scale

for cin, cout, k in [(192, 576, 1), (192, 768, 3), (768, 192, 3), (768, 192, 1), (384, 192, 3), (1536, 192, 3)]:
    B, T = 2, 300
    w, b = rnd(cout, cin, k, seed=40, scale=(cin * k) ** -0.5), rnd(cout, seed=41, scale=0.1)
    x = rnd(B, T, cin, seed=42)
    ref64 = F.conv1d(x.double().transpose(1, 2), w.double(), b.double(), padding=k // 2).transpose(1, 2)
    ref32 = F.conv1d(x.transpose(1, 2), w, b, padding=k // 2).transpose(1, 2)
    hi = x.half(); lo = (x - hi.float()).half()
    xs = torch.cat([hi, lo, hi], -1).to(DEV)
    pc = packer.pack_conv(w, b, split=True)
    # emulate packed operands in fp64
    pcd = pc.to(DEV)
    o32 = torch.empty(B, T, cout, device=DEV)
    ops.conv1d(pcd, xs, act=0, out32=o32)
    torch.cuda.synchronize()
    g = o32.cpu().double()
    xq = hi.double() + lo.double()
    refq = F.conv1d(xq.transpose(1, 2), w.double(), b.double(), padding=k // 2).transpose(1, 2)
    print(f"cin={cin} cout={cout} k={k} steps={3*cin*k//16}: |ref|max={ref64.abs().max():.3f} "
          f"gpu-ref64={(g-ref64).abs().max():.3e} cpu32-ref64={(ref32.double()-ref64).abs().max():.3e} "
          f"xquant-only={(refq-ref64).abs().max():.3e} mean_signed={(g-ref64).mean():.3e} "
          f"mean_signed_rel_sign={((g-ref64)*ref64.sign()).mean():.3e}", flush=True)
