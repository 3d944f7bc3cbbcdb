"""us per launch of conv1d_tc on the Generator's ResBlock shapes (steady state, 20 back-to-back launches per
shape, CUDA events).  A/B knobs are environment variables read by the library (B200SVS_TC_WRES, B200SVS_TC_PEEK,
B200SVS_TC_ASTAGES, B200SVS_TC_SIF), so run once per setting:
    B200SVS_TC_WRES=0 python tools/conv_shapes.py
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_generation_b200 import ops, packer

DEV = "cuda"
B, ROWS = 4, 524288            # rows at C=128; scaled so every shape moves the same bytes
N = 20


def run(C, k, dil, res, fold=False):
    """res: "c" conv1 (16-bit in, activated 16-bit out), "r" old conv2 (raw 16-bit residual in / out in place +
    activated out), "a" act16 conv2 (residual recovered from the activated tensor, activated result in place)."""
    T = ROWS * 128 // C // B
    g = torch.Generator().manual_seed(0)
    w = torch.randn(C, C, k, generator=g) * (C * k) ** -0.5
    b = torch.randn(C, generator=g) * 0.1
    pc = (packer.pack_conv_folded(w, b) if fold else packer.pack_conv(w, b, dil=dil)).to(DEV)
    f = pc.fold
    x = torch.randn(B, T // f, C * f, generator=g).half().to(DEV)
    o16 = torch.empty(B, T // f, C * f, device=DEV, dtype=torch.float16)
    kw = dict(out16=o16, post_lrelu=0.1)
    nbytes = 2 * x.numel() * 2
    if res == "r":
        r = torch.randn(B, T // f, C * f, device=DEV).half()
        kw.update(res16=r, out16r=r)
        nbytes = 4 * x.numel() * 2
    elif res == "a":
        r = torch.randn(B, T // f, C * f, device=DEV).half()
        kw.update(res16=r, res16_unlrelu=10.0, out16=r)
        nbytes = 3 * x.numel() * 2
    for _ in range(3):
        ops.conv1d(pc, x, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(N):
        ops.conv1d(pc, x, **kw)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / N
    return us, nbytes / us / 1e3


tag = " ".join(f"{k[11:]}={v}" for k, v in sorted(os.environ.items()) if k.startswith("B200SVS_TC_"))
out = []
for C in (128, 64, 32):
    for k, dil in ((3, 1), (3, 5), (7, 1), (7, 5), (11, 1), (11, 5)):
        for res in (("c", "r", "a") if dil == 1 else ("c",)):
            for fold in ((False, True) if dil == 1 and C <= 64 else (False,)):
                us, gbs = run(C, k, dil, res, fold)
                out.append(f"{C}/{k}/{dil}/{res}{'F' if fold else ''}:{us:.1f}us,{gbs:.0f}GB/s")
print(f"[{tag or 'default'}]", "\n".join(out))
