"""us per launch of b200svs_layernorm at the bench shape (64 x 2584 rows, 192 channels), fp32 + 16-bit / split outputs.
python tools/ln_shapes.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_generation_b200 import ops
B, T, C = 64, 2584, 192
g = torch.Generator().manual_seed(0)
x, res = torch.randn(B, T, C, generator=g).cuda(), torch.randn(B, T, C, generator=g).cuda()
ga, be = torch.randn(C, generator=g).cuda(), torch.randn(C, generator=g).cuda()
lens = torch.full((B,), T, dtype=torch.int32, device="cuda")
o32 = torch.empty(B, T, C, device="cuda")
for name, o16 in (("fp16", torch.empty(B, T, C, device="cuda", dtype=torch.float16)),
                  ("split", torch.empty(B, T, 3 * C, device="cuda", dtype=torch.float16))):
    kw = dict(out32=o32, out16=o16, res=res)
    if name == "split":
        kw["split16"] = True
    try:
        f = lambda: ops.layernorm(x, ga, be, lens, **kw)
        f()
    except TypeError:
        kw.pop("split16", None)
        f = lambda: ops.layernorm(x, ga, be, lens, **kw)
    for _ in range(2):
        f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        f()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 100
    nbytes = x.numel() * 4 * 3 + o16.numel() * 2
    print(f"{name}: {us:.1f} us per launch, {nbytes / us / 1e6:.2f} TB/s")
