"""us per launch of b200svs_conv_post_tanh (32 -> 1 channel, 7 taps + tanh) on a 339 MB input.   python tools/post_shapes.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_generation_b200 import ops
B, L = 8, 661504
g = torch.Generator().manual_seed(0)
x = (torch.randn(B, L, 32, generator=g) * 0.5).half().cuda()
out = torch.empty(B, L, device="cuda")
for kt in (7, 5):
    w = (torch.randn(kt, 32, generator=g) * 0.1).cuda()
    for _ in range(2):
        ops.conv_post_tanh(x, w, out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        ops.conv_post_tanh(x, w, out)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 100
    print(f"taps {kt}: {us:.1f} us per launch, {x.numel() * 2 / us / 1e6:.2f} TB/s of input")
