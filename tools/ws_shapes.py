"""conv1d_ws (weight-stationary, A operand in tensor memory) against conv1d_tc on the 64-channel dilation-1 ResBlock
convs: max abs difference and us per launch on 134 MB tensors.    python tools/ws_shapes.py [k]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_generation_b200 import ops, packer
DEV = "cuda"
ks = [int(sys.argv[1])] if len(sys.argv) > 1 else [3, 7, 11]
B, T = 4, 262144
g = torch.Generator().manual_seed(0)
x = (torch.randn(B, T, 64, generator=g) * 0.8).half().to(DEV)
src = (torch.randn(B, T, 64, generator=g) * 0.8).half().to(DEV)
mrf = (torch.randn(B, T, 64, generator=g) * 0.8).half().to(DEV)


def timeit(fn, n=10):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


for k in ks:
    w = torch.randn(64, 64, k, generator=g) * (64 * k) ** -0.5
    b = torch.randn(64, generator=g) * 0.1
    pc = packer.pack_conv(w, b).to(DEV)
    pw = packer.pack_conv_ws(w, b).to(DEV)
    for name, kw_tc, kw_ws in (
            ("conv1 (lrelu out)", dict(post_lrelu=0.1), dict()),
            ("conv2 (res)", dict(res16=src, res16_unlrelu=10.0, post_lrelu=0.1), dict(res=src)),
            ("conv2 (res + mrf, scale)", dict(res16=src, res16_unlrelu=10.0, res2=mrf, scale=1 / 3, post_lrelu=0.01),
             dict(res=src, mrf=mrf, scale=1 / 3, post_slope=0.01))):
        o_tc, o_ws = torch.empty_like(x), torch.empty_like(x)
        kw_tc = dict(kw_tc); kw_tc["out16"] = o_tc
        f_tc = lambda: ops.conv1d(pc, x, **kw_tc)
        f_ws = lambda: ops.conv1d_ws(pw, x, o_ws, **kw_ws)
        f_tc(); f_ws(); torch.cuda.synchronize()
        d = (o_tc.float() - o_ws.float()).abs()
        print(f"k={k} {name}: max abs diff {d.max().item():.3e} (mean {d.mean().item():.2e}, out max {o_tc.float().abs().max().item():.2f}) | "
              f"conv1d_tc {timeit(f_tc):.1f} us, conv1d_ws {timeit(f_ws):.1f} us")
