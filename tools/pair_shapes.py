"""us per ResBlock pair of the 32-channel stage: two conv1d_tc launches vs the fused b200svs_resblock_pair kernel
(134 MB tensors, 20 back-to-back pairs, CUDA events).   python tools/pair_shapes.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_generation_b200 import ops, packer
DEV = "cuda"
B, L = 4, 524288
N = 20
g = torch.Generator().manual_seed(0)
x = (torch.randn(B, L, 32, generator=g) * 0.8).half().to(DEV)
t16, o1, o2 = torch.empty_like(x), torch.empty_like(x), torch.empty_like(x)


def timeit(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(N):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / N


tot2 = totf = 0.0
for k in (3, 7, 11):
    for d in (1, 3, 5):
        w1, b1 = torch.randn(32, 32, k, generator=g) * (32 * k) ** -0.5, torch.randn(32, generator=g) * 0.1
        w2, b2 = torch.randn(32, 32, k, generator=g) * (32 * k) ** -0.5, torch.randn(32, generator=g) * 0.1
        c1 = (packer.pack_conv_folded(w1, b1) if d == 1 else packer.pack_conv(w1, b1, dil=d)).to(DEV)
        c2 = packer.pack_conv_folded(w2, b2).to(DEV)
        f = lambda pc, t: t.view(B, L // pc.fold, 32 * pc.fold)

        def two():
            ops.conv1d(c1, f(c1, x), out16=f(c1, t16), post_lrelu=0.1)
            ops.conv1d(c2, f(c2, t16), res16=f(c2, x), res16_unlrelu=10.0, out16=f(c2, o1), post_lrelu=0.1)
        a = timeit(two)
        b = timeit(lambda: ops.resblock_pair(c1, c2, x, o2))
        tot2 += a; totf += b
        print(f"k={k} d={d}: two launches {a:.1f} us, fused {b:.1f} us ({4 * x.numel() / b / 1e3:.0f} GB/s algorithmic)  equal={torch.equal(o1, o2)}", flush=True)
print(f"ResBlock set (9 pairs): {tot2:.0f} -> {totf:.0f} us")
