"""CTA-pair (cta_group::2) conv1d_tc vs the single-CTA path: same bits, us per launch.
    B200SVS_TC_PAIR=0 python tools/pair_check.py save /tmp/p0.pt ; B200SVS_TC_PAIR=1 python tools/pair_check.py save /tmp/p1.pt
    python tools/pair_check.py cmp /tmp/p0.pt /tmp/p1.pt"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

if sys.argv[1] == "cmp":
    a, b = torch.load(sys.argv[2]), torch.load(sys.argv[3])
    for k in a:
        same = torch.equal(a[k]["out"], b[k]["out"])
        print(f"{k}: equal={same} max|d|={float((a[k]['out'].float() - b[k]['out'].float()).abs().max()):.3e}  "
              f"{a[k]['us']:.1f} us -> {b[k]['us']:.1f} us")
    sys.exit(0)

from music_generation_b200 import ops, packer
DEV = "cuda"
C = int(os.environ.get("PC_C", "128"))
B, T = 5, (104000 + 77) * 128 // C       # odd tile count, ragged tail
res = {}
for k, dil, mode in ((3, 1, "c"), (7, 3, "c"), (11, 5, "c"), (3, 1, "a"), (7, 1, "a"), (11, 1, "a")):
    g = torch.Generator().manual_seed(k)
    gg = torch.Generator(device=DEV).manual_seed(k)
    w = torch.randn(C, C, k, generator=g) * (C * k) ** -0.5
    b = torch.randn(C, generator=g) * 0.1
    pc = packer.pack_conv(w, b, dil=dil).to(DEV)
    x = torch.randn(B, T, C, generator=gg, device=DEV).half()
    o = torch.empty(B, T, C, device=DEV, dtype=torch.float16)
    r0 = torch.randn(B, T, C, generator=gg, device=DEV).half()

    def run():
        if mode == "a":
            r = r0.clone()
            ops.conv1d(pc, x, res16=r, res16_unlrelu=10.0, out16=r, post_lrelu=0.1)
            return r
        ops.conv1d(pc, x, out16=o, post_lrelu=0.1)
        return o
    out = run().clone()
    torch.cuda.synchronize()
    kw = dict(out16=o, post_lrelu=0.1) if mode == "c" else dict(res16=r0, res16_unlrelu=10.0, out16=o, post_lrelu=0.1)
    for _ in range(3):
        ops.conv1d(pc, x, **kw)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        ops.conv1d(pc, x, **kw)
    e1.record(); torch.cuda.synchronize()
    res[f"C{C} k{k} d{dil} {mode}"] = {"out": out.cpu(), "us": e0.elapsed_time(e1) * 100}
    print(f"C{C} k{k} d{dil} {mode}: {res[f'C{C} k{k} d{dil} {mode}']['us']:.1f} us", flush=True)
torch.save(res, sys.argv[2])
