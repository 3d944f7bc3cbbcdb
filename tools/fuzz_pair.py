"""Random-shape check of b200svs_resblock_pair (32 and 64 channels) against the two conv1d_tc launches it replaces - bit for
bit, shapes large enough for several tiles per CTA, every case repeated (races).    python tools/fuzz_pair.py [cases]"""
import os, random, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_generation_b200 import ops, packer
DEV = "cuda"
rnd = random.Random(11)
n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 40
bad = 0
for it in range(n_cases):
    ch = rnd.choice([32, 64])
    k = rnd.choice([3, 7, 11] if ch == 32 else [3, 7])
    d = rnd.choice([1, 3, 5])
    B = rnd.randint(1, 5)
    L = rnd.choice([2, 4, 126, 128, 130, 244, 256, 258, 1000, 2 * 1237, 20000, 2 * rnd.randint(1, 40000)])
    mode = rnd.choice(["pair", "mrf_first", "mrf_mid", "mrf_last"])
    g = torch.Generator().manual_seed(it)
    w1, b1 = torch.randn(ch, ch, k, generator=g) * (ch * k) ** -0.5, torch.randn(ch, generator=g) * 0.1
    w2, b2 = torch.randn(ch, ch, k, generator=g) * (ch * k) ** -0.5, torch.randn(ch, generator=g) * 0.1
    if ch == 32:
        c1 = (packer.pack_conv_folded(w1, b1) if d == 1 else packer.pack_conv(w1, b1, dil=d)).to(DEV)
        c2 = packer.pack_conv_folded(w2, b2).to(DEV)
    else:
        c1, c2 = packer.pack_conv(w1, b1, dil=d).to(DEV), packer.pack_conv(w2, b2).to(DEV)
    fold = lambda pc, t: t.view(t.shape[0], t.shape[1] // pc.fold, t.shape[2] * pc.fold) if pc.fold > 1 else t
    x = (torch.randn(B, L, ch, generator=g) * 0.8).half().to(DEV)
    mrf = (torch.randn(B, L, ch, generator=g) * 0.5).half().to(DEV) if mode in ("mrf_mid", "mrf_last") else None
    t16, want = torch.empty_like(x), torch.empty_like(x)
    ops.conv1d(c1, fold(c1, x), out16=fold(c1, t16), post_lrelu=0.1)
    kw = dict(res16=fold(c2, x), res16_unlrelu=10.0)
    if mrf is not None:
        kw["res2"] = fold(c2, mrf)
    if mode == "pair":
        kw.update(out16=fold(c2, want), post_lrelu=0.1)
    elif mode in ("mrf_first", "mrf_mid"):
        kw.update(out16r=fold(c2, want))
    else:
        kw.update(scale=1.0 / 3, out16=fold(c2, want), post_lrelu=0.01)
    ops.conv1d(c2, fold(c2, t16), **kw)
    pkw = dict(pair=dict(out_lrelu=True, post_slope=0.1), mrf_first=dict(out_lrelu=False), mrf_mid=dict(out_lrelu=False),
               mrf_last=dict(out_lrelu=True, post_slope=0.01, scale=1.0 / 3))[mode]
    for rep in range(3):
        got = torch.full_like(x, 7.0)
        ops.resblock_pair(c1, c2, x, got, mrf=mrf, **pkw)
        torch.cuda.synchronize()
        if not torch.equal(got, want):
            bad += 1
            nz = (got != want).nonzero()
            print("FAIL", dict(ch=ch, k=k, d=d, B=B, L=L, mode=mode, rep=rep), "mismatches", len(nz), "first", nz[0].tolist(), "last", nz[-1].tolist())
            break
print(f"{n_cases} cases, {bad} failed")
