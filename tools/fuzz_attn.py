"""Random-shape check of b200svs_attention_tc against the CPU semantics (tests/emu_ops.attention); found the PV-barrier phase\nrace of attention_ts at T % 128 == 0.    python tools/fuzz_attn.py"""
import sys, random
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import torch, emu_ops
from music_generation_b200 import ops
random.seed(5)
worst = {False: 0.0, True: 0.0}
for it in range(36):
    B = random.randint(1, 3); T = random.choice([1, 2, 5, 63, 64, 65, 127, 128, 129, 191, 200, 257, 320, 449, 640])
    lens = [random.choice([T, random.randint(0, T), max(1, T // 2)]) for _ in range(B)]
    g = torch.Generator().manual_seed(it)
    sc = random.choice([0.5, 1.0, 2.0])
    qkv = torch.randn(B, T, 576, generator=g) * sc
    rel_k = torch.randn(9, 96, generator=g) * 96 ** -0.5; rel_v = torch.randn(9, 96, generator=g) * 96 ** -0.5
    ln = torch.tensor(lens, dtype=torch.int32)
    for split in (False, True):
        ref_in = qkv if split else qkv.half().float()
        e32 = torch.empty(B, T, 192); emu_ops.attention(ref_in, e32, None, rel_k, rel_v, ln)
        r32 = torch.full((B, T, 192), 7.0, device="cuda")
        ops.attention_tc(qkv.cuda(), r32, None, rel_k.cuda(), rel_v.cuda(), ln.cuda(), dtype=torch.float16, split=split)
        torch.cuda.synchronize()
        err = (e32 - r32.cpu()).abs().max().item()
        worst[split] = max(worst[split], err)
        tol = 2e-4 if split else 1e-2
        if not (err < tol):
            d = (e32 - r32.cpu()).abs().amax(dim=2)
            bad = [(b, int((d[b] > tol).nonzero()[0]), int((d[b] > tol).nonzero()[-1]), int((d[b] > tol).sum())) for b in range(B) if (d[b] > tol).any()]
            print("FAIL", B, T, lens, "scale", sc, "split", split, "err %.3g" % err, "bad rows (b, first, last, count):", bad)
print("worst plain %.3e split %.3e" % (worst[False], worst[True]))
