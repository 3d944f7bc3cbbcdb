"""Per-tile timeline of b200svs_resblock_pair's roles (clock64 stamps).  Slots: 0 A-load issued, 1 conv1 issue start,
2 conv1 issued, 3 conv2 issue start, 4 epi1 got acc1, 5 epi1 got the mid buffer, 6 epi1 done, 12 epi2 loop top,
7 epi2 got acc2, 8 epi2 got its slot, 9 epi2 done, 10 store got slot, 11 store done.   python tools/trace_pair.py [k] [d]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_generation_b200 import _lib, ops, packer
k = int(sys.argv[1]) if len(sys.argv) > 1 else 3
d = int(sys.argv[2]) if len(sys.argv) > 2 else 1
DEV = "cuda"
B, L = 4, 524288
g = torch.Generator().manual_seed(0)
x = (torch.randn(B, L, 32, generator=g) * 0.8).half().to(DEV)
o = torch.empty_like(x)
w1, b1 = torch.randn(32, 32, k, generator=g) * (32 * k) ** -0.5, torch.randn(32, generator=g) * 0.1
c1 = (packer.pack_conv_folded(w1, b1) if d == 1 else packer.pack_conv(w1, b1, dil=d)).to(DEV)
c2 = packer.pack_conv_folded(w1, b1).to(DEV)
for _ in range(3):
    ops.resblock_pair(c1, c2, x, o)
cap = 64
buf = torch.zeros(148, cap, 16, dtype=torch.int64, device=DEV)
_lib.call("b200svs_resblock_pair_set_trace", buf.data_ptr(), cap)
ops.resblock_pair(c1, c2, x, o)
torch.cuda.synchronize()
_lib.call("b200svs_resblock_pair_set_trace", None, 0)
t = buf.cpu()
cta = 5
base = int(t[cta, 20, 0])
names = {0: "Aload", 1: "c1beg", 2: "c1iss", 3: "c2beg", 4: "e1acc", 5: "e1mid", 6: "e1end", 12: "e2top", 13: "c2top", 14: "c2iss", 15: "c1top", 7: "e2acc", 8: "e2slot", 9: "e2end", 10: "stbeg", 11: "stend"}
for it in range(20, 26):
    print(f"tile {it}: " + " ".join(f"{names[s]}={int(t[cta, it, s]) - base}" for s in (0, 1, 2, 4, 5, 6, 3, 12, 7, 8, 9, 10, 11)))
ev = sorted((int(t[cta, it, s_]) - base, f"{names[s_]}({it})") for it in range(22, 27) for s_ in names if int(t[cta, it, s_]))
print(" ".join(f"{n}@{c}" for c, n in ev if c > 0))
print("cycles per tile (c1beg to c1beg):", [(int(t[cta, i + 1, 1]) - int(t[cta, i, 1])) for i in range(20, 30)])
