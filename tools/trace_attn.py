"""Per-tile timeline of attention_ts (clock64 stamps of CTA (1,0,0)).  Softmax warp 0: 0 loop top, 1 got S_j, 2 S in
registers, 3 max exchanged, 4 exp done, 5 P published, 6 got PV_{j-1}, 7 folded.  MMA thread: 8 S_j issue top, 9 got K_j,
10 PV_j top, 11 got P_j, 12 PV_j issued.     B200SVS_ATTN_TS=2 python tools/trace_attn.py [split] [parts]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_generation_b200 import _lib, ops
split = len(sys.argv) > 1 and sys.argv[1] == "1"
B, T = 64, 2584
g = torch.Generator().manual_seed(0)
qkv = torch.randn(B, T, 576, generator=g).cuda()
rel_k = (torch.randn(9, 96, generator=g) * 0.1).cuda()
rel_v = (torch.randn(9, 96, generator=g) * 0.1).cuda()
lens = torch.full((B,), T, dtype=torch.int32, device="cuda")
out32 = torch.empty(B, T, 192, device="cuda")
ws = {}
for _ in range(2):
    ops.attention_tc(qkv, out32, None, rel_k, rel_v, lens, dtype=torch.float16, split=split, ws=ws)
cap = 48
buf = torch.zeros(cap, 16, dtype=torch.int64, device="cuda")
_lib.call("b200svs_attention_set_trace", buf.data_ptr(), cap)
ops.attention_tc(qkv, out32, None, rel_k, rel_v, lens, dtype=torch.float16, split=split, ws=ws)
torch.cuda.synchronize()
_lib.call("b200svs_attention_set_trace", None, 0)
t = buf.cpu()
names = ["top", "gotS", "Sreg", "max", "exp", "Ppub", "gotO", "fold", "Mtop", "Siss", "early", "gotP", "PViss"]
base = int(t[20, 0])
for j in range(20, 26):
    print(f"tile {j}: " + " ".join(f"{names[s]}={int(t[j, s]) - base}" for s in range(13)))
print("cycles per tile (softmax top to top):", [int(t[j + 1, 0]) - int(t[j, 0]) for j in range(10, 38)])
print("prologue: setup done -> Q landed", int(t[1, 14]) - int(t[0, 14]), "-> band done", int(t[1, 15]) - int(t[1, 14]), "-> first top", int(t[0, 0]) - int(t[1, 15]))
print("CTA: setup done -> first tile top", int(t[0, 0]) - int(t[0, 14]), "| main loop", int(t[0, 13]) - int(t[0, 0]), "| epilogue", int(t[0, 15]) - int(t[0, 13]),
      "| tiles", sum(1 for j in range(cap) if int(t[j, 0])))
import time
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    ops.attention_tc(qkv, out32, None, rel_k, rel_v, lens, dtype=torch.float16, split=split, ws=ws)
e1.record(); torch.cuda.synchronize()
print("us per call (with prep):", e0.elapsed_time(e1) * 100)
