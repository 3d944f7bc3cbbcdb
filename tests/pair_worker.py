"""Worker of tests/test_gpu_pair.py (not a test module): runs a set of conv1d_tc launches and saves their outputs.
The library reads B200SVS_TC_PAIR* once per process, so the single-CTA and the CTA-pair runs are two processes.
    python tests/pair_worker.py out.pt"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from music_generation_b200 import ops, packer  # noqa: E402

DEV = "cuda"
out = {}


def rnd(*shape, seed, scale=1.0):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale


# (name, Cin, Cout, k, dil, B, T, kind)
CASES = [
    ("gen128_k7_conv1", 128, 128, 7, 3, 3, 1000, "c"),          # odd tile count: phantom tile
    ("gen128_k11_act16", 128, 128, 11, 1, 2, 1500, "a"),
    ("gen256_k3_conv1", 256, 256, 3, 1, 3, 700, "c"),
    ("gen256_k7_act16", 256, 256, 7, 1, 2, 900, "a"),
    ("wn_gate_k5", 192, 384, 5, 1, 4, 1111, "gate"),             # N = 192 x 2 tiles, sequence mask, ragged lengths
    ("ffn2_k3_mask", 768, 192, 3, 1, 4, 900, "mask32"),          # fp32 residual in place + mask
    ("ffn1_split", 576, 768, 3, 1, 2, 640, "split"),             # 3 N tiles of 256, [hi | lo | hi] output
    ("pre_k7_2tiles", 192, 512, 7, 1, 2, 1024, "c"),
]
for i, (name, cin, cout, k, dil, B, T, kind) in enumerate(CASES):
    w = rnd(cout, cin, k, seed=100 + i, scale=(cin * k) ** -0.5)
    b = rnd(cout, seed=200 + i, scale=0.1)
    x = rnd(B, T, cin, seed=300 + i).half().to(DEV)
    lens = torch.tensor([T - 37 * j for j in range(B)], dtype=torch.int32, device=DEV)
    if kind == "gate":
        pc = packer.pack_conv(w, b, gate=True).to(DEV)
        o16 = torch.empty(B, T, cout // 2, device=DEV, dtype=torch.float16)
        ops.conv1d(pc, x, out16=o16, lens=lens)
        res = [o16]
    elif kind == "mask32":
        pc = packer.pack_conv(w, b, dil=dil).to(DEV)
        r = rnd(B, T, cout, seed=400 + i).to(DEV)
        o16 = torch.empty(B, T, cout, device=DEV, dtype=torch.float16)
        ops.conv1d(pc, x, res1=r, out32=r, out16=o16, lens=lens)
        res = [r, o16]
    elif kind == "split":
        pc = packer.pack_conv(w, b, dil=dil).to(DEV)
        o16 = torch.empty(B, T, 3 * cout, device=DEV, dtype=torch.float16)
        ops.conv1d(pc, x, act=ops.ACT_RELU, out16=o16, split16=True, lens=lens)
        res = [o16]
    elif kind == "a":
        pc = packer.pack_conv(w, b, dil=dil).to(DEV)
        r = torch.nn.functional.leaky_relu(rnd(B, T, cout, seed=400 + i), 0.1).half().to(DEV)
        ops.conv1d(pc, x, res16=r, res16_unlrelu=10.0, out16=r, post_lrelu=0.1)
        res = [r]
    else:
        pc = packer.pack_conv(w, b, dil=dil).to(DEV)
        o16 = torch.empty(B, T, cout, device=DEV, dtype=torch.float16)
        ops.conv1d(pc, x, out16=o16, post_lrelu=0.1)
        res = [o16]
    torch.cuda.synchronize()
    out[name] = [t.cpu() for t in res]
torch.save(out, sys.argv[1])
print("ok", len(out))
