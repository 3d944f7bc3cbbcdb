"""us per launch of attn_prep + attention_tc (plain fp16 and split hi/lo) at the bench shape (B x T frames,
192 channels, 2 heads), 10 back-to-back calls, CUDA events; the prep kernel is timed separately.
    python tools/attn_shapes.py [B] [T]
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_generation_b200 import _lib, ops

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
T = int(sys.argv[2]) if len(sys.argv) > 2 else 2584
dev = "cuda"
g = torch.Generator().manual_seed(0)
qkv = torch.randn(B, T, 576, generator=g).to(dev)
rel_k = (torch.randn(9, 96, generator=g) * 0.1).to(dev)
rel_v = (torch.randn(9, 96, generator=g) * 0.1).to(dev)
lens = torch.full((B,), T, dtype=torch.int32, device=dev)
out32 = torch.empty(B, T, 192, device=dev)
N = 10


def timeit(fn):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(N):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / N


res = []
for split in (False, True):
    ws = {}
    full = timeit(lambda: ops.attention_tc(qkv, out32, None, rel_k, rel_v, lens, dtype=torch.float16, split=split, ws=ws))
    code = ops.dt_code(torch.float16)
    Tp = (T + 7) // 8 * 8
    prep = timeit(lambda: _lib.call("b200svs_attn_prep", qkv.data_ptr(), 576, ws["qk_hi"].data_ptr(),
                                    ws["qk_lo"].data_ptr() if split else None, ws["vt_hi"].data_ptr(),
                                    ws["vt_lo"].data_ptr() if split else None, code, B, T, Tp, 192, None))
    flops = B * 2 * 4 * T * T * 96 * (3 if split else 1)
    res.append(f"{'split' if split else 'plain'}: total {full:.0f} us, prep {prep:.0f}, attention {full - prep:.0f} "
               f"({flops / (full - prep) / 1e6:.0f} TFLOP/s executed)")
print(f"B={B} T={T} |", " | ".join(res))
