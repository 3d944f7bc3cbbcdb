"""Per-tile pipeline timeline of conv1d_tc (diagnostic; GPU box).  Uses b200svs_conv1d_tc_set_trace.
Stamp slots (clock64 of the CTA's SM):
  0 producer: activation load of the tile issued     1 producer: last weight load of the tile issued
  2 mma: accumulator free (tmem_empty)               3 mma: activation tile landed (a_full)
  4 mma: last MMA + commit issued                    5 epilogue: accumulator ready (tmem_full)
  6 epilogue: tile done (warp 4)                     7 epilogue: cycles waiting for staging slots (e_full)
  8 store: first chunk of the tile ready             9 store: last chunk's store issued + slot recycled
 10 epilogue: cycles waiting for tcgen05.ld
    python tools/trace_conv.py
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from music_generation_b200 import _lib, ops, packer

DEV = "cuda"
CAP = 64
B, ROWS = 4, 524288            # rows at C=128 (Generator stage 1 of 4 x 2048 frames); scaled so bytes stay equal


def run(C, k, dil, res, block_n=None, fold=False):
    T = ROWS * 128 // C // B
    g = torch.Generator().manual_seed(0)
    w = torch.randn(C, C, k, generator=g) * (C * k) ** -0.5
    b = torch.randn(C, generator=g) * 0.1
    pc = (packer.pack_conv_folded(w, b) if fold else packer.pack_conv(w, b, dil=dil, block_n=block_n)).to(DEV)
    if fold:
        T, C = T // 2, C * 2
    x = (torch.randn(B, T, C, generator=g)).half().to(DEV)
    o16 = torch.empty(B, T, C, device=DEV, dtype=torch.float16)
    kw = dict(out16=o16, post_lrelu=0.1)
    if res == 3:       # act16: residual recovered from the activated tensor, activated result in place
        r = torch.randn(B, T, C, device=DEV).half()
        kw.update(res16=r, res16_unlrelu=10.0, out16=r)
    elif res == 2:       # the Generator's 16-bit residual stream: res16 in, raw out16r in place, activated out16
        r = torch.randn(B, T, C, device=DEV).half()
        kw.update(res16=r, out16r=r)
    elif res:
        r = torch.randn(B, T, C, device=DEV)
        kw.update(res1=r, out32=r)
    for _ in range(2):
        ops.conv1d(pc, x, **kw)
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record(); ops.conv1d(pc, x, **kw); ev[1].record(); torch.cuda.synchronize()
    us = ev[0].elapsed_time(ev[1]) * 1e3
    buf = torch.zeros(148, CAP, 16, dtype=torch.int64, device=DEV)
    _lib.call("b200svs_conv1d_tc_set_trace", buf.data_ptr(), CAP)
    ops.conv1d(pc, x, **kw)
    torch.cuda.synchronize()
    _lib.call("b200svs_conv1d_tc_set_trace", None, 0)
    t = buf.cpu().double()
    n_it = int((t[0, :, 4] > 0).sum())
    print(f"\nC={C} k={k} dil={dil} res={res} block_n={pc.block_n}: {us:.1f} us, T={T}, tiles/CTA~{n_it}")
    if COMPACT:
        tt = t[:, 4:n_it - 2]
        per = float(((tt[:, -1, 4] - tt[:, 0, 4]) / (tt.shape[1] - 1)).mean())
        mma = float((tt[:, :, 4] - tt[:, :, 3]).mean()); epi = float((tt[:, :, 6] - tt[:, :, 5]).mean())
        n_mma = 2 * k * C // 16
        bytes_tile = 256 * C * (8 if res == 2 else 16 if res else 4)
        print(f"   cycles/tile {per:.0f} | mma span {mma:.0f} ({mma / n_mma:.0f}/MMA x {n_mma}) | epi span {epi:.0f} | "
              f"slot wait {float(tt[:, :, 7].mean()):.0f} | hbm floor(24B/clk) {bytes_tile / 24:.0f} | alg {B * T * C * (bytes_tile / 256 / C) / us / 1e3:.0f} GB/s")
        return
    for cta in (0, 77):
        tt = t[cta, :n_it]
        base = tt[0, 0]
        lo, hi = 4, n_it - 2          # steady state
        per_tile = (tt[hi, 4] - tt[lo, 4]) / (hi - lo)
        d = lambda a, b_: float((tt[lo:hi, a] - tt[lo:hi, b_]).mean())
        print(f"  cta {cta}: cycles/tile {per_tile:.0f} | A issue->landed(mma sees) {d(3,0):.0f} | mma waits acc-free->A landed {d(3,2):.0f}"
              f" | mma issue span {d(4,3):.0f} | mma done->epi start {d(5,4):.0f} | epi span {d(6,5):.0f}"
              f" (slot wait {float(tt[lo:hi,7].mean()):.0f}, ld wait {float(tt[lo:hi,10].mean()):.0f})"
              f" | store first->last {d(9,8):.0f} | epi done->store done {d(9,6):.0f}")
        print(f"     prev last-MMA-issue -> acc free {float((tt[lo+1:hi+1, 2] - tt[lo:hi, 4]).mean()):.0f}")
        if cta == 0:
            for i in range(3, 7):
                print("     tile", i, [int(v - base) for v in tt[i, [0, 1, 2, 3, 4, 5, 6, 8, 9]]])


import sys as _s
COMPACT = len(_s.argv) > 1 and _s.argv[1] == 'gen'
cfgs = [(256, 3, 1, False, 128), (256, 3, 1, False, 256), (256, 11, 1, False, 128), (256, 11, 1, False, 256), (256, 7, 1, True, 128), (256, 7, 1, True, 256)] if len(_s.argv) > 1 and _s.argv[1] == 'bn' else None
if len(_s.argv) > 1 and _s.argv[1] == 'hbm':
    cfgs = [(32, 3, 1, False), (32, 3, 1, 2), (64, 3, 1, False), (64, 3, 1, 2)]
if len(_s.argv) > 1 and _s.argv[1] == 'gap':
    cfgs = [(32, 11, 1, False), (64, 7, 1, False), (128, 11, 1, False), (32, 3, 1, False)]
if len(_s.argv) > 1 and _s.argv[1] == 'act':
    cfgs = [(128, 3, 1, False), (128, 3, 1, 3), (64, 3, 1, False), (64, 3, 1, 3), (32, 3, 1, False, None, True), (32, 3, 1, 3, None, True)]
if len(_s.argv) > 1 and _s.argv[1] == 'pair':
    cfgs = [(128, 11, 1, False), (128, 7, 1, False), (128, 11, 1, 3)]
if COMPACT:
    cfgs = [(C, k, 1, r) for C in (128, 64, 32) for k in (3, 7, 11) for r in (False, 2)]
for cfg in cfgs or [(128, 3, 1, False), (128, 3, 1, True), (128, 11, 5, False), (64, 3, 1, False), (64, 7, 3, True), (32, 3, 1, False), (32, 11, 5, False), (32, 7, 1, True), (256, 3, 1, False)]:
    run(*cfg)
