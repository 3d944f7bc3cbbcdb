"""Waveform error of the constructor's precision knobs on the config-4 parity case (gin 256, B=2, T=2000): which 16-bit
stream costs how much of the 2e-3 max-abs budget.  python tools/diag_err.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from common import build_model, max_abs, snr_db
from music_generation_b200 import synthetic
from oracle import visinger_oracle as vo

DEV = "cuda"
gin, B, P, T = 256, 2, 200, 2000
inp = synthetic.make_inputs(B, P, seed=4, ragged=True)
ph_dur = synthetic.make_ph_dur(inp["x_lengths"], P, T, seed=4)
noise = synthetic.make_noise(B, 192, T, seed=4)
g = torch.randn(B, gin, 1, generator=torch.Generator().manual_seed(4))
_, sd = build_model("ae", gin=gin)
taps = {}
ref = vo.infer(sd, noise=noise, ph_dur=ph_dur, g=g, taps=taps, **inp)
for kw in ({}, {"generator_residual": "fp16"}, {"generator_mrf": "fp32"}, {"generator_residual": "fp16", "generator_mrf": "fp32"},
           {"generator_residual": "fp32"}, {"wn_skip": "accumulate"}, {"f0_precision": "fp32", "text_precision": "fp32"},
           *[{k: v} for k, v in (("generator_last_fp32", True),) if "--last" in sys.argv]):
    net, _ = build_model("ae", gin=gin, device=DEV, **kw)
    i = {k: v.to(DEV) for k, v in inp.items()}
    res = net.infer(i["x"], i["x_lengths"], i["note_pitch_id"], i["note_duration"], phone_shengdiao=i["phone_shengdiao"],
                    ph_dur=ph_dur.to(DEV), noise=noise.to(DEV), g=g.to(DEV), f0_score=taps["f0_coarse"].to(DEV))
    o = res[0].cpu()
    e = (o - ref[0]).abs()
    print(kw, "snr %.2f max_abs %.3e  p99.99 %.3e  z snr %.1f spec snr %.1f" % (
        snr_db(ref[0], o), max_abs(ref[0], o), float(e.flatten().kthvalue(int(e.numel() * 0.9999)).values),
        snr_db(ref[2][0], res[2][0].cpu()), snr_db(ref[4], res[4].cpu())), flush=True)
    if not kw:
        mel = taps["mel_len"]
        for b in range(B):
            n = int(mel[b]) * 256
            eb = e[b, 0]
            k = int(eb.argmax())
            print(f"  row {b}: mel_len {int(mel[b])} argmax frame {k / 256:.1f} valid max {float(eb[:n].max()):.3e} "
                  f"padded max {float(eb[n:].max()) if n < eb.numel() else 0:.3e} |ref| there {float(ref[0][b, 0, k]):.3f} "
                  f"rms err valid {float(eb[:n].pow(2).mean().sqrt()):.3e} ref rms {float(ref[0][b, 0, :n].pow(2).mean().sqrt()):.3f}")
        top = e.flatten().topk(20).indices
        print("  top-20 error frames:", sorted(set(((top % e.shape[-1]) // 256).tolist())), "rows", sorted(set((top // e.shape[-1]).tolist())))
    if "--one" in sys.argv:
        break
