"""Diagnostic (GPU box): log_f0 / f0_coarse accuracy of the F0 predictor modes against an fp64 evaluation of
the oracle, next to the oracle's own fp32 evaluation (what the reference computes)."""
import sys, torch
sys.path.insert(0, "tests")
from common import build_model
from music_generation_b200 import synthetic
from oracle import visinger_oracle as vo

B, P, T = 2, 150, 1500
inp = synthetic.make_inputs(B, P, seed=77, ragged=True)
ph_dur = synthetic.make_ph_dur(inp["x_lengths"], P, T, seed=77)
Tm = int(ph_dur.sum(1).max())
noise = torch.zeros(B, 192, Tm)
out = {}
for prec in ("split", "fp32"):
    net, sd = build_model("ae", device="cuda", f0_precision=prec)
    i = {k: v.cuda() for k, v in inp.items()}
    net.generator_chunk_frames = 4096
    res = net.infer(i["x"], i["x_lengths"], i["note_pitch_id"], i["note_duration"], phone_shengdiao=i["phone_shengdiao"],
                    noise_scale=1., length_scale=1, ph_dur=ph_dur.cuda(), noise=noise.cuda())
    torch.cuda.synchronize()
    out[prec] = (res[3].cpu(), net.last_aux["f0_coarse"].cpu())

def f0_oracle(sd, dt):
    sdd = {k: (v.to(dt) if v.is_floating_point() else v) for k, v in sd.items()}
    h, m = vo.text_encoder(sdd, inp["x"], inp["note_pitch_id"], inp["note_duration"].to(dt), inp["x_lengths"], inp["phone_shengdiao"])
    h_lr, mel_len, _ = vo.length_regulate(h, ph_dur)
    ym = vo.sequence_mask(mel_len, h_lr.size(2)).unsqueeze(1).to(dt)
    pos = sdd["text_post_net.position_embeddings.weight"][:h_lr.shape[-1]]
    xe = (h_lr + pos.t().unsqueeze(0)) * ym
    lf = vo.f0_predictor(sdd, xe, ym)
    return lf, vo.f0_to_coarse(torch.exp(lf) - 1.0), ym

lf64, c64, ym = f0_oracle(sd, torch.float64)
lf32, c32, _ = f0_oracle(sd, torch.float32)
valid = ym[:, 0].bool()
n = int(valid.sum())
print(f"frames={n}  log_f0 range [{lf64[valid].min():.3f},{lf64[valid].max():.3f}]")
def rep(name, lf, c):
    e = (lf.double() - lf64).abs()[valid]
    print(f"{name:14s} max|dlogf0|={e.max():.3e} rms={e.pow(2).mean().sqrt():.3e}  coarse!=fp64: {int((c != c64)[valid].sum())}"
          f"  coarse!=oracle32: {int((c != c32)[valid].sum())}", flush=True)
rep("oracle fp32", lf32, c32)
for prec in out:
    rep("gpu " + prec, *out[prec])
