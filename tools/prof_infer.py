"""Profiling driver: one warm `infer`, then one inside a cudaProfiler range (ncu --profile-from-start off).

    python tools/prof_infer.py [--utts 4] [--frames 2048] [--dtype fp16]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from bench import HP  # noqa: E402
from music_generation_b200 import models_ae, synthetic  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--utts", type=int, default=4)
ap.add_argument("--frames", type=int, default=2048)
ap.add_argument("--dtype", default="fp16")
a = ap.parse_args()
dev = "cuda:0"
pos = max(2000, a.frames)
net = models_ae.SynthesizerTrn(100, 513, 48, **HP, max_position_embeddings=pos, operand_dtype=a.dtype).eval()
net.load_state_dict(synthetic.apply_recipe({k: v.clone() for k, v in net.state_dict().items()}, positions=pos))
net = net.to(dev)
ph = max(1, a.frames // 10)
inp = {k: v.to(dev) for k, v in synthetic.make_inputs(a.utts, ph).items()}
ph_dur = synthetic.make_ph_dur(inp["x_lengths"].cpu(), ph, a.frames).to(dev)


def run():
    return net.infer(inp["x"], inp["x_lengths"], inp["note_pitch_id"], inp["note_duration"],
                     phone_shengdiao=inp["phone_shengdiao"], ph_dur=ph_dur)


run()
torch.cuda.synchronize()
torch.cuda.profiler.start()
o = run()[0]
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok", tuple(o.shape), float(o.abs().max()))
