"""Latency of small (launch-bound) infer calls, eager vs cuda_graphs=True: the reference's own call pattern is one
lyric line per call at B=1 (pipline_inference_svs_v5.py:125-139).    python tools/small_batch.py"""
import os, statistics, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bench import HP, HOP, SR
from music_generation_b200 import models_ae, synthetic

dev = torch.device("cuda", 0)


def model(**kw):
    net = models_ae.SynthesizerTrn(100, 513, 48, **HP, **kw).eval()
    net.load_state_dict(synthetic.apply_recipe({k: v.clone() for k, v in net.state_dict().items()}))
    return net.to(dev)


def timeit(fn, reps=15):
    for _ in range(4):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return statistics.median(ts)


nets = {"eager": model(), "graphs": model(cuda_graphs=True)}
for B, T in ((1, 256), (1, 861), (1, 2000), (4, 861), (16, 861)):
    P = max(8, T // 10)
    inp = {k: v.to(dev) for k, v in synthetic.make_inputs(B, P, seed=5).items()}
    ph = synthetic.make_ph_dur(inp["x_lengths"].cpu(), P, T, seed=5).to(dev)
    row = []
    for name, net in nets.items():
        ms = timeit(lambda: net.infer(inp["x"], inp["x_lengths"], inp["note_pitch_id"], inp["note_duration"],
                                      phone_shengdiao=inp["phone_shengdiao"], ph_dur=ph))
        row.append(f"{name} {ms:.2f} ms ({B * T * HOP / SR / (ms / 1e3):.0f}x RT)")
    print(f"B={B} T={T}: " + " | ".join(row), flush=True)
