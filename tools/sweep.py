"""SURVEY.md 8d configs 2, 4 and 5 on one B200 (parity-test shapes, not bench lines): Generator-only z[16,192,2048]
(fp16 / bf16 operands), v6-multispeaker stand-in (gin_channels=256, 9 segments x 2000 frames), and the
frames x batch throughput sweep.  Device-resident inputs, CUDA events, 3 warm-ups, median of `--reps` calls.

    python tools/sweep.py [--reps 7] [--out gpurun_out/sweep.json]
"""
import argparse, json, os, statistics, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from bench import HP, HOP, SR, mflop_per_frame
from music_generation_b200 import models_ae, synthetic

ap = argparse.ArgumentParser()
ap.add_argument("--reps", type=int, default=7)
ap.add_argument("--out", default="gpurun_out/sweep.json")
a = ap.parse_args()
dev = torch.device("cuda", 0)


def model(gin=0, dtype="fp16", positions=8192):
    net = models_ae.SynthesizerTrn(100, 513, 48, **HP, gin_channels=gin, max_position_embeddings=positions,
                                   operand_dtype=dtype).eval()
    net.load_state_dict(synthetic.apply_recipe({k: v.clone() for k, v in net.state_dict().items()}, positions=positions))
    return net.to(dev)


def timeit(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(a.reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return statistics.median(ts)


out = {"gpu": torch.cuda.get_device_name(0), "reps": a.reps, "points": []}
# ---- config 2: Generator only
for dt in ("fp16", "bf16"):
    net = model(dtype=dt)
    net._pack()
    z = (torch.randn(16, 2048, 192, generator=torch.Generator().manual_seed(3)) * 0.14).to(net.operand_dtype).to(dev)
    ms = timeit(lambda: net._generator(z))
    sec = 16 * 2048 * HOP / SR
    out["points"].append({"config": "2: Generator only z[16,192,2048]", "dtype": dt, "ms": ms, "x_rt": sec / (ms / 1e3),
                          "tflops": 614.9e6 * 16 * 2048 / (ms * 1e-3) / 1e12})
    print(out["points"][-1], flush=True)
    del net
# ---- config 4: v6 multispeaker stand-in, 180 s song as 9 segments of 2000 frames
net = model(gin=256)
B, T = 9, 2000
inp = {k: v.to(dev) for k, v in synthetic.make_inputs(B, 200, seed=4).items()}
ph = synthetic.make_ph_dur(inp["x_lengths"].cpu(), 200, T, seed=4).to(dev)
g = torch.randn(B, 256, 1, generator=torch.Generator().manual_seed(4)).to(dev)
ms = timeit(lambda: net.infer(inp["x"], inp["x_lengths"], inp["note_pitch_id"], inp["note_duration"],
                              phone_shengdiao=inp["phone_shengdiao"], ph_dur=ph, g=g))
out["points"].append({"config": "4: v6 multispeaker stand-in (gin 256), 9 x 2000 frames", "ms": ms, "x_rt": B * T * HOP / SR / (ms / 1e3)})
print(out["points"][-1], flush=True)
del net
# ---- config 5: frames x batch sweep
net = model()
for T in (256, 512, 1024, 2048, 4096, 8192):
    for B in (1, 4, 16, 64, 256):
        if B * T > 64 * 4096:           # keep the working set bounded
            continue
        P = max(8, T // 10)
        inp = {k: v.to(dev) for k, v in synthetic.make_inputs(B, P, seed=5).items()}
        ph = synthetic.make_ph_dur(inp["x_lengths"].cpu(), P, T, seed=5).to(dev)
        ms = timeit(lambda: net.infer(inp["x"], inp["x_lengths"], inp["note_pitch_id"], inp["note_duration"],
                                      phone_shengdiao=inp["phone_shengdiao"], ph_dur=ph))
        sec = B * T * HOP / SR
        pt = {"config": "5: sweep", "B": B, "T": T, "ms": ms, "x_rt": sec / (ms / 1e3),
              "tflops": mflop_per_frame(T) * 1e6 * B * T / (ms * 1e-3) / 1e12}
        out["points"].append(pt)
        print(pt, flush=True)
json.dump(out, open(a.out, "w"), indent=1)
