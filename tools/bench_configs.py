"""BASELINE.json configs 2, 4 and 5 at 1/2/4/8 GPUs (parity-test shapes, not the bench line - bench.py is config 3).

    python tools/bench_configs.py --config 2|4|5 [--steps 5]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/bench_configs.py --config 4 --steps 5

config 2: Generator only, z [16,192,2048], the 16 utterances sharded over the ranks (strong scaling).
config 4: v6-multispeaker stand-in (gin_channels 256): 8 songs x 180 s (T = 15 504 frames each), every song ONE
          utterance synthesised by infer_long (segments of 1800 + 2 x 100 frames), songs sharded over the ranks.
config 5: frames x batch sweep of the full v5 infer; every rank runs the same point (weak scaling, per-GPU figure).
One JSON line per config on rank 0 (same timing rules as bench.py: barrier + synchronize, CUDA events, max over ranks).
"""
import argparse
import json
import os
import statistics
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

from bench import HP, HOP, SR, ClockSampler, load_peaks, mflop_per_frame
from music_generation_b200 import _lib, models_ae, sharding, synthetic

ap = argparse.ArgumentParser()
ap.add_argument("--config", type=int, required=True, choices=[2, 4, 5])
ap.add_argument("--steps", type=int, default=5)
ap.add_argument("--dtype", default="fp16")
a = ap.parse_args()
world, rank, local = (int(os.environ.get(k, d)) for k, d in (("WORLD_SIZE", "1"), ("RANK", "0"), ("LOCAL_RANK", "0")))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)


def model(gin=0, positions=2000):
    net = models_ae.SynthesizerTrn(100, 513, 48, **HP, gin_channels=gin, max_position_embeddings=positions,
                                   operand_dtype=a.dtype).eval()
    net.load_state_dict(synthetic.apply_recipe({k: v.clone() for k, v in net.state_dict().items()}, positions=positions))
    return net.to(dev)


def timed(fn, steps, warm=3):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    l0 = _lib.launches()
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
    return ms, (_lib.launches() - l0) // steps


def emit(**kw):
    if rank == 0:
        print(json.dumps(kw), flush=True)


peaks = load_peaks()
if a.config == 2:
    B, T = 16, 2048
    net = model()
    net._pack()
    z = (torch.randn(B, T, 192, generator=torch.Generator().manual_seed(3)) * 0.14).to(net.operand_dtype)
    mine = sharding.partition([T] * B, world)[rank]
    zz = z[mine].to(dev)
    sampler = ClockSampler(local); sampler.start()
    ms, launches = timed(lambda: net._generator(zz) if len(mine) else None, a.steps)
    clocks = sampler.stop()
    sec = B * T * HOP / SR
    tf = 614.9e6 * B * T / (ms * 1e-3) / 1e12
    emit(config="2: Generator only z[16,192,2048]", n_gpus=world, scaling="strong", value=sec / (ms / 1e3), unit="audio_s/s",
         ms_per_step=ms, dtype=a.dtype, utterances_per_gpu=[len(b) for b in sharding.partition([T] * B, world)],
         gpu_launches=launches, clocks=clocks,
         roofline={"bound": "tensor", "achieved": tf / world, "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
                   "frac": tf / world / peaks["tf_sustained"], "how": "614.9 MFLOP/frame x frames / step time, per GPU"})
elif a.config == 4:
    songs, T, P, gin = 8, 15504, 1550, 256
    net = model(gin=gin)
    inp = synthetic.make_inputs(songs, P, seed=4)
    ph = synthetic.make_ph_dur(inp["x_lengths"], P, T, seed=4)
    g = torch.randn(songs, gin, 1, generator=torch.Generator().manual_seed(4))
    bins = sharding.partition(ph.sum(1).tolist(), world)
    mine = bins[rank]
    sh = {k: v.to(dev) for k, v in sharding.shard_batch(dict(inp, ph_dur=ph, g=g), mine).items()}

    def step():
        if not mine:
            return
        net.infer_long(sh["x"], sh["x_lengths"], sh["note_pitch_id"], sh["note_duration"], sh["phone_shengdiao"],
                       ph_dur=sh["ph_dur"], g=sh["g"], segment_frames=1800, overlap=100)
    sampler = ClockSampler(local); sampler.start()
    ms, launches = timed(step, a.steps)
    clocks = sampler.stop()
    sec = songs * T * HOP / SR
    # algorithmic work: conv / linear 666.0 MFLOP per frame + attention 7680 x (segment context length) FLOP per frame
    ctx = 2000.0
    tf = (666.0e6 + 7680.0 * ctx) * songs * T / (ms * 1e-3) / 1e12
    emit(config="4: v6 multispeaker stand-in (gin 256), 8 songs x 180 s, infer_long segments 1800 + 2 x 100", n_gpus=world,
         scaling="strong", value=sec / (ms / 1e3), unit="audio_s/s", ms_per_step=ms, dtype=a.dtype,
         songs_per_gpu=[len(b) for b in bins], frames_per_song=T, gpu_launches=launches, clocks=clocks,
         roofline={"bound": "tensor", "achieved": tf / world, "peak": peaks["tf_sustained"], "unit": "TFLOP/s",
                   "frac": tf / world / peaks["tf_sustained"],
                   "how": "(666.0 MFLOP + 7680 x 2000 FLOP) per frame x frames / step time, per GPU (whole step, not one kernel)"})
else:
    net = model(positions=8192)
    pts = []
    for T in (256, 512, 1024, 2048, 4096, 8192):
        for B in (1, 4, 16, 64, 256):
            if B * T > 64 * 4096:
                continue
            P = max(8, T // 10)
            inp = {k: v.to(dev) for k, v in synthetic.make_inputs(B, P, seed=5 + rank).items()}
            ph = synthetic.make_ph_dur(inp["x_lengths"].cpu(), P, T, seed=5 + rank).to(dev)
            ms, launches = timed(lambda: net.infer(inp["x"], inp["x_lengths"], inp["note_pitch_id"], inp["note_duration"],
                                                   phone_shengdiao=inp["phone_shengdiao"], ph_dur=ph), a.steps)
            sec = B * T * HOP / SR * world
            tf = mflop_per_frame(T) * 1e6 * B * T / (ms * 1e-3) / 1e12
            pts.append({"B": B, "T": T, "ms": ms, "x_rt": sec / (ms / 1e3), "x_rt_per_gpu": sec / world / (ms / 1e3),
                        "tflops_per_gpu": tf, "frac": tf / peaks["tf_sustained"], "launches": launches})
            if rank == 0:
                print(pts[-1], file=sys.stderr, flush=True)
    emit(config="5: frames x batch sweep, full v5 infer", n_gpus=world, scaling="weak", unit="audio_s/s", dtype=a.dtype, points=pts)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
