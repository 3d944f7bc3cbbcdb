"""Where a B = 1 call spends its time: host enqueue time vs device time, per op vs stage API vs CUDA graph."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from common import build_model
from music_generation_b200 import _lib, synthetic
DEV = "cuda"
for T, P in ((256, 26), (861, 86)):
    inp = synthetic.make_inputs(1, P, seed=1234)
    ph = synthetic.make_ph_dur(inp["x_lengths"], P, T, seed=1234).to(DEV)
    i = {k: v.to(DEV) for k, v in inp.items()}
    for name, kw in (("per_op", dict(stage_api=False)), ("stage_api", dict(stage_api=True)), ("graph", dict(cuda_graphs=True))):
        net, _ = build_model("ae", device=DEV, **kw)
        f = lambda: net.infer(i["x"], i["x_lengths"], i["note_pitch_id"], i["note_duration"], phone_shengdiao=i["phone_shengdiao"], ph_dur=ph)
        for _ in range(5):
            f()
        torch.cuda.synchronize()
        n = 20
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = _lib.launches()
        t0 = time.perf_counter(); e0.record()
        for _ in range(n):
            f()
        e1.record(); t1 = time.perf_counter()
        torch.cuda.synchronize(); t2 = time.perf_counter()
        print(f"T={T} {name}: host enqueue {1e3*(t1-t0)/n:.2f} ms, wall {1e3*(t2-t0)/n:.2f} ms, device span {e0.elapsed_time(e1)/n:.2f} ms, launches {(_lib.launches()-l0)//n}", flush=True)
