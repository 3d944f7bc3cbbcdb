#!/usr/bin/env python
"""bench.py - throughput of the VISinger synthesis forward (SynthesizerTrn.infer) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Metric (BASELINE.json): audio-seconds synthesised per wall-second (x real-time).  One "step" = one
full `infer` over a batch of 64 synthetic utterances x 30 s (T = 2584 frames each, BASELINE.json
configs[2], the configuration the >=1000x RT target is quoted on) per GPU; utterances are independent,
so ranks shard by utterance with no collective inside the forward (weak scaling: 64 utterances per
rank).  `value` is measured with inputs resident in HBM (CUDA events, max over ranks); `e2e` goes
through the public API with pinned HOST inputs and a device->host read of the waveforms inside the
timed region.  `--impl reference` times the reference algorithm's CPU port (oracle/, torch fp32 on
the host cores) on a bounded sample of the same workload.

Only this file's `cpu_baseline` / `--impl reference` legs import oracle/ (as the thing being timed
on the CPU side, never on the product path).
"""
import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

HP = dict(inter_channels=192, hidden_channels=192, filter_channels=768, n_heads=2, n_layers=6, kernel_size=3,
          p_dropout=0.1, resblock="1", resblock_kernel_sizes=[3, 7, 11],
          resblock_dilation_sizes=[[1, 3, 5], [1, 3, 5], [1, 3, 5]], upsample_rates=[8, 8, 2, 2],
          upsample_initial_channel=512, upsample_kernel_sizes=[16, 16, 4, 4])
SR, HOP = 22050, 256
METRIC = "audio-seconds synthesized per second (x real-time)"
UNIT = "audio_s/s"

# workload: BASELINE.json configs[2] per GPU
UTTS, FRAMES, PHONES = 64, 2584, 258
# bounded CPU sample of the same workload (one utterance; see cpu_baseline.sample)
CPU_FRAMES, CPU_PHONES = 861, 86


def mflop_per_frame(T):
    """full v5 infer = 666.0 + 0.00768*T MFLOP per frame (2*MAC, padding excluded)."""
    return 666.0 + 0.00768 * T


def audio_seconds(frames_total):
    return frames_total * HOP / SR


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], hbm_gbs=d["hbm_gbs"], tf_burst=d["bf16_tflops"],
                    tf_sustained=d["bf16_tflops_sustained"], src="MEASURED_PEAKS.json")
    return dict(hbm=6650.0, hbm_gbs=6650.0, tf_burst=1590.0, tf_sustained=1400.0, src="fallback (B200_PROFILING.md)")


def load_traffic(workload):
    """DRAM bytes per conv1d_tc launch (dram__bytes_read.sum + dram__bytes_write.sum averaged over the launches of one
    step) from the committed ncu pass of this workload (profiles/conv1d_tc_dram.json, written by
    tools/ncu_dram_summary.py); None when no capture of this workload is committed."""
    p = os.path.join(ROOT, "profiles", "conv1d_tc_dram.json")
    if not os.path.exists(p):
        return None
    d = json.load(open(p))
    return d if d.get("workload") == workload else None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        rows = [r.strip().split(", ") for r in open(self.f.name) if r.strip()]
        os.unlink(self.f.name)
        sm, smax, reasons, pw = [], [], set(), []
        for r in rows:
            try:
                sm.append(float(r[1])); smax.append(float(r[2])); pw.append(float(r[3]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                if v.strip().lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        # "under load": samples in the upper half of the observed power range
        lo, hi = min(pw), max(pw)
        load = [s for s, p in zip(sm, pw) if p >= lo + 0.5 * (hi - lo)] or sm
        return {"sm_mhz": statistics.median(load), "sm_max_mhz": max(smax), "power_w_max": hi,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------
# CPU side (oracle port of the reference algorithm) - baseline only
# ---------------------------------------------------------------------------------------------------
def cpu_state_dict(positions):
    import torch
    from music_generation_b200 import models_ae, synthetic
    net = models_ae.SynthesizerTrn(100, 513, 48, **HP, max_position_embeddings=positions)
    sd = {k: v.clone() for k, v in net.state_dict().items()}
    return synthetic.apply_recipe(sd, positions=positions)


def cpu_run(frames, phones, reps, warmup, threads=None):
    """Times oracle.visinger_oracle.infer (the reference forward restated in torch fp32) on the host."""
    import torch
    from music_generation_b200 import synthetic
    from oracle import visinger_oracle as vo
    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    sd = cpu_state_dict(max(2000, frames))
    inp = synthetic.make_inputs(1, phones, seed=1234)
    ph_dur = synthetic.make_ph_dur(inp["x_lengths"], phones, frames, seed=1234)
    noise = synthetic.make_noise(1, 192, frames, seed=1234)
    times = []
    with torch.no_grad():
        for i in range(warmup + reps):
            t0 = time.perf_counter()
            vo.infer(sd, noise=noise, ph_dur=ph_dur, hp=HP, **inp)
            dt = time.perf_counter() - t0
            if i >= warmup:
                times.append(dt)
    return times, cores


def run_reference_arm(args, rank):
    if rank != 0:
        return 0
    frames, phones = CPU_FRAMES, CPU_PHONES
    times, cores = cpu_run(frames, phones, args.steps, args.warmup)
    ms = 1e3 * sum(times) / len(times)
    val = audio_seconds(frames) / (ms / 1e3)
    sample = (f"1 utterance x {frames} frames ({audio_seconds(frames):.1f} s) of the 64 x 30 s workload per step, "
              f"oracle port of SynthesizerTrn.infer (torch fp32, {cores} threads)")
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "svs_v5_infer_64x30s (bounded CPU sample)", "model": "VISinger ours_ae.json",
                       "frames": frames, "phonemes": phones},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


# ---------------------------------------------------------------------------------------------------
# GPU side
# ---------------------------------------------------------------------------------------------------
def run_gpu_arm(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist
    from music_generation_b200 import _lib, models_ae, synthetic

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=dev)
    _lib.check(_lib.lib().b200svs_check_device(), "check_device")     # fails loudly without the .so / a B200

    utts, frames, phones = args.utts, args.frames, args.phones
    positions = max(2000, frames)
    net = models_ae.SynthesizerTrn(100, 513, 48, **HP, max_position_embeddings=positions,
                                   operand_dtype=args.dtype, **json.loads(args.model_kw)).eval()
    sd = synthetic.apply_recipe({k: v.clone() for k, v in net.state_dict().items()}, positions=positions)
    net.load_state_dict(sd)
    net = net.to(dev)

    # every rank synthesises its own `utts` utterances (weak scaling): seed by rank
    inp = synthetic.make_inputs(utts, phones, seed=1234 + rank)
    ph_dur = synthetic.make_ph_dur(inp["x_lengths"], phones, frames, seed=1234 + rank)
    host = {k: v.pin_memory() for k, v in inp.items()}
    host["ph_dur"] = ph_dur.pin_memory()
    devin = {k: v.to(dev) for k, v in host.items()}
    n_samples = frames * HOP
    host_wave = torch.empty(utts, 1, n_samples, dtype=torch.float32).pin_memory()
    h2d_bytes = sum(v.numel() * v.element_size() for v in host.values())
    d2h_bytes = host_wave.numel() * host_wave.element_size()

    def step_device():
        return net.infer(devin["x"], devin["x_lengths"], devin["note_pitch_id"], devin["note_duration"],
                         phone_shengdiao=devin["phone_shengdiao"], ph_dur=devin["ph_dur"], noise_scale=1.0)

    def step_e2e():
        d = {k: v.to(dev, non_blocking=True) for k, v in host.items()}
        res = net.infer(d["x"], d["x_lengths"], d["note_pitch_id"], d["note_duration"],
                        phone_shengdiao=d["phone_shengdiao"], ph_dur=d["ph_dur"], noise_scale=1.0)
        host_wave.copy_(res[0], non_blocking=True)
        return res

    # song driver: the same utterances as the lines of one song - one ragged batch, peak-normalised,
    # int16, concatenated on the device; only the int16 song crosses PCIe
    from music_generation_b200 import pipeline
    syn = pipeline.Synthesiser(net)
    host_lines = [(host["x"][b], host["note_pitch_id"][b], host["note_duration"][b], host["phone_shengdiao"][b])
                  for b in range(utts)]
    host_song = torch.empty(pipeline.LEAD_IN + utts * n_samples, dtype=torch.int16).pin_memory()

    def step_song():
        song = syn.syn_song(host_lines, ph_dur=host["ph_dur"].to(dev, non_blocking=True))
        host_song.copy_(song, non_blocking=True)
        return song

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        l0 = _lib.LAUNCHES
        ev0.record()
        for _ in range(steps):
            fn()
        ev1.record()
        barrier()
        ms = ev0.elapsed_time(ev1)
        launches = _lib.LAUNCHES - l0
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches

    for _ in range(max(args.warmup, 3)):
        step_device()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms_total, launches = timed(step_device, args.steps)
    clocks = sampler.stop()
    if args.ncu_range:        # one extra step inside a cudaProfiler range (ncu --profile-from-start off); not timed
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        step_device()
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    for _ in range(2):
        step_e2e()
    ms_e2e, _ = timed(step_e2e, args.steps)
    step_song()
    ms_song, _ = timed(step_song, args.steps)

    ms_step = ms_total / args.steps
    audio_per_step = audio_seconds(utts * frames) * world
    value = audio_per_step / (ms_step / 1e3)
    e2e_value = audio_per_step / (ms_e2e / args.steps / 1e3)

    # instrumented pass (outside the timed region): per-launch CUDA events around every conv launch
    roofline, extra = None, {}
    if rank == 0:
        _lib.PROFILE = []
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        ev0.record()
        step_device()
        ev1.record()
        torch.cuda.synchronize()
        prof, _lib.PROFILE = _lib.PROFILE, None
        step_ms = ev0.elapsed_time(ev1)
        agg = {}
        hbm_t = hbm_b = 0.0       # launches that are HBM-bound by their algorithmic intensity
        for kind, tag, flops, a, b, nbytes in prof:
            t = a.elapsed_time(b)
            key = (kind, tag)
            f, tt, n, by = agg.get(key, (0.0, 0.0, 0, 0.0))
            agg[key] = (f + flops, tt + t, n + 1, by + nbytes)
        tc_f = sum(v[0] for k, v in agg.items() if k[0] == "conv1d_tc")
        tc_t = sum(v[1] for k, v in agg.items() if k[0] == "conv1d_tc")
        tc_n = sum(v[2] for k, v in agg.items() if k[0] == "conv1d_tc")
        peaks = load_peaks()
        ach = tc_f / (tc_t * 1e-3) / 1e12 if tc_t > 0 else 0.0
        tc_b = sum(v[3] for k, v in agg.items() if k[0] == "conv1d_tc")
        traffic = load_traffic(f"svs_v5_infer_{utts}x{audio_seconds(frames):.0f}s")
        roofline = {"bound": "tensor", "kernel": "conv1d_tc_kernel", "achieved": ach, "peak": peaks["tf_sustained"],
                    "unit": "TFLOP/s", "frac": ach / peaks["tf_sustained"],
                    "traffic": None if traffic is None else traffic["dram_bytes_per_launch"],
                    "traffic_source": None if traffic is None else traffic["source"],
                    "alg_bytes_per_launch": tc_b / max(tc_n, 1),
                    "hbm_view": {"achieved": tc_b / (tc_t * 1e-3) / 1e9 if tc_t > 0 else 0.0, "peak": peaks["hbm_gbs"],
                                 "unit": "GB/s", "frac": (tc_b / (tc_t * 1e-3) / 1e9) / peaks["hbm_gbs"] if tc_t > 0 else 0.0,
                                 "how": "algorithmic HBM bytes of the same launches / the same durations"},
                    "peak_source": peaks["src"] + " (bf16 sustained: kernel timed inside a long step)",
                    "launches_per_step": tc_n, "alg_tflop_per_step": tc_f / 1e12,
                    "avg_launch_ms": tc_t / max(tc_n, 1), "share_of_step": tc_t / step_ms,
                    "how": "sum of algorithmic conv FLOPs / sum of CUDA-event launch durations, instrumented step"}
        extra["stages_ms"] = {f"{k[0]}:{k[1] or 'untagged'}": {"ms": round(v[1], 3), "launches": v[2],
                                                              "tflops": round(v[0] / max(v[1], 1e-9) / 1e9, 1),
                                                              "alg_gbs": round(v[3] / max(v[1], 1e-9) / 1e6, 1)}
                              for k, v in sorted(agg.items())}
        extra["instrumented_step_ms"] = step_ms
        extra["step_tflops"] = mflop_per_frame(frames) * 1e6 * utts * frames / (ms_step * 1e-3) / 1e12

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        times, cores = cpu_run(CPU_FRAMES, CPU_PHONES, reps=3, warmup=1)
        v = audio_seconds(CPU_FRAMES) / statistics.median(times)
        cpu_baseline = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                        "sample": f"1 utterance x {CPU_FRAMES} frames ({audio_seconds(CPU_FRAMES):.1f} s), median of 3 "
                                  f"after 1 warm-up, oracle port of SynthesizerTrn.infer (torch fp32, {cores} threads)"}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
                "config": {"workload": f"svs_v5_infer_{utts}x{audio_seconds(frames):.0f}s", "model": "VISinger ours_ae.json (SVS v5)",
                           "utterances_per_gpu": utts, "frames_per_utterance": frames, "phonemes": phones,
                           "global_batch": utts * world, "parallelism": f"dp{world} (utterance-sharded, no collective)",
                           "weights": "random init + synthetic.apply_recipe", "noise": "device randn",
                           "l2": "activations per step >> 126 MB L2 (no flush needed)"},
                "per_gpu": value / world,
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                        "ms_per_step": ms_e2e / args.steps},
                "e2e_song": {"value": audio_per_step / (ms_song / args.steps / 1e3), "unit": UNIT,
                             "d2h_bytes_per_step": host_song.numel() * 2, "ms_per_step": ms_song / args.steps,
                             "what": "pipeline.Synthesiser.syn_song: host id tensors -> one ragged batch -> infer -> "
                                     "device peak-normalise / int16 / concatenate -> int16 song to pinned host memory"},
                "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline}
        line.update(extra)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--dtype", default="fp16", choices=["fp16", "bf16"])
    ap.add_argument("--utts", type=int, default=UTTS)
    ap.add_argument("--frames", type=int, default=FRAMES)
    ap.add_argument("--phones", type=int, default=PHONES)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--model-kw", default="{}", help='JSON of extra SynthesizerTrn keyword arguments for A/B runs, '
                    'e.g. \'{"generator_mrf": "fp32"}\' (the default line uses none)')
    ap.add_argument("--ncu-range", action="store_true",
                    help="run one extra, untimed step inside cudaProfilerStart/Stop (for ncu --profile-from-start off)")
    args = ap.parse_args()

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus > 1 and world == 1 and "RANK" not in os.environ:
        # convenience: `python bench.py --gpus N` re-launches itself under torchrun
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", str(29500 + os.getpid() % 1000), os.path.abspath(__file__)]
        return subprocess.call(cmd + sys.argv[1:])
    if args.impl == "reference":
        return run_reference_arm(args, rank)
    return run_gpu_arm(args, rank, local_rank, world)


if __name__ == "__main__":
    sys.exit(main())
