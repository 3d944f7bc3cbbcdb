#!/usr/bin/env python
"""bench.py - throughput of the VISinger synthesis forward (SynthesizerTrn.infer) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W

Metric (BASELINE.json): audio-seconds synthesised per wall-second (x real-time).  One "step" = one
pass of the hot path over BASELINE.json configs[2] - a FIXED global batch of 64 synthetic utterances
x 30 s (T = 2584 frames each), the configuration the >= 1000x RT target is quoted on.  Utterances are
independent, so with N GPUs `sharding.partition` gives every rank 64/N of them (strong scaling, no
collective inside the forward) and the int16 waveforms are gathered on rank 0 over NCCL at the end.
`value` is the whole job (64 x 30 s of audio / max-over-ranks device time) with inputs resident in
HBM; `e2e` goes through the public API from pinned HOST id tensors: H2D, `infer` on the shard, device
peak-normalise + int16 (`b200svs_song_assemble`), all_gather_into_tensor of the int16 rows, reorder,
D2H to pinned host memory on rank 0 - all inside the timed region.  `weak` (N > 1 only) keeps round
1's weak-scaling figure (64 utterances per rank).  `--impl reference` times the UNMODIFIED reference
`SynthesizerTrn.infer` (oracle/_ref, staged by oracle/stage_ref.py; the oracle port only if the stage
is missing) on the host cores on one utterance of the same workload.

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
# BASELINE.json configs[0] (the reference's CPU-runnable case): reported inside cpu_baseline
CPU_FRAMES, CPU_PHONES = 861, 86
# C-ABI entry points whose launches are tensor-core implicit-GEMM convs (the roofline kernel)
TC_KINDS = ("conv1d_tc", "resblock_pair")


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
# CPU side: the unmodified reference (oracle/_ref staged copy, or /root/reference in the build container); the oracle
# port only when neither exists - baseline only
# ---------------------------------------------------------------------------------------------------
def cpu_state_dict(positions):
    import torch
    from music_generation_b200 import models_ae, synthetic
    net = models_ae.SynthesizerTrn(100, 513, 48, **HP, max_position_embeddings=positions)
    sd = {k: v.clone() for k, v in net.state_dict().items()}
    return synthetic.apply_recipe(sd, positions=positions)


def cpu_info():
    model = "unknown"
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                model = ln.split(":", 1)[1].strip()
                break
    except OSError:
        pass
    return {"cpu_model": model, "logical_cpus": os.cpu_count()}


class CpuArm:
    """SynthesizerTrn.infer on the host cores.  kind 'reference': the reference's own modules (ref_harness);
    kind 'port': oracle.visinger_oracle (only when no reference tree / stage is present)."""

    def __init__(self, positions):
        import torch
        from oracle import ref_harness
        self.torch = torch
        self.sd = cpu_state_dict(positions)
        self.net = None
        if ref_harness.available():
            self.kind = "reference"
            self.src = "oracle/_ref (staged copy)" if ref_harness.source_kind() == "staged" else ref_harness.REF_ROOT
            self.rh = ref_harness
            self.net = ref_harness.enlarge_positions(ref_harness.build_reference("ae"), positions)
            self.net.load_state_dict(self.sd, strict=True)
        else:
            self.kind, self.src = "port", "oracle/visinger_oracle.py"

    def infer_times(self, frames, phones, reps, warmup, threads):
        from music_generation_b200 import synthetic
        torch = self.torch
        torch.set_num_threads(threads)
        inp = synthetic.make_inputs(1, phones, seed=1234)
        ph_dur = synthetic.make_ph_dur(inp["x_lengths"], phones, frames, seed=1234)
        noise = synthetic.make_noise(1, 192, frames, seed=1234)
        times = []
        with torch.no_grad():
            for i in range(warmup + reps):
                t0 = time.perf_counter()
                if self.net is not None:
                    self.rh.run_reference_infer(self.net, inp, noise, ph_dur)
                else:
                    from oracle import visinger_oracle as vo
                    vo.infer(self.sd, noise=noise, ph_dur=ph_dur, hp=HP, **inp)
                dt = time.perf_counter() - t0
                if i >= warmup:
                    times.append(dt)
        return times

    def generator_times(self, frames, reps, warmup, threads):
        """Generator.forward only (models_ae.py:327-346) on z [1,192,frames]."""
        torch = self.torch
        torch.set_num_threads(threads)
        z = torch.randn(1, 192, frames, generator=torch.Generator().manual_seed(3)) * 0.14
        times = []
        with torch.no_grad():
            for i in range(warmup + reps):
                t0 = time.perf_counter()
                if self.net is not None:
                    self.net.decoder(z)
                else:
                    from oracle import visinger_oracle as vo
                    vo.generator(self.sd, z, HP)
                dt = time.perf_counter() - t0
                if i >= warmup:
                    times.append(dt)
        return times


def run_reference_arm(args, rank):
    """`--impl reference`: the reference's own CPU forward, all host threads, one utterance (T = args.frames) of the
    arm's workload per step."""
    if rank != 0:
        return 0
    frames, phones = args.frames, args.phones
    cores = os.cpu_count() or 1
    arm = CpuArm(max(2000, frames))
    times = arm.infer_times(frames, phones, args.steps, args.warmup, cores)
    ms = 1e3 * sum(times) / len(times)
    val = audio_seconds(frames) / (ms / 1e3)
    sample = (f"1 utterance x {frames} frames ({audio_seconds(frames):.1f} s) = 1/{args.utts} of the {args.utts} x "
              f"{audio_seconds(frames):.0f} s workload per step; {arm.kind} SynthesizerTrn.infer from {arm.src}, torch fp32, "
              f"{cores} threads")
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"svs_v5_infer_{args.utts}x{audio_seconds(frames):.0f}s (bounded CPU sample: 1 utterance)",
                       "model": "VISinger ours_ae.json (SVS v5)", "frames": frames, "phonemes": phones, **cpu_info()},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": arm.kind, "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def cpu_baseline_block(args):
    """cpu_baseline of the GPU arm's line (rank 0, N = 1): ~30-60 s of CPU work in total."""
    frames, phones = args.frames, args.phones
    cores = os.cpu_count() or 1
    arm = CpuArm(max(2000, frames))
    med = statistics.median
    t_full = arm.infer_times(frames, phones, reps=3, warmup=1, threads=cores)
    v = audio_seconds(frames) / med(t_full)
    out = {"value": v, "unit": UNIT, "cores": cores, "kind": arm.kind,
           "sample": f"1 utterance x {frames} frames ({audio_seconds(frames):.1f} s) of the workload, median of 3 after 1 "
                     f"warm-up; {arm.kind} SynthesizerTrn.infer from {arm.src}, torch fp32, {cores} threads",
           **cpu_info()}
    # config 1 (B=1, T=861) on all cores and on one core; Generator only z[1,192,2048]
    t1 = arm.infer_times(CPU_FRAMES, CPU_PHONES, reps=3, warmup=1, threads=cores)
    out["config1_T861_all_cores"] = {"value": audio_seconds(CPU_FRAMES) / med(t1), "cores": cores}
    t1c = arm.infer_times(CPU_FRAMES, CPU_PHONES, reps=1, warmup=0, threads=1)
    out["config1_T861_one_core"] = {"value": audio_seconds(CPU_FRAMES) / med(t1c), "cores": 1}
    tg = arm.generator_times(2048, reps=2, warmup=1, threads=cores)
    out["generator_only_1x192x2048"] = {"value": audio_seconds(2048) / med(tg), "cores": cores}
    return out


def same_box_gpu_reference(args, batch=16):
    """the UNMODIFIED reference forward on this B200 in eager PyTorch (cuDNN / cuBLAS), fp32 with TF32 off
    and on - the 'same-box GPU reference'.  Part of the cpu_baseline leg (the only place bench.py runs oracle/): the reference's
    own SynthesizerTrn.infer, stock code path, `batch` utterances of the workload's length per call, median of 3 after 1 warm-up."""
    import torch
    frames, phones = args.frames, args.phones
    arm = CpuArm(max(2000, frames))
    if arm.net is None:
        return {"unavailable": "no reference tree / stage"}
    from music_generation_b200 import synthetic
    dev = torch.device("cuda", 0)
    net = arm.net.to(dev)
    net.text_post_net.position_ids = net.text_post_net.position_ids.to(dev)
    inp = {k: v.to(dev) for k, v in synthetic.make_inputs(batch, phones, seed=1234).items()}
    ph_dur = synthetic.make_ph_dur(inp["x_lengths"].cpu(), phones, frames, seed=1234).to(dev)
    noise = synthetic.make_noise(batch, 192, frames, seed=1234).to(dev)
    out = {"batch": batch, "frames": frames, "what": "reference SynthesizerTrn.infer (" + arm.src + "), eager PyTorch on cuda:0"}
    old = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    try:
        for name, tf32 in (("fp32", False), ("tf32", True)):
            torch.backends.cuda.matmul.allow_tf32 = tf32
            torch.backends.cudnn.allow_tf32 = tf32
            times = []
            for i in range(4):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                arm.rh.run_reference_infer(net, inp, noise, ph_dur)
                torch.cuda.synchronize()
                if i:
                    times.append(time.perf_counter() - t0)
            out[name] = {"value": batch * audio_seconds(frames) / statistics.median(times), "unit": UNIT,
                         "ms_per_call": 1e3 * statistics.median(times)}
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = old
        del net, arm
        torch.cuda.empty_cache()
    return out


# ---------------------------------------------------------------------------------------------------
# GPU side
# ---------------------------------------------------------------------------------------------------
def run_gpu_arm(args, rank, local_rank, world):
    import torch
    import torch.distributed as dist
    from music_generation_b200 import _lib, models_ae, ops, sharding, synthetic

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
    hop = net.hop

    class Job:
        """One global batch (identical on every rank: same seed), partitioned by sharding.partition; this rank's shard
        as pinned host tensors + device-resident copies."""

        def __init__(self, n_utts, seed, ragged_frames=None):
            inp = synthetic.make_inputs(n_utts, phones, seed=seed)
            if ragged_frames is None:
                ph_dur = synthetic.make_ph_dur(inp["x_lengths"], phones, frames, seed=seed)
            else:       # utterance b gets ragged_frames[b] frames over all of its phonemes
                ph_dur = torch.cat([synthetic.make_ph_dur(inp["x_lengths"][b:b + 1], phones, int(f), seed=seed + b)
                                    for b, f in enumerate(ragged_frames)], 0)
            self.frames = ph_dur.sum(1)
            self.bins = sharding.partition(self.frames.tolist(), world)
            self.mine = self.bins[rank]
            self.per = max(1, max(len(b) for b in self.bins))
            self.audio_s = audio_seconds(int(self.frames.sum()))
            sh = sharding.shard_batch(dict(inp, ph_dur=ph_dur), self.mine)
            self.host = {k: v.pin_memory() for k, v in sh.items()}
            self.dev = {k: v.to(dev) for k, v in self.host.items()}
            self.n = len(self.mine)
            self.smax = int(self.frames.max()) * hop
            # int16 path: per-rank rows [per, smax] (rows >= n stay zero), receive buffer, pinned host result on rank 0
            self.rows = torch.zeros(self.per, self.smax, device=dev, dtype=torch.int16)
            self.offsets = (torch.arange(max(self.n, 1), device=dev, dtype=torch.int64) * self.smax)[:max(self.n, 1)]
            self.recv = torch.empty(world * self.per, self.smax, device=dev, dtype=torch.int16) if world > 1 else None
            self.host_wave16 = torch.empty(n_utts, self.smax, dtype=torch.int16).pin_memory() if rank == 0 else None
            self.h2d_bytes = sum(v.numel() * v.element_size() for v in self.host.values())
            self.d2h_bytes = n_utts * self.smax * 2 if rank == 0 else 0

        def infer(self, d):
            return net.infer(d["x"], d["x_lengths"], d["note_pitch_id"], d["note_duration"],
                             phone_shengdiao=d["phone_shengdiao"], ph_dur=d["ph_dur"], noise_scale=1.0)

        def step_device(self):
            return self.infer(self.dev) if self.n else None

        def step_e2e(self):
            """host ids -> H2D -> infer -> device peak-normalise + int16 -> NCCL all-gather -> reorder -> D2H (rank 0)."""
            if self.n:
                d = {k: v.to(dev, non_blocking=True) for k, v in self.host.items()}
                res = self.infer(d)
                mel_len = d["ph_dur"].sum(1)
                ops.song_assemble(res[0], mel_len, self.offsets, self.rows.view(-1), hop=hop, lead_in=0)
            if world > 1:
                full = sharding.gather_rows(self.rows[:self.n], self.bins, dst=0, out=self.recv)
            else:
                full = self.rows[:self.n]
            if rank == 0:
                self.host_wave16.copy_(full, non_blocking=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        l0 = _lib.launches()
        ev0.record()
        for _ in range(steps):
            fn()
        ev1.record()
        barrier()
        ms = ev0.elapsed_time(ev1)
        launches = _lib.launches() - l0
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, launches

    job = Job(utts, 1234)
    warm = max(args.warmup, 3)
    for _ in range(warm):
        job.step_device()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms_total, launches = timed(job.step_device, args.steps)
    clocks = sampler.stop()
    if args.ncu_range:        # one extra step inside a cudaProfiler range (ncu --profile-from-start off); not timed
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        job.step_device()
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    for _ in range(2):
        job.step_e2e()
    ms_e2e, _ = timed(job.step_e2e, args.steps)

    ms_step = ms_total / args.steps
    value = job.audio_s / (ms_step / 1e3)
    e2e_value = job.audio_s / (ms_e2e / args.steps / 1e3)
    extra = {}

    # the same global batch with ragged lengths (20-30 s): sharding.partition has something to balance
    if not args.no_extras:
        g = torch.Generator().manual_seed(4321)
        rag = torch.randint(int(frames * 2 / 3), frames + 1, (utts,), generator=g)
        rag[0] = frames
        rjob = Job(utts, 1234, ragged_frames=rag.tolist())
        net.generator_bucket = True        # valid samples bit-identical; padded frames of short utterances are skipped
        for _ in range(2):
            rjob.step_e2e()
        ms_r, _ = timed(rjob.step_device, 3)
        ms_re, _ = timed(rjob.step_e2e, 3)
        net.generator_bucket = False
        loads = [sum(sharding.utterance_cost(int(rjob.frames[i])) for i in b) for b in rjob.bins]
        extra["ragged_20_30s"] = {"value": rjob.audio_s / (ms_r / 3 / 1e3), "e2e": rjob.audio_s / (ms_re / 3 / 1e3),
                                  "unit": UNIT, "ms_per_step": ms_r / 3, "audio_s": rjob.audio_s,
                                  "partition_imbalance": max(loads) / (sum(loads) / world),
                                  "what": "64 utterances of 20-30 s (uniform random frames), LPT partition, same e2e path, "
                                          "generator_bucket=True (length-sorted Generator groups)"}
        del rjob
    # weak scaling (round 1's SCALE figure): every rank synthesises its own `utts` utterances
    if world > 1 and not args.no_extras:
        inp = synthetic.make_inputs(utts, phones, seed=1234 + rank)
        ph = synthetic.make_ph_dur(inp["x_lengths"], phones, frames, seed=1234 + rank)
        wd = {k: v.to(dev) for k, v in dict(inp, ph_dur=ph).items()}
        wstep = lambda: job.infer(wd)
        wstep()
        ms_w, _ = timed(wstep, min(args.steps, 5))
        extra["weak"] = {"value": audio_seconds(utts * frames) * world / (ms_w / min(args.steps, 5) / 1e3), "unit": UNIT,
                         "utterances_per_gpu": utts, "ms_per_step": ms_w / min(args.steps, 5), "scaling": "weak"}
        del wd

    # song driver on this rank's shard: one ragged batch -> int16 song with lead-in -> pinned host
    if rank == 0 and world == 1 and not args.no_extras:
        from music_generation_b200 import pipeline
        syn = pipeline.Synthesiser(net)
        host_lines = [(job.host["x"][b], job.host["note_pitch_id"][b], job.host["note_duration"][b],
                       job.host["phone_shengdiao"][b]) for b in range(job.n)]
        host_song = torch.empty(pipeline.LEAD_IN + job.n * job.smax, dtype=torch.int16).pin_memory()

        def step_song():
            song = syn.syn_song(host_lines, ph_dur=job.host["ph_dur"].to(dev, non_blocking=True))
            host_song.copy_(song, non_blocking=True)
        step_song()
        ms_song, _ = timed(step_song, args.steps)
        extra["e2e_song"] = {"value": job.audio_s / (ms_song / args.steps / 1e3), "unit": UNIT,
                             "d2h_bytes_per_step": host_song.numel() * 2, "ms_per_step": ms_song / args.steps,
                             "what": "pipeline.Synthesiser.syn_song: host id tensors -> one ragged batch -> infer -> "
                                     "device peak-normalise / int16 / concatenate -> int16 song to pinned host memory"}

    # instrumented pass (outside the timed region): per-launch CUDA events around every conv launch
    roofline = None
    if rank == 0:
        _lib.PROFILE = []
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        ev0.record()
        job.step_device()
        ev1.record()
        torch.cuda.synchronize()
        prof, _lib.PROFILE = _lib.PROFILE, None
        step_ms = ev0.elapsed_time(ev1)
        agg = {}
        for kind, tag, flops, a, b, nbytes in prof:
            t = a.elapsed_time(b)
            key = (kind, tag)
            f, tt, n, by = agg.get(key, (0.0, 0.0, 0, 0.0))
            agg[key] = (f + flops, tt + t, n + 1, by + nbytes)
        is_tc = lambda k: k[0] in TC_KINDS
        tc_f = sum(v[0] for k, v in agg.items() if is_tc(k))
        tc_t = sum(v[1] for k, v in agg.items() if is_tc(k))
        tc_n = sum(v[2] for k, v in agg.items() if is_tc(k))
        peaks = load_peaks()
        ach = tc_f / (tc_t * 1e-3) / 1e12 if tc_t > 0 else 0.0
        tc_b = sum(v[3] for k, v in agg.items() if is_tc(k))
        traffic = load_traffic(f"svs_v5_infer_{job.n}x{audio_seconds(frames):.0f}s")
        roofline = {"bound": "tensor", "kernel": "conv1d_tc_kernel (+ resblock_pair_kernel, the same implicit-GEMM conv fused pairwise)",
                    "achieved": ach, "peak": peaks["tf_sustained"],
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
        extra["step_tflops"] = mflop_per_frame(frames) * 1e6 * int(job.frames.sum()) / (ms_step * 1e-3) / 1e12

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu_baseline = cpu_baseline_block(args)
        try:
            cpu_baseline["same_box_gpu_eager"] = same_box_gpu_reference(args)
        except Exception as e:                       # reported, never fatal: it is a side measurement
            cpu_baseline["same_box_gpu_eager"] = {"unavailable": f"{type(e).__name__}: {e}"[:300]}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": warm, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
                "config": {"workload": f"svs_v5_infer_{utts}x{audio_seconds(frames):.0f}s", "model": "VISinger ours_ae.json (SVS v5)",
                           "global_batch": utts, "utterances_per_gpu": [len(b) for b in job.bins],
                           "frames_per_utterance": frames, "phonemes": phones,
                           "parallelism": f"dp{world} (sharding.partition by utterance, no collective inside the forward; "
                                          f"final int16 all_gather_into_tensor over NCCL in e2e)",
                           "weights": "random init + synthetic.apply_recipe", "noise": "device randn",
                           "l2": "activations per step >> 126 MB L2 (no flush needed)"},
                "per_gpu": value / world,
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": job.h2d_bytes, "d2h_bytes_per_step": job.d2h_bytes,
                        "ms_per_step": ms_e2e / args.steps,
                        "what": "pinned host ids -> H2D -> SynthesizerTrn.infer -> device peak-normalise + int16 "
                                "(b200svs_song_assemble) -> all_gather_into_tensor (N > 1) -> int16 waveforms D2H on rank 0"},
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
    ap.add_argument("--no-extras", action="store_true", help="skip the ragged / weak / song extra measurements")
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
