"""Utterance-sharded data parallelism: one process per GPU, no collective inside the
forward, one optional gather of the waveforms at the end.

The reference synthesises a song as independent lyric lines, one `infer` per line
(pipline_inference_svs_v5.py:206-229); utterances share nothing but read-only weights, so the work
shards by utterance.  `partition` is a longest-processing-time greedy bin-packing on the per-utterance
cost  T * (666 + 0.00768 * T)  MFLOP, deterministic for a given input.
"""
from typing import List, Sequence

import torch
import torch.distributed as dist


def utterance_cost(frames: int) -> float:
    return frames * (666.0 + 0.00768 * frames)


def partition(frames: Sequence[int], world_size: int) -> List[List[int]]:
    """Returns, per rank, the indices of the utterances it synthesises (each list sorted ascending)."""
    order = sorted(range(len(frames)), key=lambda i: (-utterance_cost(int(frames[i])), i))
    loads = [0.0] * world_size
    bins: List[List[int]] = [[] for _ in range(world_size)]
    for i in order:
        r = min(range(world_size), key=lambda k: (loads[k], k))
        bins[r].append(i)
        loads[r] += utterance_cost(int(frames[i]))
    return [sorted(b) for b in bins]


def shard_batch(inputs: dict, indices: Sequence[int]) -> dict:
    idx = torch.as_tensor(list(indices), dtype=torch.long)
    return {k: v.index_select(0, idx.to(v.device)) for k, v in inputs.items()}


def gather_waveforms(local_wave: torch.Tensor, local_len: torch.Tensor, indices: Sequence[int], n_total: int,
                     max_samples: int, dst: int = 0):
    """Collects per-utterance waveforms on rank `dst` in the original utterance order.

    local_wave [n_local, S_local] fp32, local_len [n_local] valid samples.  Every rank pads to
    `max_samples`, one all_gather moves the payload (NCCL over NVLink on GPUs, gloo on CPU tests).
    Returns (wave [n_total, max_samples], lens [n_total]) on `dst`, None elsewhere."""
    world = dist.get_world_size()
    rank = dist.get_rank()
    per = (n_total + world - 1) // world
    per = max(per, max(1, len(indices)))
    per_t = torch.tensor([per], device=local_wave.device)
    dist.all_reduce(per_t, op=dist.ReduceOp.MAX)
    per = int(per_t.item())
    buf = torch.zeros(per, max_samples, device=local_wave.device, dtype=local_wave.dtype)
    meta = torch.full((per, 2), -1, device=local_wave.device, dtype=torch.int64)
    n = local_wave.shape[0]
    if n:
        s = min(max_samples, local_wave.shape[1])
        buf[:n, :s] = local_wave[:, :s]
        meta[:n, 0] = torch.as_tensor(list(indices), device=meta.device)
        meta[:n, 1] = local_len.to(torch.int64)
    bufs = [torch.empty_like(buf) for _ in range(world)]
    metas = [torch.empty_like(meta) for _ in range(world)]
    dist.all_gather(bufs, buf)
    dist.all_gather(metas, meta)
    if rank != dst:
        return None
    wave = torch.zeros(n_total, max_samples, device=buf.device, dtype=buf.dtype)
    lens = torch.zeros(n_total, device=buf.device, dtype=torch.int64)
    for b, m in zip(bufs, metas):
        ok = m[:, 0] >= 0
        wave[m[ok, 0]] = b[ok]
        lens[m[ok, 0]] = m[ok, 1]
    return wave, lens


def gather_rows(local: torch.Tensor, bins: Sequence[Sequence[int]], dst: int = 0, out: torch.Tensor = None):
    """Final gather of config 3: every rank holds the rows (e.g. int16 waveforms [n_local, S]) of the
    utterances `bins[rank]` (the deterministic `partition` every rank computes for itself, so no metadata travels);
    ONE `all_gather_into_tensor` of equally padded blocks moves the payload (NCCL over NVLink on GPUs, gloo in the
    CPU tests) and `dst` returns the rows in the original utterance order [n_total, S]; other ranks return None.
    `out` (optional, [world * per, S]) is a reusable receive buffer."""
    world, rank = dist.get_world_size(), dist.get_rank()
    if local.dtype == torch.int16:        # neither NCCL nor gloo moves int16: the payload travels as bytes
        res = gather_rows(local.view(torch.uint8), bins, dst, None if out is None else out.view(torch.uint8))
        return None if res is None else res.view(torch.int16)
    per = max(1, max(len(b) for b in bins))
    n = local.shape[0]
    if n != len(bins[rank]):
        raise ValueError(f"rank {rank} holds {n} rows but its bin has {len(bins[rank])} utterances")
    if n == per:
        send = local.contiguous()
    else:
        send = local.new_zeros((per,) + tuple(local.shape[1:]))
        send[:n] = local
    if out is None:
        out = local.new_empty((world * per,) + tuple(local.shape[1:]))
    dist.all_gather_into_tensor(out, send)
    if rank != dst:
        return None
    n_total = sum(len(b) for b in bins)
    perm = torch.empty(n_total, dtype=torch.long)
    for r, b in enumerate(bins):
        for slot, i in enumerate(b):
            perm[i] = r * per + slot
    return out.index_select(0, perm.to(out.device))
