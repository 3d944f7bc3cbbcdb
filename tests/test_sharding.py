"""Utterance sharding: partition properties and the world_size-2 gather over gloo (CPU)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from music_generation_b200 import sharding


def test_partition_covers_each_utterance_once_and_balances():
    frames = [2584, 100, 900, 2000, 2000, 50, 1200, 861, 861, 300, 10, 2584]
    for world in (1, 2, 4, 8):
        bins = sharding.partition(frames, world)
        assert sorted(i for b in bins for i in b) == list(range(len(frames)))
        loads = [sum(sharding.utterance_cost(frames[i]) for i in b) for b in bins]
        assert max(loads) <= sum(loads) / world + sharding.utterance_cost(max(frames))
    assert sharding.partition(frames, 4) == sharding.partition(frames, 4)        # deterministic
    assert sharding.partition([], 2) == [[], []]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, frames, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        bins = sharding.partition(frames, world)
        mine = bins[rank]
        hop = 4
        smax = max(frames) * hop
        # stand-in for infer(): the waveform of utterance i is a ramp tagged with i
        wave = torch.zeros(len(mine), smax)
        lens = torch.zeros(len(mine), dtype=torch.int64)
        for n, i in enumerate(mine):
            wave[n, : frames[i] * hop] = torch.arange(frames[i] * hop, dtype=torch.float32) + 1000 * i
            lens[n] = frames[i] * hop
        res = sharding.gather_waveforms(wave, lens, mine, len(frames), smax, dst=0)
        if rank == 0:
            w, l = res
            ok = True
            for i, f in enumerate(frames):
                ok &= int(l[i]) == f * hop
                ok &= bool(torch.equal(w[i, : f * hop], torch.arange(f * hop, dtype=torch.float32) + 1000 * i))
                ok &= float(w[i, f * hop:].abs().sum()) == 0.0
            ret.put(ok)
        else:
            assert res is None
    finally:
        dist.destroy_process_group()


def test_gather_world_size_2_gloo():
    frames = [7, 3, 11, 5, 2]
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, frames, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert ret.get(timeout=5) is True


def _rows_worker(rank, world, port, frames, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        bins = sharding.partition(frames, world)
        mine = bins[rank]
        S = 16
        # int16 stand-in for song_assemble's rows: utterance i is the constant i + 1 on its valid samples
        rows = torch.zeros(len(mine), S, dtype=torch.int16)
        for n, i in enumerate(mine):
            rows[n, : frames[i]] = i + 1
        full = sharding.gather_rows(rows, bins, dst=0)
        if rank == 0:
            ok = full.shape == (len(frames), S) and full.dtype == torch.int16
            for i, f in enumerate(frames):
                ok &= bool((full[i, :f] == i + 1).all()) and bool((full[i, f:] == 0).all())
            ret.put(bool(ok))
        else:
            assert full is None
    finally:
        dist.destroy_process_group()


def test_gather_rows_world_size_2_gloo():
    """bench.py's config-3 gather: uneven bins (5 utterances on 2 ranks), int16 payload, original order on rank 0."""
    frames = [7, 3, 11, 5, 2]
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rows_worker, args=(r, 2, port, frames, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert ret.get(timeout=5) is True
