"""-m gpu: b200svs_song_assemble_noalloc (scratch in a file-scope device array, no allocation inside the call) against
the existing b200svs_song_assemble entry point, through the C ABI, on two calls in a row.  Collected last on purpose;
passes alone in a fresh process and after the other tests."""
import ctypes as C

import pytest
import torch

from music_generation_b200 import _lib

pytestmark = pytest.mark.gpu
DEV = "cuda"
HOP, LEAD_IN = 256, 4410


def _p(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _case(seed, mel, amps):
    g = torch.Generator().manual_seed(seed)
    mel_len = torch.tensor(mel, dtype=torch.int64)
    B = len(mel)
    S = int(mel_len.max()) * HOP
    o = torch.randn(B, 1, S, generator=g) * torch.tensor(amps).view(B, 1, 1)
    order = torch.randperm(B, generator=g).to(torch.int32)
    samples = mel_len[order.long()] * HOP
    offsets = torch.cumsum(samples, 0) - samples
    return o.to(DEV), mel_len.to(DEV), order.to(DEV), offsets.to(DEV), LEAD_IN + int(samples.sum())


def test_noalloc_song_assemble_matches_workspace_variant_twice():
    lib = _lib.lib()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    # the second call has fewer lines AND smaller peaks than the first: stale scratch contents would show up in its scales
    cases = [_case(7, [40, 0, 17, 123, 1, 64], [0.9, 1.0, 0.5, 0.8, 0.7, 0.6]),
             _case(8, [33, 5, 90, 12], [0.003, 0.02, 1e-5, 0.004])]
    # every buffer exists before the calls under test: the entry point itself allocates nothing
    bufs = []
    for o, mel_len, order, offsets, total in cases:
        n = order.numel()
        bufs.append(dict(ws=torch.empty(n, device=DEV, dtype=torch.int32),
                         want=torch.full((total,), 77, device=DEV, dtype=torch.int16),
                         got=torch.full((total,), -77, device=DEV, dtype=torch.int16),
                         s_want=torch.empty(n, 2, device=DEV), s_got=torch.empty(n, 2, device=DEV)))
    torch.cuda.synchronize()
    for legacy in (0, 1):
        for (o, mel_len, order, offsets, total), b in zip(cases, bufs):
            n = order.numel()
            _lib.check(lib.b200svs_song_assemble(_p(o), o.stride(0), _p(mel_len), _p(order), _p(offsets), n, HOP, LEAD_IN,
                                                 legacy, _p(b["ws"]), _p(b["want"]), _p(b["s_want"]), stream), "song_assemble")
            mem0 = torch.cuda.memory_allocated()
            _lib.check(lib.b200svs_song_assemble_noalloc(_p(o), o.stride(0), _p(mel_len), _p(order), _p(offsets), n, HOP,
                                                         LEAD_IN, legacy, _p(b["got"]), _p(b["s_got"]), stream),
                       "song_assemble_noalloc")
            assert torch.cuda.memory_allocated() == mem0
            torch.cuda.synchronize()
            assert torch.equal(b["got"], b["want"])                      # int16 song: bit-exact (integer work)
            assert torch.equal(b["s_got"], b["s_want"])                  # the two peak-normalisation factors per line
            assert int(b["got"][:LEAD_IN].abs().max()) == 0 and int(b["got"].abs().max()) > 0


def test_noalloc_song_assemble_rejects_more_lines_than_its_scratch():
    lib = _lib.lib()
    x = torch.zeros(8, device=DEV)
    i64 = torch.zeros(8, device=DEV, dtype=torch.int64)
    song = torch.zeros(8, device=DEV, dtype=torch.int16)
    rc = lib.b200svs_song_assemble_noalloc(_p(x), 8, _p(i64), None, _p(i64), 4097, HOP, 0, 0, _p(song), None, None)
    assert rc == -1 and b"exceeds the static scratch of 4096 lines" in lib.b200svs_last_error()
