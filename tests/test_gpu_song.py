"""-m gpu: song assembly on the B200 (b200svs_song_assemble, pipeline.Synthesiser) vs the CPU oracle - int16 bit-exact."""
import numpy as np
import pytest
import torch

from common import SNR_DB_TOL, build_model, snr_db
from music_generation_b200 import ops, pipeline, synthetic
from oracle import song_oracle as so
from oracle import visinger_oracle as vo

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("legacy", [False, True])
def test_song_assemble_bit_exact(legacy):
    g = torch.Generator().manual_seed(7)
    B, hop = 6, 256
    mel_len = torch.tensor([40, 0, 17, 123, 1, 64], dtype=torch.int64)
    S = int(mel_len.max()) * hop
    o = torch.randn(B, 1, S, generator=g) * torch.tensor([0.3, 1.0, 0.004, 0.9, 0.05, 1e-5]).view(B, 1, 1)
    order = torch.tensor([3, 0, 5, 2, 4, 1], dtype=torch.int32)
    samples = (mel_len[order.long()] * hop)
    offsets = torch.cumsum(samples, 0) - samples
    song = torch.full((so.LEAD_IN + int(samples.sum()),), 77, dtype=torch.int16, device=DEV)
    scales = torch.empty(B, 2, device=DEV)
    ops.song_assemble(o.to(DEV), mel_len.to(DEV), offsets.to(DEV), song, order=order.to(DEV), hop=hop, lead_in=so.LEAD_IN,
                      legacy_scalar=legacy, scales_out=scales)
    torch.cuda.synchronize()
    lines = [o[b, 0, :int(mel_len[b]) * hop].numpy() for b in order.tolist()]
    want = so.assemble_song(lines, legacy_scalar=legacy)
    assert np.array_equal(song.cpu().numpy(), want)
    for i, a in enumerate(lines):
        if a.size:
            assert float(scales[i, 0]) == float(so.peak_scale(np.max(np.abs(a)), legacy))


def test_synthesiser_song_matches_oracle_pipeline():
    """All lines of a song in one ragged batch: the int16 song equals the oracle's assembly of the device waveform
    bit for bit, and matches the assembly of the fp32 reference waveform to the waveform tolerance."""
    net, sd = build_model("ae", device=DEV)
    B, P, T = 4, 24, 220
    inp = synthetic.make_inputs(B, P, seed=11, ragged=True)
    ph_dur = synthetic.make_ph_dur(inp["x_lengths"], P, T, seed=11)
    Tm = int(ph_dur.sum(1).max())
    noise = synthetic.make_noise(B, 192, Tm, seed=11)
    lines = [(inp["x"][b, :n], inp["note_pitch_id"][b, :n], inp["note_duration"][b, :n], inp["phone_shengdiao"][b, :n])
             for b, n in enumerate(inp["x_lengths"].tolist())]
    syn = pipeline.Synthesiser(net)
    song = syn.syn_song(lines, noise=noise.to(DEV), ph_dur=ph_dur.to(DEV))
    torch.cuda.synchronize()
    mel_len = ph_dur.sum(1)
    assert torch.equal(syn.last["mel_len"].cpu(), mel_len)
    o = syn.last["o"].cpu()
    own = so.assemble_song([o[b, 0, :int(mel_len[b]) * 256].numpy() for b in range(B)])
    assert song.dtype == torch.int16 and np.array_equal(song.cpu().numpy(), own)
    ref = vo.infer(sd, noise=noise, ph_dur=ph_dur, **inp)[0]
    want = so.assemble_song([ref[b, 0, :int(mel_len[b]) * 256].numpy() for b in range(B)])
    assert snr_db(torch.from_numpy(want.astype(np.float32)), song.cpu().float()) >= SNR_DB_TOL
