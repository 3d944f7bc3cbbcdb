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


def test_generator_bucketing_keeps_valid_samples_bit_identical():
    """generator_bucket=True (what syn_song uses): a ragged batch's Generator runs on length-sorted groups over their own
    longest length + halo; every VALID sample equals the unbucketed result bit for bit, the padded region is zero."""
    B, P, T = 6, 60, 700
    inp = synthetic.make_inputs(B, P, seed=51, ragged=True)
    ph_dur = synthetic.make_ph_dur(inp["x_lengths"], P, T, seed=51)
    ph_dur[2, 5:] = 0; ph_dur[4, 20:] = 0                                   # two much shorter utterances
    mel_len = ph_dur.sum(1)
    noise = synthetic.make_noise(B, 192, int(mel_len.max()), seed=51).to(DEV)
    net, _ = build_model("ae", device=DEV)
    net.generator_chunk_frames = 1500                                        # groups of two utterances
    i = {k: v.to(DEV) for k, v in inp.items()}
    call = lambda: net.infer(i["x"], i["x_lengths"], i["note_pitch_id"], i["note_duration"], phone_shengdiao=i["phone_shengdiao"],
                             ph_dur=ph_dur.to(DEV), noise=noise)[0]
    plain = call()
    net.generator_bucket = True
    bucketed = call()
    torch.cuda.synchronize()
    assert not torch.equal(plain, bucketed)                                  # the padded region differs (zeros) ...
    for b in range(B):
        n = int(mel_len[b]) * 256
        assert torch.equal(plain[b, :, :n], bucketed[b, :, :n])              # ... the valid samples do not
    # the shortest utterances sit in the last group: beyond that group's own length + halo nothing was computed
    short = int(mel_len.argmin())
    assert float(bucketed[short, :, (int(mel_len.sort().values[1]) + 40) * 256:].abs().max()) == 0.0
    assert float(plain[short, :, (int(mel_len.sort().values[1]) + 40) * 256:].abs().max()) > 0.0
