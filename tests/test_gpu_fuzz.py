"""-m gpu: seeded random shapes through the whole forward against the CPU oracle - batch, phoneme count, frame count (incl.
multiples of the 64-key and 128-query attention tiles, odd frame counts, one-tile utterances) and ragged lengths, both
through the stage-level C entry points (default) and the per-kernel path.  Tolerance as everywhere (BASELINE.json)."""
import random

import pytest
import torch

from common import MAX_ABS_TOL, SNR_DB_TOL, build_model, max_abs, report, snr_db
from music_generation_b200 import synthetic
from oracle import visinger_oracle as vo

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _infer(net, inp, **kw):
    i = {k: v.to(DEV) for k, v in inp.items()}
    return net.infer(i["x"], i["x_lengths"], i["note_pitch_id"], i["note_duration"],
                     phone_shengdiao=i["phone_shengdiao"], noise_scale=1., length_scale=1, **kw)


@pytest.mark.parametrize("stage_api", [True, False], ids=["stage_api", "per_op"])
def test_infer_random_shapes_against_oracle(stage_api):
    rnd = random.Random(2024 + stage_api)
    net, sd = build_model("ae", device=DEV, stage_api=stage_api)
    worst = 1e9
    for it in range(7):
        B = rnd.randint(1, 3)
        T = rnd.choice([64, 128, 192, 256, 333, 384, 512, 640, 41, 700])
        P = max(3, T // rnd.choice([6, 10, 17]))
        ragged = B > 1 and rnd.random() < 0.7
        inp = synthetic.make_inputs(B, P, seed=it, ragged=ragged)
        ph_dur = synthetic.make_ph_dur(inp["x_lengths"], P, T, seed=it)
        Tm = int(ph_dur.sum(1).max())
        noise = synthetic.make_noise(B, 192, Tm, seed=it)
        taps = {}
        ref = vo.infer(sd, noise=noise, ph_dur=ph_dur, taps=taps, **inp)
        res = _infer(net, inp, ph_dur=ph_dur.to(DEV), noise=noise.to(DEV), f0_score=taps["f0_coarse"].to(DEV))
        torch.cuda.synchronize()
        assert torch.equal(net.last_aux["lr_index"].long().cpu(), taps["lr_index"]), (it, B, P, T)
        assert torch.equal(res[1].cpu(), ref[1]), (it, B, P, T)
        o = res[0].cpu()
        s, m = snr_db(ref[0], o), max_abs(ref[0], o)
        worst = min(worst, s)
        assert s >= SNR_DB_TOL and m <= MAX_ABS_TOL, (it, B, P, T, ragged, s, m)
        assert snr_db(ref[5], res[5].cpu()) > 50, (it, B, P, T)          # mel_predict: the attention path
    report(f"waveform_snr_db/fuzz_worst_{'stage_api' if stage_api else 'per_op'}", round(worst, 2))


@pytest.mark.parametrize("B,P,T", [(1, 1, 1), (1, 2, 2), (2, 2, 3), (1, 3, 7), (3, 4, 17)])
def test_infer_tiny_shapes_against_oracle(B, P, T):
    """Degenerate sizes: one phoneme / one frame (256 samples), fewer frames than any tile, odd frame counts."""
    net, sd = build_model("ae", device=DEV)
    inp = synthetic.make_inputs(B, P, seed=B * 100 + T)
    ph_dur = synthetic.make_ph_dur(inp["x_lengths"], P, T, seed=T)
    Tm = int(ph_dur.sum(1).max())
    noise = synthetic.make_noise(B, 192, Tm, seed=T)
    taps = {}
    ref = vo.infer(sd, noise=noise, ph_dur=ph_dur, taps=taps, **inp)
    res = _infer(net, inp, ph_dur=ph_dur.to(DEV), noise=noise.to(DEV), f0_score=taps["f0_coarse"].to(DEV))
    torch.cuda.synchronize()
    o = res[0].cpu()
    assert o.shape == ref[0].shape == (B, 1, Tm * 256)
    assert torch.equal(res[1].cpu(), ref[1])
    assert snr_db(ref[0], o) >= SNR_DB_TOL and max_abs(ref[0], o) <= MAX_ABS_TOL
