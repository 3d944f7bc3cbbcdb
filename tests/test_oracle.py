"""The oracle (oracle/visinger_oracle.py) against the golden vectors produced by the UNMODIFIED
reference (oracle/make_golden.py), and against the live reference when /root/reference exists."""
import json
import os

import pytest
import torch

from common import GOLDEN, flat, load_golden, max_abs, state_dict_for
from oracle import ref_harness, visinger_oracle as vo
from oracle.make_golden import CASES, G_CASES, case_inputs, g_case_inputs, posterior_noise, posterior_spectrogram
from music_generation_b200 import synthetic

TOL = 2e-5     # fp32 reduction-order noise between thread counts / machines


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_matches_reference_golden(name):
    g = load_golden(name)
    variant, inp, ph_dur, seed = case_inputs(name)
    sd = state_dict_for(variant)
    T = int(g["T"])
    noise = synthetic.make_noise(inp["x"].shape[0], 192, T, seed=seed)
    taps = {}
    res = vo.infer(sd, noise=noise, ph_dur=ph_dur, variant=variant, taps=taps, **inp)
    o, y_mask, (z, z_p, m_p, logs_p), log_f0 = res[0], res[1], res[2], res[3]
    assert o.shape == g["o"].shape
    assert torch.equal(y_mask, g["y_mask"])                      # integer: bit-exact
    assert int(taps["mel_len"].max()) == T
    assert max_abs(g["o"], o) < TOL
    assert max_abs(g["log_f0"], log_f0) < 1e-4
    for key, t in (("z", z), ("z_p", z_p), ("m_p", m_p), ("logs_p", logs_p), ("mel", res[-1])):
        assert max_abs(g[key], t[:, ::4]) < 1e-4, key
    if variant == "ae":
        assert max_abs(g["spec"], res[4][:, ::16]) < 1e-4


@pytest.mark.skipif(not ref_harness.available(), reason="reference tree only exists in the build container")
@pytest.mark.parametrize("variant", ["ae", "v4"])
def test_oracle_matches_live_reference(variant):
    sd = state_dict_for(variant)
    net = ref_harness.build_reference(variant)
    net.load_state_dict(sd, strict=True)
    inp = synthetic.make_inputs(2, 9, seed=7, ragged=True)
    ph_dur = synthetic.make_ph_dur(inp["x_lengths"], 9, 40, seed=7)
    ph_dur[1, 2] = 0                                             # zero-duration phoneme vanishes
    noise = synthetic.make_noise(2, 192, int(ph_dur.sum(1).max()), seed=7)
    ref = ref_harness.run_reference_infer(net, inp, noise, ph_dur)
    out = vo.infer(sd, noise=noise, ph_dur=ph_dur, variant=variant, **inp)
    for a, b in zip(flat(ref), flat(out)):
        assert a.shape == b.shape and a.dtype == b.dtype
        assert max_abs(a, b) < TOL


@pytest.mark.parametrize("name", sorted(G_CASES))
def test_oracle_speaker_conditioning_matches_reference_golden(name):
    """The `g` path (WN.cond_layer modules.py:152-159, Generator.cond models_ae.py:329-330) of the oracle against
    outputs of the reference's own ResidualCouplingBlock / SpecDecoder / Generator called with g."""
    gold = load_golden(name)
    z_p, mel_len, g, gin = g_case_inputs(name)
    sd = state_dict_for("ae", gin=gin)
    y_mask = (torch.arange(z_p.shape[2])[None] < mel_len[:, None]).unsqueeze(1).float()
    z = vo.flow_reverse(sd, z_p * y_mask, y_mask, g=g)
    assert max_abs(gold["z"], z) < TOL
    spec = vo.spec_decoder(sd, z, mel_len, g=g)
    assert max_abs(gold["spec"], spec[:, ::8]) < 1e-4
    hid = vo.conv(sd, "spec_proj", spec) * y_mask
    assert max_abs(gold["hid"], hid) < 1e-4
    o = vo.generator(sd, gold["hid"], vo.DEFAULT_HP, g=g)
    assert o.shape == gold["o"].shape and max_abs(gold["o"], o) < TOL
    # g must matter: the unconditioned path gives a different waveform
    assert max_abs(gold["o"], vo.generator(sd, gold["hid"], vo.DEFAULT_HP)) > 1e-3
    # training-side PosteriorEncoder (models_ae.py:256-291), same g
    pz, pm, plogs, pmask, px = vo.posterior_encoder(sd, posterior_spectrogram(name), mel_len, posterior_noise(name), g=g)
    for key, t in (("post_z", pz), ("post_m", pm), ("post_logs", plogs), ("post_x", px)):
        assert max_abs(gold[key], t) < 1e-4, key
    # forward flow (training side, models_ae.py:246-253 with reverse=False) on the posterior sample, and its inverse
    zp = vo.flow_forward(sd, gold["post_z"], y_mask, g=g)
    assert max_abs(gold["post_zp"], zp) < 1e-4
    assert max_abs(gold["post_z"], vo.flow_reverse(sd, zp, y_mask, g=g)) < 1e-4


@pytest.mark.skipif(not ref_harness.available(), reason="reference tree only exists in the build container")
def test_oracle_speaker_conditioning_matches_live_reference():
    gin = 256
    sd = state_dict_for("ae", gin=gin)
    net = ref_harness.build_reference("ae", gin_channels=gin)
    net.load_state_dict(sd, strict=True)
    gen = torch.Generator().manual_seed(5)
    z_p, mel_len, g = torch.randn(2, 192, 33, generator=gen), torch.tensor([33, 20]), torch.randn(2, gin, 1, generator=gen)
    y_mask = (torch.arange(33)[None] < mel_len[:, None]).unsqueeze(1).float()
    with torch.no_grad():
        z = net.spec_flow(z_p * y_mask, y_mask, g=g, reverse=True)
        spec = net.spec_decoder(z, mel_len, g=g)
        o = net.decoder(net.spec_proj(spec) * y_mask, g=g)
    assert max_abs(z, vo.flow_reverse(sd, z_p * y_mask, y_mask, g=g)) < TOL
    assert max_abs(spec, vo.spec_decoder(sd, z, mel_len, g=g)) < 1e-4
    assert max_abs(o, vo.generator(sd, net.spec_proj(spec) * y_mask, vo.DEFAULT_HP, g=g)) < TOL


def test_length_regulator_edge_cases():
    h = torch.arange(2 * 3 * 4, dtype=torch.float32).reshape(2, 3, 4)     # [B,C,P]
    dur = torch.tensor([[2, 0, 1, 3], [0, 0, 0, 0]])
    out, mel_len, idx = vo.length_regulate(h, dur)
    assert mel_len.tolist() == [6, 0]
    assert idx[0].tolist() == [0, 0, 2, 3, 3, 3]
    assert idx[1].tolist() == [-1] * 6
    assert torch.equal(out[0, :, 2], h[0, :, 2]) and float(out[1].abs().sum()) == 0.0
    # negative durations behave like zero (vec.expand(int(<=0)) is empty / clamp)
    out2, mel2, _ = vo.length_regulate(h, torch.tensor([[1, -3, 0, 0], [1, 1, 1, 1]]))
    assert mel2.tolist() == [1, 4]


def test_f0_to_coarse_range_and_monotone():
    f0 = torch.tensor([0.0, 1.0, 49.0, 50.0, 51.0, 250.0, 1099.0, 1100.0, 5000.0])
    c = vo.f0_to_coarse(f0)
    assert c.min() >= 1 and c.max() <= 255
    assert c[0] == 1 and c[-1] == 255
    assert (c[1:] >= c[:-1]).all()


def test_position_table_limit_raises_like_reference():
    sd = state_dict_for("ae")
    x = torch.zeros(1, 192, 2001)
    with pytest.raises(RuntimeError):
        vo.text_post_net(sd, x, torch.ones(1, 1, 2001))


def test_manifest_lists_every_fixture():
    m = json.load(open(os.path.join(GOLDEN, "MANIFEST.json")))
    for name in list(m["cases"]) + list(m["g_cases"]):
        assert os.path.exists(os.path.join(GOLDEN, name + ".npz"))
