"""-m gpu: SURVEY 8f-3 on the device - the resident singer bank (three checkpoints of one architecture, selected by
singer id like pipline_inference_svs_v6_multispeaker.py:167-169,234-239), the packed-weight cache attached to a device
model, the device of the parameters deciding where the kernels run, and the fp16 numeric-range canary."""
import pytest
import torch

from common import MAX_ABS_TOL, SNR_DB_TOL, max_abs, report, snr_db, state_dict_for
from conftest import HP
from music_generation_b200 import checkpoint, models_ae, synthetic
from oracle import visinger_oracle as vo

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _net(**kw):
    return models_ae.SynthesizerTrn(100, 513, 48, **HP, **kw).eval()


def _call(net, inp, dev, **kw):
    i = {k: v.to(dev) for k, v in inp.items()}
    return net.infer(i["x"], i["x_lengths"], i["note_pitch_id"], i["note_duration"], phone_shengdiao=i["phone_shengdiao"],
                     **{k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in kw.items()})


def _case(seed=17, B=2, P=14, T=110):
    inp = synthetic.make_inputs(B, P, seed=seed, ragged=True)
    ph_dur = synthetic.make_ph_dur(inp["x_lengths"], P, T, seed=seed)
    noise = synthetic.make_noise(B, 192, int(ph_dur.sum(1).max()), seed=seed)
    return inp, ph_dur, noise


def test_singer_bank_on_device_parity_per_singer_and_cache_equivalence(tmp_path):
    seeds = {0: 1234, 1: 11, 2: 22}
    paths, sds = {}, {}
    for sid, seed in seeds.items():
        sds[sid] = state_dict_for("ae", seed=seed)
        src = _net()
        src.load_state_dict(sds[sid])
        paths[sid] = str(tmp_path / f"G_AE_{sid}.pth")
        checkpoint.save_checkpoint(src, None, 2e-4, 1000 + sid, paths[sid])
    bank = checkpoint.SingerBank(_net, paths, device=DEV, packed_cache=True)       # first load: packs + writes the caches
    assert len(bank) == 3 and bank.iterations == {0: 1000, 1: 1001, 2: 1002}
    inp, ph_dur, noise = _case()
    outs = {}
    for sid in seeds:
        net = bank.get(sid)
        assert net.device.type == "cuda"
        res = _call(net, inp, DEV, ph_dur=ph_dur, noise=noise)
        ref = vo.infer(sds[sid], noise=noise, ph_dur=ph_dur, **inp)
        assert torch.equal(res[1].cpu(), ref[1])
        s, m = snr_db(ref[0], res[0].cpu()), max_abs(ref[0], res[0].cpu())
        report(f"waveform_snr_db/singer_bank_id{sid}", round(s, 2))
        assert s >= SNR_DB_TOL and m <= MAX_ABS_TOL, (sid, s, m)
        outs[sid] = res[0]
    assert not torch.equal(outs[0], outs[1]) and not torch.equal(outs[1], outs[2])      # three different singers
    assert bank.get(7) is bank.get(0)                                                  # unknown id -> default (`else: syn`)
    # second bank: every model attaches its cache ("hit") and is bit-equal to the freshly packed one
    bank2 = checkpoint.SingerBank(_net, paths, device=DEV, packed_cache=True)
    for sid in seeds:
        net = bank2.get(sid)
        assert net._packed is not None                                                  # attached at load, not packed lazily
        assert checkpoint.load_packed(net, paths[sid] + checkpoint.CACHE_SUFFIX)        # fingerprint matches on the device model
        res = _call(net, inp, DEV, ph_dur=ph_dur, noise=noise)
        assert torch.equal(res[0], outs[sid])


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_model_on_second_device_while_first_is_current():
    """The kernels run where the parameters live, not on the current device (per-device launch state)."""
    sd = state_dict_for("ae")
    inp, ph_dur, noise = _case(seed=19)
    torch.cuda.set_device(0)
    n0 = _net(); n0.load_state_dict(sd); n0 = n0.to("cuda:0")
    n1 = _net(); n1.load_state_dict(sd); n1 = n1.to("cuda:1")
    a = _call(n0, inp, "cuda:0", ph_dur=ph_dur, noise=noise)[0]
    assert torch.cuda.current_device() == 0
    b = _call(n1, inp, "cuda:1", ph_dur=ph_dur, noise=noise)[0]
    assert b.device == torch.device("cuda:1") and torch.cuda.current_device() == 0
    assert torch.equal(a.cpu(), b.cpu())
    with pytest.raises(RuntimeError, match="is on"):
        _call(n1, inp, "cuda:0", ph_dur=ph_dur, noise=noise)


@pytest.mark.parametrize("scale,expect_sat", [(4000.0, False), (6.0e4, True)], ids=["in_range", "overflow"])
def test_fp16_numeric_range_saturates_and_reports(scale, expect_sat):
    """the Generator's activations scaled by `scale` (conv_pre x scale, conv_post / scale: the waveform stays
    O(1); 6e4 keeps every WEIGHT inside fp16 - the packer refuses weights beyond it - while activations reach ~1.5e5).  Inside fp16's range the result keeps the tolerance and the canary stays 0; beyond 65504 every 16-bit store
    saturates (finite waveform, no inf / NaN) and the canary reports it, instead of a silent NaN waveform."""
    base = state_dict_for("ae")
    sd = synthetic.apply_recipe({k: v.clone() for k, v in base.items()}, pre_gain=4.0 * scale, post_gain=1.0 / scale)
    net = _net(); net.load_state_dict(sd); net = net.to(DEV)
    inp, ph_dur, noise = _case(seed=23)
    res = _call(net, inp, DEV, ph_dur=ph_dur, noise=noise, check_range=True)
    o = res[0].cpu()
    events = net.last_aux["saturation_events"]
    report(f"saturation_events/scale_{scale:g}", events)
    assert bool(torch.isfinite(o).all()) and float(o.abs().max()) <= 1.0
    taps = {}
    ref = vo.infer(sd, noise=noise, ph_dur=ph_dur, taps=taps, **inp)
    peak = max(float(taps[k].abs().max()) for k in taps if k.startswith(("ups", "stage")))
    report(f"generator_peak_activation/scale_{scale:g}", round(peak, 1))
    if expect_sat:
        assert peak > 65504.0 and events > 0
        # bf16 operands have fp32's exponent range: same weights, no clipping, looser mantissa
        nb = _net(operand_dtype="bf16"); nb.load_state_dict(sd); nb = nb.to(DEV)
        ob = _call(nb, inp, DEV, ph_dur=ph_dur, noise=noise, check_range=True)[0].cpu()
        report(f"waveform_snr_db/bf16_scale_{scale:g}", round(snr_db(ref[0], ob), 2))
        assert bool(torch.isfinite(ob).all()) and snr_db(ref[0], ob) > 25.0
    else:
        assert peak < 65504.0 and events == 0
        s, m = snr_db(ref[0], o), max_abs(ref[0], o)
        report(f"waveform_snr_db/fp16_scale_{scale:g}", round(s, 2))
        assert s >= SNR_DB_TOL and m <= MAX_ABS_TOL, (s, m)


def test_posterior_encoder_on_device_matches_reference_golden():
    """`posterior_encode` (= reference spec_encoder, models_ae.py:256-291, with speaker conditioning) on the
    B200 against outputs of the reference's own module."""
    from common import load_golden
    from oracle.make_golden import g_case_inputs, posterior_noise, posterior_spectrogram
    name = "ae_gin256_stages"
    gold = load_golden(name)
    _, mel_len, g, gin = g_case_inputs(name)
    net = _net(gin_channels=gin)
    net.load_state_dict(state_dict_for("ae", gin=gin))
    net = net.to(DEV)
    z, m, logs, mask, x = net.posterior_encode(posterior_spectrogram(name).to(DEV), mel_len.to(DEV), g=g.to(DEV),
                                               noise=posterior_noise(name).to(DEV))
    torch.cuda.synchronize()
    s = snr_db(gold["post_z"], z.cpu())
    report("posterior_z_snr_db/gin256", round(s, 2))
    assert s > 50 and snr_db(gold["post_m"], m.cpu()) > 50 and snr_db(gold["post_x"], x.cpu()) > 45
    assert float(z[1, :, int(mel_len[1]):].abs().max()) == 0.0
    # forward flow z -> z_p (reference spec_flow(z, y_mask, g), models_ae.py:246-253) on the reference's own posterior sample
    zp = net.flow_forward(gold["post_z"].to(DEV), mel_len.to(DEV), g=g.to(DEV))
    torch.cuda.synchronize()
    s = snr_db(gold["post_zp"], zp.cpu())
    report("flow_forward_snr_db/gin256", round(s, 2))
    assert s > 50 and zp.shape == gold["post_zp"].shape
    assert float(zp[1, :, int(mel_len[1]):].abs().max()) == 0.0
