"""Host-side logic on CPU: weight packing, Flip folding, polyphase ConvTranspose, epilogue wiring and
the whole infer() orchestration, run through the emulated ops (tests/emu_ops.py) against the oracle."""
import pytest
import torch
import torch.nn.functional as F

import emu_ops
from common import MAX_ABS_TOL, SNR_DB_TOL, build_model, flat, load_golden, max_abs, snr_db, state_dict_for
from music_generation_b200 import packer, synthetic, ops
from oracle import visinger_oracle as vo
from oracle.make_golden import CASES, case_inputs


def _emu(monkeypatch):
    emu_ops.install(monkeypatch)
    monkeypatch.setattr(ops, "_btc", lambda t, n: 0 if t is None else t.stride(1))


def test_fold_weight_norm_matches_torch():
    torch.manual_seed(0)
    v, g = torch.randn(6, 5, 3), torch.rand(6, 1, 1) + 0.5
    w = packer.fold_weight_norm({"c.weight_v": v, "c.weight_g": g}, "c")
    assert torch.allclose(w, torch._weight_norm(v, g, 0), atol=1e-6)


@pytest.mark.parametrize("k,s", [(16, 8), (4, 2)])
def test_polyphase_conv_transpose(k, s):
    torch.manual_seed(1)
    cin, cout, T = 10, 6, 23
    w, b = torch.randn(cin, cout, k), torch.randn(cout)
    x = torch.randn(2, cin, T)
    ref = F.conv_transpose1d(x, w, b, stride=s, padding=(k - s) // 2)                # [B,cout,T*s]
    pc = packer.pack_conv_transpose(w, b, s, (k - s) // 2, dtype=torch.float32)
    pc.w = pc.w.float()                                                             # exact check
    out = torch.zeros(2, T * s, cout)
    emu_ops.conv1d(pc, x.transpose(1, 2).contiguous(), out32=out.view(2, T, s * cout))
    assert max_abs(ref.transpose(1, 2), out) < 1e-4


def test_polyphase_rejects_unsupported_geometry():
    with pytest.raises(ValueError):
        packer.pack_conv_transpose(torch.randn(4, 4, 32), None, 8, 4)


def test_gate_packing_equals_fused_add_tanh_sigmoid_multiply():
    torch.manual_seed(2)
    H = 192
    w, b = torch.randn(2 * H, H, 5) * 0.05, torch.randn(2 * H) * 0.1
    x = torch.randn(1, 40, H)
    y = F.conv1d(x.transpose(1, 2), w, b, padding=2)
    ref = (torch.tanh(y[:, :H]) * torch.sigmoid(y[:, H:])).transpose(1, 2)
    pc = packer.pack_conv(w, b, gate=True, dtype=torch.float32)
    pc.w = pc.w.float()
    out = torch.zeros(1, 40, H)
    emu_ops.conv1d(pc, x, out32=out)
    assert max_abs(ref, out) < 1e-4


def test_choose_block_n():
    for cout in (1, 32, 80, 96, 192, 256, 384, 512, 513, 576, 768, 1024, 2048):
        bn = packer.choose_block_n(cout)
        assert bn % 32 == 0 and 32 <= bn <= 256
        assert packer._round_up(cout, bn) - cout < bn


@pytest.mark.parametrize("name", sorted(CASES))
def test_infer_orchestration_against_reference_golden(name, monkeypatch):
    """Package infer() (emulated kernels, fp16 operand rounding) vs the reference's own outputs."""
    _emu(monkeypatch)
    g = load_golden(name)
    variant, inp, ph_dur, seed = case_inputs(name)
    net, _ = build_model(variant)
    noise = synthetic.make_noise(inp["x"].shape[0], 192, int(g["T"]), seed=seed)
    res = net.infer(inp["x"], inp["x_lengths"], inp["note_pitch_id"], inp["note_duration"],
                    phone_shengdiao=inp["phone_shengdiao"], ph_dur=ph_dur, noise_scale=1., length_scale=1, noise=noise)
    assert len(res) == (6 if variant == "ae" else 5)
    assert torch.equal(res[1], g["y_mask"]) and res[1].dtype == torch.int64
    assert res[0].shape == g["o"].shape and res[0].dtype == torch.float32
    assert snr_db(g["o"], res[0]) >= SNR_DB_TOL and max_abs(g["o"], res[0]) <= MAX_ABS_TOL
    assert max_abs(g["log_f0"], res[3]) < 1e-3
    assert snr_db(g["z"], res[2][0][:, ::4]) > 60


def test_infer_with_speaker_conditioning_and_f0_override(monkeypatch):
    _emu(monkeypatch)
    net, sd = build_model("ae", gin=256)
    inp = synthetic.make_inputs(2, 8, seed=5, ragged=True)
    ph_dur = synthetic.make_ph_dur(inp["x_lengths"], 8, 48, seed=5)
    T = int(ph_dur.sum(1).max())
    noise = synthetic.make_noise(2, 192, T, seed=5)
    gsp = torch.randn(2, 256, 1, generator=torch.Generator().manual_seed(3))
    f0 = torch.randint(1, 256, (2, T), generator=torch.Generator().manual_seed(4))
    taps = {}
    ref = vo.infer(sd, noise=noise, ph_dur=ph_dur, g=gsp, f0_score=f0, taps=taps, **inp)
    out = net.infer(inp["x"], inp["x_lengths"], inp["note_pitch_id"], inp["note_duration"], inp["phone_shengdiao"],
                    ph_dur=ph_dur, noise=noise, g=gsp, f0_score=f0)
    assert torch.equal(net.last_aux["lr_index"].long(), taps["lr_index"])
    assert torch.equal(net.last_aux["f0_coarse"], f0)
    assert snr_db(ref[0], out[0]) >= SNR_DB_TOL and max_abs(ref[0], out[0]) <= MAX_ABS_TOL


def test_resblock2_variant(monkeypatch):
    _emu(monkeypatch)
    from conftest import HP
    from music_generation_b200 import models_ae
    hp = dict(HP, resblock="2", resblock_dilation_sizes=[[1, 3], [1, 3], [1, 3]])
    net = models_ae.SynthesizerTrn(100, 513, 48, **hp).eval()
    sd = synthetic.apply_recipe({k: v.clone() for k, v in net.state_dict().items()})
    net.load_state_dict(sd)
    inp = synthetic.make_inputs(1, 6, seed=11)
    ph_dur = synthetic.make_ph_dur(inp["x_lengths"], 6, 30, seed=11)
    noise = synthetic.make_noise(1, 192, 30, seed=11)
    ref = vo.infer(sd, noise=noise, ph_dur=ph_dur, hp=hp, **inp)
    out = net.infer(inp["x"], inp["x_lengths"], inp["note_pitch_id"], inp["note_duration"], inp["phone_shengdiao"],
                    ph_dur=ph_dur, noise=noise)
    assert snr_db(ref[0], out[0]) >= SNR_DB_TOL


def test_max_len_and_lf0_quirks(monkeypatch):
    """max_len truncates the generator input; lf0 only overrides the *returned* log_f0 (models_ae.py:638-650)."""
    _emu(monkeypatch)
    net, sd = build_model("ae")
    inp = synthetic.make_inputs(1, 6, seed=12)
    ph_dur = synthetic.make_ph_dur(inp["x_lengths"], 6, 24, seed=12)
    noise = synthetic.make_noise(1, 192, 24, seed=12)
    lf0 = torch.zeros(1, 24)
    out = net.infer(inp["x"], inp["x_lengths"], inp["note_pitch_id"], inp["note_duration"], inp["phone_shengdiao"],
                    ph_dur=ph_dur, noise=noise, max_len=10, lf0=lf0)
    ref = vo.infer(sd, noise=noise, ph_dur=ph_dur, max_len=10, **inp)
    assert out[0].shape == (1, 1, 10 * 256) and out[3] is lf0
    assert snr_db(ref[0], out[0]) >= SNR_DB_TOL


def test_fused_spec_proj_equals_two_convs(monkeypatch):
    """fuse_spec_proj=True: spec_proj(proj(x) * m) * m as one masked 192x192 conv; spec_predict is None,
    every other result is unchanged (ragged batch: the frame mask has to commute exactly)."""
    _emu(monkeypatch)
    inp = synthetic.make_inputs(2, 6, seed=21, ragged=True)
    ph_dur = synthetic.make_ph_dur(inp["x_lengths"], 6, 20, seed=21)
    noise = synthetic.make_noise(2, 192, int(ph_dur.sum(1).max()), seed=21)
    outs = []
    for fused in (False, True):
        net, _ = build_model("ae", fuse_spec_proj=fused)
        outs.append(net.infer(inp["x"], inp["x_lengths"], inp["note_pitch_id"], inp["note_duration"],
                              inp["phone_shengdiao"], ph_dur=ph_dur, noise=noise))
    a, b = outs
    assert a[4] is not None and b[4] is None
    assert torch.equal(a[1], b[1]) and torch.equal(a[2][0], b[2][0]) and torch.equal(a[5], b[5])
    assert snr_db(a[0], b[0]) >= 50.0 and max_abs(a[0], b[0]) <= MAX_ABS_TOL      # two 16-bit roundings apart


def test_frames_beyond_position_table_raise(monkeypatch):
    _emu(monkeypatch)
    net, _ = build_model("ae")
    inp = synthetic.make_inputs(1, 4, seed=1)
    ph_dur = torch.tensor([[600, 600, 600, 300]])
    with pytest.raises(RuntimeError, match="must match"):
        net.infer(inp["x"], inp["x_lengths"], inp["note_pitch_id"], inp["note_duration"], inp["phone_shengdiao"],
                  ph_dur=ph_dur)


def test_cpu_inputs_fail_loudly():
    net, _ = build_model("ae")
    inp = synthetic.make_inputs(1, 4, seed=1)
    with pytest.raises(RuntimeError, match="no CPU path"):
        net.infer(inp["x"], inp["x_lengths"], inp["note_pitch_id"], inp["note_duration"], inp["phone_shengdiao"])


def test_state_dict_roundtrip_and_checkpoint_format(tmp_path):
    """utils.load_checkpoint (VISinger/utils/utils.py:33-58) iterates model.state_dict() and loads by name."""
    net, sd = build_model("ae")
    ck = tmp_path / "G_AE.pth"
    torch.save({"model": sd, "iteration": 1, "optimizer": None, "learning_rate": 2e-4}, ck)
    loaded = torch.load(ck, map_location="cpu")["model"]
    new = {k: loaded.get(k, v) for k, v in net.state_dict().items()}
    net.load_state_dict(new)
    assert len(new) == 873
    assert all(k.count("weight_g") == 0 or new[k].dim() == 3 for k in new)


def test_enlarged_position_table_loads():
    net, sd = build_model("ae")
    sd2 = synthetic.apply_recipe({k: v.clone() for k, v in sd.items()}, positions=4096)
    net.load_state_dict(sd2)
    assert net.text_post_net.position_embeddings.weight.shape[0] == 4096


def test_infer_normalises_id_dtypes_and_rejects_out_of_range_ids(monkeypatch):
    """nn.Embedding / sequence_mask accept int32 ids and lengths and raise IndexError on an id outside the table
    (ADVICE r1): int32 inputs give the same result as int64 ones, a live out-of-range id raises, one in the padded
    tail of a shorter utterance does not."""
    _emu(monkeypatch)
    net, _ = build_model("ae")
    inp = synthetic.make_inputs(2, 8, seed=5, ragged=True)
    ph_dur = synthetic.make_ph_dur(inp["x_lengths"], 8, 30, seed=5)
    noise = synthetic.make_noise(2, 192, int(ph_dur.sum(1).max()), seed=5)
    call = lambda i: net.infer(i["x"], i["x_lengths"], i["note_pitch_id"], i["note_duration"],
                               phone_shengdiao=i["phone_shengdiao"], ph_dur=ph_dur, noise=noise)
    a = call(inp)
    b = call({k: (v.to(torch.int32) if v.dtype == torch.int64 else v) for k, v in inp.items()})
    assert torch.equal(a[0], b[0]) and b[1].dtype == torch.int32          # y_mask follows x.dtype like the reference
    bad = {k: v.clone() for k, v in inp.items()}
    bad["x"][0, 0] = 100                                                    # table has 100 rows
    with pytest.raises(IndexError):
        call(bad)
    short = int(inp["x_lengths"].argmin())
    if int(inp["x_lengths"][short]) < 8:
        pad = {k: v.clone() for k, v in inp.items()}
        pad["phone_shengdiao"][short, -1] = 9999                            # padded slot: never read by the reference either
        assert torch.equal(call(pad)[0], a[0])
    with pytest.raises(IndexError):
        net.infer(inp["x"], inp["x_lengths"], inp["note_pitch_id"], inp["note_duration"],
                  phone_shengdiao=inp["phone_shengdiao"], ph_dur=ph_dur, noise=noise,
                  f0_score=torch.full((2, int(ph_dur.sum(1).max())), 300))


def test_packer_refuses_weights_beyond_fp16_range():
    w = torch.zeros(4, 4, 3); w[1, 2, 0] = 7.0e4
    with pytest.raises(ValueError, match="fp16 range"):
        packer.pack_conv(w, None, dtype=torch.float16)
    assert torch.isfinite(packer.pack_conv(w, None, dtype=torch.bfloat16).w.float()).all()


def test_posterior_encode_orchestration_against_reference_golden(monkeypatch):
    """PosteriorEncoder (models_ae.py:256-291) on the SpecDecoder's kernels - host orchestration through the
    emulated ops against outputs of the reference's own spec_encoder (tests/golden/ae_gin256_stages.npz)."""
    from oracle.make_golden import g_case_inputs, posterior_noise, posterior_spectrogram
    _emu(monkeypatch)
    name = "ae_gin256_stages"
    gold = load_golden(name)
    _, mel_len, g, gin = g_case_inputs(name)
    net, _ = build_model("ae", gin=gin)
    z, m, logs, mask, x = net.posterior_encode(posterior_spectrogram(name), mel_len, g=g, noise=posterior_noise(name))
    assert z.shape == gold["post_z"].shape and mask.shape == (2, 1, z.shape[2])
    assert snr_db(gold["post_m"], m) > 50 and snr_db(gold["post_z"], z) > 50 and snr_db(gold["post_x"], x) > 45
    assert max_abs(gold["post_logs"], logs) < 2e-2
    n1 = int(mel_len[1])
    assert float(z[1, :, n1:].abs().max()) == 0.0            # z is masked (models_ae.py:290)


def test_infer_long_orchestration_vs_oracle_and_single_segment_identity(monkeypatch):
    """segment plan + scatter of the per-segment TextPostNet statistics + full-length conv stages, through
    the emulated ops against oracle.infer_long; one segment covering everything is `infer` itself."""
    from music_generation_b200 import segments
    _emu(monkeypatch)
    assert segments.plan_segments(70, 30, 5) == [(0, 30, 0, 35), (30, 60, 25, 65), (60, 70, 55, 70)]
    assert segments.plan_segments(30, 30, 5) == [(0, 30, 0, 30)] and segments.plan_segments(0, 30, 5) == []
    net, sd = build_model("ae")
    inp = synthetic.make_inputs(2, 9, seed=13, ragged=True)
    ph_dur = synthetic.make_ph_dur(inp["x_lengths"], 9, 70, seed=13)
    T = int(ph_dur.sum(1).max())
    noise = synthetic.make_noise(2, 192, T, seed=13)
    args = (inp["x"], inp["x_lengths"], inp["note_pitch_id"], inp["note_duration"])
    res = net.infer_long(*args, phone_shengdiao=inp["phone_shengdiao"], ph_dur=ph_dur, noise=noise, segment_frames=30, overlap=5)
    plan = net.last_aux["segments"]
    assert [p[0] for p in plan] == sorted(p[0] for p in plan) and plan[0][1:] == (0, 30, 0, 35)
    ref = vo.infer_long(sd, noise=noise, ph_dur=ph_dur, plan=plan, **inp)
    assert torch.equal(res[1], ref[1])
    assert snr_db(ref[0], res[0]) >= SNR_DB_TOL and max_abs(ref[0], res[0]) <= MAX_ABS_TOL
    assert snr_db(ref[2][2], res[2][2]) > 50                                  # m_p: per-segment statistics in the right frames
    one = net.infer_long(*args, phone_shengdiao=inp["phone_shengdiao"], ph_dur=ph_dur, noise=noise, segment_frames=1000)
    full = net.infer(*args, phone_shengdiao=inp["phone_shengdiao"], ph_dur=ph_dur, noise=noise)
    for a, b in zip(flat(one), flat(full)):
        assert torch.equal(a, b)
    # segmentation changes the frame-level transformer's context, so the segmented result differs from the unsegmented one
    assert max_abs(res[0], full[0]) > 1e-4
    with pytest.raises(ValueError):
        net.infer_long(*args, phone_shengdiao=inp["phone_shengdiao"], ph_dur=ph_dur, noise=noise, segment_frames=1990, overlap=10)


@pytest.mark.parametrize("variant,gin", [("ae", 0), ("ae", 256), ("v4", 0)])
def test_state_dict_keys_match_reference(variant, gin):
    """The package's parameter tree has exactly the reference's state-dict keys and shapes (what `utils.load_checkpoint`,
    VISinger/utils/utils.py:40-55, iterates), for the shipped configs and with speaker conditioning."""
    from oracle import ref_harness
    if not ref_harness.available():
        pytest.skip("reference tree / oracle/_ref stage not present")
    ref = ref_harness.build_reference(variant, gin_channels=gin).state_dict()
    net, _ = build_model(variant, gin=gin)
    mine = net.state_dict()
    assert set(mine) == set(ref)
    for k, v in ref.items():
        assert tuple(mine[k].shape) == tuple(v.shape), k
        assert mine[k].dtype == v.dtype, k


def test_save_wav_writes_the_file_the_reference_driver_writes(tmp_path):
    """pipline_inference_svs_v5.py:256-258 ends with scipy.io.wavfile.write(path, 22050, wavs.astype(np.int16)):
    pipeline.save_wav must produce the same bytes from the int16 song (stdlib only on the product side)."""
    import numpy as np
    import scipy.io.wavfile
    from music_generation_b200 import pipeline
    rng = np.random.default_rng(7)
    song = rng.integers(-32768, 32767, size=22050 * 2 + 4410 + 3, dtype=np.int16)
    ours, ref = tmp_path / "ours.wav", tmp_path / "ref.wav"
    pipeline.save_wav(str(ours), torch.from_numpy(song))
    scipy.io.wavfile.write(str(ref), pipeline.SAMPLE_RATE, song)
    assert ours.read_bytes() == ref.read_bytes()
    sr, back = scipy.io.wavfile.read(str(ours))
    assert sr == 22050 and np.array_equal(back, song)
    with pytest.raises(ValueError):
        pipeline.save_wav(str(ours), torch.zeros(4))
