"""-m gpu: SURVEY 8f-2 / BASELINE config 4 - utterances beyond the reference's 2000-frame position table
(ext_modules/text_post_net.py:26,69-72): per-segment frame-level transformers, full-length convolutional stages."""
import pytest
import torch

from common import MAX_ABS_TOL, SNR_DB_TOL, build_model, coarse_edge_distance, max_abs, report, snr_db
from music_generation_b200 import pipeline, segments, synthetic
from oracle import song_oracle as so
from oracle import visinger_oracle as vo

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _call(fn, inp, **kw):
    i = {k: v.to(DEV) for k, v in inp.items()}
    return fn(i["x"], i["x_lengths"], i["note_pitch_id"], i["note_duration"], phone_shengdiao=i["phone_shengdiao"],
              **{k: (v.to(DEV) if torch.is_tensor(v) else v) for k, v in kw.items()})


@pytest.fixture(scope="module")
def long_case():
    """One 58 s utterance (T = 5000 frames, 500 phonemes) with speaker conditioning: 3 segments of 1800 + 2 x 100 frames."""
    gin, P, T = 256, 500, 5000
    net, sd = build_model("ae", gin=gin, device=DEV)
    inp = synthetic.make_inputs(1, P, seed=41)
    ph_dur = synthetic.make_ph_dur(inp["x_lengths"], P, T, seed=41)
    noise = synthetic.make_noise(1, 192, T, seed=41)
    g = torch.randn(1, gin, 1, generator=torch.Generator().manual_seed(41))
    res = _call(net.infer_long, inp, ph_dur=ph_dur, noise=noise, g=g, segment_frames=1800, overlap=100)
    torch.cuda.synchronize()
    return dict(net=net, sd=sd, inp=inp, ph_dur=ph_dur, noise=noise, g=g, res=res, plan=net.last_aux["segments"],
                aux={k: v.cpu() for k, v in net.last_aux.items() if torch.is_tensor(v)}, T=T)


def test_long_utterance_matches_oracle_composition(long_case):
    c = long_case
    assert c["plan"] == [(0, *s) for s in segments.plan_segments(c["T"], 1800, 100)] and len(c["plan"]) == 3
    taps = {}
    ref = vo.infer_long(c["sd"], noise=c["noise"], ph_dur=c["ph_dur"], plan=c["plan"], g=c["g"], taps=taps, **c["inp"])
    assert torch.equal(c["aux"]["lr_index"].long(), taps["lr_index"]) and torch.equal(c["res"][1].cpu(), ref[1])
    diff = c["aux"]["f0_coarse"] != taps["f0_coarse"]
    report("f0_coarse_mismatches/long_t5000_3seg", int(diff.sum()))
    o = c["res"][0].cpu()
    if diff.any():        # edge flips only (see test_gpu_infer._f0_flips_reported); reported, waveform SNR still must hold
        assert float(coarse_edge_distance(ref[3])[diff].max()) < 2e-3 and int(diff.sum()) <= diff.numel() // 250
    s, m = snr_db(ref[0], o), max_abs(ref[0], o)
    report("waveform_snr_db/long_t5000_3seg", round(s, 2))
    report("waveform_max_abs/long_t5000_3seg", m)
    if not diff.any():
        assert s >= SNR_DB_TOL and m <= 2.5e-3          # gin-conditioned 16-bit residual: see test_config4_* for the max-abs note
    else:
        assert s >= 30.0                                 # a flipped bucket swaps an embedding row for one frame
    assert snr_db(ref[2][2], c["res"][2][2].cpu()) > 50  # m_p assembled from the three segments
    assert max_abs(ref[3], c["res"][3].cpu()) < 1e-3     # log_f0


def test_time_chunked_generator_is_bit_identical(long_case):
    """The Generator cut in time with 32-frame halos equals the uncut call bit for bit (exact halos), so the waveform has
    no seam at chunk boundaries either."""
    c = long_case
    net = c["net"]
    old = net.generator_time_chunk_frames
    try:
        net.generator_time_chunk_frames = 1024
        cut = _call(net.infer_long, c["inp"], ph_dur=c["ph_dur"], noise=c["noise"], g=c["g"], segment_frames=1800, overlap=100)[0]
    finally:
        net.generator_time_chunk_frames = old
    assert torch.equal(cut, c["res"][0])


def test_no_seam_unlike_per_line_concatenation(long_case):
    """What the halo buys: the reference's drivers synthesise pieces independently and concatenate them
    (pipline_inference_svs_v6_multispeaker.py:242) - every piece's flow / SpecDecoder / Generator starts from zero padding.
    Away from the cuts both agree to the waveform tolerance; inside the receptive field of a cut they do not."""
    c = long_case
    net, T = c["net"], c["T"]
    # the same three pieces synthesised as independent `infer` calls on phoneme-aligned cuts
    cum = c["ph_dur"][0].cumsum(0)
    cuts = [0]
    for target in (1800, 3600):
        cuts.append(int((cum <= target).sum()))
    cuts.append(c["ph_dur"].shape[1])
    pieces, t_edges = [], [0]
    for a, b in zip(cuts[:-1], cuts[1:]):
        sub = {k: (v[:, a:b] if v.dim() == 2 else torch.tensor([b - a])) for k, v in c["inp"].items()}
        f0, f1 = int(cum[a - 1]) if a else 0, int(cum[b - 1])
        pieces.append(_call(net.infer, sub, ph_dur=c["ph_dur"][:, a:b], noise=c["noise"][:, :, f0:f1], g=c["g"])[0].cpu())
        t_edges.append(f1)
    naive = torch.cat(pieces, -1)
    whole = _call(net.infer_long, c["inp"], ph_dur=c["ph_dur"], noise=c["noise"], g=c["g"],
                  segment_frames=2000, overlap=0)[0].cpu()            # no context: same transformer inputs only if cuts agree
    assert naive.shape == whole.shape
    hop = 256
    for e in t_edges[1:-1]:
        seam = slice((e - 8) * hop, (e + 8) * hop)
        assert max_abs(naive[..., seam], whole[..., seam]) > 10 * MAX_ABS_TOL       # the per-line seam is audible-size
    assert bool(torch.isfinite(whole).all())


def test_syn_long_song_int16(long_case):
    c = long_case
    syn = pipeline.Synthesiser(c["net"])
    line = (c["inp"]["x"][0], c["inp"]["note_pitch_id"][0], c["inp"]["note_duration"][0], c["inp"]["phone_shengdiao"][0])
    song = syn.syn_long(line, segment_frames=1800, overlap=100, noise=c["noise"].to(DEV), ph_dur=c["ph_dur"].to(DEV), g=c["g"].to(DEV))
    torch.cuda.synchronize()
    want = so.assemble_song([c["res"][0][0, 0].cpu().numpy()])
    assert song.dtype == torch.int16 and (song.cpu().numpy() == want).all()
