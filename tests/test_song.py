"""Song assembly: the oracle against the reference's literal numpy statements, host-side collation."""
import numpy as np
import torch

from music_generation_b200 import pipeline
from oracle import song_oracle as so


def _reference_statements(lines):
    """The statements of pipline_inference_svs_v5.py:136-139, 241-243, 241-258 executed verbatim by this NumPy."""
    SAMPLE_RATE = 22050
    wavs = []
    for audio in lines:
        audio = audio.copy()
        audio *= 32767 / max(0.01, np.max(np.abs(audio)))          # Synthesiser.syn (wave_save_path is not None)
        wav = audio
        wav *= 32767 / max(0.01, np.max(np.abs(wav)))              # song loop
        wavs = np.concatenate([wavs, wav])
    delay_time = 0.2
    delay_padding = [0.] * int(delay_time * SAMPLE_RATE)
    wav_music = np.concatenate([delay_padding, wavs])
    return wav_music.astype(np.int16)


def _lines(seed, n=5):
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        m = int(rng.integers(200, 5000))
        a = (rng.standard_normal(m) * rng.uniform(0.001, 0.9)).astype(np.float32)
        out.append(a)
    out.append(np.full(300, 1e-4, dtype=np.float32))              # below the 0.01 floor
    out.append(np.zeros(64, dtype=np.float32))                    # silent line
    return out


def test_oracle_matches_reference_statements():
    legacy = int(np.__version__.split(".")[0]) < 2
    for seed in range(4):
        lines = _lines(seed)
        want = _reference_statements(lines)
        got = so.assemble_song(lines, legacy_scalar=legacy)
        assert got.dtype == np.int16 and got.shape == want.shape
        assert np.array_equal(got, want)
        assert np.all(got[:so.LEAD_IN] == 0) and so.LEAD_IN == 4410


def test_scalar_semantics_differ_only_in_the_last_bit():
    rng = np.random.default_rng(0)
    peaks = rng.uniform(0.011, 1.0, 2000).astype(np.float32)
    a = np.array([so.peak_scale(p, False) for p in peaks])
    b = np.array([so.peak_scale(p, True) for p in peaks])
    assert np.all(np.abs(a.view(np.int32) - b.view(np.int32)) <= 1)
    assert so.peak_scale(np.float32(0.001)) == np.float32(3276700.0) == so.peak_scale(np.float32(0.001), True)


def test_collate_pads_like_per_line_calls():
    lines = [(torch.arange(3, 3 + k), torch.full((k,), 60), torch.full((k,), 0.3), torch.arange(k) + 5) for k in (4, 7, 2)]
    x, lens, notes, durs, tones = pipeline.Synthesiser.collate(lines)
    assert x.shape == (3, 7) and lens.tolist() == [4, 7, 2]
    assert x[0, 4:].eq(0).all() and torch.equal(x[1], lines[1][0]) and durs.dtype == torch.float32
    assert torch.equal(tones[2, :2], lines[2][3])
