"""TEST INFRASTRUCTURE ONLY - CPU restatement of the reference's per-line / per-song audio post-processing.

Follows (behaviour, not code) pipline_inference_svs_v5.py:134-139 (`Synthesiser.syn`: the returned audio has been
scaled in place because the song loop always passes a wave_save_path), :225-243 (the song loop scales `wav` again and
concatenates) and :241-258 (0.2 s of zeros in front, `astype(np.int16)`); identical in
pipline_inference_svs_v6_multispeaker.py:240-258.

Scalar semantics: `32767 / max(0.01, np.max(np.abs(audio)))` is fp32 under NumPy >= 2 (python scalars are weak, NEP 50)
and fp64-then-rounded under NumPy < 2 (value-based casting); `legacy_scalar` selects which.  tests/test_song.py pins
this restatement against the reference's literal statements executed by the NumPy in this image.
"""
import numpy as np

SAMPLE_RATE = 22050
LEAD_IN = int(0.2 * SAMPLE_RATE)       # pipline_inference_svs_v5.py:241-242


def peak_scale(peak, legacy_scalar=False):
    """32767 / max(0.01, peak) as an fp32 multiplier."""
    peak = np.float32(peak)
    if legacy_scalar:
        d = float(peak) if float(peak) > 0.01 else 0.01
        return np.float32(32767.0 / d)
    if not (peak > np.float32(0.01)):
        return np.float32(32767 / 0.01)
    return np.float32(np.float32(32767) / peak)


def normalise_line(audio, legacy_scalar=False):
    """audio: fp32 [n] valid samples of one line -> fp32 [n] after the two in-place scalings."""
    a = np.asarray(audio, dtype=np.float32).copy()
    for _ in range(2):
        peak = np.max(np.abs(a)) if a.size else np.float32(0)
        a = (a * peak_scale(peak, legacy_scalar)).astype(np.float32)
    return a


def assemble_song(lines, legacy_scalar=False):
    """lines: list of fp32 arrays (valid samples per lyric line, in song order) -> int16 song with the lead-in."""
    parts = [np.zeros(LEAD_IN, dtype=np.float64)]
    parts += [normalise_line(a, legacy_scalar).astype(np.float64) for a in lines]
    return np.concatenate(parts).astype(np.int16)
