"""Batched song driver - the B200 counterpart of `Synthesiser.syn` + the song loop of the reference's pipelines
(pipline_inference_svs_v5.py:125-139, 206-259; pipline_inference_svs_v6_multispeaker.py:212-259).

The reference calls `infer` once per lyric line at B=1, copies every line to the host, peak-normalises it with numpy
(twice), concatenates and converts to int16.  Here all lines of a song (or of several requests) form ONE ragged batch
for `SynthesizerTrn.infer`, and normalisation + int16 conversion + concatenation + the 0.2 s lead-in run as one
device pass (`b200svs_song_assemble`); a single int16 tensor crosses PCIe.

The text front-end (`get_text`: phoneme / tone / midi id construction) is not part of this component (8f-4):
lines arrive as the id tensors `get_text` returns."""
import torch

from . import ops

SAMPLE_RATE = 22050
LEAD_IN = int(0.2 * SAMPLE_RATE)


def save_wav(path, song, sample_rate=SAMPLE_RATE):
    """Writes an int16 mono song (what `syn_song` / `syn_long` return; a device tensor is copied to the host) as a PCM
    WAV file - the file `scipy.io.wavfile.write(path, SAMPLE_RATE, wavs.astype(np.int16))` writes at the end of the
    reference's drivers (pipline_inference_svs_v5.py:139, 256-258), byte for byte, with the standard library only."""
    import wave
    if isinstance(song, torch.Tensor):
        if song.dtype != torch.int16 or song.dim() != 1:
            raise ValueError("save_wav: a 1-D int16 tensor is expected")
        data = song.detach().cpu().contiguous().numpy().tobytes()
    else:
        import numpy as np
        a = np.ascontiguousarray(song)
        if a.dtype != np.int16 or a.ndim != 1:
            raise ValueError("save_wav: a 1-D int16 array is expected")
        data = a.tobytes()
    with wave.open(path, "wb") as f:
        f.setnchannels(1)
        f.setsampwidth(2)
        f.setframerate(int(sample_rate))
        f.writeframes(data)
    return path


class Synthesiser:
    def __init__(self, net_g, legacy_scalar=False):
        """net_g: a `music_generation_b200` SynthesizerTrn (loaded, on a CUDA device).
        legacy_scalar: reproduce NumPy < 2 scalar arithmetic in the peak normalisation (see oracle/song_oracle.py)."""
        self.net_g = net_g
        self.legacy_scalar = legacy_scalar
        self.last = {}

    @staticmethod
    def collate(lines):
        """lines: list of (phones [P_i] int64, notes [P_i] int64, note_lengths [P_i] fp32, phoneme_shengdiao [P_i] int64)
        -> zero-padded batch tensors + x_lengths (what `syn` builds per line with unsqueeze(0))."""
        n = len(lines)
        P = max(int(l[0].numel()) for l in lines)
        x = torch.zeros(n, P, dtype=torch.int64)
        notes = torch.zeros(n, P, dtype=torch.int64)
        durs = torch.zeros(n, P, dtype=torch.float32)
        tones = torch.zeros(n, P, dtype=torch.int64)
        lens = torch.zeros(n, dtype=torch.int64)
        for i, (ph, nt, nl, sd) in enumerate(lines):
            k = int(ph.numel())
            x[i, :k], notes[i, :k], durs[i, :k], tones[i, :k], lens[i] = ph, nt, nl, sd, k
        return x, lens, notes, durs, tones

    @torch.no_grad()
    def syn_song(self, lines, *, noise=None, ph_dur=None, g=None):
        """Synthesise all lines in one batch and return the int16 song [LEAD_IN + sum_i samples_i] on the device.
        `noise`, `ph_dur`, `g` are forwarded to `infer` (parity runs)."""
        dev = self.net_g.device
        x, lens, notes, durs, tones = (t.to(dev) for t in self.collate(lines))
        # only the valid samples of every line reach the song: let the Generator skip the padded frames of short lines
        bucket, self.net_g.generator_bucket = self.net_g.generator_bucket, True
        try:
            res = self.net_g.infer(x, lens, notes, durs, phone_shengdiao=tones, noise_scale=1., length_scale=1,
                                   noise=noise, ph_dur=ph_dur, g=g)
        finally:
            self.net_g.generator_bucket = bucket
        o = res[0]
        mel_len = res[1].sum(dim=(1, 2)).to(torch.int64)                   # y_mask [B,1,T] -> frames per line
        samples = mel_len * self.net_g.hop
        offsets = torch.cumsum(samples, 0) - samples
        total = int(samples.sum().item())
        song = torch.empty(LEAD_IN + total, device=dev, dtype=torch.int16)
        scales = torch.empty(len(lines), 2, device=dev, dtype=torch.float32)
        with torch.cuda.device(dev):          # the library launches on the current device's stream
            ops.song_assemble(o, mel_len, offsets.contiguous(), song, hop=self.net_g.hop, lead_in=LEAD_IN,
                              legacy_scalar=self.legacy_scalar, scales_out=scales)
        self.last = {"o": o, "mel_len": mel_len, "offsets": offsets, "scales": scales}
        return song

    @torch.no_grad()
    def syn_long(self, line, *, segment_frames=None, overlap=0, noise=None, ph_dur=None, g=None):
        """One utterance of any length (a multi-minute take; BASELINE config 4) -> peak-normalised int16 audio with the
        0.2 s lead-in, on the device.  `line`: the (phones, notes, note_lengths, phoneme_shengdiao) tuple `get_text`
        returns.  Frame-level transformers per segment, everything convolutional over the full length
        (SynthesizerTrn.infer_long, segments.py): no seam between segments, unlike the reference's per-line concatenation
        (pipline_inference_svs_v6_multispeaker.py:242)."""
        dev = self.net_g.device
        x, lens, notes, durs, tones = (t.to(dev) for t in self.collate([line]))
        res = self.net_g.infer_long(x, lens, notes, durs, tones, ph_dur=ph_dur, segment_frames=segment_frames, overlap=overlap,
                                    noise=noise, g=g)
        o = res[0]
        mel_len = res[1].sum(dim=(1, 2)).to(torch.int64)
        total = int(mel_len.sum().item()) * self.net_g.hop
        song = torch.empty(LEAD_IN + total, device=dev, dtype=torch.int16)
        with torch.cuda.device(dev):
            ops.song_assemble(o, mel_len, torch.zeros(1, device=dev, dtype=torch.int64), song, hop=self.net_g.hop,
                              lead_in=LEAD_IN, legacy_scalar=self.legacy_scalar)
        self.last = {"o": o, "mel_len": mel_len}
        return song
