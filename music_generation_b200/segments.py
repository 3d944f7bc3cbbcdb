"""Segment plan for utterances longer than the frame-level transformer's position table.

The reference cannot synthesise more than 2000 frames (23.2 s) per call: TextPostNet adds a 2000-row position table
(VISinger/ext_modules/text_post_net.py:26,69-72) and its drivers cut a song into lyric lines that are synthesised
independently and concatenated with no overlap (pipline_inference_svs_v6_multispeaker.py:242), so every seam restarts
the flow / SpecDecoder / Generator convolutions from zero padding.  Here only the part that HAS the limit is segmented:
the frame-level transformers run per segment (core frames plus `overlap` frames of context on each side, positions
restarting at the segment's first context frame), their per-frame statistics are written back into one full-length
tensor, and everything convolutional (prior sample, reverse flow, SpecDecoder, Generator) runs over the full length -
no seam exists there, so the waveform has no discontinuity at segment boundaries.
"""
from typing import List, Tuple


def plan_segments(n_frames: int, segment_frames: int, overlap: int) -> List[Tuple[int, int, int, int]]:
    """Consecutive cores of `segment_frames` frames covering [0, n_frames); each with its context window clipped to
    the utterance.  Returns [(core0, core1, ctx0, ctx1)], ctx1 - ctx0 <= segment_frames + 2 * overlap."""
    if segment_frames <= 0 or overlap < 0:
        raise ValueError(f"segment_frames={segment_frames} overlap={overlap}")
    out = []
    c0 = 0
    while c0 < n_frames:
        c1 = min(n_frames, c0 + segment_frames)
        out.append((c0, c1, max(0, c0 - overlap), min(n_frames, c1 + overlap)))
        c0 = c1
    return out
