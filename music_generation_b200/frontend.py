"""Front-end tensorisation: the id construction of `Synthesiser.get_text` /
`process_input_data` (pipline_inference_svs_v5.py:56-123; same code in the v5/v6 multispeaker drivers) as pure host
code, off the per-line critical path: a whole song's lines are converted once and collated into the padded, pinned
batch tensors `SynthesizerTrn.infer` / `pipeline.Synthesiser.syn_song` consume.

In scope here: splitting syllables into initial + final (SHENG_MU), duplicating the note / note length of a split
syllable, the phoneme and tone-phoneme vocabularies (`build_phone_encoder`, VISinger/cpop_text_preprocessor.py:15-18;
TokenTextEncoder.encode, VISinger/utils/text_encoder.py:197-209), note names -> MIDI numbers (`format_note`,
cpop_text_preprocessor.py:79-87).  NOT in scope: the Chinese text normaliser / G2P
(`TextNormalizer.normalize`) - it is passed in as a callable, exactly where the reference calls it.

The vocabularies are read from the deployment's own `phone_set.json` / `phone_shengdiao_set.json` (the reference reads
them from "./VISinger"); nothing of the reference's data is vendored.
"""
import json
import os
import re

import torch

# pipline_inference_svs_v5.py:29-30: syllable initials; two-letter initials are tried first
SHENG_MU = frozenset(['b', 'p', 'm', 'f', 'd', 't', 'n', 'l', 'g', 'k', 'h', 'j', 'q', 'x', 'z', 'c', 's', 'y', 'w',
                      'zh', 'ch', 'sh', 'r'])
RESERVED = ("<pad>", "<EOS>", "<UNK>")        # VISinger/utils/text_encoder.py:5-9: ids 0, 1, 2

_NOTE_BASE = {"C": 0, "D": 2, "E": 4, "F": 5, "G": 7, "A": 9, "B": 11}
_NOTE_RE = re.compile(r"^([A-Ga-g])([#b!♯♭]*)(-?\d+)?$")


class PhoneTable:
    """`TokenTextEncoder(None, vocab_list=..., replace_oov=',')` as the drivers build it: reserved tokens take ids
    0..2, the list follows; `encode` applies the reference's ve->ue, vn->un, van->uan rewrite per token and raises
    KeyError on an out-of-vocabulary token (the reference prints a note and then fails on the dict lookup)."""

    def __init__(self, vocab_list):
        self.tokens = list(RESERVED) + [t for t in vocab_list if t not in RESERVED]
        self.index = {t: i for i, t in enumerate(self.tokens)}

    @classmethod
    def from_file(cls, data_dir, fname="phone_set.json"):
        return cls(json.load(open(os.path.join(data_dir, fname))))

    def __len__(self):
        return len(self.tokens)

    def encode(self, sentence):
        toks = [t.replace("ve", "ue").replace("vn", "un").replace("van", "uan") for t in sentence.strip().split()]
        return [self.index[t] for t in toks]


def note_to_midi(note):
    """librosa.note_to_midi for one note name ("C4" -> 60, "A#3" -> 58, "Bb3" -> 58; no octave means 0 -> 12 + pc)."""
    m = _NOTE_RE.match(note.strip())
    if not m:
        raise ValueError(f"improper note format: {note}")
    pc = _NOTE_BASE[m.group(1).upper()]
    for a in m.group(2):
        pc += 1 if a in "#♯" else -1
    octave = int(m.group(3)) if m.group(3) else 0
    return 12 * (octave + 1) + pc


def format_note(notes):
    """cpop_text_preprocessor.py:79-87: 'rest' -> 0; slurred "X/Y" keeps X; octaves 5 and 6 are folded to 4 by the
    reference's string replace (which also rewrites those digits anywhere in the token - reproduced as is)."""
    return [note_to_midi(x.split("/")[0].replace("5", "4").replace("6", "4")) if x != "rest" else 0 for x in notes]


def split_syllables(phonemes, notes, note_lengths):
    """`process_input_data` after the normaliser (pipline_inference_svs_v5.py:66-93).  `phonemes`: the normalised
    tone-annotated syllable string ("s p" / "a p" spelled out as the normaliser emits them); one note and one note
    length per syllable.  Returns (tone-phoneme string, notes, note_lengths) with split syllables duplicated."""
    ph_seq, new_notes, new_lengths = [], [], []
    sylls = phonemes.replace("s p", "SP").replace("a p", "AP").split(" ")
    for pos, ph in enumerate(sylls):
        if ph in ("SP", "AP"):
            parts = [ph]
        elif ph[:2] in SHENG_MU:
            parts = [ph[:2], ph[2:]]
        elif ph[:1] in SHENG_MU:
            parts = [ph[:1], ph[1:]]
        else:
            parts = [ph]
        ph_seq += parts
        new_notes += [notes[pos]] * len(parts)
        new_lengths += [note_lengths[pos]] * len(parts)
    return " ".join(ph_seq), new_notes, new_lengths


_TONE_DIGITS = str.maketrans("", "", "12345")


class FrontEnd:
    """get_text for a deployment: `normalizer(text) -> str` is TextNormalizer.normalize(text, word_parse=False)."""

    def __init__(self, phone_table, tone_table, normalizer):
        self.phones, self.tones, self.normalizer = phone_table, tone_table, normalizer

    @classmethod
    def from_dir(cls, data_dir, normalizer):
        return cls(PhoneTable.from_file(data_dir), PhoneTable.from_file(data_dir, "phone_shengdiao_set.json"), normalizer)

    def get_text(self, line):
        """"name|lyrics|notes|note lengths" -> (phones int64 [P], notes int64 [P], note_lengths fp32 [P],
        phoneme_shengdiao int64 [P])  (pipline_inference_svs_v5.py:95-123)."""
        metas = line.split("|")
        text = metas[1].replace(" ", "")
        notes, note_lengths = metas[2].split(" "), metas[3].split(" ")
        phonemes = self.normalizer(text).replace(".", "").strip()
        tone_str, notes, note_lengths = split_syllables(phonemes, notes, note_lengths)
        phones = self.phones.encode(tone_str.translate(_TONE_DIGITS))
        tones = self.tones.encode(tone_str)
        midi = format_note(notes)
        lengths = [float(x) for x in note_lengths]
        if not (len(phones) == len(midi) == len(lengths) == len(tones)):
            raise AssertionError(f"len(phones)={len(phones)} == len(notes)={len(midi)} == len(note_lengths)={len(lengths)} "
                                 f"== len(phoneme_shengdiao) = {len(tones)}")
        return (torch.tensor(phones, dtype=torch.int64), torch.tensor(midi, dtype=torch.int64),
                torch.tensor(lengths, dtype=torch.float64).float(), torch.tensor(tones, dtype=torch.int64))

    def tensorise(self, lines, pin=True):
        """All lines of a song at once -> (per-line tuples, padded batch dict for SynthesizerTrn.infer).  The batch
        tensors are pinned so the H2D copy of a step is one asynchronous transfer per tensor."""
        per_line = [self.get_text(l) for l in lines]
        n, P = len(per_line), max(int(t[0].numel()) for t in per_line)
        batch = dict(x=torch.zeros(n, P, dtype=torch.int64), note_pitch_id=torch.zeros(n, P, dtype=torch.int64),
                     note_duration=torch.zeros(n, P, dtype=torch.float32),
                     phone_shengdiao=torch.zeros(n, P, dtype=torch.int64), x_lengths=torch.zeros(n, dtype=torch.int64))
        for i, (ph, nt, nl, sd) in enumerate(per_line):
            k = int(ph.numel())
            batch["x"][i, :k], batch["note_pitch_id"][i, :k] = ph, nt
            batch["note_duration"][i, :k], batch["phone_shengdiao"][i, :k] = nl, sd
            batch["x_lengths"][i] = k
        if pin and torch.cuda.is_available():
            batch = {k: v.pin_memory() for k, v in batch.items()}
        return per_line, batch
