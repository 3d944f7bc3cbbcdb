"""Front-end tensorisation against the reference's own code: `process_input_data` is extracted from
pipline_inference_svs_v5.py with `ast` (the module itself imports servers' dependencies that are out of scope) and run
on the same syllable strings; the vocabularies go through the reference's TokenTextEncoder."""
import ast
import os
import random

import pytest
import torch

from music_generation_b200 import frontend
from oracle import ref_harness

REF = ref_harness.REF_ROOT
HAVE_REF = os.path.isfile(os.path.join(REF, "pipline_inference_svs_v5.py")) and \
    os.path.isfile(os.path.join(REF, "VISinger", "phone_set.json"))
needs_ref = pytest.mark.skipif(not HAVE_REF, reason="reference tree / oracle/_ref stage not present")


def _reference_functions():
    """process_input_data + SHENG_MU compiled from the reference's driver source, nothing else of the module executed."""
    tree = ast.parse(open(os.path.join(REF, "pipline_inference_svs_v5.py"), encoding="utf-8").read())
    ns = {}
    for node in tree.body:
        if isinstance(node, ast.Assign) and getattr(node.targets[0], "id", "") == "SHENG_MU":
            exec(compile(ast.Module([node], []), "ref", "exec"), ns)
        if isinstance(node, ast.ClassDef) and node.name == "Synthesiser":
            fn = next(f for f in node.body if isinstance(f, ast.FunctionDef) and f.name == "process_input_data")
            exec(compile(ast.Module([fn], []), "ref", "exec"), ns)
    return ns


class _Norm:
    """Stands in for TextNormalizer: returns a prepared syllable string (the normaliser is out of scope)."""
    def __init__(self, out):
        self.out = out
    def normalize(self, text, word_parse=False):
        return self.out
    def __call__(self, text):
        return self.out


def _random_syllables(rng, tone_vocab, n):
    initials = sorted(frontend.SHENG_MU)
    finals = [t for t in tone_vocab if t[-1:].isdigit() and t[0] not in "bpmfdtnlgkhjqxzcsywr"]
    out = []
    for _ in range(n):
        r = rng.random()
        if r < 0.1:
            out.append(rng.choice(["s p", "a p"]))
        elif r < 0.8:
            out.append(rng.choice(initials) + rng.choice(finals))
        else:
            out.append(rng.choice(finals))
    return " ".join(out)


def test_note_to_midi_known_answers():
    assert [frontend.note_to_midi(n) for n in ("C4", "A4", "C#4", "Db4", "B3", "G#2", "Bb3", "C-1", "C")] == \
        [60, 69, 61, 61, 59, 44, 58, 0, 12]
    # format_note: rest -> 0, slurs keep the first note, octaves 5 / 6 fold to 4 (cpop_text_preprocessor.py:86)
    assert frontend.format_note(["rest", "C4", "D#5/E5", "A6", "F#3"]) == [0, 60, 63, 69, 54]
    with pytest.raises(ValueError):
        frontend.note_to_midi("H2")


def test_split_syllables_hand_case():
    s, n, l = frontend.split_syllables("zhang1 a p ai4 s p yu3", ["C4", "rest", "D4", "rest", "E4"], ["0.5", "0.2", "0.3", "0.1", "0.4"])
    assert s == "zh ang1 AP ai4 SP y u3"
    assert n == ["C4", "C4", "rest", "D4", "rest", "E4", "E4"] and l == ["0.5", "0.5", "0.2", "0.3", "0.1", "0.4", "0.4"]


@needs_ref
def test_front_end_matches_reference_code():
    import sys
    if REF not in sys.path:
        sys.path.insert(0, REF)
    ref_harness.import_reference()                       # installs the import stubs
    from VISinger.utils.text_encoder import TokenTextEncoder
    import json
    ns = _reference_functions()
    assert set(ns["SHENG_MU"]) == set(frontend.SHENG_MU)
    vdir = os.path.join(REF, "VISinger")
    phones_v = json.load(open(os.path.join(vdir, "phone_set.json")))
    tones_v = json.load(open(os.path.join(vdir, "phone_shengdiao_set.json")))
    ref_ph = TokenTextEncoder(None, vocab_list=phones_v, replace_oov=",")
    ref_tn = TokenTextEncoder(None, vocab_list=tones_v, replace_oov=",")
    fe = frontend.FrontEnd.from_dir(vdir, None)
    assert len(fe.phones) == ref_ph.vocab_size and len(fe.tones) == ref_tn.vocab_size
    rng = random.Random(7)
    names = ["C4", "D4", "E4", "F#4", "G3", "A3", "Bb3", "rest", "C5", "D#5/E5"]
    checked = 0
    for case in range(40):
        n = rng.randint(1, 14)
        syl = _random_syllables(rng, tones_v, n)
        notes = [rng.choice(names) for _ in range(n)]
        lens = ["%.4f" % rng.uniform(0.1, 0.9) for _ in range(n)]
        stub = type("S", (), {"text_normalizer": _Norm(syl)})()
        want_str, want_notes, want_lens = ns["process_input_data"](stub, "x", notes, lens)
        got = frontend.split_syllables(syl.replace(".", "").strip(), notes, lens)
        assert got == (want_str, want_notes, want_lens)
        try:
            want_tones = ref_tn.encode(want_str)
            want_phones = ref_ph.encode(want_str.translate(str.maketrans("", "", "12345")))
        except KeyError:                                  # a random initial+final outside the vocabulary: both sides raise
            fe.normalizer = _Norm(syl)
            with pytest.raises(KeyError):
                fe.get_text("n|lyrics|" + " ".join(notes) + "|" + " ".join(lens))
            continue
        fe.normalizer = _Norm(syl)
        if len(want_phones) != len(want_tones):           # a digits-only final vanishes from the toneless string: the
            with pytest.raises(AssertionError):           # reference's length assert (:116-117) fires, and so does ours
                fe.get_text("n|lyrics|" + " ".join(notes) + "|" + " ".join(lens))
            continue
        ph, nt, nl, sd = fe.get_text("n|ly rics|" + " ".join(notes) + "|" + " ".join(lens))
        assert ph.tolist() == want_phones and sd.tolist() == want_tones
        assert nt.tolist() == frontend.format_note(want_notes)
        assert torch.equal(nl, torch.FloatTensor([float(x) for x in want_lens]))
        assert ph.dtype == nt.dtype == sd.dtype == torch.int64 and nl.dtype == torch.float32
        checked += 1
    assert checked >= 10


def test_tensorise_pads_and_keeps_lengths():
    tab = frontend.PhoneTable(["AP", "SP", "a1", "b", "a", "zh", "ang1", "ang"])
    fe = frontend.FrontEnd(frontend.PhoneTable(["AP", "SP", "a", "b", "zh", "ang"]), tab, None)
    outs = iter(["ba1 a p", "zhang1"])
    fe.normalizer = lambda text: next(outs)
    per_line, batch = fe.tensorise(["0|x|C4 rest|0.5 0.25", "1|y|D4|0.75"], pin=False)
    assert batch["x_lengths"].tolist() == [3, 2] and batch["x"].shape == (2, 3)
    assert batch["x"][1].tolist() == [fe.phones.index["zh"], fe.phones.index["ang"], 0]
    assert batch["note_pitch_id"].tolist() == [[60, 60, 0], [62, 62, 0]]
    assert batch["note_duration"][0].tolist() == [0.5, 0.5, 0.25]
    assert batch["phone_shengdiao"][0].tolist() == [tab.index["b"], tab.index["a1"], tab.index["AP"]]
