"""TEST / BASELINE INFRASTRUCTURE ONLY - stages the UNMODIFIED reference sources of the hot path into oracle/_ref/.

The reference is a pure-Python source tree with no setup.py, so it cannot be pip-installed and /root/reference does
not exist on the GPU box.  `__graft_entry__.build()` (run in the build container, where /root/reference exists) calls
`stage()`: the files SURVEY.md 8(a) cites are copied byte for byte into the git-ignored directory oracle/_ref/
(listed in .gitignore, NOT in .gpurunignore, so it travels to the GPU box with the snapshot like the built .so),
together with MANIFEST.json holding the sha256 of every file - the copy is never edited and never enters history.
`oracle/ref_harness.py` imports the reference from /root/reference when present, else from oracle/_ref/, so that
`bench.py --impl reference` and `cpu_baseline` time the reference's own `SynthesizerTrn.infer` on the box's host
cores (`kind: "reference"`).  Nothing under music_generation_b200/ imports this.

    python -m oracle.stage_ref            # (re)stage, prints the manifest summary
"""
import hashlib
import json
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
SRC_ROOT = os.environ.get("B200SVS_REFERENCE_ROOT", "/root/reference")
DST_ROOT = os.path.join(HERE, "_ref")

# the import closure of VISinger.models_ae / VISinger.models `SynthesizerTrn.infer` (SURVEY.md 8a + Appendix A)
FILES = [
    "VISinger/models_ae.py", "VISinger/models.py", "VISinger/modules.py", "VISinger/attentions.py",
    "VISinger/commons.py", "VISinger/transforms.py",
    "VISinger/ext_modules/__init__.py", "VISinger/ext_modules/f0_predictor.py",
    "VISinger/ext_modules/frame_prior_network.py", "VISinger/ext_modules/mel_predict_network.py",
    "VISinger/ext_modules/modules.py", "VISinger/ext_modules/phoneme_predictor.py",
    "VISinger/ext_modules/spec_decoder.py", "VISinger/ext_modules/text_post_net.py",
    "VISinger/utils/__init__.py", "VISinger/utils/utils.py", "VISinger/utils/pitch_utils.py",
    "VISinger/utils/config_loader.py", "VISinger/utils/audio.py", "VISinger/utils/logger.py",
    "VISinger/configs/ours_ae.json", "VISinger/configs/ours.json",
    # front-end tensorisation: vocabularies + the reference's encoder and driver (parsed, never imported whole)
    "VISinger/phone_set.json", "VISinger/phone_shengdiao_set.json", "VISinger/utils/text_encoder.py",
    "pipline_inference_svs_v5.py",
]


def _sha(path):
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def staged() -> bool:
    return os.path.exists(os.path.join(DST_ROOT, "MANIFEST.json"))


def stage(verbose=False) -> bool:
    """Copies FILES from the reference tree; returns False (and leaves an existing stage alone) when the reference
    tree is not present (GPU box)."""
    if not os.path.isdir(os.path.join(SRC_ROOT, "VISinger")):
        return staged()
    manifest = {}
    for rel in FILES:
        src, dst = os.path.join(SRC_ROOT, rel), os.path.join(DST_ROOT, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        manifest[rel] = _sha(dst)
        assert manifest[rel] == _sha(src)
    json.dump({"source": SRC_ROOT, "files": manifest}, open(os.path.join(DST_ROOT, "MANIFEST.json"), "w"), indent=1)
    if verbose:
        print(f"staged {len(manifest)} reference files into {DST_ROOT}")
    return True


def verify() -> bool:
    """True when every staged file still matches its manifest hash (the stage is unmodified)."""
    if not staged():
        return False
    m = json.load(open(os.path.join(DST_ROOT, "MANIFEST.json")))
    return all(os.path.exists(os.path.join(DST_ROOT, r)) and _sha(os.path.join(DST_ROOT, r)) == h
               for r, h in m["files"].items())


if __name__ == "__main__":
    print("staged:", stage(verbose=True), "verified:", verify())
