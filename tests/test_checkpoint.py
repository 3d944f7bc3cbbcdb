"""Checkpoint I/O in the reference's file format, the packed-weight cache and the resident singer bank
(VISinger/utils/utils.py:33-70, pipline_inference_svs_v6_multispeaker.py:167-169, :234-239)."""
import os

import torch

from music_generation_b200 import checkpoint, models_ae
from conftest import HP
from common import state_dict_for


def _flat(obj, out):
    if isinstance(obj, torch.Tensor):
        out.append(obj)
    elif isinstance(obj, dict):
        for k in sorted(obj):
            _flat(obj[k], out)
    elif isinstance(obj, (list, tuple)):
        for v in obj:
            _flat(v, out)
    else:
        out.append(obj)
    return out


def _net():
    return models_ae.SynthesizerTrn(100, 513, 48, **HP).eval()


def test_checkpoint_roundtrip_cache_and_bank(tmp_path):
    sd = state_dict_for("ae")
    src = _net()
    src.load_state_dict(sd)
    path = str(tmp_path / "G_AE.pth")
    checkpoint.save_checkpoint(src, None, 2e-4, 250000, path)
    blob = torch.load(path, map_location="cpu", weights_only=False)
    assert set(blob) == {"model", "iteration", "optimizer", "learning_rate"}      # the reference's layout

    # a key missing from the file keeps the model's own value (utils.py:46-51); extra keys are ignored
    dropped = "dec.conv_post.weight_v" if "dec.conv_post.weight_v" in blob["model"] else next(iter(blob["model"]))
    partial = dict(blob, model={k: v for k, v in blob["model"].items() if k != dropped})
    partial["model"]["not.a.parameter"] = torch.zeros(1)
    ppath = str(tmp_path / "partial.pth")
    torch.save(partial, ppath)
    net = _net()
    own = net.state_dict()[dropped].clone()
    _, opt, lr, it = checkpoint.load_checkpoint(ppath, net, None, packed_cache=False)
    assert (opt, lr, it) == (None, 2e-4, 250000)
    assert torch.equal(net.state_dict()[dropped], own)
    other = next(k for k in blob["model"] if k != dropped and blob["model"][k].is_floating_point())
    assert torch.equal(net.state_dict()[other], blob["model"][other])

    # first load packs and writes the cache, the second attaches it; both equal a fresh packing
    net1 = _net()
    checkpoint.load_checkpoint(path, net1, None, packed_cache=True)
    cache = path + checkpoint.CACHE_SUFFIX
    assert os.path.isfile(cache) and net1._packed is not None
    net2 = _net()
    net2.load_state_dict(sd)
    assert checkpoint.load_packed(net2, cache)
    fresh = _flat(checkpoint._encode(src._pack()), [])
    cached = _flat(checkpoint._encode(net2._packed), [])
    assert len(fresh) == len(cached) > 100
    for a, b in zip(fresh, cached):
        if isinstance(a, torch.Tensor):
            assert a.dtype == b.dtype and torch.equal(a, b)
        else:
            assert a == b

    # different weights or different precision knobs: the fingerprint rejects the cache
    net3 = _net()
    sd3 = {k: v.clone() for k, v in sd.items()}
    k0 = next(k for k in sd3 if k.endswith("conv_pre.bias"))
    sd3[k0] += 1.0
    net3.load_state_dict(sd3)
    assert not checkpoint.load_packed(net3, cache)
    net4 = models_ae.SynthesizerTrn(100, 513, 48, **HP, f0_precision="fp32").eval()
    net4.load_state_dict(sd)
    assert not checkpoint.load_packed(net4, cache)
    with open(str(tmp_path / "junk.b200pack"), "wb") as f:
        f.write(b"not a cache")
    assert not checkpoint.load_packed(net2, str(tmp_path / "junk.b200pack"))

    # resident singers: unknown ids fall back to the default model (the driver's `else: syn`)
    bank = checkpoint.SingerBank(_net, {0: path, 2: ppath}, device="cpu", packed_cache=False)
    assert len(bank) == 2 and bank.iterations == {0: 250000, 2: 250000}
    assert bank.get(2) is bank.models[2] and bank.get(1) is bank.models[0] and bank.get() is bank.models[0]
