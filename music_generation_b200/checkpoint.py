"""Checkpoint I/O and the packed-weight cache.

* `load_checkpoint` / `save_checkpoint` keep the reference's file format and semantics
  (VISinger/utils/utils.py:33-58 and :60-70): a dict with `model`, `iteration`, `optimizer`, `learning_rate`; keys
  missing from the file keep the model's current value (and are logged), extra keys are ignored.
* Loading is followed by weight packing (weight-norm folding, K-major 16-bit tiles, split hi/lo copies, gate
  interleaving - `SynthesizerTrn._pack`).  That is pure host work that depends only on the weights and the
  precision knobs, so its result is cached next to the `.pth` as `<name>.pth.b200pack` and re-used when the
  fingerprint (a BLAKE2 digest of every state-dict tensor + the knobs + the pack format version) matches.
  The cache holds tensors and plain containers only and is read with `torch.load(weights_only=True)`.
  Measured in this image (54 M parameters, 16 host cores): packing 0.11 s; reading the 149 MB cache 0.80 s plus
  0.47 s for the fingerprint - so the cache is OFF by default (`packed_cache=False`) and only pays off on hosts
  whose CPU packing is slower than their disk.
* `SingerBank` keeps several checkpoints resident and picks one by `singer_id`, the way the multi-speaker
  driver keeps `syn`, `syn_f09`, `syn_f19` (pipline_inference_svs_v6_multispeaker.py:35-36, :167-169) and selects
  per request (:234-239): ids without an entry of their own fall back to the default model.
"""
import hashlib
import logging
import os

import torch

from . import ops
from .synthesizer import _Enc

logger = logging.getLogger(__name__)

PACK_FORMAT = 5            # bump when packer.py / SynthesizerTrn._pack change what they produce
CACHE_SUFFIX = ".b200pack"


# ---------------------------------------------------------------------------------------------- reference format
def load_checkpoint(checkpoint_path, model, optimizer=None, *, packed_cache=False, allow_pickle=False):
    """Reference signature and return value: (model, optimizer, learning_rate, iteration).
    The reference format is plain tensors / dicts / numbers, so the file is read with `weights_only=True` (no arbitrary
    pickle execution); `allow_pickle=True` is the explicit opt-in for a checkpoint that carries other objects (the
    reference's own `torch.load(path, map_location='cpu')` on an old torch, VISinger/utils/utils.py:35)."""
    assert os.path.isfile(checkpoint_path), checkpoint_path
    try:
        ckpt = torch.load(checkpoint_path, map_location="cpu", weights_only=True)
    except Exception:
        if not allow_pickle:
            raise
        ckpt = torch.load(checkpoint_path, map_location="cpu", weights_only=False)
    iteration = ckpt["iteration"]
    learning_rate = ckpt["learning_rate"]
    if optimizer is not None:
        optimizer.load_state_dict(ckpt["optimizer"])
    saved = ckpt["model"]
    target = model.module if hasattr(model, "module") else model
    new_sd = {}
    for k, v in target.state_dict().items():
        if k in saved:
            new_sd[k] = saved[k]
        else:
            logger.info("%s is not in the checkpoint", k)
            new_sd[k] = v
    target.load_state_dict(new_sd)
    logger.info("Loaded checkpoint '%s' (iteration %s)", checkpoint_path, iteration)
    if packed_cache and hasattr(target, "_pack"):
        attach_packed(target, checkpoint_path + CACHE_SUFFIX)
    return model, optimizer, learning_rate, iteration


def save_checkpoint(model, optimizer, learning_rate, iteration, checkpoint_path):
    logger.info("Saving model and optimizer state at iteration %s to %s", iteration, checkpoint_path)
    target = model.module if hasattr(model, "module") else model
    torch.save({"model": target.state_dict(), "iteration": iteration,
                "optimizer": None if optimizer is None else optimizer.state_dict(),
                "learning_rate": learning_rate}, checkpoint_path)


# ---------------------------------------------------------------------------------------------- packed-weight cache
def fingerprint(model):
    """Digest of everything `_pack` depends on."""
    h = hashlib.blake2b(digest_size=20)
    gf = model.generator_fold
    knobs = (PACK_FORMAT, model.variant, str(model.operand_dtype), model.f0_precision, model.text_precision,
             model.gin_channels, model.resblock, tuple(model.upsample_rates), model.hidden_channels,
             repr(sorted(gf.items())) if isinstance(gf, dict) else gf, tuple(model.resblock_kernel_sizes),
             tuple(tuple(d) for d in model.resblock_dilation_sizes), tuple(model.upsample_kernel_sizes),
             model.upsample_initial_channel, model.n_layers, model.n_heads, model.filter_channels, model.kernel_size,
             getattr(model, "generator_pair", None))
    h.update(repr(knobs).encode())
    for k, v in model.state_dict().items():
        t = v.detach().cpu().contiguous()
        h.update(k.encode())
        h.update(repr((tuple(t.shape), str(t.dtype))).encode())
        h.update(t.view(torch.uint8).numpy().tobytes() if t.numel() else b"")
    return h.hexdigest()


def _encode(obj):
    if isinstance(obj, torch.Tensor):
        return obj.detach().cpu()
    if isinstance(obj, ops.PackedConv):
        return {"__pc__": {s: _encode(getattr(obj, s)) for s in ops.PackedConv.__slots__}}
    if isinstance(obj, _Enc):
        return {"__enc__": {s: _encode(getattr(obj, s)) for s in _Enc.__slots__}}
    if isinstance(obj, dict):      # "c" / "c_gen" / "c_enc": ctypes weight tables of the stage-level API, rebuilt on attach
        return {"__dict__": {str(k): _encode(v) for k, v in obj.items() if k not in ("c", "c_gen", "c_enc")}}
    if isinstance(obj, tuple):
        return {"__tuple__": [_encode(v) for v in obj]}
    if isinstance(obj, list):
        return {"__list__": [_encode(v) for v in obj]}
    if obj is None or isinstance(obj, (bool, int, float, str)):
        return obj
    raise TypeError(f"packed weights hold an unsupported object: {type(obj).__name__}")


def _decode(obj, dev):
    if isinstance(obj, torch.Tensor):
        return obj.to(dev)
    if isinstance(obj, dict):
        if "__pc__" in obj:
            pc = ops.PackedConv.__new__(ops.PackedConv)
            for s, v in obj["__pc__"].items():
                setattr(pc, s, _decode(v, dev))
            return pc
        if "__enc__" in obj:
            e = _Enc()
            for s, v in obj["__enc__"].items():
                setattr(e, s, _decode(v, dev))
            return e
        if "__dict__" in obj:
            return {k: _decode(v, dev) for k, v in obj["__dict__"].items()}
        if "__tuple__" in obj:
            return tuple(_decode(v, dev) for v in obj["__tuple__"])
        if "__list__" in obj:
            return [_decode(v, dev) for v in obj["__list__"]]
        raise ValueError("malformed packed-weight cache")
    return obj


def save_packed(model, path):
    """Pack (if not packed yet) and write the cache file.  Returns the fingerprint."""
    P = model._packed or model._pack()
    fp = fingerprint(model)
    tmp = f"{path}.tmp{os.getpid()}"
    torch.save({"format": PACK_FORMAT, "fingerprint": fp, "packed": _encode(P)}, tmp)
    os.replace(tmp, path)
    return fp


def load_packed(model, path):
    """Attach the cached packing if the file exists and matches the model's weights and knobs."""
    if not os.path.isfile(path):
        return False
    try:
        blob = torch.load(path, map_location="cpu", weights_only=True)
    except Exception as e:           # truncated / foreign file: fall back to packing
        logger.warning("packed-weight cache %s unreadable (%s): repacking", path, e)
        return False
    if blob.get("format") != PACK_FORMAT or blob.get("fingerprint") != fingerprint(model):
        return False
    model._packed = _decode(blob["packed"], model.device)
    model._build_stage_tables(model._packed)
    return True


def attach_packed(model, path):
    """load_packed, else pack and (best effort) write the cache.  Returns "hit" | "miss" | "miss-unwritable"."""
    if load_packed(model, path):
        return "hit"
    try:
        save_packed(model, path)
        return "miss"
    except OSError as e:
        logger.warning("cannot write packed-weight cache %s: %s", path, e)
        model._packed or model._pack()
        return "miss-unwritable"


# ---------------------------------------------------------------------------------------------- resident singers
class SingerBank:
    """Several checkpoints of one architecture resident on one device, selected by singer id.

        bank = SingerBank(lambda: SynthesizerTrn(100, 513, 32, **hps.model), {0: "G_AE.pth", 1: "G_AE_F09.pth",
                                                                              2: "G_AE_F19.pth"}, device="cuda")
        net_g = bank.get(singer_id)       # ids without an entry use the default (the reference's `else: syn`)
    """

    def __init__(self, factory, checkpoints, device="cuda", default=0, packed_cache=False):
        if default not in checkpoints:
            raise KeyError(f"default singer {default} has no checkpoint")
        self.default = default
        self.models = {}
        self.iterations = {}
        for sid, path in checkpoints.items():
            net = factory().eval().to(device)
            _, _, _, it = load_checkpoint(path, net, None, packed_cache=packed_cache)
            self.models[sid] = net
            self.iterations[sid] = it

    def get(self, singer_id=None):
        return self.models.get(singer_id, self.models[self.default])

    def __len__(self):
        return len(self.models)
