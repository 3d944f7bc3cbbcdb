"""Generates tests/golden/*.npz by running the UNMODIFIED reference (/root/reference) in the build
container.  Weights are not stored: both sides rebuild them deterministically as
`music_generation_b200` seeded init + synthetic.apply_recipe (the reference model loads that
state-dict by name, which also proves state-dict compatibility).

    python -m oracle.make_golden          # rewrites every fixture
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from music_generation_b200 import models, models_ae, synthetic  # noqa: E402
from oracle import ref_harness  # noqa: E402

HP = dict(inter_channels=192, hidden_channels=192, filter_channels=768, n_heads=2, n_layers=6, kernel_size=3,
          p_dropout=0.1, resblock="1", resblock_kernel_sizes=[3, 7, 11],
          resblock_dilation_sizes=[[1, 3, 5], [1, 3, 5], [1, 3, 5]], upsample_rates=[8, 8, 2, 2],
          upsample_initial_channel=512, upsample_kernel_sizes=[16, 16, 4, 4])

# name -> (variant, batch, phonemes, frames (None = predicted durations), ragged, seed)
CASES = {
    "ae_phdur_b2": ("ae", 2, 12, 70, True, 1234),
    "ae_pred_b1": ("ae", 1, 6, None, False, 4321),
    "v4_phdur_b2": ("v4", 2, 10, 60, True, 99),
}


# speaker-conditioned stages (the reference's infer hard-codes g=None, models_ae.py:634, so the g path of
# ResidualCouplingBlock / SpecDecoder / Generator is pinned stage by stage): name -> (batch, frames, gin, seed)
G_CASES = {"ae_gin256_stages": (2, 40, 256, 777)}


def g_case_inputs(name):
    b, t, gin, seed = G_CASES[name]
    g_ = torch.Generator().manual_seed(seed)
    z_p = torch.randn(b, 192, t, generator=g_)
    mel_len = torch.tensor([t] + [max(1, t - 13 * i) for i in range(1, b)], dtype=torch.int64)
    g = torch.randn(b, gin, 1, generator=g_)
    return z_p, mel_len, g, gin


def posterior_spectrogram(name):
    b, t, gin, seed = G_CASES[name]
    return torch.rand(b, 513, t, generator=torch.Generator().manual_seed(seed + 1)) * 2.0


def posterior_noise(name):
    b, t, gin, seed = G_CASES[name]
    return torch.randn(b, 192, t, generator=torch.Generator().manual_seed(seed + 2))


def state_dict_for(variant, seed=1234, gin=0):
    mod = models_ae if variant == "ae" else models
    net = mod.SynthesizerTrn(100, 513, 48, **HP, seed=seed, gin_channels=gin)
    sd = {k: v.clone() for k, v in net.state_dict().items()}
    return synthetic.apply_recipe(sd, seed=seed)


def case_inputs(name):
    variant, b, p, frames, ragged, seed = CASES[name]
    inp = synthetic.make_inputs(b, p, seed=seed, ragged=ragged)
    ph_dur = synthetic.make_ph_dur(inp["x_lengths"], p, frames, seed=seed) if frames else None
    return variant, inp, ph_dur, seed


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    torch.set_num_threads(1)          # fixed reduction order
    for name in CASES:
        variant, inp, ph_dur, seed = case_inputs(name)
        sd = state_dict_for(variant)
        net = ref_harness.build_reference(variant)
        missing = net.load_state_dict(sd, strict=True)
        assert not missing.missing_keys and not missing.unexpected_keys
        if ph_dur is not None:
            T = int(ph_dur.sum(1).max())
        else:   # frame count of the reference's own predicted durations
            with torch.no_grad():
                h, m = net.text_encoder(inp["x"], inp["note_pitch_id"], inp["note_duration"], inp["x_lengths"],
                                        inp["phone_shengdiao"])
                d = torch.floor(net.duration_predictor(h, m) * inp["note_duration"] / 0.0116).long()
            T = int(d.sum(1).max())
        noise = synthetic.make_noise(inp["x"].shape[0], 192, T, seed=seed)
        res = ref_harness.run_reference_infer(net, inp, noise, ph_dur)
        o, y_mask, (z, z_p, m_p, logs_p), log_f0 = res[0], res[1], res[2], res[3]
        mel = res[-1]
        rec = dict(o=o.numpy(), y_mask=y_mask.numpy(), log_f0=log_f0.numpy(), T=np.int64(T),
                   z=z[:, ::4].numpy(), z_p=z_p[:, ::4].numpy(), m_p=m_p[:, ::4].numpy(), logs_p=logs_p[:, ::4].numpy(),
                   mel=mel[:, ::4].numpy())
        if variant == "ae":
            rec["spec"] = res[4][:, ::16].numpy()
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), **rec)
        print(name, "T =", T, "o", tuple(o.shape), "|o|max %.3f" % o.abs().max())
    for name in G_CASES:
        z_p, mel_len, g, gin = g_case_inputs(name)
        sd = state_dict_for("ae", gin=gin)
        net = ref_harness.build_reference("ae", gin_channels=gin)
        missing = net.load_state_dict(sd, strict=True)
        assert not missing.missing_keys and not missing.unexpected_keys
        with torch.no_grad():
            y_mask = (torch.arange(z_p.shape[2])[None] < mel_len[:, None]).unsqueeze(1).float()
            z = net.spec_flow(z_p * y_mask, y_mask, g=g, reverse=True)          # models_ae.py:246-253, modules.py:152-159
            spec = net.spec_decoder(z, mel_len, g=g)                            # ext_modules/spec_decoder.py:41-46
            hid = net.spec_proj(spec) * y_mask
            o = net.decoder(hid, g=g)                                           # models_ae.py:327-346 with cond (:329-330)
            # training-side PosteriorEncoder (models_ae.py:256-291) on a synthetic spectrogram, with g
            y = posterior_spectrogram(name)
            pn = posterior_noise(name)
            orig = torch.randn_like
            torch.randn_like = lambda t, *a, **k: pn.to(t.dtype)
            try:
                pz, pm, plogs, pmask, px = net.spec_encoder(y, mel_len, g=g)
            finally:
                torch.randn_like = orig
            pzp = net.spec_flow(pz, y_mask, g=g)                                # forward flow (training side), models_ae.py:246-253
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), z=z.numpy(), spec=spec[:, ::8].numpy(), hid=hid.numpy(),
                            o=o.numpy(), post_z=pz.numpy(), post_m=pm.numpy(), post_logs=plogs.numpy(), post_x=px.numpy(),
                            post_zp=pzp.numpy())
        print(name, "o", tuple(o.shape), "|o|max %.3f" % o.abs().max())
    json.dump({"cases": {k: list(v) for k, v in CASES.items()}, "g_cases": {k: list(v) for k, v in G_CASES.items()},
               "torch": torch.__version__,
               "reference": "GuluOfficial/music_generation VISinger/models_ae.py + models.py (unmodified, CPU fp32)"},
              open(os.path.join(out_dir, "MANIFEST.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
