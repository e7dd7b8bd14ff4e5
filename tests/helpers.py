"""Shared test helpers: golden-fixture loading and module construction."""
import os

import numpy as np
import torch

from oracle import nsf_hifigan_oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    d = np.load(os.path.join(GOLDEN, name + ".npz"))
    cfg = O.DecoderConfig(
        initial_channel=int(d["cfg_initial_channel"]),
        resblock_kernel_sizes=tuple(int(v) for v in d["cfg_resblock_kernel_sizes"]),
        resblock_dilation_sizes=tuple(tuple(int(v) for v in row) for row in d["cfg_resblock_dilation_sizes"]),
        upsample_rates=tuple(int(v) for v in d["cfg_upsample_rates"]),
        upsample_initial_channel=int(d["cfg_upsample_initial_channel"]),
        upsample_kernel_sizes=tuple(int(v) for v in d["cfg_upsample_kernel_sizes"]),
        gin_channels=int(d["cfg_gin_channels"]),
        sr=int(d["cfg_sr"]),
    )
    return d, cfg


def build_module(cfg, seed):
    """Our drop-in module, default-initialised under `seed` exactly as the reference would be."""
    from rvc_b200 import HiFiGANNSFGenerator

    torch.manual_seed(seed)
    return HiFiGANNSFGenerator(*cfg.ctor_args(), **cfg.ctor_kwargs()).eval()


def golden_state(d, cfg):
    """state_dict for a golden case: stored weights if present, else re-created from the init seed
    (checksums of every tensor are verified against the fixture)."""
    keys = [k for k in d.files if k.startswith("sd::")]
    if keys:
        return {k[4:]: torch.from_numpy(d[k]) for k in keys}
    m = build_module(cfg, int(d["init_seed"]))
    sd = m.state_dict()
    for k in d.files:
        if k.startswith("ck::"):
            v = sd[k[4:]].double()
            got = np.array([float(v.sum()), float(v.abs().sum())])
            np.testing.assert_allclose(got, d[k], rtol=1e-12, atol=1e-12, err_msg=f"init differs from the reference for {k[4:]}")
    return sd


def tensors(d):
    x = torch.from_numpy(d["x"])
    f0 = torch.from_numpy(d["f0"])
    g = torch.from_numpy(d["g"])
    noise = torch.from_numpy(d["noise"])
    y = torch.from_numpy(d["y"])
    return x, f0, g, noise, y


# ---- round-2 helpers -----------------------------------------------------------------------------------------------
def load_real_case():
    """tests/golden/real_48k_T1350.npz (oracle/make_golden_r2.py): inputs re-created from seeds and verified against the
    fixture's checksums, the reference's real RMVPE contour, and the stored pieces of the reference's waveform."""
    d = np.load(os.path.join(GOLDEN, "real_48k_T1350.npz"))
    cfg = O.CONFIG_48K
    T = d["f0"].shape[1]
    x, _, g, _ = O.synth_inputs(cfg, 1, T, seed=int(d["in_seed"]))
    noise = O.draw_reference_noise(int(d["noise_seed"]), 1, T * cfg.upp)
    for name, t in (("x", x), ("g", g), ("noise", noise)):
        got = np.array([float(t.double().sum()), float(t.double().abs().sum())])
        np.testing.assert_allclose(got, d[name + "_ck"], rtol=1e-12, atol=1e-9, err_msg=f"re-created {name} differs from the fixture's")
    return d, cfg, x, torch.from_numpy(d["f0"]), g, noise


def check_against_real_fixture(d, y, tol):
    """y (648000,) float32 numpy against the stored segments and per-frame block sums of the reference's waveform."""
    upp = 480
    worst = 0.0
    for off, seg in zip(d["seg_off"], d["seg"]):
        worst = max(worst, float(np.abs(y[off:off + seg.shape[0]] - seg).max()))
    yb = y.astype(np.float64).reshape(-1, upp)
    blk = np.stack([yb.sum(1), np.abs(yb).sum(1)], 1)
    blk_err = float(np.abs(blk - d["blocks"]).max()) / upp  # mean error per sample of the worst 10 ms block
    assert worst <= tol, worst
    assert blk_err <= tol, blk_err
    return worst, blk_err


_PARITY = []


def record_parity(entry: dict) -> None:
    """Fast-mode / fp32 margins measured by the GPU tests; tests/conftest.py writes them to
    gpurun_out/parity_r2.json at session end (copied to profiles/ by hand)."""
    _PARITY.append(entry)
