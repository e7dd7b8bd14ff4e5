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
