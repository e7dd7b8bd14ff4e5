"""Generate tests/golden/*.npz by running the UNMODIFIED reference decoder on CPU.

TEST INFRASTRUCTURE.  Run in the build container only (it imports
/root/reference, which does not exist on the GPU box):

    python oracle/make_golden.py

Each fixture holds the inputs, the injected noise and the waveform the
reference's ``HiFiGANNSFGenerator.forward`` (hifigan_nsf.py:173-212) produced,
in fp32 on CPU.  The tiny configuration also stores its full state_dict so the
fixture does not depend on any RNG; the full-size configurations store the
init seed plus per-tensor checksums of the state_dict instead (63 MB of weights
do not belong in git) - the test re-creates them with the same seed and checks
the checksums first.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from oracle import nsf_hifigan_oracle as O  # noqa: E402
from rvc.lib.algorithm.generators.hifigan_nsf import HiFiGANNSFGenerator  # noqa: E402
from rvc.lib.algorithm.generators.hifigan import SineGenerator  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def build_reference(cfg: O.DecoderConfig, seed: int) -> HiFiGANNSFGenerator:
    torch.manual_seed(seed)
    m = HiFiGANNSFGenerator(*cfg.ctor_args(), **cfg.ctor_kwargs())
    return m.eval()


def run_reference(m, x, f0, g, noise_seed):
    torch.manual_seed(noise_seed)
    with torch.no_grad():
        return m(x, f0, g=g)


def checksums(state):
    return {k: np.array([float(v.double().sum()), float(v.double().abs().sum())]) for k, v in state.items()}


def make_case(name, cfg, B, T, init_seed, in_seed, noise_seed, store_weights):
    m = build_reference(cfg, init_seed)
    x, f0, g, _ = O.synth_inputs(cfg, B, T, seed=in_seed)
    y = run_reference(m, x, f0, g, noise_seed)
    noise = O.draw_reference_noise(noise_seed, B, T * cfg.upp)
    # sanity: the oracle with that noise reproduces the reference
    w = O.folded_weights(m.state_dict())
    y_or = O.decoder_forward(w, cfg, x, f0, g, noise)
    err = float((y - y_or).abs().max())
    print(f"{name}: out {tuple(y.shape)} rms {float(y.pow(2).mean().sqrt()):.4f} oracle-vs-reference max-abs {err:.3e}")
    assert err < 5e-6, err
    data = dict(
        x=x.numpy(), f0=f0.numpy(), g=g.numpy(), noise=noise.numpy(), y=y.numpy(),
        init_seed=np.int64(init_seed), noise_seed=np.int64(noise_seed),
        cfg_initial_channel=np.int64(cfg.initial_channel),
        cfg_resblock_kernel_sizes=np.array(cfg.resblock_kernel_sizes),
        cfg_resblock_dilation_sizes=np.array(cfg.resblock_dilation_sizes),
        cfg_upsample_rates=np.array(cfg.upsample_rates),
        cfg_upsample_initial_channel=np.int64(cfg.upsample_initial_channel),
        cfg_upsample_kernel_sizes=np.array(cfg.upsample_kernel_sizes),
        cfg_gin_channels=np.int64(cfg.gin_channels), cfg_sr=np.int64(cfg.sr),
    )
    if store_weights:
        for k, v in m.state_dict().items():
            data["sd::" + k] = v.numpy()
    else:
        for k, v in checksums(m.state_dict()).items():
            data["ck::" + k] = v
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **data)


def make_nof0_case(name, cfg, B, T, init_seed, in_seed):
    """The sibling decoder of the no-f0 models: HiFiGANGenerator.forward (hifigan.py:72-97), deterministic (no noise)."""
    from rvc.lib.algorithm.generators.hifigan import HiFiGANGenerator

    torch.manual_seed(init_seed)
    m = HiFiGANGenerator(cfg.initial_channel, list(cfg.resblock_kernel_sizes), [list(d) for d in cfg.resblock_dilation_sizes],
                         list(cfg.upsample_rates), cfg.upsample_initial_channel, list(cfg.upsample_kernel_sizes),
                         gin_channels=cfg.gin_channels).eval()
    x, _, g, _ = O.synth_inputs(cfg, B, T, seed=in_seed)
    with torch.no_grad():
        y = m(x, g=g)
    y_or = O.generator_forward_nof0(O.folded_weights(m.state_dict()), cfg, x, g)
    err = float((y - y_or).abs().max())
    print(f"{name}: out {tuple(y.shape)} rms {float(y.pow(2).mean().sqrt()):.4f} oracle-vs-reference max-abs {err:.3e}")
    assert err < 5e-6, err
    data = dict(
        x=x.numpy(), g=g.numpy(), y=y.numpy(), init_seed=np.int64(init_seed),
        cfg_initial_channel=np.int64(cfg.initial_channel),
        cfg_resblock_kernel_sizes=np.array(cfg.resblock_kernel_sizes),
        cfg_resblock_dilation_sizes=np.array(cfg.resblock_dilation_sizes),
        cfg_upsample_rates=np.array(cfg.upsample_rates),
        cfg_upsample_initial_channel=np.int64(cfg.upsample_initial_channel),
        cfg_upsample_kernel_sizes=np.array(cfg.upsample_kernel_sizes),
        cfg_gin_channels=np.int64(cfg.gin_channels), cfg_sr=np.int64(cfg.sr),
    )
    for k, v in checksums(m.state_dict()).items():
        data["ck::" + k] = v
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **data)


def make_sine_case():
    """SineGenerator alone (hifigan.py:197-228), including a long contour and the real
    RMVPE contour shipped with the reference (ios_test_data/rmvpe_f0.npy)."""
    gen = torch.Generator().manual_seed(7)
    f0 = O.synth_f0(2, 250, gen)
    real = np.load("/root/reference/ios_test_data/rmvpe_f0.npy").astype(np.float32)
    sg = SineGenerator(48000)
    torch.manual_seed(11)
    s, uv, _ = sg(f0, 480)
    noise = O.draw_reference_noise(11, 2, 250 * 480)
    torch.manual_seed(12)
    s_real, _, _ = sg(torch.from_numpy(real)[None, :200], 480)
    noise_real = O.draw_reference_noise(12, 1, 200 * 480)
    np.savez_compressed(
        os.path.join(OUT, "sine_48k.npz"),
        f0=f0.numpy(), noise=noise.numpy().astype(np.float16), sine=s[..., 0].numpy().astype(np.float32),
        f0_real=real[:200], noise_real=noise_real.numpy().astype(np.float16), sine_real=s_real[..., 0].numpy(),
    )
    # the fixture stores noise as fp16 to stay small: reproduce the expected output from the rounded noise
    d = np.load(os.path.join(OUT, "sine_48k.npz"))
    for f0k, nk, sk in (("f0", "noise", "sine"), ("f0_real", "noise_real", "sine_real")):
        f = torch.from_numpy(d[f0k]).reshape(-1, d[f0k].shape[-1])
        n_exact = noise if nk == "noise" else noise_real
        uvm = (f > 0).float().repeat_interleave(480, dim=1)
        amp = uvm * 0.003 + (1 - uvm) * (0.1 / 3)
        # expected with rounded noise = reference output - amp*(exact - rounded)
        corr = amp * (n_exact - torch.from_numpy(d[nk].astype(np.float32)))
        exp = torch.from_numpy(d[sk]).reshape(f.shape[0], -1) - corr
        got = O.sine_source(f, 480, 48000, torch.from_numpy(d[nk].astype(np.float32)))
        print(f"sine {f0k}: oracle-vs-reference max-abs {float((got - exp).abs().max()):.3e}")
    full = dict(d)
    # overwrite the stored expectation with the rounded-noise one so tests compare like with like
    for f0k, nk, sk in (("f0", "noise", "sine"), ("f0_real", "noise_real", "sine_real")):
        f = torch.from_numpy(d[f0k]).reshape(-1, d[f0k].shape[-1])
        n_exact = noise if nk == "noise" else noise_real
        uvm = (f > 0).float().repeat_interleave(480, dim=1)
        amp = uvm * 0.003 + (1 - uvm) * (0.1 / 3)
        corr = amp * (n_exact - torch.from_numpy(d[nk].astype(np.float32)))
        full[sk] = (torch.from_numpy(d[sk]).reshape(f.shape[0], -1) - corr).numpy()
    np.savez_compressed(os.path.join(OUT, "sine_48k.npz"), **full)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    make_case("tiny_24k_B2_T40", O.CONFIG_TINY, 2, 40, init_seed=1234, in_seed=5, noise_seed=6, store_weights=True)
    make_case("v2_48k_B2_T24", O.CONFIG_48K, 2, 24, init_seed=1234, in_seed=21, noise_seed=22, store_weights=False)
    make_case("v2_40k_B1_T32", O.CONFIG_40K, 1, 32, init_seed=1234, in_seed=31, noise_seed=32, store_weights=False)
    make_case("v2_32k_B1_T16", O.CONFIG_32K, 1, 16, init_seed=1234, in_seed=41, noise_seed=42, store_weights=False)
    make_sine_case()
    make_nof0_case("nof0_40k_B2_T20", O.CONFIG_40K, 2, 20, init_seed=77, in_seed=78)
