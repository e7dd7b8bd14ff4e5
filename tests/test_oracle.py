"""The oracle against the golden vectors produced by the reference (oracle/make_golden.py), and
against the live reference module when /root/reference is present (build container only)."""
import os
import sys

import numpy as np
import pytest
import torch

from oracle import nsf_hifigan_oracle as O
from helpers import golden_state, load_golden, tensors

CASES = ["tiny_24k_B2_T40", "v2_48k_B2_T24", "v2_40k_B1_T32", "v2_32k_B1_T16", "odd_rates_24k_B2_T30"]


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_golden(name):
    d, cfg = load_golden(name)
    x, f0, g, noise, y = tensors(d)
    w = O.folded_weights(golden_state(d, cfg))
    got = O.decoder_forward(w, cfg, x, f0, g, noise)
    assert got.shape == y.shape == (x.shape[0], 1, x.shape[2] * cfg.upp)
    assert float((got - y).abs().max()) <= 2e-6  # fp32 reference vs fp32 oracle: summation-order noise only


def test_oracle_fp64_close_to_fp32():
    d, cfg = load_golden("tiny_24k_B2_T40")
    x, f0, g, noise, y = tensors(d)
    w = O.folded_weights(golden_state(d, cfg))
    got = O.decoder_forward(w, cfg, x, f0, g, noise, dtype=torch.float64, source_dtype=torch.float64)
    assert float((got - y).abs().max()) <= 2e-5


def test_sine_source_matches_golden():
    d = np.load(os.path.join(os.path.dirname(__file__), "golden", "sine_48k.npz"))
    for f0k, nk, sk in (("f0", "noise", "sine"), ("f0_real", "noise_real", "sine_real")):
        f0 = torch.from_numpy(d[f0k]).reshape(-1, d[f0k].shape[-1])
        noise = torch.from_numpy(d[nk].astype(np.float32))
        exp = torch.from_numpy(d[sk]).reshape(f0.shape[0], -1)
        got = O.sine_source(f0, 480, 48000, noise)
        assert float((got - exp).abs().max()) <= 1e-6
        # the fp64 phase accumulation (what the CUDA kernel does) stays within the fp32 scan's own error
        got64 = O.sine_source(f0, 480, 48000, noise, dtype=torch.float64)
        assert float((got64 - exp).abs().max()) <= 5e-5


def test_weight_norm_fold_both_namings():
    d, cfg = load_golden("tiny_24k_B2_T40")
    sd = golden_state(d, cfg)
    legacy = {}
    for k, v in sd.items():
        legacy[k.replace(".parametrizations.weight.original0", ".weight_g").replace(".parametrizations.weight.original1", ".weight_v")] = v
    a, b = O.folded_weights(sd), O.folded_weights(legacy)
    assert a.keys() == b.keys()
    for k in a:
        assert torch.equal(a[k], b[k])
    v, g_ = sd["ups.0.parametrizations.weight.original1"], sd["ups.0.parametrizations.weight.original0"]
    assert torch.allclose(a["ups.0.weight"].reshape(v.shape[0], -1).norm(dim=1), g_.flatten(), rtol=1e-5)


def test_property_shapes_and_bounds():
    """Properties the reference's own unit tests check (tests/unit/test_generators.py:102-190)."""
    d, cfg = load_golden("tiny_24k_B2_T40")
    x, f0, g, noise, y = tensors(d)
    w = O.folded_weights(golden_state(d, cfg))
    assert cfg.upp == int(np.prod(cfg.upsample_rates))
    y_nog = O.decoder_forward(w, cfg, x, f0, None, noise)
    assert y_nog.shape == y.shape
    assert float(y.abs().max()) <= 1.0 and float(y_nog.abs().max()) <= 1.0
    # voiced mask semantics: unvoiced frames carry only the (larger) noise floor
    f = torch.tensor([[0.0, 220.0]])
    s = O.sine_source(f, cfg.upp, cfg.sr, torch.zeros(1, 2 * cfg.upp))
    assert float(s[0, : cfg.upp].abs().max()) == 0.0 and float(s[0, cfg.upp :].abs().max()) > 0.05


@pytest.mark.skipif(not os.path.isdir("/root/reference/rvc"), reason="reference checkout not present (GPU box)")
def test_oracle_matches_live_reference():
    sys.path.insert(0, "/root/reference")
    from rvc.lib.algorithm.generators.hifigan_nsf import HiFiGANNSFGenerator as Ref

    cfg = O.CONFIG_40K
    torch.manual_seed(99)
    ref = Ref(*cfg.ctor_args(), **cfg.ctor_kwargs()).eval()
    x, f0, g, _ = O.synth_inputs(cfg, 1, 20, seed=3)
    torch.manual_seed(5)
    with torch.no_grad():
        y = ref(x, f0, g=g)
    noise = O.draw_reference_noise(5, 1, 20 * cfg.upp)
    got = O.decoder_forward(O.folded_weights(ref.state_dict()), cfg, x, f0, g, noise)
    assert float((got - y).abs().max()) <= 2e-6


def test_nof0_sibling_oracle_and_module_init_match_reference_golden():
    """HiFiGANGenerator (hifigan.py:11-112, the decoder of the no-f0 models): the oracle restatement reproduces the
    reference's output on the golden inputs, and rvc_b200.HiFiGANGenerator default-initialised under the same seed has
    the reference's state_dict (names, shapes, per-tensor checksums)."""
    import numpy as np
    from helpers import load_golden
    from rvc_b200 import HiFiGANGenerator

    d, cfg = load_golden("nof0_40k_B2_T20")
    torch.manual_seed(int(d["init_seed"]))
    m = HiFiGANGenerator(cfg.initial_channel, list(cfg.resblock_kernel_sizes), [list(r) for r in cfg.resblock_dilation_sizes],
                         list(cfg.upsample_rates), cfg.upsample_initial_channel, list(cfg.upsample_kernel_sizes),
                         gin_channels=cfg.gin_channels).eval()
    sd = m.state_dict()
    assert sorted(sd.keys()) == sorted(k[4:] for k in d.files if k.startswith("ck::"))
    for k in d.files:
        if k.startswith("ck::"):
            v = sd[k[4:]].double()
            np.testing.assert_allclose(np.array([float(v.sum()), float(v.abs().sum())]), d[k], rtol=1e-12, atol=1e-12, err_msg=k)
    y = O.generator_forward_nof0(O.folded_weights(sd), cfg, torch.from_numpy(d["x"]), torch.from_numpy(d["g"]))
    assert float((y - torch.from_numpy(d["y"])).abs().max()) <= 5e-6
    # the tensors handed to the C ABI add all-zero source / noise-conv layers under the NSF names
    st = m._decoder_state()
    assert st["noise_convs.0.weight"].shape == (256, 1, 80) and st["noise_convs.3.weight"].shape == (32, 1, 1)
    assert all(float(st[k].abs().sum()) == 0.0 for k in st if k.startswith(("noise_convs.", "m_source.")))
    with pytest.raises(RuntimeError):
        m(torch.from_numpy(d["x"]), torch.from_numpy(d["g"]))  # CPU tensors: no fallback


def test_philox_oracle_known_answers_and_moments():
    """The in-kernel noise generator's oracle (oracle/philox_oracle.py) reproduces the Random123 known-answer vectors
    of Philox4x32-10, and its Box-Muller normals have the moments of N(0,1)."""
    from oracle import philox_oracle as P

    for ctr, key, out in P.KAT:
        got = P.philox4x32_10(*ctr, *key)
        assert tuple(int(v) for v in got) == out
    z = P.source_noise(0x1234567890ABCDEF, 3, 2, 500_000)
    assert z.shape == (2, 500_000)
    f = z.ravel()
    assert abs(f.mean()) < 4e-3 and abs(f.var() - 1) < 6e-3
    assert abs(((f - f.mean()) ** 4).mean() / f.var() ** 2 - 3) < 0.03
    assert abs(np.corrcoef(f[:-1], f[1:])[0, 1]) < 4e-3
    assert abs(np.corrcoef(z[0], z[1])[0, 1]) < 6e-3
    # draws are a function of (seed, global item, sample): item 4 of a call == item 1 of a call that starts at item 3
    assert np.array_equal(P.source_noise(9, 0, 5, 1000)[4], P.source_noise(9, 3, 2, 1000)[1])


def test_oracle_on_the_reference_real_contour_fixture():
    """BASELINE configs[2] input: the reference's real RMVPE contour, 13.5 s.  The oracle reproduces the stored pieces
    of the reference's own waveform (segments + per-frame block sums)."""
    from helpers import check_against_real_fixture, load_real_case

    d, cfg, x, f0, g, noise = load_real_case()
    assert f0.shape == (1, 1350) and float(f0.max()) > 160 and 0.4 < float((f0 > 0).float().mean()) < 0.6
    m_sd = golden_state(d, cfg)
    y = O.decoder_forward(O.folded_weights(m_sd), cfg, x, f0, g, noise)
    worst, blk = check_against_real_fixture(d, y[0, 0].numpy(), 5e-6)
    assert abs(float(y.pow(2).mean().sqrt()) - float(d["y_rms"])) < 1e-6
