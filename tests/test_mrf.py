"""the sibling decoder HiFiGANMRFGenerator ("MRF HiFi-GAN" vocoder, hifigan_mrf.py:223-374): oracle
restatement and module against the golden generated from the unmodified reference (oracle/make_golden_r2.py --mrf)."""
import os

import numpy as np
import pytest
import torch

from oracle import nsf_hifigan_oracle as O
from helpers import GOLDEN, record_parity

KW = dict(in_channel=192, upsample_initial_channel=512, upsample_rates=[12, 10, 2, 2], upsample_kernel_sizes=[24, 20, 4, 4],
          resblock_kernel_sizes=[3, 7, 11], resblock_dilations=[[1, 3, 5]] * 3, gin_channels=256, sample_rate=48000, harmonic_num=8)


def _case():
    d = np.load(os.path.join(GOLDEN, "mrf_48k_B2_T24.npz"))
    import rvc_b200

    torch.manual_seed(int(d["init_seed"]))
    m = rvc_b200.HiFiGANMRFGenerator(**KW).eval()
    t = {k: torch.from_numpy(d[k].astype(np.float32)) for k in ("x", "f0", "g", "rand_ini", "noise", "y")}
    return d, m, t


def test_mrf_oracle_and_module_init_match_the_reference_golden():
    d, m, t = _case()
    sd = m.state_dict()
    cks = [k for k in d.files if k.startswith("ck::")]
    assert sorted(k[4:] for k in cks) == sorted(sd.keys()) and len(cks) == 246
    for k in cks:
        v = sd[k[4:]].double()
        np.testing.assert_allclose(np.array([float(v.sum()), float(v.abs().sum())]), d[k], rtol=1e-12, atol=1e-12, err_msg=k)
    w = O.folded_weights(O.mrf_state_to_nsf_names(sd, 3))
    assert w["m_source.l_linear.weight"].shape == (1, 9) and w["conv_post.bias"].shape == (1,)
    y = O.decoder_forward_mrf(w, O.CONFIG_48K, t["x"], t["f0"], t["g"], t["rand_ini"], t["noise"])
    assert float((y - t["y"]).abs().max()) <= 2e-6
    # the module hands the library the same tensors under the NSF names
    st = m._decoder_state()
    assert set(O.folded_weights(st).keys()) == set(w.keys())
    with pytest.raises(RuntimeError):
        m(t["x"], t["f0"], t["g"])


def test_mrf_source_closed_form_equals_the_per_sample_cumsum():
    """hifigan_mrf.py:124-142 takes a per-sample cumsum of (f0 (k+1) / sr) % 1 with wrap corrections; inside a frame f0 is
    constant, so the phase is a frame-level prefix plus a linear ramp.  Checked here in fp64 against a literal restatement."""
    gen = torch.Generator().manual_seed(1)
    B, T, upp, sr, H = 2, 37, 480, 48000, 9
    f0 = O.synth_f0(B, T, gen)
    ini = torch.rand(B, H, generator=gen, dtype=torch.float64)
    ini[:, 0] = 0
    f0s = f0.double().repeat_interleave(upp, dim=1)                                   # nearest upsample
    rad = (f0s.unsqueeze(-1) * torch.arange(1, H + 1, dtype=torch.float64) / sr) % 1
    rad[:, 0, :] += ini
    lit = torch.sin(torch.cumsum(rad, 1) * 2 * np.pi) * 0.1
    uv = (f0s > 0).double().unsqueeze(-1)
    got = O.mrf_sine_source(f0, upp, sr, ini, torch.zeros(B, T * upp, H), dtype=torch.float64)
    assert float((got.double() - lit * uv).abs().max()) <= 1e-6


@pytest.mark.gpu
def test_mrf_decoder_matches_the_reference_golden_on_the_gpu():
    d, m, t = _case()
    dev = torch.device("cuda", 0)
    m = m.to(dev)
    a = [t[k].to(dev) for k in ("x", "f0", "g")]
    with torch.no_grad():
        y32 = m(*a, noise=t["noise"].to(dev), rand_ini=t["rand_ini"].to(dev), mode="fp32").cpu()
        yf = m(*a, noise=t["noise"].to(dev), rand_ini=t["rand_ini"].to(dev), mode="fast").cpu()
        torch.manual_seed(3)
        yd1 = m(*a, mode="fast").cpu()        # default: device-generator draws in the reference's order
        torch.manual_seed(3)
        ri = torch.rand(2, 9, device=dev)
        nz = torch.randn(2, 24 * 480, 9, device=dev)
        yd2 = m(*a, noise=nz, rand_ini=ri, mode="fast").cpu()
        torch.manual_seed(4)
        yp = m(*a, noise="philox", mode="fast").cpu()
        torch.manual_seed(4)
        yp2 = m(*a, noise="philox", mode="fast").cpu()
    e32 = float((y32 - t["y"]).abs().max())
    corr = O.multires_logmel_corr(yf, t["y"], 48000)
    record_parity({"test": "mrf_sibling_48k_B2_T24", "fp32_max_abs": e32, "fast_max_abs": float((yf - t["y"]).abs().max()), "fast_logmel": corr})
    assert y32.shape == t["y"].shape == (2, 1, 24 * 480)
    assert e32 <= 2e-3 and e32 <= 5e-5, e32
    assert corr["min"] >= 0.999, corr
    assert torch.equal(yd1, yd2)
    assert torch.equal(yp, yp2) and bool(torch.isfinite(yp).all()) and float(yp.abs().max()) <= 1.0 and not torch.equal(yp, yd1)


@pytest.mark.gpu
def test_synthesizer_with_the_mrf_vocoder_against_the_oracle():
    """net_g.infer with vocoder="MRF HiFi-GAN" (synthesizers.py:85-97): rvc_b200.Synthesizer builds the MRF decoder and the
    library runs front end + decoder in one call; oracle = synth_oracle front + decoder_forward_mrf."""
    import rvc_b200
    from oracle import synth_oracle as SO
    from test_synth import CFGD, SYN_48K

    dev = torch.device("cuda", 0)
    torch.manual_seed(31)
    m = rvc_b200.Synthesizer(*SYN_48K, use_f0=1, text_enc_hidden_dim=768, vocoder="MRF HiFi-GAN").eval()
    assert isinstance(m.dec, rvc_b200.HiFiGANMRFGenerator)
    with torch.no_grad():
        for i in range(4):
            m.flow.flows[2 * i].post.weight.normal_(0, 0.05)
    sd = m.state_dict()
    B, T = 2, 50
    gen = torch.Generator().manual_seed(6)
    phone = torch.randn(B, T, 768, generator=gen)
    pitch = torch.randint(1, 256, (B, T), generator=gen)
    nsff0 = O.synth_f0(B, T, gen)
    sid, lens = torch.tensor([1, 2]), torch.tensor([50, 33])
    nzp = torch.randn(B, 192, T, generator=gen)
    ini = torch.rand(B, 9, generator=gen)
    ini[:, 0] = 0
    nsrc = torch.randn(B, T * 480, 9, generator=gen)
    m = m.to(dev)
    o, x_mask, (z, *_ ) = m.infer(phone.to(dev), lens.to(dev), pitch.to(dev), nsff0.to(dev), sid.to(dev), noise_zp=nzp.to(dev),
                                  noise_src=nsrc.to(dev), rand_ini=ini.to(dev), mode="fp32")
    W = SO.folded({k: v for k, v in sd.items() if not k.startswith("enc_q.")})
    zm, g, mask, *_ = SO.synthesizer_front(W, CFGD, phone, lens, pitch, sid, nzp)
    Wd = O.folded_weights(O.mrf_state_to_nsf_names({k[4:]: v for k, v in sd.items() if k.startswith("dec.")}, 3))
    y = O.decoder_forward_mrf(Wd, O.CONFIG_48K, zm, nsff0, g, ini, nsrc)
    assert float((z.cpu() - zm).abs().max()) <= 2e-4
    assert float((o.cpu() - y).abs().max()) <= 2e-3
