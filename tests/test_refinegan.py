"""the sibling decoder RefineGANGenerator (refinegan.py:268-451, vocoder "RefineGAN"): oracle restatement
and module against the golden generated from the unmodified reference (oracle/make_golden_r2.py --refinegan)."""
import os

import numpy as np
import pytest
import torch

from oracle import nsf_hifigan_oracle as O
from oracle import refinegan_oracle as RO
from helpers import GOLDEN, record_parity

KW = dict(sample_rate=48000, downsample_rates=[2, 2, 10, 12], upsample_rates=[12, 10, 2, 2], start_channels=16, num_mels=192)
RATES = [12, 10, 2, 2]


def _draws(B, T, seed):
    """The reference's draws in execution order from a CPU generator (refinegan.py:226, :257, :98-99)."""
    gen = torch.Generator().manual_seed(seed)
    torch.rand(B, 1, generator=gen)
    src = torch.randn(B, T * 480, 1, generator=gen).reshape(B, -1)
    adain, L, ch = [], T, 512
    for r in RATES:
        L *= r
        ch //= 2
        adain.append([[torch.randn(B, ch, L, generator=gen), torch.randn(B, ch, L, generator=gen)] for _ in range(3)])
    return src, adain


def _case():
    import rvc_b200

    d = np.load(os.path.join(GOLDEN, "refinegan_48k_B2_T16.npz"))
    torch.manual_seed(int(d["init_seed"]))
    m = rvc_b200.RefineGANGenerator(**KW).eval()
    with torch.no_grad():
        for n_, p_ in m.named_parameters():
            if "blocks." in n_ and p_.dim() == 1 and (n_.endswith(".0.weight") or n_.endswith(".2.weight")):
                p_.fill_(float(d["adain_weight"]))
    t = {k: torch.from_numpy(d[k]) for k in ("x", "f0", "g", "y")}
    src, adain = _draws(2, 16, int(d["noise_seed"]))
    got = np.array([float(src.double().sum()), float(adain[3][2][1].double().abs().sum())])
    np.testing.assert_allclose(got, d["noise_ck"], rtol=1e-12, err_msg="re-created draws differ from the fixture's")
    return d, m, t, src, adain


def test_refinegan_oracle_and_module_init_match_the_reference_golden():
    d, m, t, src, adain = _case()
    sd = m.state_dict()
    cks = [k for k in d.files if k.startswith("ck::")]
    assert sorted(k[4:] for k in cks) == sorted(sd.keys()) and len(cks) == 271
    for k in cks:
        v = sd[k[4:]].double()
        np.testing.assert_allclose(np.array([float(v.sum()), float(v.abs().sum())]), d[k], rtol=1e-12, atol=1e-12, err_msg=k)
    y = RO.forward(O.folded_weights(sd), RATES, 48000, t["x"], t["f0"], t["g"], src, adain)
    assert float((y - t["y"]).abs().max()) <= 2e-6
    with pytest.raises(RuntimeError):
        m(t["x"], t["f0"], t["g"])
    assert m.noise_numel(2, 16) == src.numel() + sum(a.numel() for st in adain for blk in st for a in blk)


def test_resample_filter_and_decimation_match_torchaudio():
    torchaudio = pytest.importorskip("torchaudio")
    from rvc_b200.refinegan import sinc_kaiser_taps

    x = torch.randn(2, 16, 960, generator=torch.Generator().manual_seed(0))
    for rate in (2, 10, 12):
        k, width = RO.sinc_kaiser_kernel(rate)
        assert torch.equal(k, sinc_kaiser_taps(rate)) and k.numel() == 2 * width + rate
        ref = torchaudio.functional.resample(x, orig_freq=16 * rate, new_freq=16, lowpass_filter_width=64, rolloff=RO.ROLLOFF,
                                             resampling_method="sinc_interp_kaiser", beta=RO.BETA)
        got = RO.resample_down(x, rate)
        assert got.shape == ref.shape and float((got - ref).abs().max()) <= 1e-6


@pytest.mark.gpu
def test_refinegan_decoder_matches_the_reference_golden_on_the_gpu():
    d, m, t, src, adain = _case()
    dev = torch.device("cuda", 0)
    m = m.to(dev)
    flat = torch.cat([src.reshape(-1)] + [a.reshape(-1) for st in adain for blk in st for a in blk]).to(dev)
    a = [t[k].to(dev) for k in ("x", "f0", "g")]
    y = m(*a, noise=flat).cpu()
    err = float((y - t["y"]).abs().max())
    record_parity({"test": "refinegan_sibling_48k_B2_T16", "fp32_max_abs": err, "ref_rms": float(t["y"].pow(2).mean().sqrt()),
                   "launches": m.last_launch_count})
    assert y.shape == t["y"].shape == (2, 1, 16 * 480)
    assert err <= 2e-3 and err <= 1e-4, err
    # default noise = device-generator draws in the reference's order; in-kernel noise runs and is reproducible per seed
    torch.manual_seed(5)
    y1 = m(*a).cpu()
    torch.manual_seed(5)
    y2 = m(*a, noise=m.draw_noise(2, 16, dev)).cpu()
    assert torch.equal(y1, y2)
    torch.manual_seed(6)
    p1 = m(*a, noise="philox").cpu()
    torch.manual_seed(6)
    p2 = m(*a, noise="philox").cpu()
    assert torch.equal(p1, p2) and bool(torch.isfinite(p1).all()) and float(p1.abs().max()) <= 1.0
    # a longer, single item against the oracle (several tiles per stage, g = None)
    x, f0, g, _ = O.synth_inputs(O.CONFIG_48K, 1, 70, seed=73)
    srcL, adL = _draws_T(1, 70, 74)
    W = O.folded_weights({k: v.cpu() for k, v in m.state_dict().items()})
    yo = RO.forward(W, RATES, 48000, x, f0, None, srcL, adL)
    flatL = torch.cat([srcL.reshape(-1)] + [a_.reshape(-1) for st in adL for blk in st for a_ in blk]).to(dev)
    yl = m(x.to(dev), f0.to(dev), None, noise=flatL).cpu()
    assert float((yl - yo).abs().max()) <= 2e-3


def _draws_T(B, T, seed):
    gen = torch.Generator().manual_seed(seed)
    src = torch.randn(B, T * 480, generator=gen)
    adain, L, ch = [], T, 512
    for r in RATES:
        L *= r
        ch //= 2
        adain.append([[torch.randn(B, ch, L, generator=gen), torch.randn(B, ch, L, generator=gen)] for _ in range(3)])
    return src, adain


@pytest.mark.gpu
def test_synthesizer_with_the_refinegan_vocoder_against_the_oracle():
    """net_g.infer with vocoder="RefineGAN" (synthesizers.py:98-106): front end in the library (front-end-only rvcsyn handle),
    z * x_mask into the RefineGAN decoder; oracle = synth_oracle front + refinegan_oracle.forward."""
    import rvc_b200
    from oracle import synth_oracle as SO
    from test_synth import CFGD, SYN_48K

    dev = torch.device("cuda", 0)
    torch.manual_seed(41)
    m = rvc_b200.Synthesizer(*SYN_48K, use_f0=1, text_enc_hidden_dim=768, vocoder="RefineGAN").eval()
    assert isinstance(m.dec, rvc_b200.RefineGANGenerator)
    with torch.no_grad():
        for i in range(4):
            m.flow.flows[2 * i].post.weight.normal_(0, 0.05)
    sd = m.state_dict()
    B, T = 1, 24
    gen = torch.Generator().manual_seed(8)
    phone = torch.randn(B, T, 768, generator=gen)
    pitch = torch.randint(1, 256, (B, T), generator=gen)
    nsff0 = O.synth_f0(B, T, gen)
    sid, lens = torch.tensor([4]), torch.tensor([24])
    nzp = torch.randn(B, 192, T, generator=gen)
    src, adain = _draws_T(B, T, 75)
    flat = torch.cat([src.reshape(-1)] + [a.reshape(-1) for st in adain for blk in st for a in blk])
    m = m.to(dev)
    o, x_mask, (z, *_ ) = m.infer(phone.to(dev), lens.to(dev), pitch.to(dev), nsff0.to(dev), sid.to(dev), noise_zp=nzp.to(dev),
                                  noise_src=flat.to(dev))
    W = SO.folded({k: v for k, v in sd.items() if not k.startswith(("enc_q.", "dec."))})
    zm, g, mask, *_ = SO.synthesizer_front(W, CFGD, phone, lens, pitch, sid, nzp)
    Wd = O.folded_weights({k[4:]: v for k, v in sd.items() if k.startswith("dec.")})
    y = RO.forward(Wd, RATES, 48000, zm, nsff0, g, src, adain)
    assert float((z.cpu() - zm).abs().max()) <= 2e-4
    assert float((o.cpu() - y).abs().max()) <= 2e-3
