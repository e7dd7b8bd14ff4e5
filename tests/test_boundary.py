"""The drop-in boundary proven INSIDE the reference: the reference's own ``Synthesizer`` (synthesizers.py:108-118)
constructs ``rvc_b200.HiFiGANNSFGenerator`` in place of its decoder, loads an ``extract_model``-shaped checkpoint
(fp16 weights, legacy weight-norm names, no enc_q; extract_model.py:57-109) the way ``infer.py:461-488`` does, and the
checkpoint loader builds the same decoder from the same dict.  The reference comes from /root/reference (build
container) or the staged oracle/_ref copy (GPU box); the tests skip when neither exists."""
import sys

import numpy as np
import pytest
import torch

from oracle import nsf_hifigan_oracle as O
from oracle.stage_ref import ref_path

RP = ref_path()
needs_ref = pytest.mark.skipif(RP is None, reason="reference not importable (no /root/reference and no oracle/_ref)")

V2_48K_CONFIG = [1025, 32, 192, 192, 768, 2, 6, 3, 0, "1", [3, 7, 11], [[1, 3, 5], [1, 3, 5], [1, 3, 5]], [12, 10, 2, 2], 512,
                 [24, 20, 4, 4], 109, 256, 48000]  # rvc/configs/48000.json in Synthesizer.__init__ positional order


def _ref_modules():
    if RP not in sys.path:
        sys.path.insert(0, RP)
    import rvc.lib.algorithm.synthesizers as S

    return S


def _extract_model_dict(net_g, use_f0=1, vocoder="HiFi-GAN"):
    """What extract_model.py:57-109 writes: fp16 weights without enc_q, legacy weight-norm key names."""
    w = {k: v.half() for k, v in net_g.state_dict().items() if "enc_q" not in k}
    ren = {}
    for k, v in w.items():
        ren[k.replace(".parametrizations.weight.original1", ".weight_v").replace(".parametrizations.weight.original0", ".weight_g")] = v
    return {"weight": ren, "config": list(V2_48K_CONFIG), "f0": use_f0, "version": "v2", "vocoder": vocoder, "sr": 48000}


@pytest.fixture(scope="module")
def checkpoint():
    S = _ref_modules()
    torch.manual_seed(4321)
    ref = S.Synthesizer(*V2_48K_CONFIG, use_f0=1, text_enc_hidden_dim=768, vocoder="HiFi-GAN").eval()
    return ref, _extract_model_dict(ref)


@needs_ref
def test_reference_synthesizer_builds_and_loads_the_drop_in(checkpoint, monkeypatch):
    import rvc_b200

    S = _ref_modules()
    ref, cpt = checkpoint
    monkeypatch.setattr(S, "HiFiGANNSFGenerator", rvc_b200.HiFiGANNSFGenerator)
    cfg = list(cpt["config"])
    cfg[-3] = cpt["weight"]["emb_g.weight"].shape[0]  # infer.py:472
    net_g = S.Synthesizer(*cfg, use_f0=cpt.get("f0", 1), text_enc_hidden_dim=768, vocoder=cpt.get("vocoder", "HiFi-GAN"))
    assert isinstance(net_g.dec, rvc_b200.HiFiGANNSFGenerator)
    del net_g.enc_q                                                   # infer.py:484
    res = net_g.load_state_dict(cpt["weight"], strict=False)          # infer.py:485
    assert not [k for k in res.missing_keys if k.startswith("dec.")], res.missing_keys[:5]
    assert not [k for k in res.unexpected_keys if k.startswith("dec.")], res.unexpected_keys[:5]
    assert not res.missing_keys and not res.unexpected_keys
    net_g = net_g.float().eval()                                      # infer.py:486-488
    # same parameters as the reference decoder, after the checkpoint's fp16 rounding
    want = {k: v.half().float() for k, v in ref.dec.state_dict().items()}
    got = net_g.dec.state_dict()
    assert list(got.keys()) == list(want.keys())
    for k in want:
        assert torch.equal(got[k], want[k]), k
    # protocol the callers rely on
    assert net_g.dec.upp == 480 and net_g.dec.num_kernels == 3 and net_g.dec.num_upsamples == 4
    named = dict(net_g.dec.named_modules())
    assert any(hasattr(m, "parametrizations") for m in named.values())  # realtime/pipeline.py:409-415 walks these
    net_g._remove_weight_norm_from(net_g.dec)  # synthesizers.py:150-153: finds no WeightNorm hooks, does nothing, must not raise
    # (Synthesizer.remove_weight_norm itself touches self.enc_q, which infer.py:484 has deleted by now)
    net_g.dec.remove_weight_norm()      # the reference's raises on parametrised modules; the drop-in folds
    assert not any(hasattr(m, "parametrizations") for m in dict(net_g.dec.named_modules()).values())
    folded = O.folded_weights(want)
    sd = net_g.dec.state_dict()
    for k in ("ups.0.weight", "resblocks.5.convs1.2.weight"):
        assert torch.allclose(sd[k], folded[k], rtol=1e-6, atol=1e-7), k
    with pytest.raises(RuntimeError):   # CPU tensors: no fallback, loudly
        net_g.dec(torch.zeros(1, 192, 4), torch.zeros(1, 4), g=torch.zeros(1, 256, 1))


@needs_ref
def test_checkpoint_loader_builds_the_same_decoder_and_caches_by_content(checkpoint):
    import rvc_b200
    from rvc_b200 import loader

    ref, cpt = checkpoint
    c = loader.checkpoint_config(cpt)
    assert c["sr"] == 48000 and c["gin_channels"] == 256 and c["spk_embed_dim"] == 109 and c["text_enc_hidden_dim"] == 768
    m = rvc_b200.load_checkpoint(cpt)
    assert isinstance(m, rvc_b200.HiFiGANNSFGenerator) and not m.training
    want = {k: v.half().float() for k, v in ref.dec.state_dict().items()}
    for k, v in m.state_dict().items():
        assert torch.equal(v, want[k]), k
    # content key: same voice -> same key; one changed weight -> another key; an in-place edit drops the key
    m2 = rvc_b200.load_checkpoint(cpt)
    assert m.content_key == m2.content_key and len(m.content_key) == 64
    cpt2 = dict(cpt, weight=dict(cpt["weight"]))
    cpt2["weight"]["dec.conv_post.weight"] = cpt["weight"]["dec.conv_post.weight"] * 2
    assert rvc_b200.load_checkpoint(cpt2).content_key != m.content_key
    assert m._cache_key(0) == (m.content_key, 0)
    with torch.no_grad():
        m.conv_post.weight.mul_(1.5)
    assert m._cache_key(0) is None
    # no-f0 model -> the sibling class; other vocoders are refused loudly
    nf = dict(cpt, f0=0, weight={k: v for k, v in cpt["weight"].items() if "m_source" not in k and "noise_convs" not in k})
    assert isinstance(rvc_b200.load_checkpoint(nf), rvc_b200.HiFiGANGenerator)
    with pytest.raises(NotImplementedError):
        rvc_b200.load_checkpoint(dict(cpt, vocoder="RefineGAN", f0=0))  # no decoder in the reference either (synthesizers.py:123-126)
    with pytest.raises((KeyError, RuntimeError)):
        rvc_b200.load_checkpoint(dict(cpt, vocoder="RefineGAN"))         # NSF weights do not fit the RefineGAN decoder
    with pytest.raises(KeyError):
        rvc_b200.load_checkpoint(dict(cpt, weight={k: v for k, v in cpt["weight"].items() if k != "dec.conv_pre.bias"}))


@needs_ref
@pytest.mark.gpu
def test_net_g_infer_runs_the_drop_in_decoder_on_the_gpu(checkpoint, monkeypatch):
    """``net_g.infer`` of the reference Synthesizer with the drop-in as ``dec`` (synthesizers.py:237-241), on cuda:0:
    the decoder's inputs and its device-RNG noise are captured and the waveform is checked against the oracle."""
    import rvc_b200

    S = _ref_modules()
    ref, cpt = checkpoint
    monkeypatch.setattr(S, "HiFiGANNSFGenerator", rvc_b200.HiFiGANNSFGenerator)
    cfg = list(cpt["config"])
    net_g = S.Synthesizer(*cfg, use_f0=1, text_enc_hidden_dim=768, vocoder="HiFi-GAN")
    del net_g.enc_q
    net_g.load_state_dict(cpt["weight"], strict=False)
    dev = torch.device("cuda", 0)
    net_g = net_g.to(dev).float().eval()
    net_g.dec.mode = "fp32"
    T = 120
    gen = torch.Generator().manual_seed(5)
    phone = torch.randn(1, T, 768, generator=gen).to(dev)
    pitch = torch.randint(1, 255, (1, T), generator=gen).to(dev)
    nsff0 = O.synth_f0(1, T, gen).to(dev)
    sid = torch.tensor([3], device=dev)
    cap = {}

    def pre_hook(mod, args, kwargs):
        cap["x"], cap["f0"], cap["g"] = args[0].detach().cpu(), args[1].detach().cpu(), kwargs["g"].detach().cpu()
        torch.manual_seed(99)  # the decoder draws rand(1,1,1) then randn(B, n, 1) on the device generator next

    net_g.dec.register_forward_pre_hook(pre_hook, with_kwargs=True)
    with torch.no_grad():
        o, x_mask, _ = net_g.infer(phone, torch.tensor([T], device=dev), pitch, nsff0, sid)
    assert o.shape == (1, 1, T * 480) and bool(torch.isfinite(o).all())
    torch.manual_seed(99)
    torch.rand(1, 1, 1, device=dev)
    noise = torch.randn(1, T * 480, 1, device=dev).cpu().reshape(1, -1)
    w = O.folded_weights({k: v.half().float() for k, v in ref.dec.state_dict().items()})
    y_ref = O.decoder_forward(w, O.CONFIG_48K, cap["x"], cap["f0"], cap["g"], noise)
    assert float((o.cpu() - y_ref).abs().max()) <= 2e-3
    # the loader's module shares nothing with net_g.dec but computes the same thing, and a second load of the same
    # checkpoint re-uses the packed, device-resident weights (no second fold / upload)
    from rvc_b200 import loader

    loader.PACKED.clear()
    m1 = rvc_b200.load_checkpoint(cpt, device=dev)
    with torch.no_grad():
        y1 = m1(cap["x"].to(dev), cap["f0"].to(dev), cap["g"].to(dev), noise=noise.to(dev), mode="fp32").cpu()
    assert float((y1 - o.cpu()).abs().max()) <= 1e-6
    m2 = rvc_b200.load_checkpoint(cpt, device=dev)
    with torch.no_grad():
        y2 = m2(cap["x"].to(dev), cap["f0"].to(dev), cap["g"].to(dev), noise=noise.to(dev), mode="fp32").cpu()
    assert torch.equal(y1, y2)
    assert m2._handle is m1._handle and loader.PACKED.hits >= 1
