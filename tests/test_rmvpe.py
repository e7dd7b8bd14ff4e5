"""RMVPE f0 predictor: oracle against the reference-generated golden and the reference's own decode fixture
(CPU), module naming / seeded init against the live reference when present, GPU parity through include/rvcrmvpe.h."""
import os
import sys

import numpy as np
import pytest
import torch

from oracle import rmvpe_oracle as RO
from helpers import record_parity

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _golden():
    return np.load(os.path.join(GOLDEN, "rmvpe_e2e_B2_T64.npz"))


def _seeded_module(d):
    """rvc_b200.rmvpe.E2E built like oracle/make_golden_r2.make_rmvpe_case built the reference's: seeded default init, then
    randomised BatchNorm statistics / affine parameters."""
    from rvc_b200.rmvpe import E2E

    seed = int(d["init_seed"])
    torch.manual_seed(seed)
    m = E2E(4, 1, (2, 2)).eval()
    gen = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        for mod in m.modules():
            if isinstance(mod, torch.nn.BatchNorm2d):
                mod.running_mean.copy_(0.2 * torch.randn(mod.running_mean.shape, generator=gen))
                mod.running_var.copy_(0.5 + torch.rand(mod.running_var.shape, generator=gen))
                mod.weight.copy_(1.0 + 0.2 * torch.randn(mod.weight.shape, generator=gen))
                mod.bias.copy_(0.1 * torch.randn(mod.bias.shape, generator=gen))
    return m


def test_decode_restatement_against_the_references_own_fixture():
    """ios_test_data/rmvpe_hidden.npy -> rmvpe_f0.npy (every 4th frame, tests/golden/rmvpe_decode_real.npz)."""
    d = np.load(os.path.join(GOLDEN, "rmvpe_decode_real.npz"))
    f0 = RO.decode(d["hidden"])
    assert f0.shape == d["f0"].shape
    assert float(np.abs(f0 - d["f0"]).max()) <= 1e-9
    assert (d["f0"] > 0).sum() > 50 and (d["f0"] == 0).sum() > 10  # voiced and unvoiced frames are both in the fixture


def test_module_init_and_oracle_match_the_reference_golden():
    d = _golden()
    m = _seeded_module(d)
    sd = {k: v for k, v in m.state_dict().items()}
    cs = np.array([float(v.double().abs().sum()) for k, v in sorted(sd.items()) if v.dtype.is_floating_point])
    assert cs.shape == d["checksums"].shape and np.allclose(cs, d["checksums"], rtol=0, atol=0)  # same names, same draws
    taps = {}
    hid = RO.e2e_forward(sd, torch.from_numpy(d["mel"]), taps)
    assert float((hid - torch.from_numpy(d["hidden"])).abs().max()) <= 1e-6
    assert float((taps["cnn"] - torch.from_numpy(d["tap_cnn"])).abs().max()) <= 1e-5


def test_state_dict_names_equal_the_live_reference():
    if not os.path.isdir("/root/reference/rvc/lib/predictors"):
        pytest.skip("reference checkout not present")
    sys.path.insert(0, "/root/reference")
    from oracle.make_golden_r2 import _import_reference_rmvpe
    from rvc_b200.rmvpe import E2E

    R = _import_reference_rmvpe()
    torch.manual_seed(5)
    ref = R.E2E(4, 1, (2, 2))
    torch.manual_seed(5)
    mine = E2E(4, 1, (2, 2))
    a, b = ref.state_dict(), mine.state_dict()
    assert list(a.keys()) == list(b.keys())
    assert all(torch.equal(a[k], b[k]) for k in a)
    mine.load_state_dict(a)


def test_mel_filterbank_and_stft_restatement():
    """The filter bank restates librosa's definition (unpinned: librosa is absent); sanity: rows are non-negative triangles whose
    Slaney-normalised areas are ~1 bin wide in Hz terms, centres increase, the oracle and the module agree bit for bit."""
    from rvc_b200.rmvpe import mel_filterbank

    a, b = RO.mel_filterbank(), mel_filterbank()
    assert a.shape == (128, 513) and np.array_equal(a, b)
    assert (a >= 0).all() and (a.sum(1) > 0).all()
    centres = a.argmax(1)
    assert (np.diff(centres) >= 0).all() and centres[0] >= 1 and centres[-1] <= 512
    # each triangle integrates to ~1 over frequency (area normalisation): sum * bin width (15.625 Hz) ~ 1
    assert np.allclose(a.sum(1) * (8000.0 / 512), 1.0, atol=0.35)


# salience (0..1) / tap tolerances: fp32 mode; tensor-core mode (deep 3 x 3 convolutions with fp16 operands)
TOL = {0: dict(tap=1e-4, hidden=1e-5), 1: dict(tap=3e-3, hidden=3e-4)}


@pytest.mark.gpu
@pytest.mark.parametrize("tc", [0, 1, 2])
def test_e2e_forward_matches_the_reference_golden(tc):
    """tc: 0 = fp32; 1 = default (tensor cores, TMA-fed where Cin % 64 == 0); 2 = tensor cores with the gathered kernels only"""
    d = _golden()
    dev = torch.device("cuda", 0)
    m = _seeded_module(d).to(dev)
    m.set_option("tc", 1 if tc else 0)
    m.set_option("tma", 0 if tc == 2 else 1)
    m.set_option("keep_inter", 1)
    mel = torch.from_numpy(d["mel"]).to(dev)
    hid = m(mel)
    h = m._handle
    stream = torch.cuda.current_stream(dev).cuda_stream
    B, _, T = mel.shape
    errs = {}
    for name, shape in (("cnn", (B, T, 384)), ("gru", (B, T, 512))):
        dst = torch.empty(shape, dtype=torch.float32, device=dev)
        assert h.debug_read(name, dst, m._workspace, B, T, stream) == dst.numel()
        torch.cuda.synchronize()
        ref = torch.from_numpy(d["tap_" + name])
        errs[name] = float((dst.cpu() - ref).abs().max() / ref.abs().max())
    # the deep levels barely reach the output under random weights: hold them to the oracle where they end (B, 512, T/32, 4)
    taps = {}
    RO.e2e_forward({k: v.cpu() for k, v in m.state_dict().items()}, torch.from_numpy(d["mel"]), taps)
    dst = torch.empty(taps["inter"].shape, dtype=torch.float32, device=dev)
    assert h.debug_read("inter", dst, m._workspace, B, T, stream) == dst.numel()
    torch.cuda.synchronize()
    errs["inter"] = float((dst.cpu() - taps["inter"]).abs().max() / taps["inter"].abs().max())
    assert errs["inter"] <= (5e-3 if tc else 1e-4), errs
    if tc:
        assert errs["inter"] >= 1e-6, errs   # fp16 operands leave a trace: the tensor-core path did run
    e = float((hid.cpu() - torch.from_numpy(d["hidden"])).abs().max())
    record_parity({"test": f"rmvpe_e2e_golden_B2_T64[tc={tc}]", "hidden_max_abs": e, "tap_rel": errs, "launches": m.last_launch_count})
    assert errs["cnn"] <= TOL[min(tc, 1)]["tap"] and errs["gru"] <= TOL[min(tc, 1)]["tap"], errs
    assert e <= TOL[min(tc, 1)]["hidden"], e


@pytest.mark.gpu
def test_predictor_pipeline_against_the_oracle():
    """mel spectrogram (vs torch.stft + the same filter bank), reflect padding to 32 frames, model, decode - a 1.3 s tone
    with vibrato plus noise, T = 131 frames (not a multiple of 32); and decode on the reference's own fixture."""
    import rvc_b200

    d = _golden()
    dev = torch.device("cuda", 0)
    p = rvc_b200.RMVPE0Predictor(None, dev)
    p.model = _seeded_module(d).to(dev)
    p.model.set_option("tc", 0)
    sr, n = 16000, 20900
    t = np.arange(n) / sr
    gen = np.random.default_rng(3)
    audio = (0.4 * np.sin(2 * np.pi * (220 * t + 3 * np.sin(2 * np.pi * 5 * t))) + 0.01 * gen.standard_normal(n)).astype(np.float32)
    a = torch.from_numpy(audio).unsqueeze(0)
    mel = p.mel_extractor(a.to(dev))
    mel_ref = RO.mel_spectrogram(a, torch.from_numpy(RO.mel_filterbank()))
    assert mel.shape == mel_ref.shape == (1, 128, 1 + n // 160)
    e_mel = float((mel.cpu() - mel_ref).abs().max())
    sd = {k: v.cpu() for k, v in p.model.state_dict().items()}
    hid = p.mel2hidden(mel)
    hid_ref = RO.mel2hidden(sd, mel.cpu())
    e_hid = float((hid.cpu() - hid_ref).abs().max())
    f0 = p.decode(hid[0]).cpu().numpy()
    f0_ref = RO.decode(hid.cpu().numpy()[0])
    e_f0 = float(np.abs(f0 - f0_ref).max())
    f0_all = p.infer_from_audio(audio)
    real = np.load(os.path.join(GOLDEN, "rmvpe_decode_real.npz"))
    f0_real = p.decode(torch.from_numpy(real["hidden"]).to(dev)).cpu().numpy()
    e_real = float(np.abs(f0_real - real["f0"]).max())
    record_parity({"test": "rmvpe_pipeline", "mel_max_abs": e_mel, "hidden_max_abs": e_hid, "f0_max_abs_vs_oracle_decode": e_f0,
                   "decode_vs_reference_fixture_max_abs_hz": e_real})
    assert e_mel <= 2e-3, e_mel      # log of fp32 DFT sums (n_fft = 1024) against torch.stft's FFT
    assert e_hid <= 1e-4, e_hid
    assert e_f0 <= 1e-3, e_f0
    assert e_real <= 1e-3, e_real    # fp32 store of an fp64 result, f0 <= 170 Hz
    assert np.array_equal(f0_all, f0)
    # tensor-core mode on the same clip: salience and f0 against the fp32 mode
    p.model.set_option("tc", 1)
    hid_tc = p.mel2hidden(mel)
    f0_tc = p.decode(hid_tc[0]).cpu().numpy()
    e_hid_tc = float((hid_tc - hid).abs().max())
    voiced = (f0 > 0) & (f0_tc > 0)
    cents = float(np.abs(1200 * np.log2(f0_tc[voiced] / f0[voiced])).max()) if voiced.any() else 0.0
    record_parity({"test": "rmvpe_pipeline_tc_vs_fp32", "hidden_max_abs": e_hid_tc, "f0_max_cents": cents,
                   "voicing_flips": int(((f0 > 0) != (f0_tc > 0)).sum()), "frames": int(f0.shape[0])})
    assert e_hid_tc <= 3e-4, e_hid_tc


@pytest.mark.gpu
def test_long_clip_modes_agree():
    """T = 1024 frames, B = 2: every convolution level has several tiles and the deep levels split K; the tensor-core modes
    (TMA-fed, gathered) against the fp32 mode on the same weights."""
    d = _golden()
    dev = torch.device("cuda", 0)
    m = _seeded_module(d).to(dev)
    mel = (torch.randn(2, 128, 1024, generator=torch.Generator().manual_seed(9)) * 2 - 4).to(dev)
    outs = {}
    for name, (tc, tma) in {"fp32": (0, 0), "tma": (1, 1), "gathered": (1, 0)}.items():
        m.set_option("tc", tc)
        m.set_option("tma", tma)
        outs[name] = m(mel).clone()
    e_tma = float((outs["tma"] - outs["fp32"]).abs().max())
    e_gat = float((outs["gathered"] - outs["fp32"]).abs().max())
    record_parity({"test": "rmvpe_long_clip_modes", "frames": 1024, "tma_vs_fp32_max_abs": e_tma, "gathered_vs_fp32_max_abs": e_gat})
    assert e_tma <= 3e-4 and e_gat <= 3e-4, (e_tma, e_gat)
    assert e_tma > 0 and e_gat > 0   # the tensor-core paths did run
