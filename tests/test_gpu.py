"""GPU parity tests (run with -m gpu on the B200 box): the CUDA path, called through the drop-in module
-> torch custom op -> C ABI, against the CPU oracle and the reference-generated golden fixtures.

Bars (BASELINE.json north_star): fp32 mode max-abs <= 2e-3; fast mode multi-resolution log-mel
correlation (min over the 7 resolutions) >= 0.999.
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import nsf_hifigan_oracle as O
from helpers import build_module, golden_state, load_golden, tensors

pytestmark = pytest.mark.gpu

FP32_BAR = 2e-3
FAST_BAR = 0.999


@pytest.fixture(scope="module")
def dev():
    assert torch.cuda.is_available(), "GPU tests need CUDA"
    return torch.device("cuda", 0)


def _module_for(name, dev):
    d, cfg = load_golden(name)
    m = build_module(cfg, int(d["init_seed"]))
    sd = golden_state(d, cfg)
    m.load_state_dict(sd)
    return d, cfg, m.to(dev), O.folded_weights(sd)


def layer_taps(m, cfg, x, f0, g, noise, mode):
    """Run the CUDA path with early stops and fetch the intermediates the C ABI exposes."""
    from rvc_b200 import _lib

    dev = x.device
    B, _, T = x.shape
    mm = _lib.MODES[mode]
    out = {}
    with torch.no_grad():
        m(x, f0, g, noise=noise, mode=mode)  # builds handle / workspace
    h, ws = m._handle, m._workspace
    stream = torch.cuda.current_stream(dev).cuda_stream
    L = T
    plan = [(1, "har", (B, T * cfg.upp)), (2, "pre", (B, cfg.upsample_initial_channel, T))]
    for i, u in enumerate(cfg.upsample_rates):
        L *= u
        C = cfg.upsample_initial_channel >> (i + 1)
        plan += [(10 + i, f"stage{i}.in", (B, C, L)), (20 + i, f"stage{i}.out", (B, C, L))]
    try:
        for stop, name, shape in plan:
            h.set_stop(stop)
            with torch.no_grad():
                m(x, f0, g, noise=noise, mode=mode)
            dst = torch.empty(shape, dtype=torch.float32, device=dev)
            n = h.debug_read(name, dst, ws, B, T, mm, stream)
            assert n == dst.numel()
            torch.cuda.synchronize()
            out[name] = dst.cpu()
    finally:
        h.set_stop(0)
    return out


def oracle_taps(w, cfg, x, f0, g, noise):
    taps = {}
    y = O.decoder_forward(w, cfg, x, f0, g, noise, taps=taps)
    last = len(cfg.upsample_rates) - 1
    exp = {"har": taps["har"], "pre": F.leaky_relu(taps["pre"], 0.1)}
    for i in range(last + 1):
        exp[f"stage{i}.in"] = taps[f"stage{i}.in"]
        exp[f"stage{i}.out"] = taps[f"stage{i}.out"] if i == last else F.leaky_relu(taps[f"stage{i}.out"], 0.1)
    return exp, y


def layer_errors(name, mode, dev):
    d, cfg, m, w = _module_for(name, dev)
    x, f0, g, noise, y = tensors(d)
    got = layer_taps(m, cfg, x.to(dev), f0.to(dev), g.to(dev), noise.to(dev), mode)
    exp, _ = oracle_taps(w, cfg, x, f0, g, noise)
    errs = {}
    for k in exp:
        scale = float(exp[k].abs().max()) + 1e-12
        errs[k] = (float((got[k] - exp[k]).abs().max()), scale)
    return errs


def test_source_kernel_matches_oracle(dev):
    """One kernel: f0 upsample + fp64 phase scan + sine + injected noise + Linear + tanh, vs SineGenerator."""
    d, cfg, m, w = _module_for("v2_48k_B2_T24", dev)
    gen = torch.Generator().manual_seed(3)
    f0 = O.synth_f0(3, 333, gen)
    f0[2] = 0.0  # fully unvoiced item
    noise = torch.randn(3, 333 * cfg.upp, generator=gen)
    x = torch.zeros(3, cfg.initial_channel, 333)
    got = layer_taps(m, cfg, x.to(dev), f0.to(dev), None, noise.to(dev), "fp32")["har"]
    s = O.sine_source(f0, cfg.upp, cfg.sr, noise, dtype=torch.float64)
    exp = torch.tanh(s * w["m_source.l_linear.weight"].reshape(()) + w["m_source.l_linear.bias"].reshape(()))
    assert float((got - exp).abs().max()) <= 2e-5
    # and against the reference's own fp32 phase accumulation (its cumsum error is the difference)
    s32 = O.sine_source(f0, cfg.upp, cfg.sr, noise)
    exp32 = torch.tanh(s32 * w["m_source.l_linear.weight"].reshape(()) + w["m_source.l_linear.bias"].reshape(()))
    assert float((got - exp32).abs().max()) <= 1e-4


@pytest.mark.parametrize("name", ["tiny_24k_B2_T40", "v2_48k_B2_T24", "v2_40k_B1_T32", "v2_32k_B1_T16", "odd_rates_24k_B2_T30"])
def test_fp32_mode_matches_reference_golden(dev, name):
    d, cfg, m, w = _module_for(name, dev)
    x, f0, g, noise, y = tensors(d)
    with torch.no_grad():
        got = m(x.to(dev), f0.to(dev), g.to(dev), noise=noise.to(dev), mode="fp32").cpu()
    assert got.shape == y.shape
    err = float((got - y).abs().max())
    assert err <= FP32_BAR, err
    assert err <= 5e-5, f"fp32 path should be at rounding level, got {err}"


@pytest.mark.parametrize("name", ["tiny_24k_B2_T40", "v2_48k_B2_T24"])
def test_fp32_mode_layer_by_layer(dev, name):
    errs = layer_errors(name, "fp32", dev)
    for k, (e, scale) in errs.items():
        assert e <= 2e-5 * max(1.0, scale), (k, e, scale)


def test_fast_mode_layer_by_layer(dev):
    errs = layer_errors("v2_48k_B2_T24", "fast", dev)
    for k, (e, scale) in errs.items():
        tol = 2e-5 if k == "har" else 3e-2 * max(1.0, scale)
        assert e <= tol, (k, e, scale)


@pytest.mark.parametrize("cfg_name,B,T,seed", [("48k", 2, 150, 1), ("48k", 1, 257, 2), ("40k", 2, 120, 3), ("32k", 1, 100, 4)])
@pytest.mark.parametrize("mode", ["fp32", "fast"])
def test_against_oracle_on_seeded_inputs(dev, cfg_name, B, T, seed, mode):
    cfg = {"48k": O.CONFIG_48K, "40k": O.CONFIG_40K, "32k": O.CONFIG_32K}[cfg_name]
    m = build_module(cfg, 1000 + seed)
    w = O.folded_weights(m.state_dict())
    m = m.to(dev)
    x, f0, g, noise = O.synth_inputs(cfg, B, T, seed=seed)
    y_ref = O.decoder_forward(w, cfg, x, f0, g, noise)
    with torch.no_grad():
        y = m(x.to(dev), f0.to(dev), g.to(dev), noise=noise.to(dev), mode=mode).cpu()
    assert y.shape == (B, 1, T * cfg.upp)
    err = float((y - y_ref).abs().max())
    if mode == "fp32":
        assert err <= FP32_BAR, err
    else:
        corr = O.multires_logmel_corr(y, y_ref, cfg.sr)
        assert corr["min"] >= FAST_BAR, (corr, err)
        assert err <= 2e-2, err


@pytest.mark.parametrize("fused,zs", [(0, 1), (1, 1), (2, 1), (1, 0)])
@pytest.mark.parametrize("cfg_name,B,T,seed", [("48k", 2, 150, 11), ("40k", 1, 97, 12)])
def test_fused_mrf_variants_agree(dev, fused, zs, cfg_name, B, T, seed):
    """rvcdec_set_option("fused_mrf"): per-conv kernels (0), fused ResBlock kernel per ResBlock (1) and per
    stage (2) compute the same decoder; each meets the fast bar and they agree closely with one another
    (same operand rounding, different accumulation order).  Tile edges: T*upp/4 rows is not a multiple of
    the fused tile interior, and several tiles per item are exercised."""
    cfg = {"48k": O.CONFIG_48K, "40k": O.CONFIG_40K}[cfg_name]
    m = build_module(cfg, 2000 + seed)
    w = O.folded_weights(m.state_dict())
    m = m.to(dev)
    x, f0, g, noise = O.synth_inputs(cfg, B, T, seed=seed)
    y_ref = O.decoder_forward(w, cfg, x, f0, g, noise)
    xd, fd, gd, nd_ = x.to(dev), f0.to(dev), g.to(dev), noise.to(dev)
    with torch.no_grad():
        m.fused_mrf = 0
        y0 = m(xd, fd, gd, noise=nd_, mode="fast").cpu()
        n0 = m.last_launch_count
        m.fused_mrf = fused
        m._handle.set_option("zero_stuff", zs)  # stride-2 upsamples: zero-stuffed conv (1) or polyphase (0)
        y = m(xd, fd, gd, noise=nd_, mode="fast").cpu()
        n1 = m.last_launch_count
    if fused:
        assert n1 < n0, (n0, n1)
    corr = O.multires_logmel_corr(y, y_ref, cfg.sr)
    assert corr["min"] >= FAST_BAR, corr
    assert float((y - y0).abs().max()) <= 2e-3, float((y - y0).abs().max())


@pytest.mark.parametrize("cfg_name,B,T,seed", [("48k", 2, 150, 31), ("40k", 2, 97, 32), ("48k", 1, 3, 33), ("48k", 3, 43, 34), ("32k", 1, 61, 35)])
def test_phase_major_upsample_matches_phase_scatter(dev, cfg_name, B, T, seed):
    """rvcdec_set_option("ups_pm"): the rate-10 / 12 upsamples with phase-major GEMM columns (warp-transposed, coalesced
    epilogue that also carries stage 1's 8-tap noise conv and writes the activated 16-bit operand) against the
    one-phase-per-tile scatter epilogue + noise_add pass.  Same operand rounding and tap order, so the stage inputs agree
    to fp32 rounding; ragged lengths (T*12 not a multiple of 32 or 128) exercise the partial last tiles."""
    cfg = {"48k": O.CONFIG_48K, "40k": O.CONFIG_40K, "32k": O.CONFIG_32K}[cfg_name]
    m = build_module(cfg, 3000 + seed)
    w = O.folded_weights(m.state_dict())
    m = m.to(dev)
    x, f0, g, noise = O.synth_inputs(cfg, B, T, seed=seed)
    y_ref = O.decoder_forward(w, cfg, x, f0, g, noise)
    xd, fd, gd, nd_ = x.to(dev), f0.to(dev), g.to(dev), noise.to(dev)
    with torch.no_grad():
        m(xd, fd, gd, noise=nd_, mode="fast")
        m._handle.set_option("ups_pm", 0)
        y0 = m(xd, fd, gd, noise=nd_, mode="fast").cpu()
        n0 = m.last_launch_count
        t0 = layer_taps(m, cfg, xd, fd, gd, nd_, "fast")
        m._handle.set_option("ups_pm", 1)
        y1 = m(xd, fd, gd, noise=nd_, mode="fast").cpu()
        n1 = m.last_launch_count
        t1 = layer_taps(m, cfg, xd, fd, gd, nd_, "fast")
    if cfg_name != "32k":
        assert n1 < n0, (n0, n1)  # stage 1 lost its noise_add launch
    for k in ("stage0.in", "stage0.out", "stage1.in", "stage1.out"):
        scale = float(t0[k].abs().max())
        assert float((t1[k] - t0[k]).abs().max()) <= 2e-5 * max(scale, 1.0) + (2e-2 * scale if k.endswith(".out") else 0), k
    assert float((t1["stage1.in"] - t0["stage1.in"]).abs().max()) <= 1e-5
    corr = O.multires_logmel_corr(y1, y_ref, cfg.sr)
    assert corr["min"] >= FAST_BAR, corr
    assert float((y1 - y0).abs().max()) <= 2e-3, float((y1 - y0).abs().max())


@pytest.mark.parametrize("fused", [1])
@pytest.mark.parametrize("cfg_name,B,T,seed", [("48k", 2, 150, 21), ("40k", 1, 97, 22), ("48k", 1, 3, 23), ("32k", 3, 61, 24)])
def test_fused_conv_post_matches_separate_kernel(dev, fused, cfg_name, B, T, seed):
    """rvcdec_set_option("fused_post"): conv_post + tanh computed inside the last fused-MRF launch (tile
    margins widened by conv_post's 3 rows) vs the stand-alone post kernel on the same stage output.  Same
    fp32 products, different summation order: the waveforms agree to a few ulp of the pre-tanh sum.  Covers
    sequence ends (zero padding of conv_post), items shorter than one tile and several tiles per item."""
    cfg = {"48k": O.CONFIG_48K, "40k": O.CONFIG_40K, "32k": O.CONFIG_32K}[cfg_name]
    m = build_module(cfg, 3000 + seed).to(dev)
    x, f0, g, noise = O.synth_inputs(cfg, B, T, seed=seed)
    xd, fd, gd, nd_ = x.to(dev), f0.to(dev), g.to(dev), noise.to(dev)
    with torch.no_grad():
        m.fused_mrf = fused
        m._ensure_handle(dev).set_option("fused_post", 0)
        y0 = m(xd, fd, gd, noise=nd_, mode="fast").cpu()
        n0 = m.last_launch_count
        m._handle.set_option("fused_post", 1)
        y1 = m(xd, fd, gd, noise=nd_, mode="fast").cpu()
        n1 = m.last_launch_count
    assert n1 == n0 - 1, (n0, n1)
    assert y1.shape == y0.shape
    assert float((y1 - y0).abs().max()) <= 2e-5, float((y1 - y0).abs().max())


@pytest.mark.parametrize("cfg_name,B,T,seed", [("48k", 2, 150, 41), ("40k", 1, 97, 42), ("48k", 3, 2, 43), ("32k", 2, 61, 44)])
def test_fused_pair_kernel_matches_per_conv_path(dev, cfg_name, B, T, seed):
    """rvcdec_set_option("pair_mrf"): the 128-channel stage as one launch per (conv1, conv2) pair (fp32 residual
    in TMEM, operands in shared memory) vs one launch per conv.  Same operand rounding and fp32 accumulation;
    only the order in which bias / residual enter the sums differs.  Covers items shorter than one tile, several
    tiles per item and ragged tile ends."""
    cfg = {"48k": O.CONFIG_48K, "40k": O.CONFIG_40K, "32k": O.CONFIG_32K}[cfg_name]
    m = build_module(cfg, 5000 + seed)
    w = O.folded_weights(m.state_dict())
    m = m.to(dev)
    x, f0, g, noise = O.synth_inputs(cfg, B, T, seed=seed)
    y_ref = O.decoder_forward(w, cfg, x, f0, g, noise)
    xd, fd, gd, nd_ = x.to(dev), f0.to(dev), g.to(dev), noise.to(dev)
    with torch.no_grad():
        m._ensure_handle(dev).set_option("pair_mrf", 0)
        y0 = m(xd, fd, gd, noise=nd_, mode="fast").cpu()
        n0 = m.last_launch_count
        m._handle.set_option("pair_mrf", 1)
        y1 = m(xd, fd, gd, noise=nd_, mode="fast").cpu()
        n1 = m.last_launch_count
    # 18 conv launches -> 9 pair launches; the per-conv baseline already has its noise conv inside the phase-major
    # upsample where that applies (48k / 40k), the pair path keeps a noise_add launch unless "pair_noise" is set
    assert n1 <= n0 - 8, (n0, n1)
    assert float((y1 - y0).abs().max()) <= 2e-3, float((y1 - y0).abs().max())
    if T >= 50:
        assert O.multires_logmel_corr(y1, y_ref, cfg.sr)["min"] >= FAST_BAR


@pytest.mark.parametrize("cfg_name,B,T,seed", [("48k", 5, 60, 51), ("40k", 3, 97, 52)])
def test_item_grouped_launches_are_bit_identical(dev, cfg_name, B, T, seed):
    """rvcdec_set_option("mrf_items"): the per-conv MRF stages launched over groups of items (descending order, so
    the stage output written into a0 - zero-stuffed on the way into a stride-2 upsample - never lands on an input
    that is still to be read) compute exactly what one launch over all items computes."""
    cfg = {"48k": O.CONFIG_48K, "40k": O.CONFIG_40K}[cfg_name]
    m = build_module(cfg, 6000 + seed).to(dev)
    x, f0, g, noise = O.synth_inputs(cfg, B, T, seed=seed)
    xd, fd, gd, nd_ = x.to(dev), f0.to(dev), g.to(dev), noise.to(dev)
    with torch.no_grad():
        h = m._ensure_handle(dev)
        h.set_option("mrf_items", 0)
        ypm = m(xd, fd, gd, noise=nd_, mode="fast").cpu()  # default: stage 1's noise conv + operand copy ride in the upsample
        # grouped launches keep the separate noise_add pass (the stage hand-off buffers differ), so bit identity is
        # checked between the two launch shapes of that same arithmetic
        h.set_option("ups_pm", 0)
        y0 = m(xd, fd, gd, noise=nd_, mode="fast").cpu()
        n0 = m.last_launch_count
        h.set_option("mrf_items", 1)  # 1 MiB budget: groups of 1-2 items at these sizes
        y1 = m(xd, fd, gd, noise=nd_, mode="fast").cpu()
        n1 = m.last_launch_count
        h.set_option("ups_pm", 1)
        y2 = m(xd, fd, gd, noise=nd_, mode="fast").cpu()  # grouped launches + phase-major upsamples
        h.set_option("mrf_items", 0)
    assert n1 > n0, (n0, n1)
    assert torch.equal(y0, y1)
    assert float((y2 - ypm).abs().max()) <= 2e-3 and float((y2 - y0).abs().max()) <= 2e-3


@pytest.mark.parametrize("split_k", [7, 11])
@pytest.mark.parametrize("cfg_name,B,T,seed", [("48k", 2, 150, 61), ("40k", 1, 97, 62), ("48k", 1, 2, 63)])
def test_split_fused_chains_agree(dev, split_k, cfg_name, B, T, seed):
    """rvcdec_set_option("fused_split_k"): ResBlock chains of the fused stages with kernel size >= split_k run as two
    launches (dilations 1,3 -> fp32 -> dilation 5) instead of one.  Same operands, same fp32 accumulation; only the
    point at which the conv2 biases enter the residual differs."""
    cfg = {"48k": O.CONFIG_48K, "40k": O.CONFIG_40K}[cfg_name]
    m = build_module(cfg, 7000 + seed).to(dev)
    x, f0, g, noise = O.synth_inputs(cfg, B, T, seed=seed)
    xd, fd, gd, nd_ = x.to(dev), f0.to(dev), g.to(dev), noise.to(dev)
    with torch.no_grad():
        h = m._ensure_handle(dev)
        h.set_option("fused_split_k", -1)  # never split
        y0 = m(xd, fd, gd, noise=nd_, mode="fast").cpu()
        n0 = m.last_launch_count
        h.set_option("fused_split_k", split_k)
        y1 = m(xd, fd, gd, noise=nd_, mode="fast").cpu()
        n1 = m.last_launch_count
        h.set_option("fused_split_k", 0)  # default: automatic
    assert n1 == n0 + (2 if split_k == 11 else 4), (n0, n1)  # two fused stages x ResBlocks with k >= split_k
    assert float((y1 - y0).abs().max()) <= 2e-3, float((y1 - y0).abs().max())


@pytest.mark.parametrize("cfg_name,B,T,seed", [("48k", 2, 150, 31), ("40k", 1, 97, 32)])
def test_fused_epilogue_warp_counts_agree(dev, cfg_name, B, T, seed):
    """rvcdec_set_option("fused_epi_warps"): 8 or 16 epilogue warps split the accumulator columns differently
    but compute every element with the same arithmetic; only conv_post's partial-sum order differs."""
    cfg = {"48k": O.CONFIG_48K, "40k": O.CONFIG_40K}[cfg_name]
    m = build_module(cfg, 4000 + seed).to(dev)
    x, f0, g, noise = O.synth_inputs(cfg, B, T, seed=seed)
    xd, fd, gd, nd_ = x.to(dev), f0.to(dev), g.to(dev), noise.to(dev)
    ys = []
    with torch.no_grad():
        for ew in (8, 16):
            m._ensure_handle(dev).set_option("fused_epi_warps", ew)
            ys.append(m(xd, fd, gd, noise=nd_, mode="fast").cpu())
    assert float((ys[0] - ys[1]).abs().max()) <= 2e-6, float((ys[0] - ys[1]).abs().max())


def test_readme_shape_both_modes(dev):
    """BASELINE configs[2]: 13.5 s at the README benchmark shape (T=1350), real RMVPE-like contour shape."""
    cfg = O.CONFIG_48K
    m = build_module(cfg, 1234)
    w = O.folded_weights(m.state_dict())
    m = m.to(dev)
    x, f0, g, noise = O.synth_inputs(cfg, 1, 1350, seed=1234)
    y_ref = O.decoder_forward(w, cfg, x, f0, g, noise)
    with torch.no_grad():
        y32 = m(x.to(dev), f0.to(dev), g.to(dev), noise=noise.to(dev), mode="fp32").cpu()
        yf = m(x.to(dev), f0.to(dev), g.to(dev), noise=noise.to(dev), mode="fast").cpu()
    assert y32.shape == (1, 1, 648000)
    assert float((y32 - y_ref).abs().max()) <= FP32_BAR
    assert O.multires_logmel_corr(yf, y_ref, cfg.sr)["min"] >= FAST_BAR


@pytest.mark.parametrize("mode", ["fp32", "fast"])
def test_size_independent_properties(dev, mode):
    """Batch items are independent; ragged tile edges, micro-batching and g=None behave; output is bounded."""
    cfg = O.CONFIG_48K
    m = build_module(cfg, 7).to(dev)
    x, f0, g, noise = (t.to(dev) for t in O.synth_inputs(cfg, 3, 77, seed=9))
    with torch.no_grad():
        y = m(x, f0, g, noise=noise, mode=mode)
        y1 = m(x[1:2], f0[1:2], g[1:2], noise=noise[1:2], mode=mode)
        assert torch.equal(y[1:2], y1), "batch items must not influence each other"
        y_again = m(x, f0, g, noise=noise, mode=mode)
        assert torch.equal(y, y_again), "deterministic for fixed noise"
        # micro-batching: a workspace that only fits one item gives the same result
        full = m.max_workspace_bytes
        try:
            m.max_workspace_bytes = 1
            m._workspace = None
            y_mb = m(x, f0, g, noise=noise, mode=mode)
        finally:
            m.max_workspace_bytes = full
            m._workspace = None
        assert torch.equal(y, y_mb)
        y_nog = m(x, f0, None, noise=noise, mode=mode)
        assert y_nog.shape == y.shape and not torch.equal(y_nog, y)
        assert float(y.abs().max()) <= 1.0 and bool(torch.isfinite(y).all())
        # causality of the zero padding: a longer input with the same prefix only changes the tail region
        xl = torch.cat([x, torch.randn(3, cfg.initial_channel, 23, device=dev)], dim=2)
        f0l = torch.cat([f0, f0[:, :23]], dim=1)
        nl = torch.cat([noise, noise[:, : 23 * cfg.upp]], dim=1)
        yl = m(xl, f0l, g, noise=nl, mode=mode)
        keep = (77 - 12) * cfg.upp  # receptive field is ~11 frames
        assert float((yl[:, :, :keep] - y[:, :, :keep]).abs().max()) <= 1e-5


def test_saturation_and_bounds(dev):
    """Scaled weights drive tanh into saturation (random-init outputs never get near 1)."""
    cfg = O.CONFIG_48K
    m = build_module(cfg, 11)
    with torch.no_grad():
        m.conv_post.weight.mul_(200.0)
    w = O.folded_weights(m.state_dict())
    m = m.to(dev)
    x, f0, g, noise = O.synth_inputs(cfg, 1, 40, seed=2)
    y_ref = O.decoder_forward(w, cfg, x, f0, g, noise)
    with torch.no_grad():
        y = m(x.to(dev), f0.to(dev), g.to(dev), noise=noise.to(dev), mode="fp32").cpu()
    assert float(y_ref.abs().max()) > 0.99
    assert float(y.abs().max()) <= 1.0
    assert float((y - y_ref).abs().max()) <= FP32_BAR


def test_reference_noise_order_and_host_entry(dev):
    """Default noise = torch generator draws in the reference's order; decode_host == forward(philox) for one seed."""
    cfg = O.CONFIG_48K
    m = build_module(cfg, 5).to(dev)
    x, f0, g, _ = O.synth_inputs(cfg, 2, 30, seed=4)
    xd, fd, gd = x.to(dev), f0.to(dev), g.to(dev)
    with torch.no_grad():
        torch.manual_seed(77)
        y_default = m(xd, fd, gd, mode="fp32")
        torch.manual_seed(77)
        torch.rand(1, 1, 1, device=dev)
        nz = torch.randn(2, 30 * cfg.upp, 1, device=dev)
        y_explicit = m(xd, fd, gd, noise=nz, mode="fp32")
        assert torch.equal(y_default, y_explicit)
        torch.manual_seed(5)
        y_ph = m(xd, fd, gd, noise="philox", mode="fast")
        torch.manual_seed(5)
        y_host = m.decode_host(x, f0, g, mode="fast")
        assert y_host.shape == (2, 1, 30 * cfg.upp) and not y_host.is_cuda
        assert torch.equal(y_ph.cpu(), y_host)
        # Philox draws are standard normal: unvoiced-only input isolates them (har = tanh(w*(0.1/3)z + b))
        assert m.last_launch_count > 40


def test_long_form_chunked_decode_matches_reference_stitching(dev):
    """BASELINE configs[4] in miniature: t_pad-overlap chunks decoded independently, trimmed and concatenated
    equal the oracle run chunk by chunk (the reference restarts the sine phase per chunk too)."""
    from rvc_b200 import chunking as ch

    cfg = O.CONFIG_48K
    m = build_module(cfg, 21)
    w = O.folded_weights(m.state_dict())
    m = m.to(dev)
    n_frames = 260
    chunks = ch.plan_chunks(n_frames, split_points=[90, 170])
    total = n_frames + 200
    x, f0, g, _ = O.synth_inputs(cfg, 1, total, seed=8)
    outs, refs = [], []
    for c in chunks:
        xs, fs = x[:, :, c.start:c.stop], f0[:, c.start:c.stop]
        nz = torch.randn(1, c.frames * cfg.upp, generator=torch.Generator().manual_seed(c.index))
        refs.append(O.decoder_forward(w, cfg, xs, fs, g, nz)[0, 0].numpy())
        with torch.no_grad():
            outs.append(m(xs.to(dev).contiguous(), fs.to(dev).contiguous(), g.to(dev), noise=nz.to(dev), mode="fp32")[0, 0].cpu().numpy())
    trims = [c.trim * cfg.upp for c in chunks]
    y, y_ref = ch.stitch(outs, trims), ch.stitch(refs, trims)
    assert y.shape == (n_frames * cfg.upp,)
    assert float(np.abs(y - y_ref).max()) <= FP32_BAR


def test_batched_long_form_driver_matches_per_chunk_decode(dev):
    """chunking.decode_long: all t_pad-overlap chunks of an utterance submitted as ONE zero-padded batch, trimmed and
    concatenated, against the oracle run chunk by chunk on the same injected noise (fp32 mode), and the two-rank
    sharding of the same plan (decode_chunks with rank / world) against the single-rank result."""
    from rvc_b200 import chunking as ch

    cfg = O.CONFIG_48K
    m = build_module(cfg, 22)
    w = O.folded_weights(m.state_dict())
    m = m.to(dev)
    n_frames, pad = 300, 100
    total = n_frames + 2 * pad
    x, f0, g, _ = O.synth_inputs(cfg, 1, total, seed=9)
    noise = torch.randn(1, total * cfg.upp, generator=torch.Generator().manual_seed(3))
    splits = [70, 190]  # ragged chunk lengths: 270, 320, 310 frames
    y = ch.decode_long(m, x, f0, g, mode="fp32", noise=noise, split_points=splits)
    chunks = ch.plan_chunks(n_frames, split_points=splits)
    refs = [O.decoder_forward(w, cfg, x[:, :, c.start:c.stop], f0[:, c.start:c.stop], g,
                              noise[:, c.start * cfg.upp:c.stop * cfg.upp])[0, 0].numpy() for c in chunks]
    y_ref = ch.stitch(refs, [c.trim * cfg.upp for c in chunks])
    assert y.shape == y_ref.shape == (n_frames * cfg.upp,)
    assert float(np.abs(y - y_ref).max()) <= FP32_BAR
    parts = {}
    for r in range(2):
        parts.update(ch.decode_chunks(m, x, f0, g, chunks, mode="fp32", noise=noise, rank=r, world=2))
    y2 = ch.stitch([parts[c.index] for c in chunks], [c.trim * cfg.upp for c in chunks])
    assert float(np.abs(y2 - y).max()) <= 1e-5
    # host-buffer path (in-kernel noise): shape / range only
    yh = ch.decode_long(m, x, f0, g, mode="fast", split_points=splits, crossfade=240)
    assert yh.shape == (n_frames * cfg.upp,) and np.isfinite(yh).all() and float(np.abs(yh).max()) <= 1.0


def test_nof0_sibling_decoder_matches_reference_golden(dev):
    """rvc_b200.HiFiGANGenerator (the no-f0 models' decoder, hifigan.py:72-97) on the NSF kernels with zeroed
    source / noise-conv layers, against the reference-generated golden output: fp32 mode to the fp32 bar, fast mode by
    waveform correlation and relative max-abs error (the signal is quiet: no harmonic source feeds it)."""
    from rvc_b200 import HiFiGANGenerator

    d, cfg = load_golden("nof0_40k_B2_T20")
    torch.manual_seed(int(d["init_seed"]))
    m = HiFiGANGenerator(cfg.initial_channel, list(cfg.resblock_kernel_sizes), [list(r) for r in cfg.resblock_dilation_sizes],
                         list(cfg.upsample_rates), cfg.upsample_initial_channel, list(cfg.upsample_kernel_sizes),
                         gin_channels=cfg.gin_channels).eval().to(dev)
    x, g, y_ref = torch.from_numpy(d["x"]), torch.from_numpy(d["g"]), torch.from_numpy(d["y"])
    with torch.no_grad():
        y32 = m(x.to(dev), g.to(dev), mode="fp32").cpu()
        yf = m(x.to(dev), g.to(dev), mode="fast").cpu()
        yh = m.decode_host(x, g, mode="fp32")
        m(x.to(dev), g.to(dev), mode="fast")
        n_skip = m.last_launch_count
        # option "no_source" = 0: source kernel + noise adds run and add exactly 0 (all-zero layers)
        m._handle.set_option("no_source", 0)
        y32_z = m(x.to(dev), g.to(dev), mode="fp32").cpu()
        yf_z = m(x.to(dev), g.to(dev), mode="fast").cpu()
        n_full = m.last_launch_count
        m._handle.set_option("no_source", 1)
    assert n_skip < n_full, (n_skip, n_full)
    assert torch.equal(y32_z, y32) and torch.equal(yf_z, yf)
    assert y32.shape == y_ref.shape
    assert float((y32 - y_ref).abs().max()) <= 1e-5
    assert torch.equal(yh, y32)
    scale = float(y_ref.abs().max())
    cc = float(torch.corrcoef(torch.stack([yf.flatten(), y_ref.flatten()]))[0, 1])
    assert cc >= FAST_BAR, cc
    assert float((yf - y_ref).abs().max()) <= 0.05 * scale, (float((yf - y_ref).abs().max()), scale)


@pytest.mark.parametrize("cfg_name,B,T,seed", [("48k", 2, 75, 71), ("40k", 1, 33, 72)])
def test_conv_epilogue_shape_options_are_bit_identical(dev, cfg_name, B, T, seed):
    """rvcdec_set_option("conv_epi_warps" / "conv_epi_warps_small" / "pm_a_stages"): 8 epilogue warps over 32-column
    items or 16 over 16-column items, two or three activation stages - launch shapes of the same arithmetic, so the
    waveforms must be bit-identical (per-conv MRF path everywhere so that every tiling is exercised)."""
    cfg = {"48k": O.CONFIG_48K, "40k": O.CONFIG_40K}[cfg_name]
    m = build_module(cfg, 8000 + seed).to(dev)
    x, f0, g, noise = O.synth_inputs(cfg, B, T, seed=seed)
    xd, fd, gd, nd_ = x.to(dev), f0.to(dev), g.to(dev), noise.to(dev)
    outs = []
    with torch.no_grad():
        m.fused_mrf = 0
        h = m._ensure_handle(dev)
        for ew, ews, pas in ((16, 16, 2), (8, 8, 2), (16, 8, 3), (8, 16, 3)):
            h.set_option("conv_epi_warps", ew)
            h.set_option("conv_epi_warps_small", ews)
            h.set_option("pm_a_stages", pas)
            outs.append(m(xd, fd, gd, noise=nd_, mode="fast").cpu())
        h.set_option("conv_epi_warps", 16); h.set_option("conv_epi_warps_small", 16); h.set_option("pm_a_stages", 2)
    for y in outs[1:]:
        assert torch.equal(y, outs[0])
