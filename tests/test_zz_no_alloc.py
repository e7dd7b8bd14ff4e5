"""The allocation-free entry point rvcdec_source_noalloc (include/rvcdec.h; scratch = a static device array inside the
library) against the existing entry point's harmonic source (rvcdec_forward stopped after the source, "har" tap), through
the C ABI, on two calls in a row with different inputs (the shared scratch keeps the first call's contents).

Collected last on purpose (file name): a failure here must not hide other tests."""
import ctypes as C

import pytest
import torch

from oracle import nsf_hifigan_oracle as O
from helpers import build_module

pytestmark = pytest.mark.gpu


def _har_existing(m, cfg, f0, noise, dev):
    """har through the existing entry point: rvcdec_forward with the test hook "stop after the source"."""
    from rvc_b200 import _lib

    B, T = f0.shape
    x = torch.zeros(B, cfg.initial_channel, T, device=dev)
    with torch.no_grad():
        m(x, f0, None, noise=noise, mode="fp32")  # builds handle / workspace
    h, ws = m._handle, m._workspace
    stream = torch.cuda.current_stream(dev).cuda_stream
    try:
        h.set_stop(1)
        with torch.no_grad():
            m(x, f0, None, noise=noise, mode="fp32")
        dst = torch.empty(B, T * cfg.upp, dtype=torch.float32, device=dev)
        n = h.debug_read("har", dst, ws, B, T, _lib.MODES["fp32"], stream)
        assert n == dst.numel()
    finally:
        h.set_stop(0)
    torch.cuda.synchronize()
    return dst


def test_no_alloc_source_matches_the_existing_entry_point_twice_in_a_row():
    from rvc_b200 import _lib

    if not torch.cuda.is_available():
        pytest.fail("rvcdec_source_noalloc needs a CUDA device: there is no CPU fallback")
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    cfg = O.CONFIG_48K
    m = build_module(cfg, 4242).to(dev)
    sd = m.state_dict()
    lin_w = float(sd["m_source.l_linear.weight"].reshape(()))
    lin_b = float(sd["m_source.l_linear.bias"].reshape(()))
    L = _lib.lib()
    stream = torch.cuda.current_stream(dev).cuda_stream
    gen = torch.Generator().manual_seed(11)
    # second call: fewer, shorter items than the first, so stale scratch entries of call 1 lie right behind its own
    for B, T in ((3, 333), (2, 97)):
        f0 = O.synth_f0(B, T, gen)
        f0[B - 1, : T // 2] = 0.0  # an unvoiced stretch
        noise = torch.randn(B, T * cfg.upp, generator=gen)
        f0d, nd = f0.to(dev).contiguous(), noise.to(dev).contiguous()
        want = _har_existing(m, cfg, f0d, nd, dev)
        got = torch.full((B, T * cfg.upp), float("nan"), dtype=torch.float32, device=dev)
        torch.cuda.synchronize()
        rc = L.rvcdec_source_noalloc(f0d.data_ptr(), nd.data_ptr(), C.c_uint64(0), got.data_ptr(), B, T, cfg.upp, C.c_float(cfg.sr),
                                     C.c_float(lin_w), C.c_float(lin_b), stream)
        assert rc == 0, (rc, _lib.last_error())
        torch.cuda.synchronize()
        assert bool(torch.isfinite(got).all())
        err = float((got - want).abs().max())
        assert err <= 2e-5, (B, T, err)  # the tolerance of test_source_kernel_matches_oracle
        # and against the fp64 oracle itself
        s = O.sine_source(f0, cfg.upp, cfg.sr, noise, dtype=torch.float64)
        exp = torch.tanh(s * lin_w + lin_b).to(torch.float32)
        assert float((got.cpu() - exp).abs().max()) <= 2e-5
    # beyond the static scratch: a clear error, nothing launched
    B, T = 8, 9000
    f0d = torch.zeros(B, T, device=dev)
    out = torch.zeros(B, T * cfg.upp, device=dev)
    rc = L.rvcdec_source_noalloc(f0d.data_ptr(), None, C.c_uint64(0), out.data_ptr(), B, T, cfg.upp, C.c_float(cfg.sr), C.c_float(1.0),
                                 C.c_float(0.0), stream)
    assert rc < 0 and "exceeds the static scratch" in _lib.last_error()
