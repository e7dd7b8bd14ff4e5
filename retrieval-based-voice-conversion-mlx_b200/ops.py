"""torch custom op ``rvc_b200::nsf_hifigan_decode`` - the thin layer between torch tensors and the
C ABI (include/rvcdec.h).  The op only forwards device pointers and torch's current stream to
``rvcdec_forward``; all arithmetic happens in librvcdec.so."""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib


@torch.library.custom_op("rvc_b200::nsf_hifigan_decode", mutates_args=("workspace",), device_types="cuda")
def nsf_hifigan_decode(x: torch.Tensor, f0: torch.Tensor, g: Optional[torch.Tensor], noise: Optional[torch.Tensor],
                       workspace: torch.Tensor, handle: int, mode: int, upp: int, seed: int) -> torch.Tensor:
    """x (B,C,T) fp32, f0 (B,T) fp32, g (B,gin) fp32 | None, noise (B,T*upp) fp32 | None (in-kernel Philox4x32-10 keyed by seed),
    workspace uint8 scratch, handle = rvcdec_handle* as int, mode = rvcdec_mode.  Returns (B, T*upp) fp32."""
    for t in (x, f0, g, noise):
        if t is not None and (not t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous()):
            raise RuntimeError("nsf_hifigan_decode needs contiguous fp32 CUDA tensors")
    B, _, T = x.shape
    L = _lib.lib()
    n = T * upp
    if noise is not None and (noise.shape[0] != B or noise.numel() % (B * n) != 0):
        raise RuntimeError(f"noise must be (B, T*upp[, harmonics]) with B = {B}, T*upp = {n}")
    out = torch.empty(B, n, dtype=torch.float32, device=x.device)
    stream = torch.cuda.current_stream(x.device).cuda_stream
    with torch.cuda.device(x.device):
        rc = L.rvcdec_forward(C.c_void_p(handle), x.data_ptr(), f0.data_ptr(), g.data_ptr() if g is not None else None,
                              noise.data_ptr() if noise is not None else None, C.c_uint64(seed & (2**64 - 1)), out.data_ptr(), workspace.data_ptr(), workspace.numel(),
                              B, T, mode, C.c_void_p(stream))
    _lib.check(rc, "rvcdec_forward")
    return out


@nsf_hifigan_decode.register_fake
def _(x, f0, g, noise, workspace, handle, mode, upp, seed):
    return x.new_empty(x.shape[0], x.shape[2] * upp)
