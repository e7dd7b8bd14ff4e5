"""ctypes binding of librvcdec.so (include/rvcdec.h).

The library is built in-tree by ``__graft_entry__.build()`` / ``csrc/build.sh``.
There is no fallback of any kind: if the shared object is missing or a call
fails, a ``RuntimeError`` carrying ``rvcdec_last_error()`` is raised.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Sequence

HERE = os.path.dirname(os.path.abspath(__file__))
# (RVC_B200_LIB: A/B runs against another build of the library, e.g. an older one that lacks the newer entry points)
LIB_PATH = os.environ.get("RVC_B200_LIB") or os.path.join(HERE, "librvcdec.so")

MAX_UPS = 8
MAX_KERNELS = 8
MAX_DIL = 8

MODE_FP32 = 0
MODE_FAST = 1          # bf16 operands up to stage 1, fp16 after (default fast plan)
MODE_FAST_F16 = 2      # fp16 operands everywhere
MODE_FAST_BF16 = 3     # bf16 operands everywhere
MODES = {"fp32": MODE_FP32, "fast": MODE_FAST, "fast_f16": MODE_FAST_F16, "fast_bf16": MODE_FAST_BF16}


class Config(C.Structure):
    _fields_ = [
        ("initial_channel", C.c_int32),
        ("upsample_initial_channel", C.c_int32),
        ("gin_channels", C.c_int32),
        ("sr", C.c_int32),
        ("num_upsamples", C.c_int32),
        ("upsample_rates", C.c_int32 * MAX_UPS),
        ("upsample_kernel_sizes", C.c_int32 * MAX_UPS),
        ("num_kernels", C.c_int32),
        ("resblock_kernel_sizes", C.c_int32 * MAX_KERNELS),
        ("num_dilations", C.c_int32),
        ("resblock_dilation_sizes", (C.c_int32 * MAX_DIL) * MAX_KERNELS),
        ("harmonic_num", C.c_int32),
    ]


class SynConfig(C.Structure):
    """rvcsyn_config (include/rvcsyn.h)."""
    _fields_ = [(n, C.c_int32) for n in ("inter_channels", "hidden_channels", "filter_channels", "n_heads", "n_layers", "kernel_size",
                                         "text_dim", "spk_embed_dim", "gin_channels", "use_f0", "window_size", "flow_kernel",
                                         "flow_layers", "n_flows")] + [("dec", Config)]


class RefineConfig(C.Structure):
    """rvcrefine_config (include/rvcrefine.h)."""
    _fields_ = [("sample_rate", C.c_int32), ("num_upsamples", C.c_int32), ("upsample_rates", C.c_int32 * MAX_UPS), ("num_mels", C.c_int32),
                ("start_channels", C.c_int32), ("upsample_initial_channel", C.c_int32), ("gin_channels", C.c_int32)]


class Tensor(C.Structure):
    _fields_ = [("name", C.c_char_p), ("data", C.c_void_p), ("numel", C.c_int64)]


class HubertHandle:
    """Owns one rvchubert_handle (the HuBERT / ContentVec content embedder; include/rvchubert.h)."""

    def __init__(self, device: int):
        self._h = C.c_void_p()
        self._check(lib().rvchubert_create(int(device), C.byref(self._h)), "rvchubert_create")
        self.device = int(device)

    @staticmethod
    def _check(rc: int, what: str) -> None:
        if rc < 0:
            raise RuntimeError(f"{what} failed ({rc}): {(lib().rvchubert_last_error() or b'').decode()}")

    def load_weights(self, state: Dict[str, "object"]) -> None:
        arr, keep = Handle._tensor_array(None, state)
        self._check(lib().rvchubert_load_weights(self._h, arr, len(state)), "rvchubert_load_weights")

    @staticmethod
    def frames(n: int) -> int:
        return int(lib().rvchubert_frames(int(n)))

    def workspace_bytes(self, B: int, n: int) -> int:
        return int(lib().rvchubert_workspace_bytes(self._h, B, n))

    def forward(self, audio, hidden, workspace, B, n, stream) -> None:
        self._check(lib().rvchubert_forward(self._h, audio, hidden, workspace.data_ptr(), workspace.numel(), B, n, stream), "rvchubert_forward")

    def final_proj(self, x, out, rows, stream) -> None:
        self._check(lib().rvchubert_final_proj(self._h, x, out, rows, stream), "rvchubert_final_proj")

    def set_option(self, key: str, value: int) -> None:
        self._check(lib().rvchubert_set_option(self._h, key.encode(), int(value)), "rvchubert_set_option")

    def debug_read(self, name: str, dst, workspace, B: int, n: int, stream) -> int:
        k = lib().rvchubert_debug_read(self._h, name.encode(), dst.data_ptr(), dst.numel(), workspace.data_ptr(), B, n, stream)
        self._check(k, "rvchubert_debug_read")
        return int(k)

    def launch_count(self) -> int:
        return int(lib().rvchubert_last_launch_count(self._h))

    def close(self) -> None:
        if self._h:
            lib().rvchubert_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class ProfileEntry(C.Structure):
    _fields_ = [("name", C.c_char * 32), ("ms", C.c_double), ("flops", C.c_double), ("bytes", C.c_double), ("spans", C.c_int64)]


EXPORTS = [
    "rvcdec_create", "rvcdec_load_weights", "rvcdec_workspace_bytes", "rvcdec_forward", "rvcdec_forward_host",
    "rvcdec_set_option", "rvcdec_set_initial_phases", "rvcdec_last_launch_count", "rvcdec_profile_enable", "rvcdec_profile_read", "rvcdec_debug_set_stop", "rvcdec_debug_read", "rvcdec_debug_pack_host",
    "rvcdec_debug_packed", "rvcdec_debug_fused_timing", "rvcdec_debug_stat", "rvcdec_source_noalloc", "rvcdec_destroy", "rvcdec_last_error", "rvcdec_version",
    # include/rvcsyn.h
    "rvcsyn_create", "rvcsyn_load_weights", "rvcsyn_workspace_bytes", "rvcsyn_infer", "rvcsyn_decoder", "rvcsyn_last_launch_count",
    "rvcsyn_destroy", "rvcsyn_last_error", "rvcsyn_debug_set_stop", "rvcsyn_debug_read", "rvcsyn_set_option",
    # include/rvcrefine.h
    "rvcrefine_create", "rvcrefine_load_weights", "rvcrefine_workspace_bytes", "rvcrefine_forward", "rvcrefine_noise_floats",
    "rvcrefine_last_launch_count", "rvcrefine_destroy", "rvcrefine_last_error",
    "rvcrmvpe_create", "rvcrmvpe_load_weights", "rvcrmvpe_mel_workspace_bytes", "rvcrmvpe_mel", "rvcrmvpe_workspace_bytes", "rvcrmvpe_forward",
    "rvcrmvpe_decode", "rvcrmvpe_debug_read", "rvcrmvpe_set_option", "rvcrmvpe_last_launch_count", "rvcrmvpe_destroy", "rvcrmvpe_last_error",
    # include/rvchubert.h
    "rvchubert_create", "rvchubert_load_weights", "rvchubert_frames", "rvchubert_workspace_bytes", "rvchubert_forward", "rvchubert_final_proj",
    "rvchubert_debug_read", "rvchubert_set_option", "rvchubert_last_launch_count", "rvchubert_destroy", "rvchubert_last_error",
]

_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU or PyTorch fallback for the decoder)")
    L = C.CDLL(LIB_PATH)
    if os.environ.get("RVC_B200_LIB"):
        have = L

        class _Tolerant:  # an older build: entry points it lacks get a stub that raises when called
            def __getattr__(self, name):
                try:
                    return getattr(have, name)
                except AttributeError:
                    class _Missing:
                        argtypes = restype = None

                        def __call__(self, *a, **k):
                            raise RuntimeError(f"{name} is not exported by {LIB_PATH}")
                    m = _Missing()
                    setattr(self, name, m)
                    return m
        L = _Tolerant()
    vp, i32, i64, u64, sz = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_size_t
    L.rvcdec_create.argtypes = [C.POINTER(Config), C.c_int, C.POINTER(vp)]
    L.rvcdec_create.restype = C.c_int
    L.rvcdec_load_weights.argtypes = [vp, C.POINTER(Tensor), i32]
    L.rvcdec_load_weights.restype = C.c_int
    L.rvcdec_workspace_bytes.argtypes = [vp, i32, i32, i32]
    L.rvcdec_workspace_bytes.restype = sz
    L.rvcdec_forward.argtypes = [vp, vp, vp, vp, vp, u64, vp, vp, sz, i32, i32, i32, vp]
    L.rvcdec_forward.restype = C.c_int
    L.rvcdec_forward_host.argtypes = [vp, vp, vp, vp, vp, u64, vp, i32, i32, i32, sz]
    L.rvcdec_forward_host.restype = C.c_int
    L.rvcdec_set_option.argtypes = [vp, C.c_char_p, i64]
    L.rvcdec_set_option.restype = C.c_int
    L.rvcdec_set_initial_phases.argtypes = [vp, vp]
    L.rvcdec_set_initial_phases.restype = C.c_int
    L.rvcdec_last_launch_count.argtypes = [vp]
    L.rvcdec_last_launch_count.restype = i64
    L.rvcdec_profile_enable.argtypes = [vp, i32]
    L.rvcdec_profile_enable.restype = C.c_int
    L.rvcdec_profile_read.argtypes = [vp, C.POINTER(ProfileEntry), i32]
    L.rvcdec_profile_read.restype = i32
    L.rvcdec_source_noalloc.argtypes = [vp, vp, u64, vp, i32, i32, i32, C.c_float, C.c_float, C.c_float, vp]
    L.rvcdec_source_noalloc.restype = C.c_int
    L.rvcdec_debug_set_stop.argtypes = [vp, i32]
    L.rvcdec_debug_set_stop.restype = C.c_int
    L.rvcdec_debug_read.argtypes = [vp, C.c_char_p, vp, i64, vp, i32, i32, i32, vp]
    L.rvcdec_debug_read.restype = i64
    L.rvcdec_debug_pack_host.argtypes = [vp, C.POINTER(Tensor), i32]
    L.rvcdec_debug_pack_host.restype = C.c_int
    L.rvcdec_debug_packed.argtypes = [vp, C.c_char_p, i32, vp, i64]
    L.rvcdec_debug_packed.restype = i64
    L.rvcdec_debug_fused_timing.argtypes = [vp, vp, i64]
    L.rvcdec_debug_fused_timing.restype = i64
    L.rvcdec_debug_stat.argtypes = [vp, C.c_char_p]
    L.rvcdec_debug_stat.restype = C.c_double
    L.rvcdec_destroy.argtypes = [vp]
    L.rvcdec_destroy.restype = None
    L.rvcdec_last_error.argtypes = []
    L.rvcdec_last_error.restype = C.c_char_p
    L.rvcdec_version.argtypes = []
    L.rvcdec_version.restype = C.c_char_p
    L.rvcrefine_create.argtypes = [C.POINTER(RefineConfig), C.c_int, C.POINTER(vp)]
    L.rvcrefine_create.restype = C.c_int
    L.rvcrefine_load_weights.argtypes = [vp, C.POINTER(Tensor), i32]
    L.rvcrefine_load_weights.restype = C.c_int
    L.rvcrefine_workspace_bytes.argtypes = [vp, i32, i32]
    L.rvcrefine_workspace_bytes.restype = sz
    L.rvcrefine_forward.argtypes = [vp, vp, vp, vp, vp, u64, vp, vp, sz, i32, i32, vp]
    L.rvcrefine_forward.restype = C.c_int
    L.rvcrefine_noise_floats.argtypes = [vp, i32, i32]
    L.rvcrefine_noise_floats.restype = i64
    L.rvcrefine_last_launch_count.argtypes = [vp]
    L.rvcrefine_last_launch_count.restype = i64
    L.rvcrefine_destroy.argtypes = [vp]
    L.rvcrefine_destroy.restype = None
    L.rvcrefine_last_error.argtypes = []
    L.rvcrefine_last_error.restype = C.c_char_p
    L.rvcrmvpe_create.argtypes = [C.c_int, C.POINTER(vp)]
    L.rvcrmvpe_create.restype = C.c_int
    L.rvcrmvpe_load_weights.argtypes = [vp, C.POINTER(Tensor), i32]
    L.rvcrmvpe_load_weights.restype = C.c_int
    L.rvcrmvpe_mel_workspace_bytes.argtypes = [vp, i32, i32]
    L.rvcrmvpe_mel_workspace_bytes.restype = sz
    L.rvcrmvpe_mel.argtypes = [vp, vp, vp, vp, sz, i32, i32, vp]
    L.rvcrmvpe_mel.restype = C.c_int
    L.rvcrmvpe_workspace_bytes.argtypes = [vp, i32, i32]
    L.rvcrmvpe_workspace_bytes.restype = sz
    L.rvcrmvpe_forward.argtypes = [vp, vp, vp, vp, sz, i32, i32, vp]
    L.rvcrmvpe_forward.restype = C.c_int
    L.rvcrmvpe_decode.argtypes = [vp, vp, C.c_float, vp, i32, vp]
    L.rvcrmvpe_decode.restype = C.c_int
    L.rvcrmvpe_debug_read.argtypes = [vp, C.c_char_p, vp, i64, vp, i32, i32, vp]
    L.rvcrmvpe_debug_read.restype = i64
    L.rvcrmvpe_set_option.argtypes = [vp, C.c_char_p, i64]
    L.rvcrmvpe_set_option.restype = C.c_int
    L.rvcrmvpe_last_launch_count.argtypes = [vp]
    L.rvcrmvpe_last_launch_count.restype = i64
    L.rvcrmvpe_destroy.argtypes = [vp]
    L.rvcrmvpe_destroy.restype = None
    L.rvcrmvpe_last_error.argtypes = []
    L.rvcrmvpe_last_error.restype = C.c_char_p
    L.rvchubert_create.argtypes = [C.c_int, C.POINTER(vp)]
    L.rvchubert_create.restype = C.c_int
    L.rvchubert_load_weights.argtypes = [vp, C.POINTER(Tensor), i32]
    L.rvchubert_load_weights.restype = C.c_int
    L.rvchubert_frames.argtypes = [i32]
    L.rvchubert_frames.restype = i32
    L.rvchubert_workspace_bytes.argtypes = [vp, i32, i32]
    L.rvchubert_workspace_bytes.restype = sz
    L.rvchubert_forward.argtypes = [vp, vp, vp, vp, sz, i32, i32, vp]
    L.rvchubert_forward.restype = C.c_int
    L.rvchubert_final_proj.argtypes = [vp, vp, vp, i64, vp]
    L.rvchubert_final_proj.restype = C.c_int
    L.rvchubert_debug_read.argtypes = [vp, C.c_char_p, vp, i64, vp, i32, i32, vp]
    L.rvchubert_debug_read.restype = i64
    L.rvchubert_set_option.argtypes = [vp, C.c_char_p, i64]
    L.rvchubert_set_option.restype = C.c_int
    L.rvchubert_last_launch_count.argtypes = [vp]
    L.rvchubert_last_launch_count.restype = i64
    L.rvchubert_destroy.argtypes = [vp]
    L.rvchubert_destroy.restype = None
    L.rvchubert_last_error.argtypes = []
    L.rvchubert_last_error.restype = C.c_char_p
    L.rvcsyn_create.argtypes = [C.POINTER(SynConfig), C.c_int, C.POINTER(vp)]
    L.rvcsyn_create.restype = C.c_int
    L.rvcsyn_load_weights.argtypes = [vp, C.POINTER(Tensor), i32]
    L.rvcsyn_load_weights.restype = C.c_int
    L.rvcsyn_workspace_bytes.argtypes = [vp, i32, i32, i32]
    L.rvcsyn_workspace_bytes.restype = sz
    L.rvcsyn_infer.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, u64, vp, vp, vp, sz, i32, i32, i32, vp]
    L.rvcsyn_infer.restype = C.c_int
    L.rvcsyn_set_option.argtypes = [vp, C.c_char_p, i64]
    L.rvcsyn_set_option.restype = C.c_int
    L.rvcsyn_debug_set_stop.argtypes = [vp, i32]
    L.rvcsyn_debug_set_stop.restype = C.c_int
    L.rvcsyn_debug_read.argtypes = [vp, C.c_char_p, vp, i64, vp, i32, i32, vp]
    L.rvcsyn_debug_read.restype = i64
    L.rvcsyn_decoder.argtypes = [vp]
    L.rvcsyn_decoder.restype = vp
    L.rvcsyn_last_launch_count.argtypes = [vp]
    L.rvcsyn_last_launch_count.restype = i64
    L.rvcsyn_destroy.argtypes = [vp]
    L.rvcsyn_destroy.restype = None
    L.rvcsyn_last_error.argtypes = []
    L.rvcsyn_last_error.restype = C.c_char_p
    _lib = L
    return L


def last_error() -> str:
    return (lib().rvcdec_last_error() or b"").decode("utf-8", "replace")


def check(rc: int, what: str) -> None:
    if rc < 0:
        raise RuntimeError(f"{what} failed ({rc}): {last_error()}")


def make_config(initial_channel: int, resblock_kernel_sizes: Sequence[int], resblock_dilation_sizes,
                upsample_rates: Sequence[int], upsample_initial_channel: int, upsample_kernel_sizes: Sequence[int],
                gin_channels: int, sr: int, harmonic_num: int = 0) -> Config:
    if len(upsample_rates) > MAX_UPS or len(resblock_kernel_sizes) > MAX_KERNELS:
        raise ValueError("too many upsample / resblock stages")
    nd = {len(d) for d in resblock_dilation_sizes}
    if len(nd) != 1 or nd.pop() > MAX_DIL:
        raise ValueError("every resblock must have the same number (<= 8) of dilations")
    c = Config()
    c.initial_channel = initial_channel
    c.upsample_initial_channel = upsample_initial_channel
    c.gin_channels = gin_channels
    c.sr = sr
    c.harmonic_num = harmonic_num
    c.num_upsamples = len(upsample_rates)
    for i, (u, k) in enumerate(zip(upsample_rates, upsample_kernel_sizes)):
        c.upsample_rates[i] = u
        c.upsample_kernel_sizes[i] = k
    c.num_kernels = len(resblock_kernel_sizes)
    c.num_dilations = len(resblock_dilation_sizes[0])
    for j, k in enumerate(resblock_kernel_sizes):
        c.resblock_kernel_sizes[j] = k
        for m, d in enumerate(resblock_dilation_sizes[j]):
            c.resblock_dilation_sizes[j][m] = d
    return c


class RefineHandle:
    """Owns one rvcrefine_handle (the RefineGAN sibling decoder; include/rvcrefine.h)."""

    def __init__(self, cfg: RefineConfig, device: int):
        self._h = C.c_void_p()
        self._check(lib().rvcrefine_create(C.byref(cfg), int(device), C.byref(self._h)), "rvcrefine_create")
        self.device = int(device)

    @staticmethod
    def _check(rc: int, what: str) -> None:
        if rc < 0:
            raise RuntimeError(f"{what} failed ({rc}): {(lib().rvcrefine_last_error() or b'').decode()}")

    def load_weights(self, state: Dict[str, "object"]) -> None:
        arr, keep = Handle._tensor_array(None, state)
        self._check(lib().rvcrefine_load_weights(self._h, arr, len(state)), "rvcrefine_load_weights")

    def workspace_bytes(self, B: int, T: int) -> int:
        return int(lib().rvcrefine_workspace_bytes(self._h, B, T))

    def noise_floats(self, B: int, T: int) -> int:
        return int(lib().rvcrefine_noise_floats(self._h, B, T))

    def forward(self, mel, f0, g, noise, seed, out, workspace, B, T, stream) -> None:
        self._check(lib().rvcrefine_forward(self._h, mel, f0, g, noise, C.c_uint64(seed & (2**64 - 1)), out, workspace.data_ptr(),
                                            workspace.numel(), B, T, stream), "rvcrefine_forward")

    def launch_count(self) -> int:
        return int(lib().rvcrefine_last_launch_count(self._h))

    def close(self) -> None:
        if self._h:
            lib().rvcrefine_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class SynHandle:
    """Owns one rvcsyn_handle (the whole Synthesizer.infer; include/rvcsyn.h)."""

    def __init__(self, cfg: SynConfig, device: int):
        self._h = C.c_void_p()
        rc = lib().rvcsyn_create(C.byref(cfg), int(device), C.byref(self._h))
        if rc < 0:
            raise RuntimeError(f"rvcsyn_create failed ({rc}): {(lib().rvcsyn_last_error() or b'').decode()}")
        self.device = int(device)

    def _check(self, rc: int, what: str) -> None:
        if rc < 0:
            raise RuntimeError(f"{what} failed ({rc}): {(lib().rvcsyn_last_error() or b'').decode()}")

    def load_weights(self, state: Dict[str, "object"]) -> None:
        arr, keep = Handle._tensor_array(None, state)
        self._check(lib().rvcsyn_load_weights(self._h, arr, len(state)), "rvcsyn_load_weights")

    def workspace_bytes(self, B: int, T: int, mode: int) -> int:
        return int(lib().rvcsyn_workspace_bytes(self._h, B, T, mode))

    def infer(self, phone, lengths, pitch, nsff0, sid, noise_zp, noise_src, seed, out, z_out, workspace, B, T, mode, stream) -> None:
        self._check(lib().rvcsyn_infer(self._h, phone, lengths, pitch, nsff0, sid, noise_zp, noise_src, C.c_uint64(seed & (2**64 - 1)), out,
                                       z_out, workspace.data_ptr(), workspace.numel(), B, T, mode, stream), "rvcsyn_infer")

    def set_stop(self, stop: int) -> None:
        self._check(lib().rvcsyn_debug_set_stop(self._h, int(stop)), "rvcsyn_debug_set_stop")

    def set_option(self, key: str, value: int):
        self._check(lib().rvcsyn_set_option(self._h, key.encode(), int(value)), "rvcsyn_set_option")

    def debug_read(self, name: str, dst, workspace, B: int, T: int, stream) -> int:
        n = lib().rvcsyn_debug_read(self._h, name.encode(), dst.data_ptr(), dst.numel(), workspace.data_ptr(), B, T, stream)
        self._check(n, "rvcsyn_debug_read")
        return int(n)

    def launch_count(self) -> int:
        return int(lib().rvcsyn_last_launch_count(self._h))

    def close(self) -> None:
        if self._h:
            lib().rvcsyn_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Handle:
    """Owns one rvcdec_handle."""

    def __init__(self, cfg: Config, device: int):
        self._h = C.c_void_p()
        check(lib().rvcdec_create(C.byref(cfg), int(device), C.byref(self._h)), "rvcdec_create")
        self.device = int(device)
        self._keep = None

    @property
    def ptr(self) -> C.c_void_p:
        return self._h

    def _tensor_array(self, state: Dict[str, "object"]):
        import torch

        keep = []
        arr = (Tensor * len(state))()
        for i, (k, v) in enumerate(state.items()):
            t = v.detach().to(device="cpu", dtype=torch.float32).contiguous()
            name = k.encode()
            keep.append((t, name))
            arr[i].name = name
            arr[i].data = t.data_ptr()
            arr[i].numel = t.numel()
        return arr, keep

    def load_weights(self, state: Dict[str, "object"]) -> None:
        arr, keep = self._tensor_array(state)
        check(lib().rvcdec_load_weights(self._h, arr, len(state)), "rvcdec_load_weights")

    def pack_host(self, state: Dict[str, "object"]) -> None:
        arr, keep = self._tensor_array(state)
        check(lib().rvcdec_debug_pack_host(self._h, arr, len(state)), "rvcdec_debug_pack_host")

    def packed(self, layer: str, kind: int):
        import numpy as np

        n = lib().rvcdec_debug_packed(self._h, layer.encode(), kind, None, 0)
        check(n, "rvcdec_debug_packed")
        dt = {0: np.float32, 1: np.uint16, 2: np.int32, 3: np.uint32, 4: np.uint16, 5: np.int32, 6: np.float32}[kind]
        out = np.empty(n // np.dtype(dt).itemsize, dtype=dt)
        check(lib().rvcdec_debug_packed(self._h, layer.encode(), kind, out.ctypes.data, n), "rvcdec_debug_packed")
        return out

    def workspace_bytes(self, B: int, T: int, mode: int) -> int:
        return int(lib().rvcdec_workspace_bytes(self._h, B, T, mode))

    def forward(self, x, f0, g, noise, seed, out, workspace, B, T, mode, stream) -> None:
        check(lib().rvcdec_forward(self._h, x, f0, g, noise, C.c_uint64(seed & (2**64 - 1)), out, workspace.data_ptr(),
                                   workspace.numel(), B, T, mode, stream), "rvcdec_forward")

    def forward_host(self, x, f0, g, noise, seed, out, B, T, mode, max_device_bytes=0) -> None:
        check(lib().rvcdec_forward_host(self._h, x, f0, g, noise, C.c_uint64(seed & (2**64 - 1)), out, B, T, mode,
                                        max_device_bytes), "rvcdec_forward_host")

    def set_initial_phases(self, ptr) -> None:
        check(lib().rvcdec_set_initial_phases(self._h, ptr), "rvcdec_set_initial_phases")

    def set_option(self, key: str, value: int) -> None:
        check(lib().rvcdec_set_option(self._h, key.encode(), int(value)), "rvcdec_set_option")

    def fused_timing(self, num_sms: int = 148):
        import numpy as np

        buf = np.zeros(32 * num_sms * 16, dtype=np.int64)
        n = lib().rvcdec_debug_fused_timing(self._h, buf.ctypes.data, buf.size)
        check(n, "rvcdec_debug_fused_timing")
        return buf[: n * num_sms * 16].reshape(n, num_sms, 16)

    def stat(self, name: str) -> float:
        return float(lib().rvcdec_debug_stat(self._h, name.encode()))

    def launch_count(self) -> int:
        return int(lib().rvcdec_last_launch_count(self._h))

    def profile_enable(self, on: bool) -> None:
        check(lib().rvcdec_profile_enable(self._h, 1 if on else 0), "rvcdec_profile_enable")

    def profile_read(self):
        arr = (ProfileEntry * 64)()
        n = lib().rvcdec_profile_read(self._h, arr, 64)
        check(n, "rvcdec_profile_read")
        return [dict(name=arr[i].name.decode(), ms=arr[i].ms, flops=arr[i].flops, bytes=arr[i].bytes, spans=arr[i].spans)
                for i in range(n)]

    def set_stop(self, stop: int) -> None:
        check(lib().rvcdec_debug_set_stop(self._h, stop), "rvcdec_debug_set_stop")

    def debug_read(self, name: str, dst, workspace, B, T, mode, stream) -> int:
        n = lib().rvcdec_debug_read(self._h, name.encode(), dst.data_ptr(), dst.numel(), workspace.data_ptr(), B, T, mode, stream)
        check(n, "rvcdec_debug_read")
        return int(n)

    def close(self) -> None:
        if self._h:
            lib().rvcdec_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class RmvpeHandle:
    """Owns one rvcrmvpe_handle (the RMVPE f0 predictor; include/rvcrmvpe.h)."""

    def __init__(self, device: int):
        self._h = C.c_void_p()
        self._check(lib().rvcrmvpe_create(int(device), C.byref(self._h)), "rvcrmvpe_create")
        self.device = int(device)

    @staticmethod
    def _check(rc: int, what: str) -> None:
        if rc < 0:
            raise RuntimeError(f"{what} failed ({rc}): {(lib().rvcrmvpe_last_error() or b'').decode()}")

    def load_weights(self, state: Dict[str, "object"]) -> None:
        arr, keep = Handle._tensor_array(None, state)
        self._check(lib().rvcrmvpe_load_weights(self._h, arr, len(state)), "rvcrmvpe_load_weights")

    def workspace_bytes(self, B: int, T: int) -> int:
        return int(lib().rvcrmvpe_workspace_bytes(self._h, B, T))

    def mel_workspace_bytes(self, B: int, n: int) -> int:
        return int(lib().rvcrmvpe_mel_workspace_bytes(self._h, B, n))

    def forward(self, mel, hidden, workspace, B, T, stream) -> None:
        self._check(lib().rvcrmvpe_forward(self._h, mel, hidden, workspace.data_ptr(), workspace.numel(), B, T, stream), "rvcrmvpe_forward")

    def mel(self, audio, mel, workspace, B, n, stream) -> None:
        self._check(lib().rvcrmvpe_mel(self._h, audio, mel, workspace.data_ptr(), workspace.numel(), B, n, stream), "rvcrmvpe_mel")

    def decode(self, hidden, thred, f0, T, stream) -> None:
        self._check(lib().rvcrmvpe_decode(self._h, hidden, C.c_float(thred), f0, T, stream), "rvcrmvpe_decode")

    def debug_read(self, name: str, dst, workspace, B: int, T: int, stream) -> int:
        n = lib().rvcrmvpe_debug_read(self._h, name.encode(), dst.data_ptr(), dst.numel(), workspace.data_ptr(), B, T, stream)
        self._check(n, "rvcrmvpe_debug_read")
        return int(n)

    def set_option(self, key: str, value: int) -> None:
        self._check(lib().rvcrmvpe_set_option(self._h, key.encode(), int(value)), "rvcrmvpe_set_option")

    def launch_count(self) -> int:
        return int(lib().rvcrmvpe_last_launch_count(self._h))

    def close(self) -> None:
        if self._h:
            lib().rvcrmvpe_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
