"""Drop-in replacement for the reference's NSF-HiFiGAN decoder.

``HiFiGANNSFGenerator`` mirrors the constructor, parameter names and
``forward(x, f0, g) -> waveform`` of
``rvc/lib/algorithm/generators/hifigan_nsf.py:55-235`` so that
``rvc/lib/algorithm/synthesizers.py:108-118`` can construct it and
``load_state_dict(cpt["weight"], strict=False)`` (``rvc/infer/infer.py:479-488``)
fills it unchanged.  The torch submodules below only HOLD parameters: every
flop of ``forward`` runs in librvcdec.so (hand-written sm_100a CUDA) behind the
torch custom op ``rvc_b200::nsf_hifigan_decode``.  There is no PyTorch or CPU
fallback; calling ``forward`` with CPU tensors raises.
"""
from __future__ import annotations

import math
import os
from typing import Optional

import torch
from torch.nn.utils.parametrizations import weight_norm
from torch.nn.utils import parametrize

from . import _lib

LRELU_SLOPE = 0.1  # residuals.py:10


def _get_padding(kernel_size: int, dilation: int = 1) -> int:  # commons.py:19-27
    return int((kernel_size * dilation - dilation) / 2)


def _burn_init_weights(m: torch.nn.Module) -> None:
    """The reference calls ``init_weights`` (commons.py:5-16) on weight-normed convs; on a
    parametrised module that only re-draws a temporary (``m.weight`` is computed), i.e. it has no
    effect on the parameters but DOES consume the global RNG.  Same draws here so that
    ``torch.manual_seed(s); HiFiGANNSFGenerator(...)`` yields bit-identical random-init weights."""
    if m.__class__.__name__.find("Conv") != -1:
        m.weight.data.normal_(0.0, 0.01)


class SourceModuleHnNSF(torch.nn.Module):
    """Parameter holder for hifigan_nsf.py:14-52 (``l_linear`` = Linear(1, 1))."""

    def __init__(self, sample_rate: int, harmonic_num: int = 0, sine_amp: float = 0.1, add_noise_std: float = 0.003,
                 voiced_threshod: float = 0):
        super().__init__()
        if harmonic_num != 0:
            raise ValueError("the NSF decoder uses harmonic_num=0 (hifigan_nsf.py:93)")
        self.sine_amp = sine_amp
        self.noise_std = add_noise_std
        self.sampling_rate = sample_rate
        self.l_linear = torch.nn.Linear(harmonic_num + 1, 1)
        self.l_tanh = torch.nn.Tanh()


class ResBlock(torch.nn.Module):
    """Parameter holder for residuals.py:34-84."""

    def __init__(self, channels: int, kernel_size: int = 3, dilations=(1, 3, 5)):
        super().__init__()
        self.convs1 = self._create_convs(channels, kernel_size, dilations)
        self.convs2 = self._create_convs(channels, kernel_size, [1] * len(dilations))

    @staticmethod
    def _create_convs(channels, kernel_size, dilations):
        layers = torch.nn.ModuleList([
            weight_norm(torch.nn.Conv1d(channels, channels, kernel_size, 1, dilation=d, padding=_get_padding(kernel_size, d)))
            for d in dilations
        ])
        layers.apply(_burn_init_weights)
        return layers


def _strip(module: torch.nn.Module) -> None:
    if parametrize.is_parametrized(module, "weight"):
        parametrize.remove_parametrizations(module, "weight", leave_parametrized=True)


class HiFiGANNSFGenerator(torch.nn.Module):
    """B200-native NSF-HiFiGAN generator with the reference's interface (hifigan_nsf.py:75-86).

    Extra, optional, keyword-only arguments of ``forward``:
      noise  (B, T*upp) or (B, T*upp, 1) standard-normal draws to inject; default = drawn with the torch
             generator of x's device in the reference's order (``rand(1,1,1)`` then ``randn(B, T*upp, 1)``,
             hifigan.py:189,223), so a seeded call reproduces the reference's noise on the same device.
             "philox" draws them inside the source kernel instead (no noise tensor in HBM).
      mode   "fp32" | "fast" | "fast_f16" | "fast_bf16"; default ``self.mode`` ("fast", or $RVC_B200_MODE).
    """

    def __init__(self, initial_channel: int, resblock_kernel_sizes: list, resblock_dilation_sizes: list,
                 upsample_rates: list, upsample_initial_channel: int, upsample_kernel_sizes: list, gin_channels: int,
                 sr: int, checkpointing: bool = False):
        super().__init__()
        self.num_kernels = len(resblock_kernel_sizes)
        self.num_upsamples = len(upsample_rates)
        self.checkpointing = checkpointing  # training-only in the reference; accepted and ignored
        self.f0_upsamp = torch.nn.Upsample(scale_factor=math.prod(upsample_rates))
        self.m_source = SourceModuleHnNSF(sample_rate=sr, harmonic_num=0)
        self.conv_pre = torch.nn.Conv1d(initial_channel, upsample_initial_channel, 7, 1, padding=3)
        self.ups = torch.nn.ModuleList()
        self.noise_convs = torch.nn.ModuleList()
        channels = [upsample_initial_channel // (2 ** (i + 1)) for i in range(len(upsample_rates))]
        stride_f0s = [math.prod(upsample_rates[i + 1:]) if i + 1 < len(upsample_rates) else 1
                      for i in range(len(upsample_rates))]
        for i, (u, k) in enumerate(zip(upsample_rates, upsample_kernel_sizes)):
            padding = (k - u) // 2 if u % 2 == 0 else u // 2 + u % 2  # hifigan_nsf.py:113-117
            self.ups.append(weight_norm(torch.nn.ConvTranspose1d(upsample_initial_channel // (2 ** i), channels[i], k, u,
                                                                 padding=padding, output_padding=u % 2)))
            stride = stride_f0s[i]
            kernel = 1 if stride == 1 else stride * 2 - stride % 2
            padding = 0 if stride == 1 else (kernel - stride) // 2
            self.noise_convs.append(torch.nn.Conv1d(1, channels[i], kernel_size=kernel, stride=stride, padding=padding))
        self.resblocks = torch.nn.ModuleList([
            ResBlock(channels[i], k, d)
            for i in range(len(self.ups))
            for k, d in zip(resblock_kernel_sizes, resblock_dilation_sizes)
        ])
        self.conv_post = torch.nn.Conv1d(channels[-1], 1, 7, 1, padding=3, bias=False)
        self.ups.apply(_burn_init_weights)
        if gin_channels != 0:
            self.cond = torch.nn.Conv1d(gin_channels, upsample_initial_channel, 1)
        self.upp = math.prod(upsample_rates)
        self.lrelu_slope = LRELU_SLOPE

        self.mode = os.environ.get("RVC_B200_MODE", "fast")
        self.noise_source = os.environ.get("RVC_B200_NOISE", "torch")  # "torch" | "philox"
        # scratch cap per forward; 0 = automatic: 70 % of the device memory that is free when the scratch is (re)allocated,
        # so a batch that fits runs in ONE pass (4 passes over 128 x 30 s chunks cost 2 %, profiles/r1am_*)
        self.max_workspace_bytes = int(os.environ.get("RVC_B200_MAX_WORKSPACE", "0"))
        self._cfg_args = (initial_channel, list(resblock_kernel_sizes), [list(d) for d in resblock_dilation_sizes],
                          list(upsample_rates), upsample_initial_channel, list(upsample_kernel_sizes), gin_channels, sr)
        self._gin = gin_channels
        self._cin = initial_channel
        # fast-mode tuning knob (rvcdec_set_option "fused_mrf"): None = library default
        self.fused_mrf: Optional[int] = int(os.environ["RVC_B200_FUSED_MRF"]) if "RVC_B200_FUSED_MRF" in os.environ else None
        self._applied_fused = None
        self._handle: Optional[_lib.Handle] = None
        self._packed_key = None
        self._workspace: Optional[torch.Tensor] = None
        self._ws_event = None
        self._ws_stream = None
        self.last_launch_count = 0

    # -- reference protocol ------------------------------------------------
    def remove_weight_norm(self):
        """Fold weight-norm into plain weights (the reference's version raises on parametrised
        modules, hifigan_nsf.py:214-218; rvc/realtime/pipeline.py:405-415 shows the intent)."""
        for l in self.ups:
            _strip(l)
        for rb in self.resblocks:
            for l in list(rb.convs1) + list(rb.convs2):
                _strip(l)

    def __prepare_scriptable__(self):
        return self

    # -- packed weights ----------------------------------------------------
    def _decoder_state(self):
        """The tensors rvcdec_load_weights receives (reference state_dict names)."""
        return self.state_dict()

    def _weights_key(self):
        return tuple((p.data_ptr(), p._version, p.device.index) for p in self.parameters())

    def set_content_key(self, key: Optional[str]) -> None:
        """Content hash of the current weights (loader.content_hash): handles that hold the packed images of these
        weights are shared through ``loader.PACKED``.  The key is dropped as soon as a parameter is modified in place."""
        self.content_key = key
        self._content_versions = tuple(p._version for p in self.parameters())

    def _cache_key(self, idx: int):
        ck = getattr(self, "content_key", None)
        if ck is None or getattr(self, "_content_versions", None) != tuple(p._version for p in self.parameters()):
            return None
        return (ck, idx)

    def _ensure_handle(self, device: torch.device) -> _lib.Handle:
        idx = device.index if device.index is not None else torch.cuda.current_device()
        if self._handle is not None and self._handle.device != idx:
            self._handle = None  # (closed by its last reference: the packed-weight cache may share it)
        if self._handle is None:
            ckey = self._cache_key(idx)
            if ckey is not None:
                from . import loader

                cached = loader.PACKED.get(ckey)
                if cached is not None:  # folded + repacked + uploaded once already for these very weights
                    self._handle = cached
                    self._packed_key = self._weights_key()
                    self._applied_fused = None
        if self._handle is None:
            self._handle = _lib.Handle(_lib.make_config(*self._cfg_args), idx)
            self._packed_key = None
            self._applied_fused = None
            # tuning knobs for A/B runs: RVC_B200_OPTS="fused_epi_warps=8,fused_post=0" -> rvcdec_set_option
            for kv in filter(None, os.environ.get("RVC_B200_OPTS", "").split(",")):
                k, v = kv.split("=")
                self._handle.set_option(k.strip(), int(v))
        if self.fused_mrf is not None and self.fused_mrf != self._applied_fused:
            self._handle.set_option("fused_mrf", self.fused_mrf)
            self._applied_fused = self.fused_mrf
        key = self._weights_key()
        if key != self._packed_key:
            ckey = self._cache_key(idx)
            if ckey is None and getattr(self, "content_key", None) is not None:
                # weights changed after load_checkpoint: this module needs its own handle from now on
                from . import loader

                if any(v is self._handle for v in loader.PACKED._d.values()):
                    self._handle = _lib.Handle(_lib.make_config(*self._cfg_args), idx)
                    self._applied_fused = None
                self.content_key = None
            self._handle.load_weights(self._decoder_state())
            self._packed_key = key
            if ckey is not None:
                from . import loader

                loader.PACKED.put(ckey, self._handle)
        return self._handle

    def _get_workspace(self, nbytes: int, device: torch.device) -> torch.Tensor:
        ws = self._workspace
        if ws is None or ws.device != device or ws.numel() < nbytes:
            self._workspace = None
            ws = torch.empty(nbytes, dtype=torch.uint8, device=device)
            self._workspace = ws
        return ws

    # -- forward -----------------------------------------------------------
    def _draw_seed(self) -> int:
        """One 63-bit Philox key per call from the torch CPU generator (so ``torch.manual_seed`` reproduces a decode)."""
        return int(torch.randint(0, 2 ** 62, (1,)).item())

    def forward(self, x: torch.Tensor, f0: torch.Tensor, g: Optional[torch.Tensor] = None, *,
                noise: Optional[torch.Tensor] = None, mode: Optional[str] = None):
        from . import ops  # registers the custom op

        if not x.is_cuda:
            raise RuntimeError("rvc_b200.HiFiGANNSFGenerator runs only on a CUDA (sm_100a) device; there is no CPU fallback")
        if torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()) and self.training:
            raise RuntimeError("the B200 decoder is inference-only (call .eval() / use torch.no_grad())")
        B, Cx, T = x.shape
        if Cx != self._cin or f0.shape != (B, T):
            raise ValueError(f"expected x (B,{self._cin},T) and f0 (B,T); got {tuple(x.shape)} and {tuple(f0.shape)}")
        m = _lib.MODES[mode or self.mode]
        h = self._ensure_handle(x.device)
        n = T * self.upp
        seed = 0
        if noise is None:
            noise = self.noise_source
        if isinstance(noise, str):
            if noise == "torch":
                # reference draw order on this device's generator: hifigan.py:189 then :223
                torch.rand(1, 1, 1, device=x.device)
                noise = torch.randn(B, n, 1, device=x.device, dtype=torch.float32)
            elif noise == "philox":
                # in-kernel draws: one 63-bit key from the torch CPU generator per call
                noise = None
                seed = self._draw_seed()
            else:
                raise ValueError("noise must be a tensor, 'torch' or 'philox'")
        if noise is not None:
            noise = noise.reshape(B, -1).contiguous().float()  # (B, n), or (B, n * harmonics) for the MRF sibling
        if g is not None:
            if self._gin == 0:
                g = None
            else:
                g = g.reshape(B, self._gin).contiguous().float()
        need = h.workspace_bytes(B, T, m)
        cap = self.max_workspace_bytes
        if cap <= 0:
            have = self._workspace.numel() if (self._workspace is not None and self._workspace.device == x.device) else 0
            cap = need if have >= need else int(0.7 * (torch.cuda.mem_get_info(x.device)[0] + have))
        ws = self._get_workspace(min(need, max(cap, h.workspace_bytes(1, T, m))), x.device)
        # the cached scratch is shared by every forward of this module: a call on another stream must wait for the
        # previous call's kernels before it overwrites it
        cur = torch.cuda.current_stream(x.device)
        if self._ws_event is not None and self._ws_stream != cur.cuda_stream:
            cur.wait_event(self._ws_event)
        out = ops.nsf_hifigan_decode(x.contiguous().float(), f0.contiguous().float(), g, noise, ws, h.ptr.value, m, self.upp, seed)
        if self._ws_event is None:
            self._ws_event = torch.cuda.Event()
        self._ws_event.record(cur)
        self._ws_stream = cur.cuda_stream
        self.last_launch_count = h.launch_count()
        return out.unsqueeze(1)

    def decode_host(self, x: torch.Tensor, f0: torch.Tensor, g: Optional[torch.Tensor] = None, *,
                    mode: Optional[str] = None, out: Optional[torch.Tensor] = None, device: Optional[int] = None,
                    max_device_bytes: int = 0) -> torch.Tensor:
        """Host-buffer entry point (``rvcdec_forward_host``): x, f0, g are CPU tensors (pinned memory
        recommended); copies in, decodes with in-kernel noise, copies the waveform back into ``out``
        (CPU, (B,1,T*upp)) and returns it.  This is the per-chunk call of the reference's pipeline
        (rvc/infer/pipeline.py:366-371) with the host<->device hops inside."""
        if x.is_cuda or f0.is_cuda:
            raise ValueError("decode_host takes CPU tensors; use forward() for device tensors")
        B, Cx, T = x.shape
        m = _lib.MODES[mode or self.mode]
        dev = torch.device("cuda", torch.cuda.current_device() if device is None else device)
        h = self._ensure_handle(dev)
        x = x.contiguous().float()
        f0 = f0.contiguous().float()
        gp = None
        if g is not None and self._gin:
            g = g.reshape(B, self._gin).contiguous().float()
            gp = g.data_ptr()
        n = T * self.upp
        if Cx != self._cin or tuple(f0.shape) != (B, T):
            raise ValueError(f"expected x (B,{self._cin},T) and f0 (B,T); got {tuple(x.shape)} and {tuple(f0.shape)}")
        if out is None:
            out = torch.empty(B, 1, n, dtype=torch.float32, pin_memory=True)
        elif out.is_cuda or out.dtype != torch.float32 or not out.is_contiguous() or out.numel() != B * n:
            # the library DMAs B*T*upp floats into this buffer: anything else would be overrun
            raise ValueError(f"out must be a contiguous CPU float32 tensor of {B * n} elements (B*T*upp); got "
                             f"{tuple(out.shape)} {out.dtype} on {out.device}")
        seed = self._draw_seed()
        h.forward_host(x.data_ptr(), f0.data_ptr(), gp, None, seed, out.data_ptr(), B, T, m, max_device_bytes)
        self.last_launch_count = h.launch_count()
        return out


class HiFiGANGenerator(HiFiGANNSFGenerator):
    """B200-native sibling decoder of the no-f0 models: ``HiFiGANGenerator``
    (rvc/lib/algorithm/generators/hifigan.py:11-112, constructed at synthesizers.py:127-135) with the
    reference's constructor, parameter names and ``forward(x, g) -> waveform``.

    It is the NSF decoder without the harmonic source and the noise convs, and runs on the same kernels
    through the same C ABI: the library is handed zero ``noise_convs`` / ``m_source`` tensors, so every
    ``x + noise_convs[i](har)`` would add exactly 0, and the handle option ``no_source`` makes it skip the
    source kernel and those additions altogether (hifigan.py:72-97).  Sub-modules are created in the reference's order, so a seeded default
    init is bit-identical to the reference's."""

    def __init__(self, initial_channel: int, resblock_kernel_sizes: list, resblock_dilation_sizes: list,
                 upsample_rates: list, upsample_initial_channel: int, upsample_kernel_sizes: list, gin_channels: int = 0):
        torch.nn.Module.__init__(self)
        self.num_kernels = len(resblock_kernel_sizes)
        self.num_upsamples = len(upsample_rates)
        self.conv_pre = torch.nn.Conv1d(initial_channel, upsample_initial_channel, 7, 1, padding=3)
        self.ups = torch.nn.ModuleList()
        self.resblocks = torch.nn.ModuleList()
        ch = upsample_initial_channel
        for i, (u, k) in enumerate(zip(upsample_rates, upsample_kernel_sizes)):
            if u % 2 != 0:
                # hifigan.py:56 pads (k-u)//2 whatever the rate: an odd rate makes every stage one sample longer than u x its
                # input, which no released configuration uses
                raise ValueError("odd upsample rates are not supported by the no-f0 decoder")
            ch = upsample_initial_channel // (2 ** (i + 1))
            self.ups.append(weight_norm(torch.nn.ConvTranspose1d(upsample_initial_channel // (2 ** i), ch, k, u, padding=(k - u) // 2)))
            for kk, d in zip(resblock_kernel_sizes, resblock_dilation_sizes):
                self.resblocks.append(ResBlock(ch, kk, d))
        self.conv_post = torch.nn.Conv1d(ch, 1, 7, 1, padding=3, bias=False)
        self.ups.apply(_burn_init_weights)
        if gin_channels != 0:
            self.cond = torch.nn.Conv1d(gin_channels, upsample_initial_channel, 1)
        self.upp = math.prod(upsample_rates)
        self.lrelu_slope = LRELU_SLOPE
        self.checkpointing = False
        self.mode = os.environ.get("RVC_B200_MODE", "fast")
        self.noise_source = "philox"  # never used: the source output is multiplied by zero weights
        self.max_workspace_bytes = int(os.environ.get("RVC_B200_MAX_WORKSPACE", "0"))
        # the library's config carries a sample rate for the (unused) source; any valid value does
        self._cfg_args = (initial_channel, list(resblock_kernel_sizes), [list(d) for d in resblock_dilation_sizes],
                          list(upsample_rates), upsample_initial_channel, list(upsample_kernel_sizes), gin_channels, 48000)
        self._gin = gin_channels
        self._cin = initial_channel
        self._rates = list(upsample_rates)
        self._chs = [upsample_initial_channel // (2 ** (i + 1)) for i in range(len(upsample_rates))]
        self.fused_mrf = int(os.environ["RVC_B200_FUSED_MRF"]) if "RVC_B200_FUSED_MRF" in os.environ else None
        self._applied_fused = None
        self._handle = None
        self._packed_key = None
        self._workspace = None
        self._ws_event = None
        self._ws_stream = None
        self.last_launch_count = 0

    def _draw_seed(self) -> int:
        # the reference's no-f0 decoder draws nothing (hifigan.py:72-97): leave the caller's RNG stream alone
        return 0

    def _decoder_state(self):
        sd = dict(self.state_dict())
        sd["m_source.l_linear.weight"] = torch.zeros(1, 1)
        sd["m_source.l_linear.bias"] = torch.zeros(1)
        for i, C in enumerate(self._chs):
            stride = math.prod(self._rates[i + 1:]) if i + 1 < len(self._rates) else 1
            kernel = 1 if stride == 1 else stride * 2 - stride % 2
            sd[f"noise_convs.{i}.weight"] = torch.zeros(C, 1, kernel)
            sd[f"noise_convs.{i}.bias"] = torch.zeros(C)
        return sd

    def _ensure_handle(self, device: torch.device) -> _lib.Handle:
        fresh = self._handle is None or self._handle.device != (device.index if device.index is not None else torch.cuda.current_device())
        h = HiFiGANNSFGenerator._ensure_handle(self, device)
        if fresh:
            h.set_option("no_source", 1)  # skip the source kernel and the noise adds (they would add exactly 0)
        return h

    def forward(self, x: torch.Tensor, g: Optional[torch.Tensor] = None, *, mode: Optional[str] = None):
        f0 = torch.zeros(x.shape[0], x.shape[2], device=x.device, dtype=torch.float32)
        return HiFiGANNSFGenerator.forward(self, x, f0, g, noise="philox", mode=mode)

    def decode_host(self, x: torch.Tensor, g: Optional[torch.Tensor] = None, *, mode: Optional[str] = None,
                    out: Optional[torch.Tensor] = None, device: Optional[int] = None, max_device_bytes: int = 0) -> torch.Tensor:
        f0 = torch.zeros(x.shape[0], x.shape[2], dtype=torch.float32)
        return HiFiGANNSFGenerator.decode_host(self, x, f0, g, mode=mode, out=out, device=device, max_device_bytes=max_device_bytes)


class _MRFLayer(torch.nn.Module):  # hifigan_mrf.py:13-57
    def __init__(self, channels, kernel_size, dilation):
        super().__init__()
        self.conv1 = weight_norm(torch.nn.Conv1d(channels, channels, kernel_size, padding=(kernel_size * dilation - dilation) // 2,
                                                 dilation=dilation))
        self.conv2 = weight_norm(torch.nn.Conv1d(channels, channels, kernel_size, padding=kernel_size // 2, dilation=1))


class _MRFBlock(torch.nn.Module):  # hifigan_mrf.py:60-85
    def __init__(self, channels, kernel_size, dilations):
        super().__init__()
        self.layers = torch.nn.ModuleList([_MRFLayer(channels, kernel_size, d) for d in dilations])


class _MRFSource(torch.nn.Module):  # hifigan_mrf.py:179-220 (l_sin_gen has no parameters)
    def __init__(self, sampling_rate: int, harmonic_num: int = 0):
        super().__init__()
        self.sampling_rate, self.harmonic_num = sampling_rate, harmonic_num
        self.l_linear = torch.nn.Linear(harmonic_num + 1, 1)
        self.l_tanh = torch.nn.Tanh()


class HiFiGANMRFGenerator(HiFiGANNSFGenerator):
    """B200-native sibling decoder ``HiFiGANMRFGenerator`` (rvc/lib/algorithm/generators/hifigan_mrf.py:223-374, the
    "MRF HiFi-GAN" vocoder selected at synthesizers.py:85-97) with the reference's keyword constructor, parameter names
    (``upsamples``, ``mrfs.i.j.layers.m.conv1``, weight-normed ``conv_pre`` / ``conv_post`` with bias) and
    ``forward(x, f0, g)``.  It is the NSF conv stack with a 9-sine harmonic source (overtones with random initial phases
    merged by Linear(9, 1)), so it runs on the same kernels behind the same C ABI: ``rvcdec_config.harmonic_num`` selects
    the source, the state dict is handed to the library under the NSF names.

    The reference integrates the phase with a per-SAMPLE fp32 cumsum (hifigan_mrf.py:137-142); f0 is constant inside a
    frame, so the source kernel keeps the closed form it already used - a frame-level fp64 prefix plus (j+1) f0/sr inside
    the frame, times (k+1) per overtone - which equals the reference's sum in exact arithmetic.

    Extra keyword-only arguments of ``forward``: ``noise`` (B, T*upp, H) standard-normal draws and ``rand_ini`` (B, H)
    initial phases in [0, 1) (column 0 ignored: the fundamental starts at 0); default = drawn with the torch generator
    of x's device in the reference's order (``rand(B, H)`` then ``randn(B, T*upp, H)``); "philox" = in-kernel draws."""

    def __init__(self, in_channel: int, upsample_initial_channel: int, upsample_rates: list, upsample_kernel_sizes: list,
                 resblock_kernel_sizes: list, resblock_dilations: list, gin_channels: int, sample_rate: int, harmonic_num: int,
                 checkpointing: bool = False):
        torch.nn.Module.__init__(self)
        self.num_kernels = len(resblock_kernel_sizes)
        self.num_upsamples = len(upsample_rates)
        self.checkpointing = checkpointing
        self.f0_upsample = torch.nn.Upsample(scale_factor=math.prod(upsample_rates))
        self.m_source = _MRFSource(sample_rate, harmonic_num)
        self.conv_pre = weight_norm(torch.nn.Conv1d(in_channel, upsample_initial_channel, kernel_size=7, stride=1, padding=3))
        self.upsamples = torch.nn.ModuleList()
        self.noise_convs = torch.nn.ModuleList()
        stride_f0s = [math.prod(upsample_rates[i + 1:]) if i + 1 < len(upsample_rates) else 1 for i in range(len(upsample_rates))]
        for i, (u, k) in enumerate(zip(upsample_rates, upsample_kernel_sizes)):
            padding = (k - u) // 2 if u % 2 == 0 else u // 2 + u % 2  # hifigan_mrf.py:279-284
            self.upsamples.append(weight_norm(torch.nn.ConvTranspose1d(upsample_initial_channel // (2 ** i),
                                                                       upsample_initial_channel // (2 ** (i + 1)), kernel_size=k,
                                                                       stride=u, padding=padding, output_padding=u % 2)))
            stride = stride_f0s[i]
            kernel = 1 if stride == 1 else stride * 2 - stride % 2
            padding = 0 if stride == 1 else (kernel - stride) // 2
            self.noise_convs.append(torch.nn.Conv1d(1, upsample_initial_channel // (2 ** (i + 1)), kernel_size=kernel, stride=stride,
                                                    padding=padding))
        self.mrfs = torch.nn.ModuleList()
        channel = upsample_initial_channel
        for i in range(len(self.upsamples)):
            channel = upsample_initial_channel // (2 ** (i + 1))
            self.mrfs.append(torch.nn.ModuleList([_MRFBlock(channel, kernel_size=k, dilations=d)
                                                  for k, d in zip(resblock_kernel_sizes, resblock_dilations)]))
        self.conv_post = weight_norm(torch.nn.Conv1d(channel, 1, kernel_size=7, stride=1, padding=3))
        if gin_channels != 0:
            self.cond = torch.nn.Conv1d(gin_channels, upsample_initial_channel, 1)
        self.upp = math.prod(upsample_rates)
        self.lrelu_slope = LRELU_SLOPE
        self.harmonics = harmonic_num + 1
        self.mode = os.environ.get("RVC_B200_MODE", "fast")
        self.noise_source = os.environ.get("RVC_B200_NOISE", "torch")
        self.max_workspace_bytes = int(os.environ.get("RVC_B200_MAX_WORKSPACE", "0"))
        self._cfg_args = (in_channel, list(resblock_kernel_sizes), [list(d) for d in resblock_dilations], list(upsample_rates),
                          upsample_initial_channel, list(upsample_kernel_sizes), gin_channels, sample_rate, harmonic_num)
        self._gin = gin_channels
        self._cin = in_channel
        self.fused_mrf = int(os.environ["RVC_B200_FUSED_MRF"]) if "RVC_B200_FUSED_MRF" in os.environ else None
        self._applied_fused = None
        self._handle = None
        self._packed_key = None
        self._workspace = None
        self._ws_event = None
        self._ws_stream = None
        self.last_launch_count = 0

    def remove_weight_norm(self):
        for l in [self.conv_pre, self.conv_post] + list(self.upsamples):
            _strip(l)
        for stage in self.mrfs:
            for blk in stage:
                for lay in blk.layers:
                    _strip(lay.conv1)
                    _strip(lay.conv2)

    def _decoder_state(self):
        """The library's (NSF) names: upsamples.i -> ups.i, mrfs.i.j.layers.m.conv1/2 -> resblocks.(i*nk+j).convs1/2.m."""
        out = {}
        nk = self.num_kernels
        for k, v in self.state_dict().items():
            if k.startswith("upsamples."):
                k = "ups." + k[len("upsamples."):]
            elif k.startswith("mrfs."):
                p = k.split(".")
                i, j, m, which = int(p[1]), int(p[2]), int(p[4]), p[5]
                k = f"resblocks.{i * nk + j}.convs{1 if which == 'conv1' else 2}.{m}." + ".".join(p[6:])
            out[k] = v
        return out

    def forward(self, x: torch.Tensor, f0: torch.Tensor, g: Optional[torch.Tensor] = None, *,
                noise=None, rand_ini: Optional[torch.Tensor] = None, mode: Optional[str] = None):
        if not x.is_cuda:
            raise RuntimeError("rvc_b200.HiFiGANMRFGenerator runs only on a CUDA (sm_100a) device; there is no CPU fallback")
        B, _, T = x.shape
        n, H = T * self.upp, self.harmonics
        if noise is None:
            noise = self.noise_source
        if isinstance(noise, str) and noise == "torch":
            # reference draw order on this device's generator: hifigan_mrf.py:129 then :172
            rand_ini = torch.rand(B, H, device=x.device)
            noise = torch.randn(B, n, H, device=x.device, dtype=torch.float32)
        h = self._ensure_handle(x.device)
        keep = None
        if isinstance(noise, torch.Tensor):
            if rand_ini is None:
                raise ValueError("injected noise needs rand_ini (B, harmonics) as well")
            keep = rand_ini.reshape(B, H).contiguous().float().clone()
            keep[:, 0] = 0
            h.set_initial_phases(keep.data_ptr())
            noise = noise.reshape(B, n * H)
        try:
            return HiFiGANNSFGenerator.forward(self, x, f0, g, noise=noise, mode=mode)
        finally:
            h.set_initial_phases(None)
