"""Drop-in for the reference's ``RefineGANGenerator`` (rvc/lib/algorithm/generators/refinegan.py:268-451, built by
``Synthesizer`` for vocoder "RefineGAN", synthesizers.py:98-106): same keyword constructor, parameter names and
``forward(mel, f0, g) -> waveform``; every flop runs in librvcdec.so behind ``include/rvcrefine.h`` (fp32).

The torch sub-modules only HOLD parameters; they are created in the reference's order with the reference's
initialisers (``init_weights`` re-draws plain convolutions and only burns RNG on weight-normed ones, commons.py:5-16), so
a seeded construction yields the reference's state dict bit for bit.
"""
from __future__ import annotations

import math
from typing import Optional

import torch
from torch.nn.utils.parametrizations import weight_norm

from . import _lib

LPW, ROLLOFF, BETA = 64, 0.9475937167399596, 14.769656459379492  # refinegan.py:429-433


def sinc_kaiser_taps(rate: int) -> torch.Tensor:
    """The filter ``torchaudio.functional.resample(x, rate*k, k, lowpass_filter_width=64, rolloff=0.9476,
    resampling_method="sinc_interp_kaiser", beta=14.77)`` applies (refinegan.py:424-433): with new_freq reduced to 1 it is
    ONE FIR of 2*width + rate taps used at stride ``rate`` on the input zero-padded by ``width`` in front.  Same
    sequence of fp32 operations as torchaudio's ``_get_sinc_resample_kernel`` for an fp32 waveform."""
    orig = rate
    base = 1 * ROLLOFF  # min(orig, new) * rolloff with new = 1
    width = math.ceil(LPW * orig / base)
    idx = torch.arange(-width, width + orig, dtype=torch.float32)[None, None] / orig
    t = torch.arange(0, -1, -1, dtype=torch.float32)[:, None, None] / 1 + idx
    t *= base
    t = t.clamp_(-LPW, LPW)
    beta = torch.tensor(float(BETA))
    window = torch.i0(beta * torch.sqrt(1 - (t / LPW) ** 2)) / torch.i0(beta)
    t *= math.pi
    k = torch.where(t == 0, torch.tensor(1.0).to(t), t.sin() / t)
    k *= window * (base / orig)
    return k.reshape(-1).float()


def _init_weights(m: torch.nn.Module) -> None:  # commons.py:5-16
    if m.__class__.__name__.find("Conv") != -1:
        m.weight.data.normal_(0.0, 0.01)


class _ResBlock(torch.nn.Module):  # refinegan.py:28-75
    def __init__(self, channels: int, kernel_size: int = 7, dilation=(1, 3, 5), leaky_relu_slope: float = 0.2):
        super().__init__()
        self.leaky_relu_slope = leaky_relu_slope
        self.convs1 = torch.nn.ModuleList([
            weight_norm(torch.nn.Conv1d(channels, channels, kernel_size, stride=1, dilation=d, padding=int((kernel_size * d - d) / 2)))
            for d in dilation])
        self.convs1.apply(_init_weights)
        self.convs2 = torch.nn.ModuleList([
            weight_norm(torch.nn.Conv1d(channels, channels, kernel_size, stride=1, dilation=1, padding=int((kernel_size - 1) / 2)))
            for _ in dilation])
        self.convs2.apply(_init_weights)


class _AdaIN(torch.nn.Module):  # refinegan.py:88-100
    def __init__(self, *, channels: int, leaky_relu_slope: float = 0.2):
        super().__init__()
        self.weight = torch.nn.Parameter(torch.ones(channels) * 1e-4)
        self.activation = torch.nn.LeakyReLU(leaky_relu_slope)


class _ParallelResBlock(torch.nn.Module):  # refinegan.py:103-160
    def __init__(self, *, in_channels: int, out_channels: int, kernel_sizes=(3, 7, 11), dilation=(1, 3, 5), leaky_relu_slope: float = 0.2):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.input_conv = torch.nn.Conv1d(in_channels=in_channels, out_channels=out_channels, kernel_size=7, stride=1, padding=3)
        self.input_conv.apply(_init_weights)
        self.blocks = torch.nn.ModuleList([
            torch.nn.Sequential(_AdaIN(channels=out_channels),
                                _ResBlock(out_channels, kernel_size=k, dilation=dilation, leaky_relu_slope=leaky_relu_slope),
                                _AdaIN(channels=out_channels))
            for k in kernel_sizes])


class _SineGenerator(torch.nn.Module):  # refinegan.py:180-216
    def __init__(self, samp_rate, harmonic_num=0):
        super().__init__()
        self.sampling_rate, self.harmonic_num, self.dim = samp_rate, harmonic_num, harmonic_num + 1
        self.merge = torch.nn.Sequential(torch.nn.Linear(self.dim, 1, bias=False), torch.nn.Tanh())


class RefineGANGenerator(torch.nn.Module):
    """B200-native RefineGAN generator.  Extra keyword-only arguments of ``forward``: ``noise`` = a flat fp32 tensor with
    every standard-normal draw of the reference in its order (source (B, T*upp), then per stage and parallel block the two
    AdaIN draws (B, C, L); ``noise_numel(B, T)`` gives the size), or "philox" for in-kernel draws; default = drawn with
    the torch generator of mel's device in the reference's order (``rand(B, 1)`` first, refinegan.py:226)."""

    def __init__(self, *, sample_rate: int = 44100, downsample_rates=(2, 2, 8, 8), upsample_rates=(8, 8, 2, 2), leaky_relu_slope: float = 0.2,
                 num_mels: int = 128, start_channels: int = 16, gin_channels: int = 256, checkpointing: bool = False,
                 upsample_initial_channel=512):
        super().__init__()
        self.upsample_rates = list(upsample_rates)
        self.leaky_relu_slope = leaky_relu_slope
        self.checkpointing = checkpointing
        self.upp = int(math.prod(upsample_rates))
        self.m_source = _SineGenerator(sample_rate)
        self.pre_conv = weight_norm(torch.nn.Conv1d(1, 16, 7, 1, padding=3))
        channels = start_channels
        self.downsample_blocks = torch.nn.ModuleList([])
        for _ in upsample_rates:
            self.downsample_blocks.append(weight_norm(torch.nn.Conv1d(channels, channels * 2, 7, 1, padding=3)))
            channels *= 2
        channels = upsample_initial_channel
        self.mel_conv = weight_norm(torch.nn.Conv1d(num_mels, channels // 2, 7, 1, padding=3))
        self.mel_conv.apply(_init_weights)
        if gin_channels != 0:
            self.cond = torch.nn.Conv1d(256, channels // 2, 1)
        self.upsample_blocks = torch.nn.ModuleList([])
        self.upsample_conv_blocks = torch.nn.ModuleList([])
        for rate in upsample_rates:
            new_channels = channels // 2
            self.upsample_blocks.append(torch.nn.Upsample(scale_factor=rate, mode="linear"))
            self.upsample_conv_blocks.append(_ParallelResBlock(in_channels=channels + channels // 4, out_channels=new_channels,
                                                               kernel_sizes=(3, 7, 11), dilation=(1, 3, 5), leaky_relu_slope=leaky_relu_slope))
            channels = new_channels
        self.conv_post = weight_norm(torch.nn.Conv1d(channels, 1, 7, 1, padding=3, bias=False))
        self.conv_post.apply(_init_weights)
        if abs(leaky_relu_slope - 0.2) > 1e-12:
            raise ValueError("the B200 RefineGAN decoder is built for leaky_relu_slope = 0.2 (the reference's only value)")
        self._cfg = dict(sample_rate=sample_rate, rates=list(upsample_rates), num_mels=num_mels, start_channels=start_channels,
                         upsample_initial_channel=upsample_initial_channel, gin_channels=256 if gin_channels != 0 else 0)
        self._gin = self._cfg["gin_channels"]
        self.noise_source = "torch"
        self._handle: Optional[_lib.RefineHandle] = None
        self._packed_key = None
        self._workspace: Optional[torch.Tensor] = None
        self.last_launch_count = 0

    def remove_weight_norm(self):
        from torch.nn.utils import parametrize

        for m in list(self.modules()):
            if parametrize.is_parametrized(m, "weight"):
                parametrize.remove_parametrizations(m, "weight", leave_parametrized=True)

    def _library_state(self):
        sd = dict(self.state_dict())
        n = len(self.upsample_rates)
        for i in range(n):
            sd[f"resample_taps.{i}"] = sinc_kaiser_taps(self.upsample_rates[n - 1 - i])
        return sd

    def _ensure_handle(self, device: torch.device) -> _lib.RefineHandle:
        idx = device.index if device.index is not None else torch.cuda.current_device()
        if self._handle is not None and self._handle.device != idx:
            self._handle = None
        if self._handle is None:
            c = _lib.RefineConfig()
            c.sample_rate = self._cfg["sample_rate"]
            c.num_upsamples = len(self._cfg["rates"])
            for i, r in enumerate(self._cfg["rates"]):
                c.upsample_rates[i] = r
            c.num_mels = self._cfg["num_mels"]
            c.start_channels = self._cfg["start_channels"]
            c.upsample_initial_channel = self._cfg["upsample_initial_channel"]
            c.gin_channels = self._cfg["gin_channels"]
            self._handle = _lib.RefineHandle(c, idx)
            self._packed_key = None
        key = tuple((p.data_ptr(), p._version, p.device.index) for p in self.parameters())
        if key != self._packed_key:
            self._handle.load_weights(self._library_state())
            self._packed_key = key
        return self._handle

    def noise_numel(self, B: int, T: int) -> int:
        n = B * T * self.upp
        L, ch = T, self._cfg["upsample_initial_channel"]
        for r in self.upsample_rates:
            L *= r
            ch //= 2
            n += 6 * B * ch * L
        return n

    def draw_noise(self, B: int, T: int, device, generator=None) -> torch.Tensor:
        """The reference's draws in its order (refinegan.py:226, :257, :98-99) as one flat tensor."""
        torch.rand(B, 1, device=device, generator=generator)  # rand_ini, zeroed for the fundamental
        parts = [torch.randn(B, T * self.upp, 1, device=device, generator=generator).reshape(-1)]
        L, ch = T, self._cfg["upsample_initial_channel"]
        for r in self.upsample_rates:
            L *= r
            ch //= 2
            for _ in range(6):
                parts.append(torch.randn(B, ch, L, device=device, generator=generator).reshape(-1))
        return torch.cat(parts)

    @torch.no_grad()
    def forward(self, mel: torch.Tensor, f0: torch.Tensor, g: Optional[torch.Tensor] = None, *, noise=None):
        if not mel.is_cuda:
            raise RuntimeError("rvc_b200.RefineGANGenerator runs only on a CUDA (sm_100a) device; there is no CPU fallback")
        B, C, T = mel.shape
        if C != self._cfg["num_mels"] or tuple(f0.shape) != (B, T):
            raise ValueError(f"expected mel (B,{self._cfg['num_mels']},T) and f0 (B,T); got {tuple(mel.shape)} and {tuple(f0.shape)}")
        dev = mel.device
        h = self._ensure_handle(dev)
        seed = 0
        if noise is None:
            noise = self.noise_source
        if isinstance(noise, str):
            if noise == "torch":
                noise = self.draw_noise(B, T, dev)
            elif noise == "philox":
                noise = None
                seed = int(torch.randint(0, 2 ** 62, (1,)).item())
            else:
                raise ValueError("noise must be a tensor, 'torch' or 'philox'")
        if noise is not None:
            noise = noise.reshape(-1).contiguous().float()
            if noise.numel() != self.noise_numel(B, T):
                raise ValueError(f"noise must hold {self.noise_numel(B, T)} draws, got {noise.numel()}")
        gp = None
        if g is not None and self._gin:
            g = g.reshape(B, self._gin).contiguous().float()
            gp = g.data_ptr()
        need = h.workspace_bytes(B, T)
        if self._workspace is None or self._workspace.device != dev or self._workspace.numel() < need:
            self._workspace = None
            self._workspace = torch.empty(need, dtype=torch.uint8, device=dev)
        out = torch.empty(B, T * self.upp, dtype=torch.float32, device=dev)
        mel, f0 = mel.contiguous().float(), f0.contiguous().float()
        h.forward(mel.data_ptr(), f0.data_ptr(), gp, noise.data_ptr() if noise is not None else None, seed, out.data_ptr(),
                  self._workspace, B, T, torch.cuda.current_stream(dev).cuda_stream)
        self.last_launch_count = h.launch_count()
        return out.unsqueeze(1)
