"""CPU oracle for the NSF-HiFiGAN decoder hot path.  TEST INFRASTRUCTURE ONLY.

This file is a CPU restatement (torch CPU tensors, fp32 or fp64 arithmetic) of
the reference algorithm for the one hot path this repo accelerates:
``HiFiGANNSFGenerator.forward`` and everything it calls.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl
reference`` legs may import it; the product (``rvc_b200``) never does.

Parity status: PINNED.  The reference has no golden vectors for this path
(SURVEY.md section 8c), so the oracle is pinned against outputs of the
reference module itself, run in the build container by
``oracle/make_golden.py`` (which imports ``/root/reference``) and committed
under ``tests/golden/``; ``tests/test_oracle.py`` re-checks the oracle against
those fixtures on every run and, when ``/root/reference`` is present, against
the live reference module as well.

Reference files restated (paths relative to the reference repo root):
  rvc/lib/algorithm/generators/hifigan_nsf.py   (SourceModuleHnNSF :14-52, HiFiGANNSFGenerator :55-212)
  rvc/lib/algorithm/generators/hifigan.py       (SineGenerator :115-228)
  rvc/lib/algorithm/residuals.py                (ResBlock :34-84, LRELU_SLOPE :10)
  rvc/lib/algorithm/commons.py                  (get_padding :19-27)
The arithmetic itself (conv1d / conv_transpose1d / sin / tanh) lives in the
third-party tensor runtime the reference calls (PyTorch ATen, pinned by the
reference at torch==2.4.0, pyproject.toml:39; torch 2.11 in this image); the
oracle calls the same published operators on CPU.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn.functional as F

LRELU_SLOPE = 0.1  # residuals.py:10


@dataclass
class DecoderConfig:
    """The nine constructor values of HiFiGANNSFGenerator (hifigan_nsf.py:75-86)."""

    initial_channel: int = 192
    resblock_kernel_sizes: Sequence[int] = (3, 7, 11)
    resblock_dilation_sizes: Sequence[Sequence[int]] = ((1, 3, 5), (1, 3, 5), (1, 3, 5))
    upsample_rates: Sequence[int] = (12, 10, 2, 2)
    upsample_initial_channel: int = 512
    upsample_kernel_sizes: Sequence[int] = (24, 20, 4, 4)
    gin_channels: int = 256
    sr: int = 48000

    @property
    def upp(self) -> int:  # hifigan_nsf.py:170
        return math.prod(self.upsample_rates)

    def ctor_args(self):
        return (
            self.initial_channel,
            list(self.resblock_kernel_sizes),
            [list(d) for d in self.resblock_dilation_sizes],
            list(self.upsample_rates),
            self.upsample_initial_channel,
            list(self.upsample_kernel_sizes),
        )

    def ctor_kwargs(self):
        return dict(gin_channels=self.gin_channels, sr=self.sr)


CONFIG_48K = DecoderConfig()
CONFIG_40K = DecoderConfig(upsample_rates=(10, 10, 2, 2), upsample_kernel_sizes=(16, 16, 4, 4), sr=40000)
CONFIG_32K = DecoderConfig(upsample_rates=(10, 8, 2, 2), upsample_kernel_sizes=(20, 16, 4, 4), sr=32000)
# A narrow decoder with the same topology, small enough to commit its weights as a fixture.
CONFIG_TINY = DecoderConfig(
    initial_channel=24,
    upsample_rates=(6, 4, 2, 2),
    upsample_initial_channel=128,
    upsample_kernel_sizes=(12, 8, 4, 4),
    gin_channels=16,
    sr=24000,
)


# Odd upsample rates (hifigan_nsf.py:113-117 padding u//2 + 1 / output_padding 1, noise convs with odd strides :142-143).
CONFIG_ODD = DecoderConfig(initial_channel=32, upsample_rates=(5, 3, 2, 2), upsample_initial_channel=128,
                           upsample_kernel_sizes=(10, 6, 4, 4), gin_channels=16, sr=24000)


def get_padding(kernel_size: int, dilation: int = 1) -> int:
    """commons.py:19-27."""
    return int((kernel_size * dilation - dilation) / 2)


def fold_weight_norm(g: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
    """w = v * (g / ||v||) with the norm over every dim but 0, no epsilon.

    torch.nn.utils.parametrizations.weight_norm(dim=0) as used at
    hifigan_nsf.py:120 and residuals.py:14.  For ConvTranspose1d dim 0 is C_in.
    """
    norm = v.reshape(v.shape[0], -1).norm(dim=1).reshape(-1, *([1] * (v.dim() - 1)))
    return v * (g / norm)


def folded_weights(state: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Turn a decoder state_dict (either naming convention) into plain conv weights.

    Accepts ``*.parametrizations.weight.original0/1`` (torch>=2.1 naming) and the
    legacy ``*.weight_g`` / ``*.weight_v`` written by the reference's model
    extractor (rvc/train/process/extract_model.py:57-109, rvc/train/utils.py:49-55).
    """
    out: Dict[str, torch.Tensor] = {}
    for k, t in state.items():
        t = t.detach().to(torch.float32).cpu()
        if k.endswith(".parametrizations.weight.original0"):
            base = k[: -len(".parametrizations.weight.original0")]
            out[base + ".weight"] = fold_weight_norm(t, state[base + ".parametrizations.weight.original1"].detach().float().cpu())
        elif k.endswith(".weight_g"):
            base = k[: -len(".weight_g")]
            out[base + ".weight"] = fold_weight_norm(t, state[base + ".weight_v"].detach().float().cpu())
        elif k.endswith(".parametrizations.weight.original1") or k.endswith(".weight_v"):
            continue
        else:
            out[k] = t
    return out


def sine_source(
    f0: torch.Tensor,
    upp: int,
    sr: int,
    noise: torch.Tensor,
    dtype: torch.dtype = torch.float32,
    sine_amp: float = 0.1,
    noise_std: float = 0.003,
) -> torch.Tensor:
    """SineGenerator.forward (hifigan.py:197-228) with harmonic_num = 0.

    f0: (B, T) Hz, 0 = unvoiced.  noise: (B, T*upp) standard-normal draws, the
    values ``torch.randn_like(sine_waves)`` would have produced (hifigan.py:223).
    ``dtype`` float32 follows the reference bit for bit; float64 is the
    higher-precision phase accumulation the CUDA kernel uses.
    Returns (B, T*upp).
    """
    B, T = f0.shape
    f0 = f0.to(dtype)
    grid = torch.arange(1, upp + 1, dtype=dtype)  # hifigan.py:167-169
    inc = (f0 / sr).unsqueeze(-1) * grid  # (B, T, upp)  :172
    rem = torch.fmod(inc[:, :-1, -1:] + 0.5, 1.0) - 0.5  # :173
    cum = rem.cumsum(dim=1).fmod(1.0).to(dtype)  # :174
    inc = inc + F.pad(cum, (0, 0, 1, 0))  # :175-177
    phase = inc.reshape(B, -1)  # :180 (harmonic scale 1, random phase of the fundamental zeroed :183-191)
    sine = torch.sin(2 * math.pi * phase) * sine_amp  # :194, :204
    uv = (f0 > 0).to(dtype).repeat_interleave(upp, dim=1)  # :146-154, :208-215 nearest upsample
    amp = uv * noise_std + (1 - uv) * (sine_amp / 3)  # :218-220
    out = sine * uv + amp * noise.to(dtype)  # :223-226
    return out.to(torch.float32)


def draw_reference_noise(seed: int, B: int, n: int) -> torch.Tensor:
    """The noise the reference draws after ``torch.manual_seed(seed)``.

    Order per forward on the CPU generator: ``torch.rand(1,1,1)`` (hifigan.py:189)
    then ``torch.randn_like(sine_waves)`` of shape (B, n, 1) (hifigan.py:223).
    """
    gen = torch.Generator().manual_seed(seed)
    torch.rand(1, 1, 1, generator=gen)
    return torch.randn(B, n, 1, generator=gen).reshape(B, n)


def decoder_forward(
    w: Dict[str, torch.Tensor],
    cfg: DecoderConfig,
    x: torch.Tensor,
    f0: torch.Tensor,
    g: Optional[torch.Tensor],
    noise: torch.Tensor,
    dtype: torch.dtype = torch.float32,
    taps: Optional[Dict[str, torch.Tensor]] = None,
    source_dtype: Optional[torch.dtype] = None,
) -> torch.Tensor:
    """HiFiGANNSFGenerator.forward (hifigan_nsf.py:173-212), inference branch.

    ``w`` are folded weights (see folded_weights).  ``taps`` if given receives the
    intermediate tensors the CUDA path is checked against layer by layer.
    """
    W = {k: v.to(dtype) for k, v in w.items()}
    x = x.to(dtype)
    nk = len(cfg.resblock_kernel_sizes)
    # m_source: SineGenerator then Linear(1,1) + tanh (hifigan_nsf.py:48-52), transposed to (B,1,n) :177
    s = sine_source(f0, cfg.upp, cfg.sr, noise, dtype=source_dtype or torch.float32).to(dtype)
    har = torch.tanh(s * W["m_source.l_linear.weight"].reshape(()) + W["m_source.l_linear.bias"].reshape(())).unsqueeze(1)
    if taps is not None:
        taps["har"] = har[:, 0].float()
    x = F.conv1d(x, W["conv_pre.weight"], W["conv_pre.bias"], padding=3)  # :179
    if g is not None:
        x = x + F.conv1d(g.to(dtype), W["cond.weight"], W["cond.bias"])  # :181-182
    if taps is not None:
        taps["pre"] = x.float()
    for i, (u, k) in enumerate(zip(cfg.upsample_rates, cfg.upsample_kernel_sizes)):
        x = F.leaky_relu(x, LRELU_SLOPE)  # :185
        if u % 2 == 0:  # :113-117
            pad, opad = (k - u) // 2, 0
        else:
            pad, opad = u // 2 + u % 2, u % 2
        x = F.conv_transpose1d(x, W[f"ups.{i}.weight"], W[f"ups.{i}.bias"], stride=u, padding=pad, output_padding=opad)  # :198
        stride = math.prod(cfg.upsample_rates[i + 1 :]) if i + 1 < len(cfg.upsample_rates) else 1  # :106-109
        nkern = 1 if stride == 1 else stride * 2 - stride % 2  # :142-143
        npad = 0 if stride == 1 else (nkern - stride) // 2
        x = x + F.conv1d(har, W[f"noise_convs.{i}.weight"], W[f"noise_convs.{i}.bias"], stride=stride, padding=npad)  # :199
        if taps is not None:
            taps[f"stage{i}.in"] = x.float()
        xs = None
        for j in range(nk):  # :200-206
            y = resblock_forward(W, f"resblocks.{i * nk + j}", x, cfg.resblock_kernel_sizes[j], cfg.resblock_dilation_sizes[j])
            if taps is not None:
                taps[f"stage{i}.rb{j}"] = y.float()
            xs = y if xs is None else xs + y
        x = xs / nk  # :207
        if taps is not None:
            taps[f"stage{i}.out"] = x.float()
    x = F.leaky_relu(x)  # default slope 0.01  :209
    x = torch.tanh(F.conv1d(x, W["conv_post.weight"], None, padding=3))  # :210 (bias=False :164)
    return x.float()


def generator_forward_nof0(w: Dict[str, torch.Tensor], cfg: DecoderConfig, x: torch.Tensor, g: Optional[torch.Tensor],
                           dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """HiFiGANGenerator.forward (rvc/lib/algorithm/generators/hifigan.py:72-97): the sibling decoder of the no-f0
    models (selected at synthesizers.py:127-135) - the same conv stack without the harmonic source and noise convs."""
    W = {k: v.to(dtype) for k, v in w.items()}
    nk = len(cfg.resblock_kernel_sizes)
    x = F.conv1d(x.to(dtype), W["conv_pre.weight"], W["conv_pre.bias"], padding=3)  # :74
    if g is not None:
        x = x + F.conv1d(g.to(dtype), W["cond.weight"], W["cond.bias"])  # :76-77
    for i, (u, k) in enumerate(zip(cfg.upsample_rates, cfg.upsample_kernel_sizes)):
        x = F.leaky_relu(x, LRELU_SLOPE)  # :80
        x = F.conv_transpose1d(x, W[f"ups.{i}.weight"], W[f"ups.{i}.bias"], stride=u, padding=(k - u) // 2)  # :81, :56
        xs = None
        for j in range(nk):  # :82-87
            y = resblock_forward(W, f"resblocks.{i * nk + j}", x, cfg.resblock_kernel_sizes[j], cfg.resblock_dilation_sizes[j])
            xs = y if xs is None else xs + y
        x = xs / nk  # :88
    x = F.leaky_relu(x)  # :90 default slope 0.01
    return torch.tanh(F.conv1d(x, W["conv_post.weight"], None, padding=3)).float()  # :91-93


def resblock_forward(W, prefix: str, x: torch.Tensor, k: int, dilations: Sequence[int]) -> torch.Tensor:
    """ResBlock.forward (residuals.py:71-80) with x_mask=None."""
    for m, d in enumerate(dilations):
        res = x
        xt = F.leaky_relu(x, LRELU_SLOPE)
        xt = F.conv1d(xt, W[f"{prefix}.convs1.{m}.weight"], W[f"{prefix}.convs1.{m}.bias"], dilation=d, padding=get_padding(k, d))
        xt = F.leaky_relu(xt, LRELU_SLOPE)
        xt = F.conv1d(xt, W[f"{prefix}.convs2.{m}.weight"], W[f"{prefix}.convs2.{m}.bias"], dilation=1, padding=get_padding(k, 1))
        x = xt + res
    return x


# --------------------------------------------------------------------------
# Synthetic inputs of the named shapes.
# --------------------------------------------------------------------------
def synth_f0(B: int, T: int, gen: torch.Generator) -> torch.Tensor:
    """Piecewise-smooth f0 contour in Hz with unvoiced runs set to exactly 0."""
    t = torch.arange(T, dtype=torch.float32) / 100.0
    median = 90 + 210 * torch.rand(B, 1, generator=gen)
    vib = (0.5 / 12.0) * torch.sin(2 * math.pi * 5.0 * t + 6.28 * torch.rand(B, 1, generator=gen))
    drift = torch.cumsum(torch.randn(B, T, generator=gen) * 0.004, dim=1)
    f0 = (median * torch.pow(2.0, vib + drift)).clamp(50.0, 1100.0)
    # unvoiced runs of 10-60 frames covering roughly a quarter of the frames
    for b in range(B):
        pos = int(torch.randint(0, 40, (1,), generator=gen))
        while pos < T:
            run = int(torch.randint(10, 61, (1,), generator=gen))
            f0[b, pos : pos + run] = 0.0
            pos += run + int(torch.randint(60, 240, (1,), generator=gen))
    return f0


def synth_inputs(cfg: DecoderConfig, B: int, T: int, seed: int = 1234):
    gen = torch.Generator().manual_seed(seed)
    x = torch.randn(B, cfg.initial_channel, T, generator=gen)
    g = torch.randn(B, cfg.gin_channels, 1, generator=gen) if cfg.gin_channels else None
    f0 = synth_f0(B, T, gen)
    noise = torch.randn(B, T * cfg.upp, generator=gen)
    return x, f0, g, noise


# --------------------------------------------------------------------------
# Parity metrics.
# --------------------------------------------------------------------------
def _mel_fbank(n_freqs: int, n_mels: int, sr: int) -> torch.Tensor:
    """Slaney-scale, slaney-normalised triangular mel bank (librosa.filters.mel defaults)."""
    def hz_to_mel(f):
        f = torch.as_tensor(f, dtype=torch.float64)
        lin = f / (200.0 / 3)
        log = 15.0 + torch.log(f.clamp(min=1e-10) / 1000.0) / (math.log(6.4) / 27.0)
        return torch.where(f >= 1000.0, log, lin)

    def mel_to_hz(m):
        lin = m * (200.0 / 3)
        log = 1000.0 * torch.exp((math.log(6.4) / 27.0) * (m - 15.0))
        return torch.where(m >= 15.0, log, lin)

    freqs = torch.linspace(0, sr / 2, n_freqs, dtype=torch.float64)
    mpts = torch.linspace(float(hz_to_mel(0.0)), float(hz_to_mel(sr / 2)), n_mels + 2, dtype=torch.float64)
    fpts = mel_to_hz(mpts)
    fdiff = fpts[1:] - fpts[:-1]
    ramps = fpts.unsqueeze(1) - freqs.unsqueeze(0)
    lower = -ramps[:-2] / fdiff[:-1].unsqueeze(1)
    upper = ramps[2:] / fdiff[1:].unsqueeze(1)
    fb = torch.clamp(torch.minimum(lower, upper), min=0)
    fb = fb * (2.0 / (fpts[2:] - fpts[:-2])).unsqueeze(1)
    return fb.float()  # (n_mels, n_freqs)


def multires_logmel_corr(y: torch.Tensor, y_ref: torch.Tensor, sr: int) -> Dict[str, float]:
    """Pearson correlation of log-mel spectrograms at the 7 resolutions of
    MultiScaleMelSpectrogramLoss (rvc/train/mel_processing.py:114-188)."""
    y = y.reshape(-1, y.shape[-1]).float().cpu()
    y_ref = y_ref.reshape(-1, y_ref.shape[-1]).float().cpu()
    res = {}
    vals = []
    for n_mels, win in zip([5, 10, 20, 40, 80, 160, 320], [32, 64, 128, 256, 512, 1024, 2048]):
        fb = _mel_fbank(win // 2 + 1, n_mels, sr)
        window = torch.hann_window(win)

        def lm(a):
            s = torch.stft(a, n_fft=win, hop_length=win // 4, win_length=win, window=window, center=True, return_complex=True)
            mag = torch.sqrt(s.real**2 + s.imag**2 + 1e-6)
            return torch.log10(torch.clamp(fb @ mag, min=1e-5)).flatten()

        a, b = lm(y).double(), lm(y_ref).double()
        a, b = a - a.mean(), b - b.mean()
        c = float((a @ b) / (a.norm() * b.norm() + 1e-30))
        res[f"win{win}"] = c
        vals.append(c)
    res["min"] = min(vals)
    res["mean"] = sum(vals) / len(vals)
    return res


# --------------------------------------------------------------------------
# Sibling decoder HiFiGANMRFGenerator (rvc/lib/algorithm/generators/hifigan_mrf.py:223-374)
# --------------------------------------------------------------------------
def mrf_sine_source(f0: torch.Tensor, upp: int, sr: int, rand_ini: torch.Tensor, noise: torch.Tensor, dtype: torch.dtype = torch.float64,
                    sine_amp: float = 0.1, noise_std: float = 0.003) -> torch.Tensor:
    """SineGenerator.forward of the MRF variant (hifigan_mrf.py:86-176) in closed form.

    The reference upsamples f0 to the sample rate (nearest, :338), forms rad = (f0 * (k+1) / sr) % 1 per harmonic k, adds a
    random initial phase to sample 0 of every overtone (:129-134) and takes a per-sample cumsum with an explicit wrap
    correction (:137-142); sin(2 pi x) only sees the cumsum modulo 1.  Because f0 is constant inside a frame the cumsum is
    closed-form there:  phase_k[l*upp + j] = ini_k + (k+1) * (sum_{l'<l} upp * f0[l']/sr + (j+1) * f0[l]/sr)  (mod 1),
    so only a FRAME-level prefix is needed.  dtype float64 is the accumulation the CUDA kernel uses; the reference's fp32
    per-sample cumsum differs from it by its own rounding (grows with the number of samples).
    f0 (B, T); rand_ini (B, H) uniform [0,1) with column 0 = 0; noise (B, T*upp, H).  Returns (B, T*upp, H)."""
    B, T = f0.shape
    H = rand_ini.shape[1]
    f = f0.to(dtype) / sr
    prefix = torch.cumsum(f * upp, dim=1) - f * upp                       # (B, T): advance before frame l
    j = torch.arange(1, upp + 1, dtype=dtype)
    base = (prefix.unsqueeze(-1) + f.unsqueeze(-1) * j).reshape(B, T * upp)  # fundamental phase (cycles)
    k = torch.arange(1, H + 1, dtype=dtype)
    phase = base.unsqueeze(-1) * k + rand_ini.to(dtype).unsqueeze(1)
    sines = torch.sin(2 * math.pi * (phase - torch.floor(phase))) * sine_amp
    uv = (f0 > 0).to(dtype).repeat_interleave(upp, dim=1).unsqueeze(-1)
    amp = uv * noise_std + (1 - uv) * sine_amp / 3
    return (sines * uv + amp * noise.to(dtype)).to(torch.float32)


def mrf_state_to_nsf_names(state: Dict[str, torch.Tensor], nk: int) -> Dict[str, torch.Tensor]:
    """HiFiGANMRFGenerator state dict -> the NSF decoder's names (same layers under other names):
    upsamples.i -> ups.i, mrfs.i.j.layers.m.conv1/conv2 -> resblocks.(i*nk+j).convs1/convs2.m."""
    out = {}
    for k, v in state.items():
        if k.startswith("upsamples."):
            k = "ups." + k[len("upsamples."):]
        elif k.startswith("mrfs."):
            p = k.split(".")
            i, j, m, which = int(p[1]), int(p[2]), int(p[4]), p[5]
            k = f"resblocks.{i * nk + j}.convs{1 if which == 'conv1' else 2}.{m}." + ".".join(p[6:])
        out[k] = v
    return out


def decoder_forward_mrf(w: Dict[str, torch.Tensor], cfg: DecoderConfig, x, f0, g, rand_ini, noise, dtype=torch.float32,
                        source_dtype=torch.float64) -> torch.Tensor:
    """HiFiGANMRFGenerator.forward (hifigan_mrf.py:335-367).  ``w`` = folded weights under the NSF names
    (mrf_state_to_nsf_names + folded_weights); conv_post has a bias here (:327-329), l_linear is (1, H)."""
    W = {k: v.to(dtype) for k, v in w.items()}
    nk = len(cfg.resblock_kernel_sizes)
    s = mrf_sine_source(f0, cfg.upp, cfg.sr, rand_ini, noise, dtype=source_dtype).to(dtype)   # (B, n, H)
    har = torch.tanh(F.linear(s, W["m_source.l_linear.weight"], W["m_source.l_linear.bias"])).transpose(1, 2)  # (B, 1, n)
    x = F.conv1d(x.to(dtype), W["conv_pre.weight"], W["conv_pre.bias"], padding=3)
    if g is not None:
        x = x + F.conv1d(g.to(dtype), W["cond.weight"], W["cond.bias"])
    for i, (u, k) in enumerate(zip(cfg.upsample_rates, cfg.upsample_kernel_sizes)):
        x = F.leaky_relu(x, LRELU_SLOPE)
        pad, opad = ((k - u) // 2, 0) if u % 2 == 0 else (u // 2 + u % 2, u % 2)  # hifigan_mrf.py:279-296
        x = F.conv_transpose1d(x, W[f"ups.{i}.weight"], W[f"ups.{i}.bias"], stride=u, padding=pad, output_padding=opad)
        stride = math.prod(cfg.upsample_rates[i + 1:]) if i + 1 < len(cfg.upsample_rates) else 1
        nkern = 1 if stride == 1 else stride * 2 - stride % 2
        npad = 0 if stride == 1 else (nkern - stride) // 2
        x = x + F.conv1d(har, W[f"noise_convs.{i}.weight"], W[f"noise_convs.{i}.bias"], stride=stride, padding=npad)
        xs = None
        for j in range(nk):
            y = resblock_forward(W, f"resblocks.{i * nk + j}", x, cfg.resblock_kernel_sizes[j], cfg.resblock_dilation_sizes[j])
            xs = y if xs is None else xs + y
        x = xs / nk
    x = F.leaky_relu(x)
    return torch.tanh(F.conv1d(x, W["conv_post.weight"], W["conv_post.bias"], padding=3)).float()
