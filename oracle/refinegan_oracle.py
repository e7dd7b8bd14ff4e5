"""CPU oracle for the sibling decoder RefineGANGenerator (rvc/lib/algorithm/generators/refinegan.py:268-451, selected by
vocoder "RefineGAN" at synthesizers.py:98-106).  TEST INFRASTRUCTURE ONLY.

Restated with torch CPU operators; every random draw of the reference is an explicit input, in the order the reference
makes them: ``rand_ini`` (B, 1) [SineGenerator, :226-229, zeroed for the fundamental], ``noise_src`` (B, n, 1) [:257], then
per upsample stage and per parallel block the two AdaIN draws (B, C, L) [:98-99] - ``adain[stage][block][0 / 1]``.

  SineGenerator.forward      refinegan.py:245-263   f0 linearly interpolated to the sample rate (:419), per-sample cumsum (:232-238)
  downsample path            :422-436               leaky_relu(0.2), torchaudio sinc_interp_kaiser resample, weight-normed k=7 convs
  ParallelResBlock           :105-178               input conv k=7, 3 x [AdaIN -> ResBlock(k, (1,3,5), slope 0.2) -> AdaIN], mean
  upsample path              :438-455               leaky_relu(0.2), linear Upsample, concat with the skip, ParallelResBlock
  tail                       :457-460               leaky_relu(0.2), conv_post (no bias), tanh

The per-sample phase is evaluated in closed form (fp64): linear interpolation with align_corners=False is piecewise
linear between frame centres, so the cumsum inside a piece is an arithmetic series; the reference's fp32 cumsum differs
from it by its own rounding.  The resampling filter is restated by ``sinc_kaiser_kernel`` (the published torchaudio
algorithm, torchaudio.functional.resample with the reference's arguments); tests pin both against the live libraries.
Parity status: PINNED by tests/golden/refinegan_*.npz (oracle/make_golden_r2.py --refinegan).
"""
from __future__ import annotations

import math
from typing import Dict, List, Sequence

import torch
import torch.nn.functional as F

LPW, ROLLOFF, BETA = 64, 0.9475937167399596, 14.769656459379492  # refinegan.py:429-433


def sinc_kaiser_kernel(rate: int) -> tuple:
    """torchaudio.functional.resample(orig_freq=rate*k, new_freq=k, lowpass_filter_width=64, rolloff, sinc_interp_kaiser,
    beta) reduces to ONE filter of 2*width + rate taps applied with stride `rate` (fp32 arithmetic as torchaudio does for
    an fp32 waveform).  Returns (kernel (taps,), width)."""
    orig, new = rate, 1
    base = min(orig, new) * ROLLOFF
    width = math.ceil(LPW * orig / base)
    idx = torch.arange(-width, width + orig, dtype=torch.float32)[None, None] / orig
    t = torch.arange(0, -new, -1, dtype=torch.float32)[:, None, None] / new + idx
    t *= base
    t = t.clamp_(-LPW, LPW)
    beta = torch.tensor(float(BETA))
    window = torch.i0(beta * torch.sqrt(1 - (t / LPW) ** 2)) / torch.i0(beta)
    t *= math.pi
    scale = base / orig
    k = torch.where(t == 0, torch.tensor(1.0).to(t), t.sin() / t)
    k *= window * scale
    return k.reshape(-1).float(), width


def resample_down(x: torch.Tensor, rate: int) -> torch.Tensor:
    """(B, C, L) -> (B, C, L / rate): pad (width, width + rate), depthwise FIR, stride rate (torchaudio _apply_sinc_resample_kernel)."""
    k, width = sinc_kaiser_kernel(rate)
    B, C, L = x.shape
    xp = F.pad(x.reshape(B * C, 1, L), (width, width + rate))
    y = F.conv1d(xp, k.reshape(1, 1, -1).to(x.dtype), stride=rate)
    return y[..., : math.ceil(L / rate)].reshape(B, C, -1)


def sine_source(f0: torch.Tensor, upp: int, sr: int, noise: torch.Tensor, dtype=torch.float64) -> torch.Tensor:
    """SineGenerator.forward (refinegan.py:245-263) on f0 linearly interpolated to the sample rate (:419).
    f0 (B, T); noise (B, T*upp).  Returns sine_waves (B, T*upp) BEFORE the merge (Linear(1,1,bias=False) + tanh)."""
    B, T = f0.shape
    n = T * upp
    f = F.interpolate(f0.to(dtype).unsqueeze(1), size=n, mode="linear")[:, 0]   # (B, n), align_corners=False
    # closed-form prefix of f / sr: pieces of `upp` samples between frame centres, constant half-pieces at both ends
    fd = f0.to(dtype) / sr
    s = torch.arange(n, dtype=dtype)
    src = ((s + 0.5) / upp - 0.5).clamp(min=0)
    i0 = src.floor().long().clamp(max=T - 1)
    i1 = (i0 + 1).clamp(max=T - 1)
    half = upp // 2
    # advance accumulated before piece p (piece 0 = the first half frame, piece p >= 1 starts at sample half + (p-1) upp)
    a, b = fd[:, :-1], fd[:, 1:]
    piece_sum = torch.cat([fd[:, :1] * half, upp * (a + b) / 2], dim=1)           # (B, T): sums of pieces 0 .. T-1
    before = torch.cat([torch.zeros(B, 1, dtype=dtype), torch.cumsum(piece_sum, dim=1)], dim=1)   # (B, T+1): advance before piece p
    p = torch.where(s < half, torch.zeros_like(s), ((s - half) / upp).floor() + 1).long()        # piece of sample s
    j = torch.where(s < half, s, s - half - (p - 1).to(dtype) * upp)                             # index inside the piece
    fa = torch.where(p == 0, fd[:, :1].expand(B, n), fd.gather(1, (p - 1).clamp(min=0).expand(B, n)))
    fb = torch.where(p == 0, fd[:, :1].expand(B, n), fd.gather(1, p.clamp(max=T - 1).expand(B, n)))
    inside = (j + 1) * fa + (fb - fa) * (j + 1) ** 2 / (2 * upp)
    phase = before.gather(1, p.expand(B, n)) + inside
    sines = torch.sin(2 * math.pi * (phase - phase.floor())) * 0.1
    uv = (f > 0).to(dtype)
    amp = uv * 0.003 + (1 - uv) * 0.1 / 3
    return (sines * uv + amp * noise.to(dtype)).float()


def resblock(W, pre: str, x: torch.Tensor, k: int, dils: Sequence[int], slope: float) -> torch.Tensor:
    for m, d in enumerate(dils):  # refinegan.py:78-85
        xt = F.leaky_relu(x, slope)
        xt = F.conv1d(xt, W[f"{pre}.convs1.{m}.weight"], W[f"{pre}.convs1.{m}.bias"], dilation=d, padding=(k * d - d) // 2)
        xt = F.leaky_relu(xt, slope)
        xt = F.conv1d(xt, W[f"{pre}.convs2.{m}.weight"], W[f"{pre}.convs2.{m}.bias"], padding=(k - 1) // 2)
        x = xt + x
    return x


def forward(W: Dict[str, torch.Tensor], rates: Sequence[int], sr: int, mel, f0, g, noise_src, adain: List[List[List[torch.Tensor]]],
            slope: float = 0.2) -> torch.Tensor:
    """RefineGANGenerator.forward (refinegan.py:415-462).  W = folded weights (oracle.nsf_hifigan_oracle.folded_weights)."""
    upp = math.prod(rates)
    s = sine_source(f0, upp, sr, noise_src)
    har = torch.tanh(s * W["m_source.merge.0.weight"].reshape(())).unsqueeze(1)           # (B, 1, n)
    x = F.conv1d(har, W["pre_conv.weight"], W["pre_conv.bias"], padding=3)
    downs = []
    for i in range(len(rates)):
        x = F.leaky_relu(x, slope)
        downs.append(x)
        x = resample_down(x, rates[-i - 1])
        x = F.conv1d(x, W[f"downsample_blocks.{i}.weight"], W[f"downsample_blocks.{i}.bias"], padding=3)
    m = F.conv1d(mel, W["mel_conv.weight"], W["mel_conv.bias"], padding=3)
    if g is not None:
        m = m + F.conv1d(g, W["cond.weight"], W["cond.bias"])
    x = torch.cat([m, x], dim=1)
    for i, (rate, down) in enumerate(zip(rates, reversed(downs))):
        x = F.leaky_relu(x, slope)
        x = F.interpolate(x, scale_factor=float(rate), mode="linear")
        x = torch.cat([x, down], dim=1)
        pre = f"upsample_conv_blocks.{i}"
        x = F.conv1d(x, W[pre + ".input_conv.weight"], W[pre + ".input_conv.bias"], padding=3)
        outs = []
        for j, k in enumerate((3, 7, 11)):
            a = F.leaky_relu(x + adain[i][j][0] * W[f"{pre}.blocks.{j}.0.weight"][None, :, None], slope)
            y = resblock(W, f"{pre}.blocks.{j}.1", a, k, (1, 3, 5), slope)
            outs.append(F.leaky_relu(y + adain[i][j][1] * W[f"{pre}.blocks.{j}.2.weight"][None, :, None], slope))
        x = torch.stack(outs, 0).mean(0)
    x = F.leaky_relu(x, slope)
    return torch.tanh(F.conv1d(x, W["conv_post.weight"], None, padding=3))
