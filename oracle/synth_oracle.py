"""CPU oracle for the rest of ``Synthesizer.infer``.  TEST INFRASTRUCTURE ONLY.

Restates, with torch CPU operators, everything ``rvc/lib/algorithm/synthesizers.py:205-243`` does before the decoder:

  emb_g lookup                      synthesizers.py:226
  TextEncoder.forward               encoders.py:127-144   (emb_phone + emb_pitch, * sqrt(H), leaky_relu 0.1, mask)
  Encoder.forward                   encoders.py:72-85     (6 x [rel-pos MHA -> LayerNorm -> FFN -> LayerNorm])
  MultiHeadAttention.attention      attentions.py:89-125  (window-10 relative keys / values, -1e4 mask fill)
  FFN.forward                       attentions.py:221-226 (k = 3 "same" padding, relu)
  LayerNorm                         normalization.py:19-26 (over channels, eps 1e-5)
  z_p sampling                      synthesizers.py:228   ((m + exp(logs) * noise * 0.66666) * mask)
  ResidualCouplingBlock(reverse)    residuals.py:153-165  (Flip, mean-only coupling layers in reverse order)
  ResidualCouplingLayer.forward     residuals.py:231-256
  WaveNet.forward                   modules.py:84-117 with fused_add_tanh_sigmoid_multiply commons.py:88-103

The relative-position terms are written in their direct form (score[t, s] += q_t . E_k[s - t + w] for |s - t| <= w) rather
than through the reference's pad / reshape skewing (attentions.py:156-180); tests pin the equality on goldens generated
from the unmodified reference.  Parity status: PINNED by tests/golden/synth_*.npz (oracle/make_golden_r2.py).
"""
from __future__ import annotations

import math
from typing import Dict, Optional

import torch
import torch.nn.functional as F

from .nsf_hifigan_oracle import fold_weight_norm


def folded(state: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Synthesizer state dict (either weight-norm naming) -> plain fp32 weights."""
    out = {}
    for k, t in state.items():
        t = t.detach().float().cpu()
        for g_sfx, v_sfx in ((".parametrizations.weight.original0", ".parametrizations.weight.original1"), (".weight_g", ".weight_v")):
            if k.endswith(g_sfx):
                base = k[: -len(g_sfx)]
                out[base + ".weight"] = fold_weight_norm(t, state[base + v_sfx].detach().float().cpu())
                break
        else:
            if not (k.endswith(".parametrizations.weight.original1") or k.endswith(".weight_v")):
                out[k] = t
    return out


def attention(W, pre: str, x: torch.Tensor, mask: torch.Tensor, n_heads: int, window: int) -> torch.Tensor:
    """MultiHeadAttention.forward (attentions.py:79-125) for self-attention; mask (B, 1, T) of 0/1."""
    B, C, T = x.shape
    kc = C // n_heads
    q = F.conv1d(x, W[pre + ".conv_q.weight"], W[pre + ".conv_q.bias"])
    k = F.conv1d(x, W[pre + ".conv_k.weight"], W[pre + ".conv_k.bias"])
    v = F.conv1d(x, W[pre + ".conv_v.weight"], W[pre + ".conv_v.bias"])
    q = q.view(B, n_heads, kc, T).transpose(2, 3) / math.sqrt(kc)   # (B, h, T, kc)
    k = k.view(B, n_heads, kc, T).transpose(2, 3)
    v = v.view(B, n_heads, kc, T).transpose(2, 3)
    scores = q @ k.transpose(-2, -1)                                # (B, h, T, T)
    ek, ev = W[pre + ".emb_rel_k"][0], W[pre + ".emb_rel_v"][0]     # (2w+1, kc), heads_share
    idx = torch.arange(T)
    rel = idx[None, :] - idx[:, None]                               # s - t
    band = rel.abs() <= window
    rl = q @ ek.t()                                                 # (B, h, T, 2w+1): q_t . E_k[r]
    gather = (rel + window).clamp(0, 2 * window)
    scores = scores + torch.where(band, rl.gather(-1, gather.expand(B, n_heads, T, T)), torch.zeros(()))
    am = mask.unsqueeze(2) * mask.unsqueeze(-1)                     # (B, 1, T, T)
    scores = scores.masked_fill(am == 0, -1e4)
    p = F.softmax(scores, dim=-1)
    out = p @ v
    # relative values: out[t] += sum_{|s-t|<=w} p[t, s] E_v[s - t + w]
    pw = torch.zeros(B, n_heads, T, 2 * window + 1)
    for r in range(2 * window + 1):
        d = r - window
        lo, hi = max(0, -d), min(T, T - d)
        if lo < hi:
            t_idx = torch.arange(lo, hi)
            pw[:, :, t_idx, r] = p[:, :, t_idx, t_idx + d]
    out = out + pw @ ev
    out = out.transpose(2, 3).contiguous().view(B, C, T)
    return F.conv1d(out, W[pre + ".conv_o.weight"], W[pre + ".conv_o.bias"])


def layer_norm(W, pre: str, x: torch.Tensor) -> torch.Tensor:
    return F.layer_norm(x.transpose(1, -1), (x.shape[1],), W[pre + ".gamma"], W[pre + ".beta"], 1e-5).transpose(1, -1)


def ffn(W, pre: str, x: torch.Tensor, mask: torch.Tensor, k: int) -> torch.Tensor:
    p = (k - 1) // 2
    y = F.conv1d(F.pad(x * mask, (p, p)), W[pre + ".conv_1.weight"], W[pre + ".conv_1.bias"])
    y = torch.relu(y)
    y = F.conv1d(F.pad(y * mask, (p, p)), W[pre + ".conv_2.weight"], W[pre + ".conv_2.bias"])
    return y * mask


def text_encoder(W, phone, pitch, lengths, hidden: int, out_ch: int, n_heads: int, n_layers: int, kernel: int, window: int = 10):
    """TextEncoder.forward (encoders.py:127-144) + Encoder.forward (:72-85) -> m, logs, mask."""
    x = F.linear(phone, W["enc_p.emb_phone.weight"], W["enc_p.emb_phone.bias"])
    if pitch is not None and "enc_p.emb_pitch.weight" in W:
        x = x + W["enc_p.emb_pitch.weight"][pitch]
    x = F.leaky_relu(x * math.sqrt(hidden), 0.1).transpose(1, -1)
    T = x.shape[2]
    mask = (torch.arange(T)[None, :] < lengths[:, None]).unsqueeze(1).to(x.dtype)
    x = x * mask
    for i in range(n_layers):
        y = attention(W, f"enc_p.encoder.attn_layers.{i}", x, mask, n_heads, window)
        x = layer_norm(W, f"enc_p.encoder.norm_layers_1.{i}", x + y)
        y = ffn(W, f"enc_p.encoder.ffn_layers.{i}", x, mask, kernel)
        x = layer_norm(W, f"enc_p.encoder.norm_layers_2.{i}", x + y)
    x = x * mask
    stats = F.conv1d(x, W["enc_p.proj.weight"], W["enc_p.proj.bias"]) * mask
    m, logs = torch.split(stats, out_ch, dim=1)
    return m, logs, mask


def wavenet(W, pre: str, x, mask, g, hidden: int, kernel: int, n_layers: int, dilation_rate: int = 1):
    """WaveNet.forward (modules.py:84-117)."""
    out = torch.zeros_like(x)
    gc = F.conv1d(g, W[pre + ".cond_layer.weight"], W[pre + ".cond_layer.bias"]) if g is not None else None
    for i in range(n_layers):
        d = dilation_rate ** i
        a = F.conv1d(x, W[f"{pre}.in_layers.{i}.weight"], W[f"{pre}.in_layers.{i}.bias"], dilation=d, padding=(kernel * d - d) // 2)
        if gc is not None:
            a = a + gc[:, i * 2 * hidden:(i + 1) * 2 * hidden]
        acts = torch.tanh(a[:, :hidden]) * torch.sigmoid(a[:, hidden:])
        rs = F.conv1d(acts, W[f"{pre}.res_skip_layers.{i}.weight"], W[f"{pre}.res_skip_layers.{i}.bias"])
        if i < n_layers - 1:
            x = (x + rs[:, :hidden]) * mask
            out = out + rs[:, hidden:]
        else:
            out = out + rs
    return out * mask


def flow_reverse(W, z, mask, g, channels: int, hidden: int, kernel: int = 5, n_layers: int = 3, n_flows: int = 4):
    """ResidualCouplingBlock.forward(reverse=True) (residuals.py:162-165): reversed([RCL0, Flip, ..., RCL3, Flip])."""
    half = channels // 2
    x = z
    for i in reversed(range(n_flows)):
        x = torch.flip(x, [1])                                   # Flip
        x0, x1 = x[:, :half], x[:, half:]
        pre = f"flow.flows.{2 * i}"
        h = F.conv1d(x0, W[pre + ".pre.weight"], W[pre + ".pre.bias"]) * mask
        h = wavenet(W, pre + ".enc", h, mask, g, hidden, kernel, n_layers)
        m = F.conv1d(h, W[pre + ".post.weight"], W[pre + ".post.bias"]) * mask   # mean_only
        x1 = (x1 - m) * mask
        x = torch.cat([x0, x1], 1)
    return x


def synthesizer_front(W, cfg: dict, phone, lengths, pitch, sid, noise_zp):
    """Everything ``Synthesizer.infer`` computes before ``self.dec`` (rate=None): returns (z * mask, g, mask, z_p, m_p, logs_p).
    ``noise_zp`` (B, inter, T) = the values ``torch.randn_like(m_p)`` would have produced (synthesizers.py:228)."""
    g = W["emb_g.weight"][sid].unsqueeze(-1)
    m_p, logs_p, mask = text_encoder(W, phone, pitch, lengths, cfg["hidden_channels"], cfg["inter_channels"], cfg["n_heads"],
                                     cfg["n_layers"], cfg["kernel_size"])
    z_p = (m_p + torch.exp(logs_p) * noise_zp * 0.66666) * mask
    z = flow_reverse(W, z_p, mask, g, cfg["inter_channels"], cfg["hidden_channels"])
    return z * mask, g, mask, z_p, m_p, logs_p
