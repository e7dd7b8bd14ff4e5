"""Bring-up diagnostic (not a pytest file): stage-by-stage max-abs error of the Synthesizer front end (rvcsyn_infer with
debug stops) against the oracle on the golden fixture.  Usage on the GPU box: python tests/synth_diag.py"""
import math
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import synth_oracle as SO  # noqa: E402
from test_synth import CFGD, _golden, _inputs, _module  # noqa: E402

d = _golden()
dev = torch.device("cuda", 0)
m = _module(d)
sd = m.state_dict()
W = SO.folded({k: v for k, v in sd.items() if not k.startswith("enc_q.")})
m = m.to(dev)
phone, lens, pitch, nsff0, sid, nzp, nsrc = _inputs(d)
B, T = phone.shape[:2]
args = [t.to(dev) for t in (phone, lens, pitch, nsff0, sid)]

# oracle intermediates of layer 0
x = F.linear(phone, W["enc_p.emb_phone.weight"], W["enc_p.emb_phone.bias"]) + W["enc_p.emb_pitch.weight"][pitch]
x = F.leaky_relu(x * math.sqrt(192), 0.1).transpose(1, -1)
mask = (torch.arange(T)[None, :] < lens[:, None]).unsqueeze(1).float()
x0 = x * mask
pre = "enc_p.encoder.attn_layers.0"
qkv = torch.cat([F.conv1d(x0, W[pre + f".conv_{n}.weight"], W[pre + f".conv_{n}.bias"]) for n in "qkv"], 1)
y = SO.attention(W, pre, x0, mask, 2, 10)
x1 = SO.layer_norm(W, "enc_p.encoder.norm_layers_1.0", x0 + y) * mask
y2 = SO.ffn(W, "enc_p.encoder.ffn_layers.0", x1, mask, 3)
x2 = SO.layer_norm(W, "enc_p.encoder.norm_layers_2.0", x1 + y2) * mask
zm, g, _, z_p, m_p, logs_p = SO.synthesizer_front(W, CFGD, phone, lens, pitch, sid, nzp)
exp = {1: ("x", x0), 2: ("qkv", qkv), 4: ("x", x1), 5: ("x", x2), 6: ("stats", torch.cat([m_p, logs_p], 1)), 7: ("xf", z_p), 9: ("xf", zm)}
h = m._ensure_handle(dev)
for stop, (name, ref) in exp.items():
    h.set_stop(stop)
    m.infer(*args, noise_zp=nzp.to(dev), noise_src=nsrc.to(dev), mode="fp32")
    dst = torch.empty(ref.shape, dtype=torch.float32, device=dev)
    h.debug_read(name, dst, m._workspace, B, T, torch.cuda.current_stream(dev).cuda_stream)
    torch.cuda.synchronize()
    got = dst.cpu()
    valid = mask.expand_as(ref) > 0 if name != "qkv" else torch.ones_like(ref, dtype=torch.bool)
    err = (got - ref).abs()
    print(f"stop {stop} {name}: max-abs {float(err.max()):.3e} (valid rows {float(err[valid].max()):.3e}), ref max {float(ref.abs().max()):.3f}")
h.set_stop(0)
