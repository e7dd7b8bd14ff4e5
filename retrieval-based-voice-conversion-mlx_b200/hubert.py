"""HuBERT / ContentVec content embedder on the B200 library: drop-in for the reference's
``HubertModelWithFinalProj`` (rvc/lib/utils.py:31-34, a subclass of ``transformers.HubertModel``) as the inference
pipeline calls it (rvc/infer/pipeline.py:326-334)::

    feats = model(feats)["last_hidden_state"]
    feats = model.final_proj(feats[0]).unsqueeze(0) if version == "v1" else feats

The module holds the parameters under transformers' names (``feature_extractor.conv_layers.0.conv.weight``,
``encoder.pos_conv_embed.conv.parametrizations.weight.original1``, ``encoder.layers.11.final_layer_norm.bias``,
``final_proj.weight`` ...), so the published ``pytorch_model.bin`` of an embedder loads with ``load_state_dict``
(the older ``weight_g`` / ``weight_v`` names of the positional conv are mapped on load, as transformers does), and hands
them to ``include/rvchubert.h``.  Only the published base configuration is implemented.  There is no CPU fallback: CPU
tensors raise.
"""
from __future__ import annotations

import json
import os
from typing import Dict, Optional

import torch
from torch import nn

from . import _lib

CONV_DIM, HIDDEN, LAYERS, HEADS, FFN, PROJ, POS_K, POS_GROUPS = 512, 768, 12, 12, 3072, 256, 128, 16
CONV_KERNEL = (10, 3, 3, 3, 3, 2, 2)
CONV_STRIDE = (5, 2, 2, 2, 2, 2, 2)

_BASE_CONFIG = {
    "hidden_size": HIDDEN, "num_hidden_layers": LAYERS, "num_attention_heads": HEADS, "intermediate_size": FFN,
    "conv_dim": [CONV_DIM] * 7, "conv_kernel": list(CONV_KERNEL), "conv_stride": list(CONV_STRIDE), "conv_bias": False,
    "feat_extract_norm": "group", "feat_extract_activation": "gelu", "hidden_act": "gelu", "do_stable_layer_norm": False,
    "num_conv_pos_embeddings": POS_K, "num_conv_pos_embedding_groups": POS_GROUPS, "classifier_proj_size": PROJ,
    "layer_norm_eps": 1e-5, "feat_proj_layer_norm": True,
}


def check_config(config: Optional[dict]) -> None:
    """Raise unless ``config`` (a config.json dict or an object with those attributes) is the base configuration."""
    if config is None:
        return
    get = config.get if isinstance(config, dict) else (lambda k, d=None: getattr(config, k, d))
    for k, want in _BASE_CONFIG.items():
        have = get(k, want)
        if isinstance(have, tuple):
            have = list(have)
        if have != want:
            raise NotImplementedError(f"the B200 HuBERT path implements the base configuration only: {k} = {have!r}, expected {want!r}")


class _FinalProj(nn.Linear):
    """``model.final_proj`` (rvc/lib/utils.py:34): a Linear whose CUDA forward runs on the library's GEMM."""

    def __init__(self, owner: "HubertModelWithFinalProj"):
        super().__init__(HIDDEN, PROJ)
        object.__setattr__(self, "_owner", owner)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not x.is_cuda:
            raise RuntimeError("rvc_b200.hubert runs only on a CUDA (sm_100a) device; there is no CPU fallback")
        owner = self._owner
        h = owner._ensure_handle(x.device)
        x2 = x.contiguous().float().view(-1, HIDDEN)
        out = torch.empty(x2.shape[0], PROJ, dtype=torch.float32, device=x.device)
        if x2.shape[0]:
            h.final_proj(x2.data_ptr(), out.data_ptr(), x2.shape[0], torch.cuda.current_stream(x.device).cuda_stream)
        return out.view(*x.shape[:-1], PROJ)


class HubertModelWithFinalProj(nn.Module):
    """``HubertModelWithFinalProj(config)`` (rvc/lib/utils.py:31-34).  ``config`` may be ``None`` (base configuration), a
    config.json dict or a ``transformers.HubertConfig``; anything but the base configuration raises."""

    def __init__(self, config=None):
        super().__init__()
        check_config(config)
        self.masked_spec_embed = nn.Parameter(torch.empty(HIDDEN).uniform_())  # training only; kept for load_state_dict
        fe = nn.Module()
        fe.conv_layers = nn.ModuleList()
        for i, k in enumerate(CONV_KERNEL):
            layer = nn.Module()
            layer.conv = nn.Conv1d(1 if i == 0 else CONV_DIM, CONV_DIM, k, CONV_STRIDE[i], bias=False)
            if i == 0:
                layer.layer_norm = nn.GroupNorm(CONV_DIM, CONV_DIM, affine=True)
            fe.conv_layers.append(layer)
        self.feature_extractor = fe
        fp = nn.Module()
        fp.layer_norm = nn.LayerNorm(CONV_DIM, eps=1e-5)
        fp.projection = nn.Linear(CONV_DIM, HIDDEN)
        self.feature_projection = fp
        enc = nn.Module()
        pce = nn.Module()
        pce.conv = nn.utils.parametrizations.weight_norm(nn.Conv1d(HIDDEN, HIDDEN, POS_K, padding=POS_K // 2, groups=POS_GROUPS), name="weight", dim=2)
        enc.pos_conv_embed = pce
        enc.layer_norm = nn.LayerNorm(HIDDEN, eps=1e-5)
        enc.layers = nn.ModuleList()
        for _ in range(LAYERS):
            lyr = nn.Module()
            att = nn.Module()
            att.k_proj, att.v_proj, att.q_proj, att.out_proj = (nn.Linear(HIDDEN, HIDDEN) for _ in range(4))
            lyr.attention = att
            lyr.layer_norm = nn.LayerNorm(HIDDEN, eps=1e-5)
            ff = nn.Module()
            ff.intermediate_dense = nn.Linear(HIDDEN, FFN)
            ff.output_dense = nn.Linear(FFN, HIDDEN)
            lyr.feed_forward = ff
            lyr.final_layer_norm = nn.LayerNorm(HIDDEN, eps=1e-5)
            enc.layers.append(lyr)
        self.encoder = enc
        self.final_proj = _FinalProj(self)
        self._register_load_state_dict_pre_hook(self._map_old_names)
        self._handle: Optional[_lib.HubertHandle] = None
        self._sig = None
        self._workspace: Optional[torch.Tensor] = None
        self._options: Dict[str, int] = {}
        self.last_launch_count = 0

    def set_option(self, key: str, value: int) -> None:
        """``rvchubert_set_option``: "tc" (1, default: GEMMs on the tensor cores with fp16 operands and fp32 accumulation;
        0: fp32 CUDA-core GEMMs), "keep_layer0" (test hook)."""
        self._options[key] = int(value)
        if self._handle is not None:
            self._handle.set_option(key, int(value))

    @staticmethod
    def _map_old_names(state_dict, prefix, *args):
        """checkpoints written before torch's parametrised weight norm name the positional conv's tensors weight_g / weight_v"""
        p = prefix + "encoder.pos_conv_embed.conv."
        for old, new in (("weight_g", "parametrizations.weight.original0"), ("weight_v", "parametrizations.weight.original1")):
            if p + old in state_dict:
                state_dict[p + new] = state_dict.pop(p + old)

    @classmethod
    def from_pretrained(cls, path: str) -> "HubertModelWithFinalProj":
        """Directory with ``config.json`` and ``pytorch_model.bin`` (what rvc/lib/utils.py:106-170 downloads per embedder)."""
        cfg = None
        if os.path.exists(os.path.join(path, "config.json")):
            with open(os.path.join(path, "config.json")) as f:
                cfg = json.load(f)
        model = cls(cfg)
        sd = torch.load(os.path.join(path, "pytorch_model.bin"), map_location="cpu", weights_only=True)
        missing, unexpected = model.load_state_dict(sd, strict=False)
        missing = [k for k in missing if not k.startswith("final_proj.") and k != "masked_spec_embed"]
        if missing:
            raise RuntimeError(f"{path}: missing tensors {missing[:5]} ...")
        return model.eval()

    # ------------------------------------------------------------------
    def _apply(self, fn, *args, **kwargs):
        self._sig_tensors = None  # .to() / .cuda() may replace tensor objects
        return super()._apply(fn, *args, **kwargs)

    def _tensor_sig(self):
        """(version, address) of every parameter / buffer: the handle re-uploads when one changed.  The tensor list is cached
        (walking 100+ submodules costs ~1 ms per call, more than a short clip's GPU time)."""
        ts = getattr(self, "_sig_tensors", None)
        if ts is None:
            ts = self._sig_tensors = list(self.parameters()) + list(self.buffers())
        return tuple((v._version, v.data_ptr()) for v in ts)

    def _ensure_handle(self, dev: torch.device) -> "_lib.HubertHandle":
        sig = (dev.index or 0, self._tensor_sig())
        if self._handle is None or self._sig != sig:
            sd = self.state_dict()
            if self._handle is None or self._handle.device != (dev.index or 0):
                self._handle = _lib.HubertHandle(dev.index or 0)
                for k, v in self._options.items():
                    self._handle.set_option(k, v)
            state = {k: v.detach().to("cpu", torch.float32).contiguous() for k, v in sd.items() if v.dtype.is_floating_point}
            self._handle.load_weights(state)
            self._sig = sig
        return self._handle

    def _ws(self, need: int, dev: torch.device) -> torch.Tensor:
        if self._workspace is None or self._workspace.device != dev or self._workspace.numel() < need:
            self._workspace = None
            self._workspace = torch.empty(need, dtype=torch.uint8, device=dev)
        return self._workspace

    def forward(self, input_values: torch.Tensor, attention_mask=None, **kwargs) -> Dict[str, torch.Tensor]:
        """(B, n) 16 kHz waveform -> {"last_hidden_state": (B, T, 768)} (HubertModel.forward, modeling_hubert.py:889-966,
        with the arguments the reference passes: no mask, eval mode)."""
        if attention_mask is not None or any(v for v in kwargs.values()):
            raise NotImplementedError("the B200 HuBERT path implements model(input_values) as the reference calls it (pipeline.py:331)")
        if not input_values.is_cuda:
            raise RuntimeError("rvc_b200.hubert runs only on a CUDA (sm_100a) device; there is no CPU fallback")
        if input_values.dim() != 2:
            raise ValueError(f"expected (B, n) audio, got {tuple(input_values.shape)}")
        dev = input_values.device
        B, n = input_values.shape
        T = _lib.HubertHandle.frames(n)
        if T <= 0:
            raise ValueError(f"{n} samples are shorter than one frame (400 samples)")
        h = self._ensure_handle(dev)
        ws = self._ws(h.workspace_bytes(B, n), dev)
        audio = input_values.contiguous().float()
        out = torch.empty(B, T, HIDDEN, dtype=torch.float32, device=dev)
        h.forward(audio.data_ptr(), out.data_ptr(), ws, B, n, torch.cuda.current_stream(dev).cuda_stream)
        self.last_launch_count = h.launch_count()
        return {"last_hidden_state": out}
