"""Checkpoint loader: build the B200 decoder from an extracted RVC model (``.pth``).

Restates what ``rvc/infer/infer.py:461-488`` does with a checkpoint written by
``rvc/train/process/extract_model.py:57-109``:

  * ``cpt["config"]`` is the POSITIONAL argument list of ``Synthesizer.__init__`` (synthesizers.py:40-58):
    [spec_channels, segment_size, inter_channels, hidden_channels, filter_channels, n_heads, n_layers, kernel_size,
     p_dropout, resblock, resblock_kernel_sizes, resblock_dilation_sizes, upsample_rates, upsample_initial_channel,
     upsample_kernel_sizes, spk_embed_dim, gin_channels, sr]; ``infer.py:472`` overwrites ``config[-3]`` with the
    row count of ``emb_g.weight``;
  * ``cpt["weight"]`` holds fp16 tensors (``value.half()``, extract_model.py:59-61) without ``enc_q``, weight-norm
    pairs under the legacy names ``*.weight_g`` / ``*.weight_v`` (extract_model.py:97-105);
  * ``cpt["f0"]`` (default 1), ``cpt["version"]`` (default "v1"), ``cpt["vocoder"]`` (default "HiFi-GAN") select the
    decoder class (synthesizers.py:84-135).

Only the decoder (``dec.*``) is built here - the path this repo accelerates.  Packed weights are cached by CONTENT
hash (sha256 over names, shapes and fp32 bytes of the decoder tensors + the decoder configuration): loading the
same voice again - the reference re-creates ``net_g`` on every model switch, infer.py:490-520 - re-uses the handle
that already holds the folded, repacked, device-resident images instead of folding and uploading 63 MB again.
"""
from __future__ import annotations

import hashlib
from collections import OrderedDict
from typing import Dict, Optional, Tuple

import torch

from . import _lib
from .decoder import HiFiGANGenerator, HiFiGANMRFGenerator, HiFiGANNSFGenerator
from .refinegan import RefineGANGenerator

CONFIG_FIELDS = ("spec_channels", "segment_size", "inter_channels", "hidden_channels", "filter_channels", "n_heads",
                 "n_layers", "kernel_size", "p_dropout", "resblock", "resblock_kernel_sizes", "resblock_dilation_sizes",
                 "upsample_rates", "upsample_initial_channel", "upsample_kernel_sizes", "spk_embed_dim", "gin_channels", "sr")

_LEGACY = ((".weight_g", ".parametrizations.weight.original0"), (".weight_v", ".parametrizations.weight.original1"))


def checkpoint_config(cpt: dict) -> Dict[str, object]:
    """Named view of ``cpt["config"]`` plus the selector keys, with the reference's defaults (infer.py:471-477)."""
    cfg = list(cpt["config"])
    if len(cfg) != len(CONFIG_FIELDS):
        raise ValueError(f'cpt["config"] has {len(cfg)} entries, expected {len(CONFIG_FIELDS)} (Synthesizer.__init__ positional order)')
    out = dict(zip(CONFIG_FIELDS, cfg))
    w = cpt.get("weight", {})
    if "emb_g.weight" in w:
        out["spk_embed_dim"] = int(w["emb_g.weight"].shape[0])  # infer.py:472
    out["use_f0"] = int(cpt.get("f0", 1))
    out["version"] = cpt.get("version", "v1")
    out["vocoder"] = cpt.get("vocoder", "HiFi-GAN")
    out["text_enc_hidden_dim"] = 768 if out["version"] == "v2" else 256
    return out


def decoder_state(weights: Dict[str, torch.Tensor], prefix: str = "dec.") -> "OrderedDict[str, torch.Tensor]":
    """The decoder's tensors out of a Synthesizer state dict: ``dec.`` prefix dropped, fp16 -> fp32 (``.float()``,
    infer.py:487), legacy weight-norm names mapped to the parametrised ones the modules register."""
    out: "OrderedDict[str, torch.Tensor]" = OrderedDict()
    for k, v in weights.items():
        if not k.startswith(prefix):
            continue
        name = k[len(prefix):]
        for old, new in _LEGACY:
            if name.endswith(old):
                name = name[: -len(old)] + new
        out[name] = v.detach().to(torch.float32)
    return out


def content_hash(cfg_args: Tuple, state: Dict[str, torch.Tensor]) -> str:
    h = hashlib.sha256()
    h.update(repr(cfg_args).encode())
    for k in sorted(state):
        t = state[k].detach().to(device="cpu", dtype=torch.float32).contiguous()
        h.update(k.encode())
        h.update(repr(tuple(t.shape)).encode())
        h.update(t.numpy().tobytes())
    return h.hexdigest()


class PackedCache:
    """LRU of library handles keyed by (content hash, device): a handle owns the folded + repacked weight images on
    its device (``rvcdec_load_weights``), so a hit skips the fold, the repack and the upload."""

    def __init__(self, capacity: int = 8):
        self.capacity = capacity
        self._d: "OrderedDict[Tuple[str, int], _lib.Handle]" = OrderedDict()
        self.hits = 0
        self.misses = 0

    def get(self, key):
        h = self._d.get(key)
        if h is not None:
            self._d.move_to_end(key)
            self.hits += 1
        else:
            self.misses += 1
        return h

    def put(self, key, handle):
        self._d[key] = handle
        self._d.move_to_end(key)
        while len(self._d) > self.capacity:
            self._d.popitem(last=False)  # modules that still use the handle keep it alive; the last reference closes it

    def clear(self):
        self._d.clear()


PACKED = PackedCache()


def load_checkpoint(cpt: dict, device: Optional[torch.device] = None, strict: bool = True):
    """Build the decoder of an extracted model and fill it from ``cpt["weight"]``.

    Returns an ``HiFiGANNSFGenerator`` (f0 models, vocoder "HiFi-GAN"), ``HiFiGANMRFGenerator`` (f0 models, vocoder
    "MRF HiFi-GAN") or ``HiFiGANGenerator`` (no-f0 models) in eval
    mode, moved to ``device`` when given.  ``module.content_key`` carries the packed-weight cache key; the first
    forward on a device looks the handle up in ``loader.PACKED`` before folding anything."""
    c = checkpoint_config(cpt)
    voc = c["vocoder"]
    if voc not in ("HiFi-GAN", "MRF HiFi-GAN", "RefineGAN") or (voc != "HiFi-GAN" and not c["use_f0"]):
        raise NotImplementedError(f'vocoder "{voc}" with f0={c["use_f0"]} has no decoder (synthesizers.py:85-135): "HiFi-GAN" (with or '
                                  'without f0), "MRF HiFi-GAN" and "RefineGAN" (f0 models) are supported')
    if voc == "RefineGAN":  # synthesizers.py:98-106
        m = RefineGANGenerator(sample_rate=c["sr"], downsample_rates=list(c["upsample_rates"])[::-1], upsample_rates=list(c["upsample_rates"]),
                               start_channels=16, num_mels=c["inter_channels"])
        missing, unexpected = m.load_state_dict(decoder_state(cpt["weight"]), strict=False)
        if strict and (missing or unexpected):
            raise KeyError(f"decoder state mismatch: missing {list(missing)[:4]}..., unexpected {list(unexpected)[:4]}...")
        m.eval()
        return m.to(device) if device is not None else m
    args = (c["inter_channels"], list(c["resblock_kernel_sizes"]), [list(d) for d in c["resblock_dilation_sizes"]],
            list(c["upsample_rates"]), c["upsample_initial_channel"], list(c["upsample_kernel_sizes"]))
    if voc == "MRF HiFi-GAN":  # synthesizers.py:86-97
        m = HiFiGANMRFGenerator(in_channel=c["inter_channels"], upsample_initial_channel=c["upsample_initial_channel"],
                                upsample_rates=list(c["upsample_rates"]), upsample_kernel_sizes=list(c["upsample_kernel_sizes"]),
                                resblock_kernel_sizes=list(c["resblock_kernel_sizes"]),
                                resblock_dilations=[list(d) for d in c["resblock_dilation_sizes"]], gin_channels=c["gin_channels"],
                                sample_rate=c["sr"], harmonic_num=8)
    elif c["use_f0"]:
        m = HiFiGANNSFGenerator(*args, gin_channels=c["gin_channels"], sr=c["sr"])
    else:
        m = HiFiGANGenerator(*args, gin_channels=c["gin_channels"])
    sd = decoder_state(cpt["weight"])
    missing, unexpected = m.load_state_dict(sd, strict=False)
    if strict and (missing or unexpected):
        raise KeyError(f"decoder state mismatch: missing {list(missing)[:4]}..., unexpected {list(unexpected)[:4]}...")
    m.eval()
    m.set_content_key(content_hash(m._cfg_args, m._decoder_state()))
    if device is not None:
        m = m.to(device)
    return m
