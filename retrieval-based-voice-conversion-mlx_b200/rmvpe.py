"""RMVPE f0 predictor on the B200 library: drop-in for ``rvc/lib/predictors/RMVPE.py``.

``E2E`` holds the parameters under the reference's names (``unet.encoder.layers.0.conv.0.conv.0.weight``,
``unet.decoder.layers.4.conv1.1.running_var``, ``fc.0.gru.weight_hh_l0_reverse``, ``fc.1.bias`` ...; RMVPE.py:13-339), so the
published ``rmvpe.pt`` loads with ``load_state_dict`` and a seeded construction consumes the RNG like the reference's;
``forward`` hands them to ``include/rvcrmvpe.h`` (BatchNorm folded at load).  ``RMVPE0Predictor`` mirrors RMVPE.py:420-541:
mel spectrogram -> reflect-pad to a multiple of 32 frames -> model -> decode, every stage on the GPU.  There is no CPU
fallback: CPU tensors raise.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Optional

import numpy as np
import torch
from torch import nn

from . import _lib

N_MELS, N_CLASS = 128, 360


def _conv_block_res(cin: int, cout: int, momentum: float = 0.01) -> nn.Module:
    """Parameter holder with ConvBlockRes' names: conv.{0,1,3,4}, shortcut (RMVPE.py:23-51)."""
    m = nn.Module()
    m.conv = nn.Sequential(nn.Conv2d(cin, cout, 3, 1, 1, bias=False), nn.BatchNorm2d(cout, momentum=momentum), nn.ReLU(),
                           nn.Conv2d(cout, cout, 3, 1, 1, bias=False), nn.BatchNorm2d(cout, momentum=momentum), nn.ReLU())
    if cin != cout:
        m.shortcut = nn.Conv2d(cin, cout, 1)
    return m


def _res_block(cin: int, cout: int, n_blocks: int) -> nn.Module:
    m = nn.Module()
    m.conv = nn.ModuleList([_conv_block_res(cin if j == 0 else cout, cout) for j in range(n_blocks)])
    return m


class E2E(nn.Module):
    """``E2E(n_blocks, n_gru, kernel_size, en_de_layers=5, inter_layers=4, in_channels=1, en_out_channels=16)`` (RMVPE.py:303-333).
    The library implements the published configuration ``E2E(4, 1, (2, 2))`` (RMVPE.py:431)."""

    def __init__(self, n_blocks: int = 4, n_gru: int = 1, kernel_size=(2, 2), en_de_layers: int = 5, inter_layers: int = 4,
                 in_channels: int = 1, en_out_channels: int = 16):
        super().__init__()
        if (n_blocks, n_gru, tuple(kernel_size), en_de_layers, inter_layers, in_channels, en_out_channels) != (4, 1, (2, 2), 5, 4, 1, 16):
            raise NotImplementedError("the B200 RMVPE path implements E2E(4, 1, (2, 2)) with the default layer counts (RMVPE.py:431)")
        unet = nn.Module()
        enc = nn.Module()
        enc.bn = nn.BatchNorm2d(in_channels, momentum=0.01)
        enc.layers = nn.ModuleList()
        cin, cout = in_channels, en_out_channels
        for _ in range(en_de_layers):
            enc.layers.append(_res_block(cin, cout, n_blocks))
            cin, cout = cout, cout * 2
        unet.encoder = enc
        inter = nn.Module()
        inter.layers = nn.ModuleList([_res_block(cin if i == 0 else cout, cout, n_blocks) for i in range(inter_layers)])
        unet.intermediate = inter
        dec = nn.Module()
        dec.layers = nn.ModuleList()
        c = cout
        for _ in range(en_de_layers):
            layer = nn.Module()
            layer.conv1 = nn.Sequential(nn.ConvTranspose2d(c, c // 2, 3, 2, 1, output_padding=1, bias=False),
                                        nn.BatchNorm2d(c // 2, momentum=0.01), nn.ReLU())
            layer.conv2 = nn.ModuleList([_conv_block_res(c if j == 0 else c // 2, c // 2) for j in range(n_blocks)])
            dec.layers.append(layer)
            c //= 2
        unet.decoder = dec
        self.unet = unet
        self.cnn = nn.Conv2d(en_out_channels, 3, 3, padding=1)
        gru = nn.Module()
        gru.gru = nn.GRU(3 * N_MELS, 256, num_layers=n_gru, batch_first=True, bidirectional=True)
        self.fc = nn.Sequential(gru, nn.Linear(512, N_CLASS), nn.Dropout(0.25), nn.Sigmoid())
        self._handle: Optional[_lib.RmvpeHandle] = None
        self._sig = None
        self._workspace: Optional[torch.Tensor] = None
        self._mel_basis: Optional[np.ndarray] = None
        self._options = {}
        self.last_launch_count = 0

    def set_option(self, key: str, value: int) -> None:
        """``rvcrmvpe_set_option``: "tc" (1, default: the deep levels' 3 x 3 convolutions on the tensor cores with fp16
        operands; 0: every layer in fp32)."""
        self._options[key] = int(value)
        if self._handle is not None:
            self._handle.set_option(key, int(value))

    # ------------------------------------------------------------------
    def set_mel_basis(self, basis) -> None:
        """(128, 513) filter bank of the mel front end (MelSpectrogram's ``mel_basis`` buffer, RMVPE.py:368-378)."""
        self._mel_basis = np.ascontiguousarray(np.asarray(basis, dtype=np.float32))
        self._sig = None

    def _apply(self, fn, *args, **kwargs):
        self._sig_tensors = None  # .to() / .cuda() replace the buffer objects
        return super()._apply(fn, *args, **kwargs)

    def _tensor_sig(self):
        """(version, address) of every parameter / buffer; the tensor list is cached (walking the 741-entry state dict costs
        5 ms per call, as much as the GPU needs for a short clip)."""
        ts = getattr(self, "_sig_tensors", None)
        if ts is None:
            ts = self._sig_tensors = list(self.parameters()) + list(self.buffers())
        return tuple((v._version, v.data_ptr()) for v in ts)

    def _ensure_handle(self, dev: torch.device) -> "_lib.RmvpeHandle":
        sig = (dev.index or 0, self._tensor_sig())
        if self._handle is None or self._sig != sig:
            if self._handle is None or self._handle.device != (dev.index or 0):
                self._handle = _lib.RmvpeHandle(dev.index or 0)
                for k, v in self._options.items():
                    self._handle.set_option(k, v)
            state = {k: v.detach().to("cpu", torch.float32).contiguous() for k, v in self.state_dict().items()
                     if v.dtype.is_floating_point}
            basis = self._mel_basis if self._mel_basis is not None else mel_filterbank()
            state["mel_basis"] = torch.from_numpy(np.ascontiguousarray(basis, dtype=np.float32))
            self._handle.load_weights(state)
            self._sig = sig
        return self._handle

    def _ws(self, need: int, dev: torch.device) -> torch.Tensor:
        if self._workspace is None or self._workspace.device != dev or self._workspace.numel() < need:
            self._workspace = None
            self._workspace = torch.empty(need, dtype=torch.uint8, device=dev)
        return self._workspace

    def forward(self, mel: torch.Tensor) -> torch.Tensor:
        """mel (B, 128, T), T a multiple of 32 -> salience (B, T, 360) (RMVPE.py:335-339)."""
        if not mel.is_cuda:
            raise RuntimeError("rvc_b200.rmvpe.E2E runs only on a CUDA (sm_100a) device; there is no CPU fallback")
        dev = mel.device
        B, M, T = mel.shape
        if M != N_MELS:
            raise ValueError(f"expected {N_MELS} mel bins, got {M}")
        h = self._ensure_handle(dev)
        ws = self._ws(h.workspace_bytes(B, T), dev)
        mel = mel.contiguous().float()
        out = torch.empty(B, T, N_CLASS, dtype=torch.float32, device=dev)
        h.forward(mel.data_ptr(), out.data_ptr(), ws, B, T, torch.cuda.current_stream(dev).cuda_stream)
        self.last_launch_count = h.launch_count()
        return out


def mel_filterbank(sr: int = 16000, n_fft: int = 1024, n_mels: int = N_MELS, fmin: float = 30.0, fmax: float = 8000.0) -> np.ndarray:
    """What ``librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax, htk=True)`` returns (RMVPE.py:368-376; Slaney-normalised
    triangles on the HTK mel scale), computed here because librosa is not a dependency."""
    hz2mel = lambda f: 2595.0 * np.log10(1.0 + np.asarray(f, dtype=np.float64) / 700.0)  # noqa: E731
    mel2hz = lambda m: 700.0 * (10.0 ** (np.asarray(m, dtype=np.float64) / 2595.0) - 1.0)  # noqa: E731
    fft_f = np.linspace(0.0, sr / 2.0, n_fft // 2 + 1)
    pts = mel2hz(np.linspace(hz2mel(fmin), hz2mel(fmax), n_mels + 2))
    width = np.diff(pts)
    ramps = pts[:, None] - fft_f[None, :]
    w = np.maximum(0.0, np.minimum(-ramps[:-2] / width[:-1, None], ramps[2:] / width[1:, None]))
    w *= (2.0 / (pts[2:] - pts[:-2]))[:, None]
    return w.astype(np.float32)


class RMVPE0Predictor:
    """``RMVPE0Predictor(model_path, device)`` (RMVPE.py:420-541).  ``model_path`` may also be a state dict or ``None``
    (random weights: tests, benchmarks)."""

    def __init__(self, model_path=None, device=None):
        self.device = torch.device(device if device is not None else "cuda")
        if self.device.type != "cuda":
            raise RuntimeError("rvc_b200.RMVPE0Predictor runs only on a CUDA (sm_100a) device; there is no CPU fallback")
        self.model = E2E(4, 1, (2, 2))
        if model_path is not None:
            ckpt = model_path if isinstance(model_path, dict) else torch.load(model_path, map_location="cpu", weights_only=True)
            self.model.load_state_dict(ckpt)
        self.model.eval()
        self._mel_ws: Optional[torch.Tensor] = None

    def mel_extractor(self, audio: torch.Tensor) -> torch.Tensor:
        """MelSpectrogram(128, 16000, 1024, 160, None, 30, 8000).forward(audio, center=True) (RMVPE.py:388-417): (B, n) -> (B, 128, 1 + n // 160)."""
        dev = audio.device
        h = self.model._ensure_handle(dev)
        B, n = audio.shape
        need = h.mel_workspace_bytes(B, n)
        if self._mel_ws is None or self._mel_ws.device != dev or self._mel_ws.numel() < need:
            self._mel_ws = torch.empty(need, dtype=torch.uint8, device=dev)
        audio = audio.contiguous().float()
        mel = torch.empty(B, N_MELS, 1 + n // 160, dtype=torch.float32, device=dev)
        h.mel(audio.data_ptr(), mel.data_ptr(), self._mel_ws, B, n, torch.cuda.current_stream(dev).cuda_stream)
        return mel

    def mel2hidden(self, mel: torch.Tensor, chunk_size: int = 32000) -> torch.Tensor:
        """RMVPE.py:445-482: reflect-pad the frames to a multiple of 32, run the model in chunks, crop."""
        n_frames = mel.shape[-1]
        mel = torch.nn.functional.pad(mel, (0, 32 * ((n_frames - 1) // 32 + 1) - n_frames), mode="reflect")
        outs = [self.model(mel[..., s: min(s + chunk_size, mel.shape[-1])]) for s in range(0, mel.shape[-1], chunk_size)]
        return torch.cat(outs, dim=1)[:, :n_frames]

    def decode(self, hidden: torch.Tensor, thred: float = 0.03) -> torch.Tensor:
        """salience (T, 360) on the device -> f0 (T,) in Hz on the device (RMVPE.py:484-541)."""
        dev = hidden.device
        h = self.model._ensure_handle(dev)
        hidden = hidden.contiguous().float()
        f0 = torch.empty(hidden.shape[0], dtype=torch.float32, device=dev)
        h.decode(hidden.data_ptr(), float(thred), f0.data_ptr(), hidden.shape[0], torch.cuda.current_stream(dev).cuda_stream)
        return f0

    def infer_from_audio(self, audio, thred: float = 0.03) -> np.ndarray:
        """RMVPE.py:497-513: 16 kHz mono audio (numpy) -> f0 per 10 ms frame (numpy)."""
        a = torch.from_numpy(np.asarray(audio)).float().to(self.device).unsqueeze(0)
        hidden = self.mel2hidden(self.mel_extractor(a))
        return self.decode(hidden[0], thred).cpu().numpy()
