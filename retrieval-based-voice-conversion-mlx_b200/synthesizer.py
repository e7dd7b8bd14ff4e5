"""``Synthesizer`` with the reference's constructor, parameter names and ``infer`` signature
(``rvc/lib/algorithm/synthesizers.py:12-243``) whose ``infer`` runs entirely in librvcdec.so (``rvcsyn_infer``, include/rvcsyn.h):
speaker embedding, TextEncoder, z_p sampling, reverse flow and the decoder behind one call.

The torch sub-modules below only HOLD parameters - they are created in the reference's order with the reference's
initialisers, so a seeded construction yields the reference's state dict bit for bit (tests check per-tensor checksums),
``load_state_dict(cpt["weight"], strict=False)`` (infer.py:485) fills them unchanged and ``del net_g.enc_q`` (infer.py:484)
works.  There is no PyTorch fallback: ``infer`` raises on CPU tensors, ``forward`` (training) is not provided.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch.nn.utils.parametrizations import weight_norm

from . import _lib
from .decoder import HiFiGANGenerator, HiFiGANMRFGenerator, HiFiGANNSFGenerator
from .refinegan import RefineGANGenerator


class _LayerNorm(torch.nn.Module):  # normalization.py:4-26
    def __init__(self, channels: int, eps: float = 1e-5):
        super().__init__()
        self.eps = eps
        self.gamma = torch.nn.Parameter(torch.ones(channels))
        self.beta = torch.nn.Parameter(torch.zeros(channels))


class _MultiHeadAttention(torch.nn.Module):  # attentions.py:22-77
    def __init__(self, channels: int, out_channels: int, n_heads: int, p_dropout: float = 0.0, window_size: int = None):
        super().__init__()
        self.channels, self.out_channels, self.n_heads = channels, out_channels, n_heads
        self.k_channels = channels // n_heads
        self.window_size = window_size
        self.conv_q = torch.nn.Conv1d(channels, channels, 1)
        self.conv_k = torch.nn.Conv1d(channels, channels, 1)
        self.conv_v = torch.nn.Conv1d(channels, channels, 1)
        self.conv_o = torch.nn.Conv1d(channels, out_channels, 1)
        self.drop = torch.nn.Dropout(p_dropout)
        if window_size:
            rel_stddev = self.k_channels ** -0.5
            self.emb_rel_k = torch.nn.Parameter(torch.randn(1, 2 * window_size + 1, self.k_channels) * rel_stddev)
            self.emb_rel_v = torch.nn.Parameter(torch.randn(1, 2 * window_size + 1, self.k_channels) * rel_stddev)
        for c in (self.conv_q, self.conv_k, self.conv_v, self.conv_o):
            torch.nn.init.xavier_uniform_(c.weight)


class _FFN(torch.nn.Module):  # attentions.py:201-219
    def __init__(self, in_channels: int, out_channels: int, filter_channels: int, kernel_size: int, p_dropout: float = 0.0):
        super().__init__()
        self.conv_1 = torch.nn.Conv1d(in_channels, filter_channels, kernel_size)
        self.conv_2 = torch.nn.Conv1d(filter_channels, out_channels, kernel_size)
        self.drop = torch.nn.Dropout(p_dropout)


class _Encoder(torch.nn.Module):  # encoders.py:24-70
    def __init__(self, hidden_channels, filter_channels, n_heads, n_layers, kernel_size=1, p_dropout=0.0, window_size=10):
        super().__init__()
        self.hidden_channels, self.n_layers, self.window_size = hidden_channels, n_layers, window_size
        self.drop = torch.nn.Dropout(p_dropout)
        self.attn_layers = torch.nn.ModuleList([_MultiHeadAttention(hidden_channels, hidden_channels, n_heads, p_dropout=p_dropout,
                                                                    window_size=window_size) for _ in range(n_layers)])
        self.norm_layers_1 = torch.nn.ModuleList([_LayerNorm(hidden_channels) for _ in range(n_layers)])
        self.ffn_layers = torch.nn.ModuleList([_FFN(hidden_channels, hidden_channels, filter_channels, kernel_size, p_dropout=p_dropout)
                                               for _ in range(n_layers)])
        self.norm_layers_2 = torch.nn.ModuleList([_LayerNorm(hidden_channels) for _ in range(n_layers)])


class _TextEncoder(torch.nn.Module):  # encoders.py:104-125
    def __init__(self, out_channels, hidden_channels, filter_channels, n_heads, n_layers, kernel_size, p_dropout, embedding_dim, f0=True):
        super().__init__()
        self.hidden_channels, self.out_channels = hidden_channels, out_channels
        self.emb_phone = torch.nn.Linear(embedding_dim, hidden_channels)
        self.lrelu = torch.nn.LeakyReLU(0.1, inplace=True)
        self.emb_pitch = torch.nn.Embedding(256, hidden_channels) if f0 else None
        self.encoder = _Encoder(hidden_channels, filter_channels, n_heads, n_layers, kernel_size, p_dropout)
        self.proj = torch.nn.Conv1d(hidden_channels, out_channels * 2, 1)


class _WaveNet(torch.nn.Module):  # modules.py:18-82
    def __init__(self, hidden_channels, kernel_size, dilation_rate, n_layers, gin_channels=0, p_dropout=0):
        super().__init__()
        self.hidden_channels, self.n_layers, self.gin_channels = hidden_channels, n_layers, gin_channels
        self.kernel_size, self.dilation_rate = kernel_size, dilation_rate
        self.in_layers = torch.nn.ModuleList()
        self.res_skip_layers = torch.nn.ModuleList()
        self.drop = torch.nn.Dropout(p_dropout)
        if gin_channels:
            self.cond_layer = weight_norm(torch.nn.Conv1d(gin_channels, 2 * hidden_channels * n_layers, 1), name="weight")
        for i in range(n_layers):
            d = dilation_rate ** i
            self.in_layers.append(weight_norm(torch.nn.Conv1d(hidden_channels, 2 * hidden_channels, kernel_size, dilation=d,
                                                              padding=(kernel_size * d - d) // 2), name="weight"))
            rs = hidden_channels if i == n_layers - 1 else 2 * hidden_channels
            self.res_skip_layers.append(weight_norm(torch.nn.Conv1d(hidden_channels, rs, 1), name="weight"))


class _PosteriorEncoder(torch.nn.Module):  # encoders.py:161-182 (training only; infer.py:484 deletes it)
    def __init__(self, in_channels, out_channels, hidden_channels, kernel_size, dilation_rate, n_layers, gin_channels=0):
        super().__init__()
        self.out_channels = out_channels
        self.pre = torch.nn.Conv1d(in_channels, hidden_channels, 1)
        self.enc = _WaveNet(hidden_channels, kernel_size, dilation_rate, n_layers, gin_channels=gin_channels)
        self.proj = torch.nn.Conv1d(hidden_channels, out_channels * 2, 1)


class _Flip(torch.nn.Module):  # residuals.py:87-101 (folded into the coupling layers' weights by rvcsyn_load_weights)
    pass


class _ResidualCouplingLayer(torch.nn.Module):  # residuals.py:196-229
    def __init__(self, channels, hidden_channels, kernel_size, dilation_rate, n_layers, p_dropout=0, gin_channels=0, mean_only=False):
        super().__init__()
        self.half_channels = channels // 2
        self.mean_only = mean_only
        self.pre = torch.nn.Conv1d(self.half_channels, hidden_channels, 1)
        self.enc = _WaveNet(hidden_channels, kernel_size, dilation_rate, n_layers, p_dropout=p_dropout, gin_channels=gin_channels)
        self.post = torch.nn.Conv1d(hidden_channels, self.half_channels * (2 - mean_only), 1)
        self.post.weight.data.zero_()
        self.post.bias.data.zero_()


class _ResidualCouplingBlock(torch.nn.Module):  # residuals.py:118-151
    def __init__(self, channels, hidden_channels, kernel_size, dilation_rate, n_layers, n_flows=4, gin_channels=0):
        super().__init__()
        self.n_flows, self.kernel_size, self.n_layers = n_flows, kernel_size, n_layers
        self.flows = torch.nn.ModuleList()
        for _ in range(n_flows):
            self.flows.append(_ResidualCouplingLayer(channels, hidden_channels, kernel_size, dilation_rate, n_layers,
                                                     gin_channels=gin_channels, mean_only=True))
            self.flows.append(_Flip())


class Synthesizer(torch.nn.Module):
    """B200-native ``Synthesizer`` (synthesizers.py:40-153): same positional constructor, same sub-module / parameter
    names; ``infer(phone, phone_lengths, pitch, nsff0, sid, rate=None)`` returns ``(o, x_mask, (z, z_p, m_p, logs_p))`` with
    ``o`` and ``x_mask`` as in the reference; the intermediates stay on the device inside the library, so the third element
    carries ``z * x_mask`` and ``None`` for the others unless they are asked for through the debug taps.

    Extra keyword-only arguments of ``infer``: ``noise_zp`` / ``noise_src`` inject the standard-normal draws of
    synthesizers.py:228 and hifigan.py:223 (default: drawn with the torch generator of the input's device in the reference's
    order - ``randn_like(m_p)``, then ``rand(1,1,1)`` and ``randn(B, T*upp, 1)``); ``mode`` as for the decoder."""

    def __init__(self, spec_channels: int, segment_size: int, inter_channels: int, hidden_channels: int, filter_channels: int,
                 n_heads: int, n_layers: int, kernel_size: int, p_dropout: float, resblock: str, resblock_kernel_sizes: list,
                 resblock_dilation_sizes: list, upsample_rates: list, upsample_initial_channel: int, upsample_kernel_sizes: list,
                 spk_embed_dim: int, gin_channels: int, sr: int, use_f0: bool, text_enc_hidden_dim: int = 768,
                 vocoder: str = "HiFi-GAN", randomized: bool = True, checkpointing: bool = False, **kwargs):
        super().__init__()
        if vocoder not in ("HiFi-GAN", "MRF HiFi-GAN", "RefineGAN") or (vocoder != "HiFi-GAN" and not use_f0):
            raise NotImplementedError(f'vocoder "{vocoder}" without pitch guidance has no decoder in the reference either (synthesizers.py:120-126)'
                                      if vocoder in ("MRF HiFi-GAN", "RefineGAN") else f'unknown vocoder "{vocoder}"')
        self.segment_size, self.use_f0, self.randomized = segment_size, use_f0, randomized
        self.enc_p = _TextEncoder(inter_channels, hidden_channels, filter_channels, n_heads, n_layers, kernel_size, p_dropout,
                                  text_enc_hidden_dim, f0=use_f0)
        if use_f0 and vocoder == "MRF HiFi-GAN":  # synthesizers.py:86-97
            self.dec = HiFiGANMRFGenerator(in_channel=inter_channels, upsample_initial_channel=upsample_initial_channel,
                                           upsample_rates=upsample_rates, upsample_kernel_sizes=upsample_kernel_sizes,
                                           resblock_kernel_sizes=resblock_kernel_sizes, resblock_dilations=resblock_dilation_sizes,
                                           gin_channels=gin_channels, sample_rate=sr, harmonic_num=8, checkpointing=checkpointing)
        elif use_f0 and vocoder == "RefineGAN":  # synthesizers.py:98-106
            self.dec = RefineGANGenerator(sample_rate=sr, downsample_rates=upsample_rates[::-1], upsample_rates=upsample_rates,
                                          start_channels=16, num_mels=inter_channels, checkpointing=checkpointing)
        elif use_f0:
            self.dec = HiFiGANNSFGenerator(inter_channels, resblock_kernel_sizes, resblock_dilation_sizes, upsample_rates,
                                           upsample_initial_channel, upsample_kernel_sizes, gin_channels=gin_channels, sr=sr,
                                           checkpointing=checkpointing)
        else:
            self.dec = HiFiGANGenerator(inter_channels, resblock_kernel_sizes, resblock_dilation_sizes, upsample_rates,
                                        upsample_initial_channel, upsample_kernel_sizes, gin_channels=gin_channels)
        self.enc_q = _PosteriorEncoder(spec_channels, inter_channels, hidden_channels, 5, 1, 16, gin_channels=gin_channels)
        self.flow = _ResidualCouplingBlock(inter_channels, hidden_channels, 5, 1, 3, gin_channels=gin_channels)
        self.emb_g = torch.nn.Embedding(spk_embed_dim, gin_channels)
        self._syn_args = dict(inter_channels=inter_channels, hidden_channels=hidden_channels, filter_channels=filter_channels,
                              n_heads=n_heads, n_layers=n_layers, kernel_size=kernel_size, text_dim=text_enc_hidden_dim,
                              spk_embed_dim=spk_embed_dim, gin_channels=gin_channels, use_f0=1 if use_f0 else 0, window_size=10,
                              flow_kernel=5, flow_layers=3, n_flows=4)
        self.mode = getattr(self.dec, "mode", "fp32")
        self._handle: Optional[_lib.SynHandle] = None
        self._packed_key = None
        self._workspace: Optional[torch.Tensor] = None
        self.last_launch_count = 0

    # -- reference protocol ------------------------------------------------
    def remove_weight_norm(self):
        """synthesizers.py:155-157 looks for legacy WeightNorm hooks and finds none on parametrised modules; the decoder's
        own ``remove_weight_norm`` folds (see decoder.py).  Nothing to do for the library: it folds at load time."""
        return None

    def __prepare_scriptable__(self):
        return self

    def forward(self, *args, **kwargs):
        raise RuntimeError("rvc_b200.Synthesizer is inference-only: call infer() (training runs the reference's modules)")

    # -- packed weights ----------------------------------------------------
    def _state_for_library(self):
        sd = {k: v for k, v in self.state_dict().items() if not k.startswith("enc_q.")}
        if isinstance(self.dec, RefineGANGenerator):  # front end only: the decoder has its own handle (include/rvcrefine.h)
            return {k: v for k, v in sd.items() if not k.startswith("dec.")}
        if isinstance(self.dec, HiFiGANMRFGenerator):  # the library knows the decoder's layers under the NSF names
            sd = {k: v for k, v in sd.items() if not k.startswith("dec.")}
            for k, v in self.dec._decoder_state().items():
                sd["dec." + k] = v
        if not self.use_f0:  # the library receives all-zero source / noise-conv layers for the no-f0 decoder (decoder.py)
            for k, v in self.dec._decoder_state().items():
                sd["dec." + k] = v
        return sd

    def _weights_key(self):
        return tuple((p.data_ptr(), p._version, p.device.index) for n, p in self.named_parameters() if not n.startswith("enc_q."))

    def _ensure_handle(self, device: torch.device) -> _lib.SynHandle:
        idx = device.index if device.index is not None else torch.cuda.current_device()
        if self._handle is not None and self._handle.device != idx:
            self._handle = None
        if self._handle is None:
            c = _lib.SynConfig()
            for k, v in self._syn_args.items():
                setattr(c, k, int(v))
            if not isinstance(self.dec, RefineGANGenerator):
                c.dec = _lib.make_config(*self.dec._cfg_args)  # (left zeroed = front end only for the RefineGAN decoder)
            self._handle = _lib.SynHandle(c, idx)
            self._packed_key = None
            if not self.use_f0:  # the handle's own decoder skips the source kernel and the noise adds (decoder.py)
                dec_ptr = _lib.C.c_void_p(_lib.lib().rvcsyn_decoder(self._handle._h))
                _lib.check(_lib.lib().rvcdec_set_option(dec_ptr, b"no_source", 1), "rvcdec_set_option")
        key = self._weights_key()
        if key != self._packed_key:
            self._handle.load_weights(self._state_for_library())
            self._packed_key = key
        return self._handle

    # -- inference ---------------------------------------------------------
    @torch.no_grad()
    def infer(self, phone: torch.Tensor, phone_lengths: torch.Tensor, pitch: Optional[torch.Tensor] = None,
              nsff0: Optional[torch.Tensor] = None, sid: torch.Tensor = None, rate: Optional[torch.Tensor] = None, *,
              noise_zp: Optional[torch.Tensor] = None, noise_src: Optional[torch.Tensor] = None,
              rand_ini: Optional[torch.Tensor] = None, mode: Optional[str] = None):
        if not phone.is_cuda:
            raise RuntimeError("rvc_b200.Synthesizer.infer runs only on a CUDA (sm_100a) device; there is no CPU fallback")
        if rate is not None:
            raise NotImplementedError("rate (head trimming, synthesizers.py:230-234) is not supported: trim the inputs instead")
        dev = phone.device
        B, T, D = phone.shape
        if isinstance(self.dec, RefineGANGenerator):
            return self._infer_refinegan(phone, phone_lengths, pitch, nsff0, sid, noise_zp, noise_src)
        m = _lib.MODES[mode or self.mode]
        h = self._ensure_handle(dev)
        upp = self.dec.upp
        inter = self._syn_args["inter_channels"]
        if noise_zp is None:  # reference draw order on this device's generator: synthesizers.py:228, then hifigan.py:189,223
            noise_zp = torch.randn(B, inter, T, device=dev, dtype=torch.float32)
        H = getattr(self.dec, "harmonics", 1)
        if noise_src is None and self.use_f0:
            if H > 1:  # hifigan_mrf.py:129,172
                rand_ini = torch.rand(B, H, device=dev)
                noise_src = torch.randn(B, T * upp, H, device=dev, dtype=torch.float32)
            else:
                torch.rand(1, 1, 1, device=dev)
                noise_src = torch.randn(B, T * upp, 1, device=dev, dtype=torch.float32)
        ini = None
        if H > 1:
            if rand_ini is None:
                raise ValueError("injected source noise needs rand_ini (B, harmonics) for the MRF decoder")
            ini = rand_ini.reshape(B, H).to(dev).contiguous().float().clone()
            ini[:, 0] = 0
        lengths = phone_lengths.to(device=dev, dtype=torch.int32).contiguous()
        phone = phone.contiguous().float()
        pitch_p = pitch.to(device=dev, dtype=torch.int64).contiguous() if (self.use_f0 and pitch is not None) else None
        f0 = nsff0.contiguous().float() if (self.use_f0 and nsff0 is not None) else torch.zeros(B, T, device=dev)
        sid = sid.to(device=dev, dtype=torch.int64).contiguous()
        nz = noise_zp.reshape(B, inter, T).contiguous().float()
        ns = noise_src.reshape(B, T * upp * H).contiguous().float() if noise_src is not None else None
        need = h.workspace_bytes(B, T, m)
        if self._workspace is None or self._workspace.device != dev or self._workspace.numel() < need:
            self._workspace = None
            self._workspace = torch.empty(need, dtype=torch.uint8, device=dev)
        out = torch.empty(B, T * upp, dtype=torch.float32, device=dev)
        z = torch.empty(B, inter, T, dtype=torch.float32, device=dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        dec_ptr = _lib.C.c_void_p(_lib.lib().rvcsyn_decoder(h._h))
        _lib.check(_lib.lib().rvcdec_set_initial_phases(dec_ptr, ini.data_ptr() if ini is not None else None), "rvcdec_set_initial_phases")
        h.infer(phone.data_ptr(), lengths.data_ptr(), pitch_p.data_ptr() if pitch_p is not None else None, f0.data_ptr(),
                sid.data_ptr(), nz.data_ptr(), ns.data_ptr() if ns is not None else None, 0, out.data_ptr(), z.data_ptr(),
                self._workspace, B, T, m, stream)
        self.last_launch_count = h.launch_count()
        x_mask = (torch.arange(T, device=dev)[None, :] < phone_lengths.to(dev)[:, None]).unsqueeze(1).to(phone.dtype)
        return out.unsqueeze(1), x_mask, (z, None, None, None)

    def _infer_refinegan(self, phone, phone_lengths, pitch, nsff0, sid, noise_zp, noise_dec):
        """vocoder "RefineGAN": the library's front end (rvcsyn_infer on a front-end-only handle) hands z * x_mask to the
        RefineGAN decoder's own call (rvcrefine_forward).  ``noise_src`` of ``infer`` carries the decoder's flat draw
        buffer here (``dec.noise_numel``)."""
        dev = phone.device
        B, T, _ = phone.shape
        h = self._ensure_handle(dev)
        inter = self._syn_args["inter_channels"]
        if noise_zp is None:
            noise_zp = torch.randn(B, inter, T, device=dev, dtype=torch.float32)
        lengths = phone_lengths.to(device=dev, dtype=torch.int32).contiguous()
        phone = phone.contiguous().float()
        pitch_p = pitch.to(device=dev, dtype=torch.int64).contiguous()
        sid = sid.to(device=dev, dtype=torch.int64).contiguous()
        nz = noise_zp.reshape(B, inter, T).contiguous().float()
        need = h.workspace_bytes(B, T, 0)
        if self._workspace is None or self._workspace.device != dev or self._workspace.numel() < need:
            self._workspace = None
            self._workspace = torch.empty(need, dtype=torch.uint8, device=dev)
        z = torch.empty(B, inter, T, dtype=torch.float32, device=dev)
        h.infer(phone.data_ptr(), lengths.data_ptr(), pitch_p.data_ptr(), None, sid.data_ptr(), nz.data_ptr(), None, 0, None, z.data_ptr(),
                self._workspace, B, T, 0, torch.cuda.current_stream(dev).cuda_stream)
        g = self.emb_g.weight.detach()[sid].unsqueeze(-1)  # row gather of the speaker table (synthesizers.py:226)
        o = self.dec(z, nsff0.to(dev).float(), g, noise=noise_dec)
        self.last_launch_count = h.launch_count() + self.dec.last_launch_count
        x_mask = (torch.arange(T, device=dev)[None, :] < phone_lengths.to(dev)[:, None]).unsqueeze(1).to(phone.dtype)
        return o, x_mask, (z, None, None, None)
