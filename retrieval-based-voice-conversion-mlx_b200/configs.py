"""Decoder configurations of the reference's v2 models (rvc/configs/{48000,40000,32000}.json ``model`` block,
positional order of ``HiFiGANNSFGenerator.__init__``, hifigan_nsf.py:75-86).  Product-side constants: bench.py
and the checkpoint loader read them here, nothing in the package imports ``oracle/``."""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Sequence


@dataclass(frozen=True)
class DecoderConfig:
    initial_channel: int = 192                      # "inter_channels"
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
        return (self.initial_channel, list(self.resblock_kernel_sizes), [list(d) for d in self.resblock_dilation_sizes],
                list(self.upsample_rates), self.upsample_initial_channel, list(self.upsample_kernel_sizes))

    def ctor_kwargs(self):
        return dict(gin_channels=self.gin_channels, sr=self.sr)


V2_48K = DecoderConfig()
V2_40K = DecoderConfig(upsample_rates=(10, 10, 2, 2), upsample_kernel_sizes=(16, 16, 4, 4), sr=40000)
V2_32K = DecoderConfig(upsample_rates=(10, 8, 2, 2), upsample_kernel_sizes=(20, 16, 4, 4), sr=32000)
BY_SR = {48000: V2_48K, 40000: V2_40K, 32000: V2_32K}

# Algorithmic MACs per input frame (10 ms of audio), biases excluded
MMAC_PER_FRAME = {48000: 550.760448e6, 40000: 457.24672e6, 32000: 384.442368e6}
