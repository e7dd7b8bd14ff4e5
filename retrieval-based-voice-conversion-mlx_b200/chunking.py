"""Host-side chunk geometry, sharding and stitching around the decoder.

Restates only the arithmetic the reference's pipeline applies around ``net_g.infer``
(``rvc/infer/pipeline.py``): split a long utterance every ``x_center`` seconds when it is
longer than ``x_max`` (:441-454), give every chunk ``x_pad`` seconds of context on each side
(:458, :478-544), decode chunks independently, trim ``t_pad_tgt`` samples per side (:494, :527)
and concatenate (:545).  Chunks are the unit of data parallelism: they share no state, so they
are dealt to GPUs with no collective on the data path.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Sequence

import numpy as np

FRAMES_PER_SECOND = 100  # hop = sr/100 at every rate (rvc/configs/*.json "hop_length")


@dataclass(frozen=True)
class Chunk:
    index: int
    start: int  # first frame of the padded chunk, in frames of the PADDED utterance
    stop: int   # one past the last frame
    trim: int   # frames of context to drop on each side after decoding

    @property
    def frames(self) -> int:
        return self.stop - self.start


def plan_chunks(n_frames: int, x_pad: int = 1, x_query: int = 6, x_center: int = 38, x_max: int = 41,
                split_points: Sequence[int] | None = None) -> List[Chunk]:
    """Chunk plan for an utterance of ``n_frames`` (unpadded) frames.

    The reference pads the utterance by ``t_pad`` on both sides (pipeline.py:458) and, when the
    padded... unpadded+window signal exceeds ``t_max`` (:443), cuts at ``opt_ts``: one point per
    ``t_center`` (:446), each moved to the quietest sample within +-``t_query`` (:447-454).  The
    quietest-sample search needs the waveform; callers that have it pass ``split_points`` (frames,
    unpadded coordinates).  Without it the nominal multiples of ``t_center`` are used (synthetic
    latents have no waveform).  Chunk k covers padded frames [s_k, t_k + 2*pad) (:485-487), the last
    one [t_last, end) (:519-521); every decoded chunk is trimmed by ``pad`` frames per side.
    """
    pad = x_pad * FRAMES_PER_SECOND
    total = n_frames + 2 * pad
    if split_points is None:
        split_points = []
        if n_frames > x_max * FRAMES_PER_SECOND:
            split_points = list(range(x_center * FRAMES_PER_SECOND, n_frames, x_center * FRAMES_PER_SECOND))
    chunks: List[Chunk] = []
    s = 0
    for k, t in enumerate(split_points):
        if not (s < t < n_frames):
            raise ValueError("split points must be increasing and inside the utterance")
        chunks.append(Chunk(k, s, t + 2 * pad, pad))
        s = t
    chunks.append(Chunk(len(chunks), s, total, pad))
    return chunks


def shard_units(lengths: Sequence[int], n_ranks: int) -> List[List[int]]:
    """Deal independent units (chunks / utterances) to ranks: longest-processing-time first, ties and
    equal lengths in input order, so equal-length batches become contiguous equal shares."""
    order = sorted(range(len(lengths)), key=lambda i: (-lengths[i], i))
    load = [0] * n_ranks
    out: List[List[int]] = [[] for _ in range(n_ranks)]
    for i in order:
        r = min(range(n_ranks), key=lambda q: (load[q], q))
        out[r].append(i)
        load[r] += lengths[i]
    for r in range(n_ranks):
        out[r].sort()
    return out


def stitch(decoded: Sequence[np.ndarray], trims: Sequence[int], crossfade: int = 0) -> np.ndarray:
    """Join decoded chunks (1-D sample arrays).  ``trims`` = samples to drop per side.

    crossfade == 0 reproduces the reference exactly: trim then concatenate (pipeline.py:494,:527,:545).
    crossfade > 0 keeps ``crossfade`` extra samples of context on both sides of every joint and blends
    them with the reference's only crossfade convention, the sin^2 windows of the realtime path
    (rvc/realtime/core.py:383-398: fade_in = sin^2(pi/2 * linspace(0,1,N)), fade_out = 1 - fade_in).
    """
    if len(decoded) != len(trims):
        raise ValueError("one trim per chunk")
    if crossfade == 0:
        return np.concatenate([d[t: len(d) - t] if t else d for d, t in zip(decoded, trims)])
    if any(t < crossfade for t in trims[:-1]) or any(t < crossfade for t in trims[1:]):
        raise ValueError("crossfade longer than the available context")
    fade_in = np.sin(0.5 * np.pi * np.linspace(0.0, 1.0, 2 * crossfade, dtype=np.float32)) ** 2
    fade_out = 1.0 - fade_in
    # one pass into a preallocated buffer (growing the output by concatenation copied it once per chunk)
    n = 2 * crossfade
    spans = []
    for k, (d, t) in enumerate(zip(decoded, trims)):
        lo = t - (crossfade if k > 0 else 0)
        hi = len(d) - t + (crossfade if k < len(decoded) - 1 else 0)
        spans.append((lo, hi))
    total = sum(hi - lo for lo, hi in spans) - n * (len(decoded) - 1)
    out = np.empty(total, dtype=np.float32)
    pos = 0
    for k, (d, (lo, hi)) in enumerate(zip(decoded, spans)):
        seg = np.asarray(d[lo:hi], dtype=np.float32)
        if k == 0:
            out[: len(seg)] = seg
            pos = len(seg)
            continue
        out[pos - n: pos] = out[pos - n: pos] * fade_out + seg[:n] * fade_in
        out[pos: pos + len(seg) - n] = seg[n:]
        pos += len(seg) - n
    return out


def pinned(t):
    """Page-locked copy of a host tensor when a CUDA driver is present (staging buffers of the host-buffer calls);
    the tensor itself otherwise (CPU-only tests of the host logic)."""
    import torch

    return t.pin_memory() if torch.cuda.is_available() else t


def decode_chunks(model, x, f0, g, chunks: Sequence[Chunk], *, mode=None, noise=None, device=None,
                  rank: int = 0, world: int = 1):
    """Decode this rank's share of ``chunks`` of one padded utterance in ONE batched call and return
    ``{chunk.index: 1-D float32 waveform of chunk.frames * upp samples}`` (untrimmed).

    The reference walks ``opt_ts`` sequentially and calls ``net_g.infer`` once per chunk
    (pipeline.py:478-544); chunks share no state, so here they are dealt to ``world`` ranks
    (``shard_units``: longest first) and each rank submits its chunks as one batch.  Chunks shorter than
    the rank's longest are zero-padded at the end: the decoder's receptive field (about 12 frames) is far
    below the ``t_pad`` context (100 frames) that ``stitch`` trims, so the kept samples do not see the padding.

    x (1, C, n) / f0 (1, n): CPU tensors of the padded utterance; g (1, gin, 1) or None.
    noise: None = in-kernel Philox draws through the host-buffer entry point (``decode_host``); or a CPU tensor
    (1, n * upp) of standard-normal draws, cut per chunk and injected through the device-pointer entry
    point (``forward``) - what the parity tests use."""
    import torch

    mine = [chunks[i] for i in shard_units([c.frames for c in chunks], world)[rank]]
    if not mine:
        return {}
    upp = model.upp
    B, Tmax = len(mine), max(c.frames for c in mine)
    xb = pinned(torch.zeros(B, x.shape[1], Tmax, dtype=torch.float32))
    fb = pinned(torch.zeros(B, Tmax, dtype=torch.float32))
    for k, c in enumerate(mine):
        xb[k, :, :c.frames] = x[0, :, c.start:c.stop]
        fb[k, :c.frames] = f0[0, c.start:c.stop]
    gb = None if g is None else g.reshape(1, -1, 1).expand(B, -1, 1).contiguous()
    if noise is None:
        y = model.decode_host(xb, fb, gb, mode=mode, device=device)
    else:
        nb = torch.zeros(B, Tmax * upp, dtype=torch.float32)
        for k, c in enumerate(mine):
            nb[k, :c.frames * upp] = noise[0, c.start * upp:c.stop * upp]
        dev = torch.device("cuda", torch.cuda.current_device() if device is None else device)
        with torch.no_grad():
            y = model(xb.to(dev), fb.to(dev), None if gb is None else gb.to(dev), noise=nb.to(dev), mode=mode).cpu()
    return {c.index: y[k, 0, :c.frames * upp].numpy() for k, c in enumerate(mine)}


def decode_long(model, x, f0, g, *, mode=None, crossfade: int = 0, noise=None, device=None, split_points=None,
                x_pad: int = 1, x_query: int = 6, x_center: int = 38, x_max: int = 41) -> np.ndarray:
    """Single-process long-form decode: plan the reference's ``t_pad``-overlap chunks for a PADDED utterance
    (x: (1, C, n_frames + 2 * pad)), decode them as one batch, trim and join (``stitch``)."""
    pad = x_pad * FRAMES_PER_SECOND
    chunks = plan_chunks(x.shape[2] - 2 * pad, x_pad, x_query, x_center, x_max, split_points)
    dec = decode_chunks(model, x, f0, g, chunks, mode=mode, noise=noise, device=device)
    return stitch([dec[c.index] for c in chunks], [c.trim * model.upp for c in chunks], crossfade)


def gather_chunks(local: dict, chunks: Sequence[Chunk], upp: int, crossfade: int = 0, dst: int = 0):
    """``torchrun`` deployments (one process per GPU): every rank passes the ``{chunk.index: waveform}`` dict its
    ``decode_chunks(rank=, world=)`` call returned; rank ``dst`` receives all of them (``gather_object`` - the only
    cross-rank step of the path, a result gather), stitches and returns the utterance, other ranks get None.
    Works on any initialised backend (nccl on the GPU box, gloo in the CPU tests); without a process group it just
    stitches ``local``."""
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        parts = [None] * dist.get_world_size() if dist.get_rank() == dst else None
        dist.gather_object(local, parts, dst=dst)
        if dist.get_rank() != dst:
            return None
        merged = {}
        for p in parts:
            merged.update(p)
    else:
        merged = dict(local)
    missing = [c.index for c in chunks if c.index not in merged]
    if missing:
        raise RuntimeError(f"chunks {missing} were decoded by no rank")
    return stitch([merged[c.index] for c in chunks], [c.trim * upp for c in chunks], crossfade)
