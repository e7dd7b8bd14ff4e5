"""Single-process multi-GPU runner for independent units (padded chunks / utterances).

BASELINE configs[4]: a long utterance is cut into ``t_pad``-overlap chunks (``chunking.plan_chunks`` restates
``rvc/infer/pipeline.py:441-458,478-545``), the chunks are decoded on every GPU of the box and the host trims and joins
them (``chunking.stitch``).  The reference walks the chunks sequentially through one ``net_g.infer`` per chunk; chunks
share no state, so here they are dealt longest-first to one WORKER PER DEVICE:

  * a worker owns one library handle (``rvcdec_create`` + ``rvcdec_load_weights`` on its device), hence its own
    packed weights, scratch, compute stream and two copy streams (``rvcdec_forward_host``); workers are threads and
    every call into the library releases the GIL, so the devices run concurrently;
  * each worker submits its share as ONE zero-padded batch from a pinned staging buffer and the library copies the
    waveforms straight into that worker's rows of ONE pinned output buffer shared by all workers - the gather is the
    D2H copy itself, there is no device-to-device traffic and no collective;
  * a worker that fails (launch error, out of memory, lost device) hands its units back: they are re-dealt to the
    workers that are still healthy, and the run only fails when no healthy worker is left.
"""
from __future__ import annotations

import ctypes as C
import threading
import time
from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence

import numpy as np
import torch

from . import _lib, chunking


@dataclass
class Unit:
    """One independent decode: ``frames`` frames starting at ``start`` of a shared (1, C, n) latent / (1, n) f0 pair."""
    index: int
    start: int
    frames: int


@dataclass
class _Worker:
    device: int
    handle: Optional[_lib.Handle] = None
    healthy: bool = True
    error: Optional[str] = None
    units_done: int = 0
    seconds: float = 0.0
    x_stage: Optional[torch.Tensor] = None
    f_stage: Optional[torch.Tensor] = None
    g_stage: Optional[torch.Tensor] = None
    tmp: Optional[torch.Tensor] = None      # pinned scratch of ragged shares (grow-only: a pinned allocation costs ~20 ms per 50 MB)
    nz_stage: Optional[torch.Tensor] = None


class MultiGpuDecoder:
    """One handle per device behind one ``decode_units`` call.

    ``module`` is an ``rvc_b200.HiFiGANNSFGenerator`` (or the no-f0 sibling) holding the weights on any device; its
    state dict is packed once per device, in parallel.  ``fail_hook(device, units)`` is a test seam: when it returns
    True for a worker's batch the worker behaves as if its launch had failed."""

    def __init__(self, module, devices: Optional[Sequence[int]] = None, mode: str = "fast",
                 fail_hook: Optional[Callable[[int, List[Unit]], bool]] = None):
        if devices is None:
            devices = list(range(torch.cuda.device_count()))
        if not devices:
            raise RuntimeError("MultiGpuDecoder needs at least one CUDA device (there is no CPU fallback)")
        self.mode = _lib.MODES[mode]
        self.upp = module.upp
        self.cin = module._cin
        self.gin = module._gin
        self.fail_hook = fail_hook
        self.workers = [_Worker(int(d)) for d in devices]
        state = {k: v.detach().cpu() for k, v in module._decoder_state().items()}
        cfg_args = module._cfg_args
        no_source = 1 if type(module).__name__ == "HiFiGANGenerator" else 0

        def setup(w: _Worker):
            try:
                w.handle = _lib.Handle(_lib.make_config(*cfg_args), w.device)
                if no_source:
                    w.handle.set_option("no_source", 1)
                w.handle.load_weights(state)
            except Exception as e:  # noqa: BLE001
                w.healthy, w.error = False, f"setup: {e}"

        ts = [threading.Thread(target=setup, args=(w,)) for w in self.workers]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        if not any(w.healthy for w in self.workers):
            raise RuntimeError("no device could be set up: " + "; ".join(f"cuda:{w.device}: {w.error}" for w in self.workers))
        self.requeued = 0
        self._out_flat: Optional[torch.Tensor] = None

    # ------------------------------------------------------------------
    def _run_share(self, w: _Worker, share: List[Unit], x, f0, g, out_flat: torch.Tensor, offsets: Dict[int, int], seed: int, noise=None):
        """Decode ``share`` on worker ``w`` as one zero-padded batch; waveforms land in ``out_flat`` at ``offsets``."""
        t0 = time.perf_counter()
        B, Tmax = len(share), max(u.frames for u in share)
        n = Tmax * self.upp
        if w.x_stage is None or tuple(w.x_stage.shape) != (B, self.cin, Tmax):
            w.x_stage = chunking.pinned(torch.zeros(B, self.cin, Tmax, dtype=torch.float32))
            w.f_stage = chunking.pinned(torch.zeros(B, Tmax, dtype=torch.float32))
            w.g_stage = chunking.pinned(torch.zeros(B, max(self.gin, 1), dtype=torch.float32))
        xb, fb, gb = w.x_stage, w.f_stage, w.g_stage
        same_len = all(u.frames == Tmax for u in share)
        for k, u in enumerate(share):
            xb[k, :, :u.frames] = x[0, :, u.start:u.start + u.frames]
            fb[k, :u.frames] = f0[0, u.start:u.start + u.frames]
            if u.frames < Tmax:
                xb[k, :, u.frames:] = 0
                fb[k, u.frames:] = 0
        if g is not None and self.gin:
            gb[:] = g.reshape(1, -1)
        nz_ptr = None
        if noise is not None:  # injected source noise (parity tests): (1, n_total * upp) standard-normal draws
            if w.nz_stage is None or w.nz_stage.numel() < B * n:
                w.nz_stage = chunking.pinned(torch.zeros(B * n, dtype=torch.float32))
            nb = w.nz_stage[: B * n].view(B, n)
            nb.zero_()
            for k, u in enumerate(share):
                nb[k, :u.frames * self.upp] = noise[0, u.start * self.upp:(u.start + u.frames) * self.upp]
            nz_ptr = nb.data_ptr()
        if self.fail_hook is not None and self.fail_hook(w.device, share):
            raise RuntimeError(f"injected failure on cuda:{w.device}")
        contiguous_out = same_len and all(offsets[share[k + 1].index] - offsets[share[k].index] == n for k in range(B - 1))
        if contiguous_out:  # the library's D2H copy is the gather
            dst = out_flat[offsets[share[0].index]: offsets[share[0].index] + B * n]
            w.handle.forward_host(xb.data_ptr(), fb.data_ptr(), gb.data_ptr() if (g is not None and self.gin) else None, nz_ptr,
                                  seed + share[0].index, dst.data_ptr(), B, Tmax, self.mode, 0)
        else:  # ragged share: decode into a pinned scratch, then copy each unit's valid samples
            if w.tmp is None or w.tmp.numel() < B * n:
                w.tmp = chunking.pinned(torch.empty(B * n, dtype=torch.float32))
            tmp = w.tmp[: B * n].view(B, n)
            w.handle.forward_host(xb.data_ptr(), fb.data_ptr(), gb.data_ptr() if (g is not None and self.gin) else None, nz_ptr,
                                  seed + share[0].index, tmp.data_ptr(), B, Tmax, self.mode, 0)
            for k, u in enumerate(share):
                o = offsets[u.index]
                out_flat[o: o + u.frames * self.upp] = tmp[k, : u.frames * self.upp]
        w.units_done += B
        w.seconds += time.perf_counter() - t0

    def decode_units(self, x: torch.Tensor, f0: torch.Tensor, g: Optional[torch.Tensor], units: Sequence[Unit],
                     out_flat: Optional[torch.Tensor] = None, seed: int = 0, noise: Optional[torch.Tensor] = None):
        """Decode every unit; returns ``(out_flat, offsets)`` with unit ``u``'s ``u.frames * upp`` samples at
        ``out_flat[offsets[u.index]:]``.  When ``out_flat`` is not given the decoder's own pinned buffer is used (grow-only,
        OVERWRITTEN by the next call: copy what must outlive it, as ``decode_long`` does)."""
        units = list(units)
        offsets, total = {}, 0
        for u in units:
            offsets[u.index] = total
            total += u.frames * self.upp
        if out_flat is None or out_flat.numel() < total:
            cached = getattr(self, "_out_flat", None)
            if cached is None or cached.numel() < total:
                cached = self._out_flat = chunking.pinned(torch.empty(total, dtype=torch.float32))
            out_flat = cached
        pending = units
        while pending:
            healthy = [w for w in self.workers if w.healthy]
            if not healthy:
                raise RuntimeError("every worker failed: " + "; ".join(f"cuda:{w.device}: {w.error}" for w in self.workers))
            shares = chunking.shard_units([u.frames for u in pending], len(healthy))
            failed: List[Unit] = []
            lock = threading.Lock()

            def work(w: _Worker, share: List[Unit]):
                try:
                    self._run_share(w, share, x, f0, g, out_flat, offsets, seed, noise)
                except Exception as e:  # noqa: BLE001 - any failure of this device: hand the units back
                    w.healthy, w.error = False, str(e)
                    with lock:
                        failed.extend(share)

            ts = []
            for w, idxs in zip(healthy, shares):
                if idxs:
                    t = threading.Thread(target=work, args=(w, [pending[i] for i in idxs]))
                    t.start()
                    ts.append(t)
            for t in ts:
                t.join()
            self.requeued += len(failed)
            pending = sorted(failed, key=lambda u: u.index)
        return out_flat, offsets

    def decode_long(self, x: torch.Tensor, f0: torch.Tensor, g: Optional[torch.Tensor], *, crossfade: int = 0,
                    split_points=None, x_pad: int = 1, x_query: int = 6, x_center: int = 38, x_max: int = 41,
                    seed: int = 0, noise: Optional[torch.Tensor] = None) -> np.ndarray:
        """Long-form decode of a PADDED utterance (x: (1, C, n_frames + 2*pad)) over all healthy devices:
        plan -> deal -> decode -> trim + concat (reference semantics) or sin^2 crossfade."""
        pad = x_pad * chunking.FRAMES_PER_SECOND
        chunks = chunking.plan_chunks(x.shape[2] - 2 * pad, x_pad, x_query, x_center, x_max, split_points)
        units = [Unit(c.index, c.start, c.frames) for c in chunks]
        out, offs = self.decode_units(x, f0, g, units, seed=seed, noise=noise)
        dec = [out[offs[c.index]: offs[c.index] + c.frames * self.upp].numpy() for c in chunks]
        return chunking.stitch(dec, [c.trim * self.upp for c in chunks], crossfade)

    def stats(self):
        return [{"device": w.device, "healthy": w.healthy, "units": w.units_done, "seconds": w.seconds, "error": w.error}
                for w in self.workers]

    def close(self):
        for w in self.workers:
            if w.handle is not None:
                w.handle.close()
                w.handle = None
