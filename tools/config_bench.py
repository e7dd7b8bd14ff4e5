"""Timing records of the other BASELINE.json configurations on ONE B200 (bench.py measures configs[3], the one the
metric is quoted on).  Prints one JSON line per configuration:

  configs[1]  RVCv2 40 kHz decoder, batch 16 x 10 s chunks, fast mode, device-resident inputs
  configs[2]  RVCv2 48 kHz, one 13.5 s utterance (T = 1350, the README benchmark shape): fp32 parity mode vs fast mode
  configs[4]  RVCv2 48 kHz, 5 min long-form audio cut into the reference's t_pad-overlap chunks (8 chunks), decoded as
              one batch through the host-buffer entry point and stitched on the host (trim + concat), end to end
              (chunk slicing into pinned memory, H2D, decode, D2H, stitch) - single GPU here; chunking.decode_chunks
              takes rank / world for the sharded form.

    python tools/config_bench.py [--reps 5]

Synthetic latents / f0, random-init weights (seed 1234); CUDA events for the device-resident numbers, wall clock for
the end-to-end long-form number (it contains host work by definition)."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from bench import synth_batch  # noqa: E402
from rvc_b200 import HiFiGANNSFGenerator, chunking  # noqa: E402


class Cfg:
    def __init__(self, rates, kernels, sr):
        self.initial_channel, self.gin_channels, self.sr = 192, 256, sr
        self.rates, self.kernels = rates, kernels
        self.upp = 1
        for r in rates:
            self.upp *= r


CFG48 = Cfg([12, 10, 2, 2], [24, 20, 4, 4], 48000)
CFG40 = Cfg([10, 10, 2, 2], [16, 16, 4, 4], 40000)


def build(cfg, dev):
    torch.manual_seed(1234)
    m = HiFiGANNSFGenerator(192, [3, 7, 11], [[1, 3, 5]] * 3, cfg.rates, 512, cfg.kernels, gin_channels=256, sr=cfg.sr).eval().to(dev)
    m.noise_source = "philox"
    return m


def timed(fn, reps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)

    # configs[1]
    m = build(CFG40, dev)
    x, f0, g = (t.to(dev) for t in synth_batch(CFG40, 16, 1000, 1234, pin=False))
    with torch.no_grad():
        ms = timed(lambda: m(x, f0, g, mode="fast"), a.reps)
    print(json.dumps({"config": "configs[1]: RVCv2 40kHz decoder, batch 16 x 10 s chunks, 1 B200, fast mode", "ms_per_batch": ms,
                      "audio_s_per_s": 160.0 / (ms * 1e-3), "tflops": 2 * 457.24672e6 * 16 * 1000 / (ms * 1e-3) / 1e12,
                      "launches": m.last_launch_count}))
    del m

    # configs[2]
    m = build(CFG48, dev)
    x, f0, g = (t.to(dev) for t in synth_batch(CFG48, 1, 1350, 1234, pin=False))
    rec = {"config": "configs[2]: RVCv2 48kHz decoder, one 13.5 s utterance (T=1350), fp32 parity mode vs fast mode"}
    with torch.no_grad():
        for mode in ("fp32", "fast"):
            ms = timed(lambda: m(x, f0, g, mode=mode), a.reps)
            rec[mode] = {"ms": ms, "audio_s_per_s": 13.5 / (ms * 1e-3), "launches": m.last_launch_count}
    print(json.dumps(rec))

    # configs[4]
    n_frames, pad = 30000, 100
    x, f0, g = synth_batch(CFG48, 1, n_frames + 2 * pad, 99, pin=True)
    chunks = chunking.plan_chunks(n_frames)
    for _ in range(2):
        y = chunking.decode_long(m, x, f0, g, mode="fast")
    torch.cuda.synchronize()
    ts = []
    for _ in range(a.reps):
        t0 = time.perf_counter()
        y = chunking.decode_long(m, x, f0, g, mode="fast")
        ts.append(time.perf_counter() - t0)
    assert y.shape == (n_frames * CFG48.upp,)
    best, mean = min(ts), sum(ts) / len(ts)
    print(json.dumps({"config": "configs[4]: RVCv2 48kHz, 5 min long-form, t_pad-overlap chunks decoded as one batch on 1 B200, host trim+concat, end to end",
                      "chunks": len(chunks), "frames_per_chunk": [c.frames for c in chunks], "wall_s_mean": mean, "wall_s_best": best,
                      "audio_s_per_s_mean": 300.0 / mean, "audio_s_per_s_best": 300.0 / best,
                      "includes": "chunk slicing into pinned memory, H2D, decode, D2H, stitch"}))


if __name__ == "__main__":
    main()
