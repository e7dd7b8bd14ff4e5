"""RMVPE f0 predictor timing (rvc_b200.RMVPE0Predictor, random weights): infer_from_audio end to end from a host numpy array to
the host f0 array, and the stages on the device, for 5 s / 60 s / 300 s of 16 kHz audio (the reference README quotes
0.18 - 18.7 s (MLX) and 0.30 - 27.0 s (MPS) for 5 s - 5 min on an M3 Max).  python tools/rmvpe_bench.py > gpurun_out/rmvpe.json"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import rvc_b200  # noqa: E402


def ev(fn, n=3):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def main():
    dev = torch.device("cuda", 0)
    torch.manual_seed(1)
    p = rvc_b200.RMVPE0Predictor(None, dev)
    p.model.to(dev)
    out = {}
    for secs in (5, 60, 300):
        n = secs * 16000
        audio = (0.3 * np.sin(2 * np.pi * 200 * np.arange(n) / 16000) + 0.01 * np.random.default_rng(0).standard_normal(n)).astype(np.float32)
        p.infer_from_audio(audio)
        ts = []
        for _ in range(3):
            t0 = time.perf_counter()
            f0 = p.infer_from_audio(audio)
            ts.append(time.perf_counter() - t0)
        a = torch.from_numpy(audio).to(dev).unsqueeze(0)
        mel = p.mel_extractor(a)
        hid = p.mel2hidden(mel)
        out[f"{secs}s"] = {"frames": int(f0.shape[0]), "infer_from_audio_s": min(ts), "audio_s_per_s": secs / min(ts),
                           "mel_ms": ev(lambda: p.mel_extractor(a)), "model_ms": ev(lambda: p.mel2hidden(mel)),
                           "decode_ms": ev(lambda: p.decode(hid[0])), "launches_model": p.model.last_launch_count}
    # batched model call (the U-Net and the GRU clusters of different items run side by side): 8 x 30 s
    B, T = 8, 3008
    melb = torch.randn(B, 128, T, device=dev) * 2 - 4
    ms = ev(lambda: p.model(melb))
    out["batch_8x30s"] = {"B": B, "frames": T, "model_ms": ms, "audio_s_per_s": B * T / 100.0 / (ms * 1e-3)}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
