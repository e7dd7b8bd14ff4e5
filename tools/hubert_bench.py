"""HuBERT / ContentVec embedder timing (rvc_b200.HubertModelWithFinalProj, seeded random weights): model(audio) on the device
for 5 s / 13.5 s / 30 s clips at B = 1 and 30 s at B = 16, end to end from a pinned host array to the host features at B = 1.
python tools/hubert_bench.py > gpurun_out/hubert.json"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import rvc_b200  # noqa: E402


def ev(fn, n=5):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def flops(B, n):
    """multiply-adds x 2 of one forward (convs 1-6, projection, positional conv, 12 layers incl. attention)"""
    L = n
    Ls = []
    for k, s in zip((10, 3, 3, 3, 3, 2, 2), (5, 2, 2, 2, 2, 2, 2)):
        L = (L - k) // s + 1
        Ls.append(L)
    T = Ls[-1]
    mac = Ls[0] * 512 * 10
    for L_, k in zip(Ls[1:], (3, 3, 3, 3, 2, 2)):
        mac += L_ * 512 * 512 * k
    mac += T * 512 * 768 + T * 768 * 48 * 128
    mac += 12 * (T * (4 * 768 * 768 + 2 * 768 * 3072) + 2 * T * T * 768)
    return 2.0 * B * mac, T


def main():
    dev = torch.device("cuda", 0)
    torch.manual_seed(1)
    m = rvc_b200.HubertModelWithFinalProj().eval()
    out = {}
    for B, secs in ((1, 5.0), (1, 13.5), (1, 30.0), (16, 30.0)):
        n = int(secs * 16000)
        host = (0.3 * torch.randn(B, n)).pin_memory()
        audio = host.to(dev)
        m.set_option("tc", 0)
        ms32 = ev(lambda: m(audio))
        m.set_option("tc", 1)
        ms = ev(lambda: m(audio))
        fl, T = flops(B, n)
        rec = {"B": B, "seconds": secs, "frames": T, "model_ms": ms, "audio_s_per_s": B * secs / (ms * 1e-3), "tflops": fl / (ms * 1e-3) / 1e12,
               "model_ms_fp32_mode": ms32, "launches": m.last_launch_count}
        if B == 1:
            ts = []
            for _ in range(3):
                t0 = time.perf_counter()
                y = m(host.to(dev, non_blocking=True))["last_hidden_state"].cpu()
                ts.append(time.perf_counter() - t0)
            rec["host_to_host_ms"] = min(ts) * 1e3
        out[f"B{B}_{secs}s"] = rec
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
