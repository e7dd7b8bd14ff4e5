"""net_g.infer throughput through rvc_b200.Synthesizer (front end + decoder in one rvcsyn_infer call) against the decoder
alone, device-resident inputs, CUDA events.  Usage on the GPU box: python tools/synth_bench.py [B] [T] > gpurun_out/synth.json"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import rvc_b200  # noqa: E402

SYN_48K = [1025, 32, 192, 192, 768, 2, 6, 3, 0, "1", [3, 7, 11], [[1, 3, 5]] * 3, [12, 10, 2, 2], 512, [24, 20, 4, 4], 109, 256, 48000]


def timed(fn, n):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 3000
    dev = torch.device("cuda", 0)
    torch.manual_seed(1234)
    m = rvc_b200.Synthesizer(*SYN_48K, use_f0=1, text_enc_hidden_dim=768, vocoder="HiFi-GAN").eval().to(dev)
    gen = torch.Generator().manual_seed(1)
    phone = torch.randn(B, T, 768, generator=gen).to(dev)
    pitch = torch.randint(1, 256, (B, T), generator=gen).to(dev)
    f0 = (100 + 100 * torch.rand(B, T, generator=gen)).to(dev)
    sid = torch.zeros(B, dtype=torch.long, device=dev)
    lens = torch.full((B,), T, dtype=torch.long, device=dev)
    nzp = torch.randn(B, 192, T, device=dev)
    nsrc = torch.randn(B, T * 480, device=dev)
    res = {"B": B, "T": T, "audio_s": B * T / 100.0}
    for mode in ("fast", "fp32"):
        ms_all = timed(lambda: m.infer(phone, lens, pitch, f0, sid, noise_zp=nzp, noise_src=nsrc, mode=mode), 3)
        x = torch.randn(B, 192, T, device=dev)
        g = torch.randn(B, 256, 1, device=dev)
        with torch.no_grad():
            ms_dec = timed(lambda: m.dec(x, f0, g, noise=nsrc, mode=mode), 3)
        # front end alone: stop before the decoder (debug stop 9 = flow done)
        m._handle.set_stop(9)
        ms_front = timed(lambda: m.infer(phone, lens, pitch, f0, sid, noise_zp=nzp, noise_src=nsrc, mode=mode), 3)
        m._handle.set_stop(0)
        res[mode] = {"infer_ms": ms_all, "decoder_ms": ms_dec, "front_ms": ms_front, "infer_audio_s_per_s": B * T / 100.0 / ms_all * 1e3,
                     "decoder_audio_s_per_s": B * T / 100.0 / ms_dec * 1e3, "launches": m.last_launch_count}
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
