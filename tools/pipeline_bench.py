"""The three device stages of one voice conversion chained as the reference's pipeline chains them (rvc/infer/pipeline.py
:300-375 for a clip shorter than t_max: one chunk), on the README's 13.5 s clip length - a MEASUREMENT of the published
"full pipeline" row (MLX 1.27 s, PyTorch-MPS 11.08 s for HuBERT + RMVPE + synthesis on an M3 Max), not a
product path: file I/O, resampling, index retrieval, chunking of long clips and the CLI are out of scope.

    audio 16 kHz, reflect-padded by 1 s per side (pipeline.py:458)
      -> RMVPE f0 (get_f0, :200-291: f0 -> coarse pitch 1..255 on the mel scale)
      -> HuBERT features (:326-334), x 2 nearest upsampling (:344), p_len (:348)
      -> net_g.infer (:367), crop of the padding (:527)

Random weights (no network for checkpoints).  Host waveform in, host waveform out, wall clock.
python tools/pipeline_bench.py [seconds] [batch] > gpurun_out/pipeline.json
With batch > 1 the same clip is converted `batch` times in one pass (the throughput form: every stage takes a batch).
The f0 predictor and the embedder both read the waveform and nothing of each other: the tool runs them on two CUDA streams
(the reference runs them one after the other; `sequential_pipeline_s` is that order here)."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import rvc_b200  # noqa: E402

SYN_48K = [1025, 32, 192, 192, 768, 2, 6, 3, 0, "1", [3, 7, 11], [[1, 3, 5]] * 3, [12, 10, 2, 2], 512, [24, 20, 4, 4], 109, 256, 48000]
WINDOW, SR, TGT_SR, X_PAD = 160, 16000, 48000, 1
F0_MEL_MIN, F0_MEL_MAX = 1127 * np.log(1 + 50 / 700), 1127 * np.log(1 + 1100 / 700)


def coarse_pitch(f0: torch.Tensor) -> torch.Tensor:
    """pipeline.py:282-289 on the device"""
    mel = 1127 * torch.log1p(f0 / 700)
    mel = torch.where(mel > 0, (mel - F0_MEL_MIN) * 254 / (F0_MEL_MAX - F0_MEL_MIN) + 1, mel)
    return torch.clamp(mel, 1, 255).round().long()


def main():
    secs = float(sys.argv[1]) if len(sys.argv) > 1 else 13.5
    NB = int(sys.argv[2]) if len(sys.argv) > 2 else 1
    dev = torch.device("cuda", 0)
    torch.manual_seed(1234)
    hubert = rvc_b200.HubertModelWithFinalProj().eval()
    rmvpe = rvc_b200.RMVPE0Predictor(None, dev)
    rmvpe.model.to(dev)
    net_g = rvc_b200.Synthesizer(*SYN_48K, use_f0=1, text_enc_hidden_dim=768, vocoder="HiFi-GAN").eval().to(dev)
    n = int(secs * SR)
    t = np.arange(n) / SR
    audio = (0.4 * np.sin(2 * np.pi * (180 * t + 4 * np.sin(2 * np.pi * 4 * t))) + 0.02 * np.random.default_rng(0).standard_normal(n)).astype(np.float32)
    sid = torch.zeros(NB, dtype=torch.long, device=dev)
    stages = {}

    s_f0, s_ft = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def convert(audio_np, record=False, overlap=True):
        def mark(name, t0):
            if record:
                torch.cuda.synchronize()
                stages[name] = (time.perf_counter() - t0) * 1e3
        t_pad, t_pad_tgt = SR * X_PAD, TGT_SR * X_PAD
        t0 = time.perf_counter()
        audio_pad = np.pad(audio_np, (t_pad, t_pad), mode="reflect")                       # pipeline.py:458
        p_len = audio_pad.shape[0] // WINDOW
        a = torch.from_numpy(audio_pad).to(dev, non_blocking=True).unsqueeze(0).expand(NB, -1).contiguous()
        mark("pad_h2d", t0)
        def f0_stage():
            hidden = rmvpe.mel2hidden(rmvpe.mel_extractor(a))                             # get_f0 (:200-246), RMVPE.py:497-513
            f0 = torch.stack([rmvpe.decode(hidden[i], 0.03)[:p_len] for i in range(NB)])
            return coarse_pitch(f0), f0                                                   # :282-291

        def feature_stage():
            feats = hubert(a)["last_hidden_state"]                                        # :331 (v2: no final_proj)
            return torch.nn.functional.interpolate(feats.permute(0, 2, 1), scale_factor=2).permute(0, 2, 1)  # :344

        if overlap and not record:
            cur = torch.cuda.current_stream(dev)
            s_f0.wait_stream(cur); s_ft.wait_stream(cur)
            with torch.cuda.stream(s_f0):
                pitch, pitchf = f0_stage()
            with torch.cuda.stream(s_ft):
                feats = feature_stage()
            cur.wait_stream(s_f0); cur.wait_stream(s_ft)
        else:
            t0 = time.perf_counter()
            pitch, pitchf = f0_stage()
            mark("rmvpe_f0", t0)
            t0 = time.perf_counter()
            feats = feature_stage()
            mark("hubert", t0)
        p = min(audio_pad.shape[0] // WINDOW, feats.shape[1])                             # :348
        feats, pitch, pitchf = feats[:, :p].contiguous(), pitch[:, :p].contiguous(), pitchf[:, :p].contiguous()
        t0 = time.perf_counter()
        lens = torch.full((NB,), p, device=dev).long()
        out = net_g.infer(feats.float(), lens, pitch, pitchf.float(), sid)[0][:, 0]       # :367
        y = out[:, t_pad_tgt:-t_pad_tgt].cpu().numpy()                                    # :527
        y = y[0] if NB == 1 else y
        mark("net_g_infer_d2h", t0)
        return y

    y = convert(audio)
    convert(audio)
    ts = []
    for _ in range(5):
        t0 = time.perf_counter()
        y = convert(audio)
        ts.append(time.perf_counter() - t0)
    seq = []
    for _ in range(3):
        t0 = time.perf_counter()
        convert(audio, overlap=False)
        seq.append(time.perf_counter() - t0)
    convert(audio, record=True)
    assert np.isfinite(y).all() and abs(y.shape[-1] - int(secs * TGT_SR)) <= 3 * 480, y.shape  # HuBERT yields 2 frames fewer than audio // 160, as in the reference
    best = min(ts)
    print(json.dumps({"clip_seconds": secs, "batch": NB, "output_samples": int(y.shape[-1]), "pipeline_s": best, "realtime_factor": NB * secs / best,
                      "sequential_pipeline_s": min(seq),
                      "stage_ms_with_syncs": stages,
                      "published_reference": {"mlx_s": 1.27, "pytorch_mps_s": 11.08, "hardware": "M3 Max", "source": "BASELINE.md / reference README.md:22-29"},
                      "data": "synthetic tone + noise, random weights"}, indent=1))


if __name__ == "__main__":
    main()
