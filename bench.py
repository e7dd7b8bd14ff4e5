"""bench.py - NSF-HiFiGAN decode throughput (audio-seconds per second) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--mode fast|fp32]

Workload (BASELINE.json configs[3], the configuration the metric is quoted on): RVC v2 48 kHz decoder,
30 s padded chunks (T = 3000 frames), 128 chunks per GPU per step - at 8 GPUs that is exactly the
"1024 x 30 s chunks sharded across 8 B200" case; per-GPU work is fixed as N grows (weak scaling), chunks
are independent so there is no data-path collective.  Synthetic latents / f0 contours, random-init
weights of the named architecture (torch.manual_seed(1234), the reference's default init).

One step = one pass of the decoder over this rank's 128 chunks.
  value : device-resident inputs, CUDA events on the launching stream, barrier + synchronize on both
          sides, max over ranks.  Inputs of a step are 295 MB (> the 126 MB L2) and every layer streams
          GBs of activations, so no L2 flush is needed between iterations (stated in config).
  e2e   : same metric through the host-buffer C-ABI call (rvcdec_forward_host via
          HiFiGANNSFGenerator.decode_host): pinned host inputs -> device, decode, waveform -> pinned host,
          every step.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

SR = 48000
T_FRAMES = 3000          # 30 s chunks
CHUNKS_PER_GPU = 128     # 1024 chunks / 8 GPUs
MMAC_PER_FRAME = 550.760448e6   # SURVEY.md 8a, 48 kHz v2 (algorithmic, biases excluded)
METRIC = "NSF-HiFiGAN decode audio-sec/sec @48kHz v2"


# ------------------------------------------------------------------ distributed helpers
def rank_units(n_units_per_rank: int, rank: int, world: int):
    """Global unit ids owned by `rank` (weak scaling: every rank owns n_units_per_rank independent chunks)."""
    return list(range(rank * n_units_per_rank, (rank + 1) * n_units_per_rank))


def _reduce(value: float, device: torch.device, op) -> float:
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=op)
    return float(t.item())


def max_over_ranks(value: float, device: torch.device) -> float:
    import torch.distributed as dist

    return _reduce(value, device, dist.ReduceOp.MAX)


def sum_over_ranks(value: float, device: torch.device) -> float:
    import torch.distributed as dist

    return _reduce(value, device, dist.ReduceOp.SUM)


def barrier():
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized():
        dist.barrier()


# ------------------------------------------------------------------ clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, smax, reasons = [], None, set()
        for ln in self.lines:
            f = [s.strip() for s in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax = float(f[2])
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------ model / inputs
def build_model(device):
    from rvc_b200 import HiFiGANNSFGenerator
    from rvc_b200.configs import V2_48K as cfg

    torch.manual_seed(1234)
    m = HiFiGANNSFGenerator(*cfg.ctor_args(), **cfg.ctor_kwargs()).eval()
    m.noise_source = "philox"
    return m, cfg


def synth_batch(cfg, B, T, seed, pin):
    """Synthetic latents / f0 of the named shape, generated on the host (so the e2e leg has real host buffers)."""
    gen = torch.Generator().manual_seed(seed)
    x = torch.randn(B, cfg.initial_channel, T, generator=gen)
    g = torch.randn(B, cfg.gin_channels, 1, generator=gen)
    t = torch.arange(T, dtype=torch.float32) / 100.0
    med = 90 + 210 * torch.rand(B, 1, generator=gen)
    f0 = med * torch.pow(2.0, (0.5 / 12) * torch.sin(2 * 3.14159265 * 5.0 * t + 6.28 * torch.rand(B, 1, generator=gen)))
    uv = (torch.rand(B, T // 50 + 1, generator=gen) < 0.25).repeat_interleave(50, dim=1)[:, :T]
    f0 = torch.where(uv, torch.zeros_like(f0), f0).contiguous()
    if pin:
        x, g, f0 = x.pin_memory(), g.pin_memory(), f0.pin_memory()
    return x, f0, g


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def measured_traffic(B, T, steps, spans):
    """DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture
    (profiles/roofline_traffic.json, written by tools/launch_share.py --traffic).  The capture runs fewer chunks
    than a bench launch covers (ncu replays every kernel ~40x); traffic of these kernels is linear in the number of
    rows, so it is scaled by frames per launch (bench) / frames per launch (capture).  A bench step may take several
    passes over the batch (workspace cap), hence frames per launch = B*T*steps*launches_per_pass / launches timed."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json")))
        frames_per_launch = B * T * steps * t["launches"] / float(spans)
        scale = frames_per_launch / float(t["chunks"] * t["frames_per_chunk"])
        return t["dram_bytes_per_launch"] * scale, (f"ncu dram__bytes_read+write per launch, {t['source']} ({t['chunks']} chunks per launch), "
                                                    f"x{scale:.3g} ({frames_per_launch:.0f} frames per launch here)")
    except Exception as e:  # noqa: BLE001
        return None, f"no capture: {e}"


def workload_config(B, T, mode):
    """The `config` object of the product arm."""
    return {"workload": f"RVCv2 48kHz HiFiGAN-NSF decoder, {B} x {T / 100:.0f} s padded chunks per GPU (BASELINE configs[3]: 1024 x 30 s over 8 GPUs), "
                        f"mode={mode}, random-init weights seed 1234, in-kernel Philox source noise",
            "workload_family": WORKLOAD_FAMILY,
            "chunks_per_gpu": B, "frames_per_chunk": T, "l2": "inputs (295 MB/step) and activations (GBs/layer) exceed the 126 MB L2; no flush needed"}


WORKLOAD_FAMILY = "RVCv2 48kHz HiFiGAN-NSF decoder forward, random-init weights seed 1234, audio-s/s (size-normalised)"


# ------------------------------------------------------------------ CPU arms
def _reference_module():
    """The UNMODIFIED reference class (rvc/lib/algorithm/generators/hifigan_nsf.py) from the live checkout or from
    oracle/_ref (staged by oracle/stage_ref.py); None when neither is there."""
    from oracle.stage_ref import ref_path

    rp = ref_path()
    if rp is None:
        return None, None
    if rp not in sys.path:
        sys.path.insert(0, rp)
    from rvc.lib.algorithm.generators.hifigan_nsf import HiFiGANNSFGenerator as Ref

    return Ref, rp


def _time_cpu(fn, warm, reps):
    times = []
    for i in range(warm + reps):
        t0 = time.perf_counter()
        fn()
        dt = time.perf_counter() - t0
        if i >= warm:
            times.append(dt)
    return times


def _ref_inputs(B, T, real_f0_path=None):
    from rvc_b200.configs import V2_48K as cfg

    x, f0, g = synth_batch(cfg, B, T, seed=1234, pin=False)
    if real_f0_path and os.path.exists(real_f0_path) and B == 1:
        import numpy as np

        real = torch.from_numpy(np.load(real_f0_path).astype("float32"))
        if real.numel() >= T:
            f0 = real[:T][None].contiguous()
    return x, f0, g


def cpu_reference_runs(steps: int, warm: int, full: bool):
    """The reference's own CPU implementation of the path, timed on this host.

    Step workload: BASELINE configs[0] - a single 5 s utterance (B = 1, T = 500), the module AS SHIPPED (weight-norm
    re-evaluated in every forward, rvc/infer/infer.py:479-488 never folds), all host threads.  (Batching does not help
    the CPU: B = 8 x 5 s measured 2.2 audio-s/s against 3.4 for B = 1 on the 8-thread build container; the batched leg
    is reported under `detail`.)  `full` adds BASELINE.md section 3's other legs once each: weight-norm folded
    (rvc/realtime/pipeline.py:405-415), T = 1350 on the real RMVPE contour, a batch of 4, and one thread.
    Falls back to the oracle port (kind "port") when the reference is not importable."""
    from rvc_b200.configs import V2_48K as cfg

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    Ref, rp = _reference_module()
    B, T = 1, 500
    detail = {}
    if Ref is None:
        from oracle import nsf_hifigan_oracle as O

        m, _ = build_model(None)
        w = O.folded_weights(m.state_dict())
        x, f0, g, noise = O.synth_inputs(O.CONFIG_48K, B, T, seed=1234)
        with torch.no_grad():
            times = _time_cpu(lambda: O.decoder_forward(w, O.CONFIG_48K, x, f0, g, noise), warm, steps)
        kind, what = "port", "oracle port (torch CPU fp32 restatement of hifigan_nsf.py:173-212, weight-norm folded)"
    else:
        torch.manual_seed(1234)
        ref = Ref(*cfg.ctor_args(), **cfg.ctor_kwargs()).eval()
        x, f0, g = _ref_inputs(B, T)
        with torch.no_grad():
            times = _time_cpu(lambda: ref(x, f0, g=g), warm, steps)
            kind = "reference"
            what = (f"UNMODIFIED reference HiFiGANNSFGenerator ({'live checkout' if rp == '/root/reference' else 'oracle/_ref staged copy'}), "
                    "as shipped (weight-norm re-evaluated per forward)")
            if full:
                real = os.path.join(rp, "ios_test_data", "rmvpe_f0.npy") if rp == "/root/reference" else os.path.join(rp, "rmvpe_f0.npy")
                legs = [("as_shipped_B4_T500_all_threads", 4, 500, cores), ("as_shipped_B1_T1350_real_contour_all_threads", 1, 1350, cores),
                        ("as_shipped_B1_T500_1_thread", 1, 500, 1)]
                for name, b, t, nt in legs:
                    torch.set_num_threads(nt)
                    xi, fi, gi = _ref_inputs(b, t, real)
                    dt = min(_time_cpu(lambda: ref(xi, fi, g=gi), 0 if nt == 1 else 1, 1))
                    detail[name] = {"audio_s_per_s": b * t / 100.0 / dt, "seconds": dt, "threads": nt}
                torch.set_num_threads(cores)
                # folded: what rvc/realtime/pipeline.py:405-415 does before inference
                from torch.nn.utils import parametrize

                for mod in list(ref.modules()):
                    if parametrize.is_parametrized(mod, "weight"):
                        parametrize.remove_parametrizations(mod, "weight", leave_parametrized=True)
                dt = min(_time_cpu(lambda: ref(x, f0, g=g), 1, 1))
                detail[f"folded_B{B}_T500_all_threads"] = {"audio_s_per_s": B * T / 100.0 / dt, "seconds": dt, "threads": cores}
                xi, fi, gi = _ref_inputs(1, 1350, real)
                dt = min(_time_cpu(lambda: ref(xi, fi, g=gi), 1, 1))
                detail["folded_B1_T1350_real_contour_all_threads"] = {"audio_s_per_s": 13.5 / dt, "seconds": dt, "threads": cores}
    sec = sum(times) / len(times)
    thr = B * T / 100.0 / sec
    sample = (f"{what}; one step = one forward over B={B} chunks of {T / 100:.0f} s (T={T} frames, BASELINE configs[0] shape, "
              f"all {cores} host threads), torch {torch.__version__} CPU fp32, mean of {len(times)} after {warm} warm-up: {sec:.2f} s per step")
    return {"value": thr, "unit": "audio-s/s", "cores": torch.get_num_threads(), "kind": kind, "sample": sample,
            "B": B, "T": T, "seconds_per_step": sec, "detail": detail}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = cpu_reference_runs(max(1, args.steps), max(0, min(args.warmup, 1)), full=True)
    cfg = {"workload": f"CPU arm: RVCv2 48kHz HiFiGAN-NSF decoder, {r['B']} x {r['T'] / 100:.0f} s utterance per step on the host cores "
                       f"(BASELINE configs[0]; a bounded sample of the product arm's workload family, NOT the 128 x 30 s GPU step), fp32, "
                       f"random-init weights seed 1234, torch RNG source noise",
           "workload_family": WORKLOAD_FAMILY, "chunks_per_step": r["B"], "frames_per_chunk": r["T"]}
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": "audio-s/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": r["seconds_per_step"] * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": cfg,
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "detail")},
        "e2e": {"value": r["value"], "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def longform_leg(model, cfg, n_dev, reps=3):
    """BASELINE configs[4]: 5 min of 48 kHz audio (30 000 frames + 2 x 100 frames of t_pad context) split with the
    reference's chunk geometry (x_center = 38 s -> 8 chunks of <= 4000 frames), decoded by the single-process runner
    (one worker thread + library handle + stream set per device, waveforms gathered by the D2H copies into one pinned
    buffer), trimmed and joined on the host.  End to end from host latents to the stitched host waveform: slicing into
    pinned staging, H2D, decode, D2H, stitch.  Reported for trim + concat (reference semantics) and with the sin^2
    crossfade; wall clock (it spans several devices and the host)."""
    from rvc_b200 import MultiGpuDecoder

    n_frames, pad = 30000, 100
    x, f0, g = synth_batch(cfg, 1, n_frames + 2 * pad, seed=4321, pin=False)
    dec = MultiGpuDecoder(model, devices=list(range(n_dev)), mode="fast")
    out = {}
    try:
        for name, cf in (("trim_concat", 0), ("crossfade_sin2_10ms", 480)):
            dec.decode_long(x, f0, g, crossfade=cf)  # warm-up: handles, staging buffers, scratch
            ts = []
            for _ in range(reps):
                t0 = time.perf_counter()
                y = dec.decode_long(x, f0, g, crossfade=cf)
                ts.append(time.perf_counter() - t0)
            assert y.shape == (n_frames * cfg.upp,) and bool((abs(y) <= 1.0).all())
            out[name] = {"seconds": min(ts), "audio_s_per_s": n_frames / 100.0 / min(ts)}
        st = dec.stats()
    finally:
        dec.close()
    return {"workload": "5 min utterance, 8 t_pad-overlap chunks (reference geometry x_pad=1, x_center=38, x_max=41), single-process runner",
            "n_devices": n_dev, "chunks": 8, "units_per_device": [s["units"] // (2 * (reps + 1)) for s in st], "best_of": reps, **out}


def rmvpe_leg(dev, secs=60, reps=3):
    """the RMVPE f0 predictor (rvc_b200.RMVPE0Predictor, random weights, default mode: the 3 x 3 convolutions
    from level 1 down on the tensor cores with fp16 operands) end to end from a host numpy array of 16 kHz audio to the host f0
    array: H2D, mel spectrogram, U-Net + BiGRU, decode, D2H.  Wall clock."""
    import numpy as np
    import rvc_b200

    torch.manual_seed(1)
    p = rvc_b200.RMVPE0Predictor(None, dev)
    p.model.to(dev)
    n = secs * 16000
    audio = (0.3 * np.sin(2 * np.pi * 200 * np.arange(n) / 16000) + 0.01 * np.random.default_rng(0).standard_normal(n)).astype(np.float32)
    f0 = p.infer_from_audio(audio)
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        f0 = p.infer_from_audio(audio)
        ts.append(time.perf_counter() - t0)
    assert f0.shape == (1 + n // 160,) and bool(np.isfinite(f0).all())
    return {"workload": f"RMVPE0Predictor.infer_from_audio, {secs} s of 16 kHz audio, host array -> host f0, random weights",
            "seconds": min(ts), "audio_s_per_s": secs / min(ts), "frames": int(f0.shape[0]), "launches_model": p.model.last_launch_count}


def hubert_leg(dev, secs=30, reps=3):
    """SURVEY 8f row 4, second half: the HuBERT / ContentVec embedder (rvc_b200.HubertModelWithFinalProj, random weights,
    default mode: GEMMs and attention on the tensor cores) from a pinned host waveform to the host features: H2D, conv feature
    extractor, 12 transformer layers, D2H.  Wall clock."""
    import rvc_b200

    torch.manual_seed(1)
    m = rvc_b200.HubertModelWithFinalProj().eval()
    n = secs * 16000
    host = (0.3 * torch.randn(1, n)).pin_memory()
    y = m(host.to(dev, non_blocking=True))["last_hidden_state"].cpu()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        y = m(host.to(dev, non_blocking=True))["last_hidden_state"].cpu()
        ts.append(time.perf_counter() - t0)
    assert y.shape[0] == 1 and y.shape[2] == 768 and bool(torch.isfinite(y).all())
    return {"workload": f"HubertModelWithFinalProj(audio)['last_hidden_state'], {secs} s of 16 kHz audio, host waveform -> host features, random weights",
            "seconds": min(ts), "audio_s_per_s": secs / min(ts), "frames": int(y.shape[1]), "launches": m.last_launch_count}


def synthesizer_leg(dev, B=16, T=3000, reps=3):
    """the whole net_g.infer (speaker embedding, TextEncoder, z_p sampling, reverse flow, decoder) through
    rvc_b200.Synthesizer -> rvcsyn_infer, device-resident inputs, CUDA events; fast mode (front-end convolutions and attention on the tensor cores)."""
    import rvc_b200

    SYN_48K = [1025, 32, 192, 192, 768, 2, 6, 3, 0, "1", [3, 7, 11], [[1, 3, 5]] * 3, [12, 10, 2, 2], 512, [24, 20, 4, 4], 109, 256, 48000]
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

    def timed(stop):
        m._handle.set_stop(stop)
        m.infer(phone, lens, pitch, f0, sid, noise_zp=nzp, noise_src=nsrc, mode="fast")
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            m.infer(phone, lens, pitch, f0, sid, noise_zp=nzp, noise_src=nsrc, mode="fast")
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    m.infer(phone, lens, pitch, f0, sid, noise_zp=nzp, noise_src=nsrc, mode="fast")
    ms_all, ms_front = timed(0), timed(9)
    m._handle.set_stop(0)
    return {"workload": f"net_g.infer, {B} x {T / 100:.0f} s, fast mode (decoder, front-end convolutions and attention on the tensor cores; softmax, LayerNorm, gate, z_p fp32), injected noise, device-resident inputs",
            "infer_ms": ms_all, "front_end_ms": ms_front, "front_end_share": ms_front / ms_all,
            "audio_s_per_s": B * T / 100.0 / ms_all * 1e3, "launches": m.last_launch_count}


def parity_check(model, cfg, dev):
    """One untimed 30 s chunk (the bench's chunk shape) with injected noise against the oracle: the numbers the
    north_star bars are stated on, measured in the same process as the throughput."""
    from oracle import nsf_hifigan_oracle as O

    w = O.folded_weights(model.state_dict())
    x, f0, g, noise = O.synth_inputs(O.CONFIG_48K, 1, T_FRAMES, seed=77)
    with torch.no_grad():
        y_ref = O.decoder_forward(w, O.CONFIG_48K, x, f0, g, noise)
        y32 = model(x.to(dev), f0.to(dev), g.to(dev), noise=noise.to(dev), mode="fp32").cpu()
        yf = model(x.to(dev), f0.to(dev), g.to(dev), noise=noise.to(dev), mode="fast").cpu()
    corr = O.multires_logmel_corr(yf, y_ref, cfg.sr)
    return {"fp32_max_abs": float((y32 - y_ref).abs().max()), "fast_min_corr": corr["min"], "fast_mean_corr": corr["mean"],
            "fast_max_abs": float((yf - y_ref).abs().max()), "bars": {"fp32_max_abs": 2e-3, "fast_min_corr": 0.999},
            "sample": "one 30 s chunk (B=1, T=3000), injected noise, vs the CPU oracle (pinned to reference-generated goldens)"}


def run_ours(args):
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback for the product arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # stdout carries exactly one JSON line: NCCL writes its "NCCL version ..." banner (NCCL_DEBUG=VERSION and above)
        # and any warning to stdout by default, so its log is pointed at stderr
        # (at level VERSION the banner is a bare printf to stdout; from WARN up it goes through NCCL's log file)
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=dev)
    B, T = args.chunks, args.frames
    model, cfg = build_model(dev)
    units = rank_units(B, rank, world)
    x_h, f0_h, g_h = synth_batch(cfg, B, T, seed=1234 + units[0], pin=True)
    x, f0, g = x_h.to(dev), f0_h.to(dev), g_h.to(dev)
    n = T * cfg.upp
    audio_s = B * T / 100.0

    # ---- device-resident timing --------------------------------------------------------
    with torch.no_grad():
        for _ in range(args.warmup):
            y = model(x, f0, g, mode=args.mode)
        h = model._handle
        h.profile_enable(True)
        sampler = ClockSampler(local)
        torch.cuda.synchronize(); barrier(); torch.cuda.synchronize()
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        launches = 0
        for _ in range(args.steps):
            y = model(x, f0, g, mode=args.mode)
            launches += model.last_launch_count
        e1.record()
        torch.cuda.synchronize(); barrier(); torch.cuda.synchronize()
        clocks = sampler.stop()
        ms = e0.elapsed_time(e1)
        prof = h.profile_read()
        h.profile_enable(False)
    ms_max = max_over_ranks(ms, dev)
    total_audio = sum_over_ranks(audio_s * args.steps, dev)
    value = total_audio / (ms_max / 1e3)
    assert bool(torch.isfinite(y).all()) and float(y.abs().max()) <= 1.0

    # ---- end-to-end through the host-buffer C ABI --------------------------------------
    out_h = torch.empty(B, 1, n, dtype=torch.float32, pin_memory=True)
    for _ in range(1):
        model.decode_host(x_h, f0_h, g_h, mode=args.mode, out=out_h, device=local)
    torch.cuda.synchronize(); barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        model.decode_host(x_h, f0_h, g_h, mode=args.mode, out=out_h, device=local)
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0, dev)
    barrier()
    e2e_value = total_audio / e2e_s
    h2d = x_h.numel() * 4 + f0_h.numel() * 4 + g_h.numel() * 4
    d2h = out_h.numel() * 4

    # ---- BASELINE configs[4]: long-form across all GPUs of the job, single-process runner on rank 0 -----------------
    longform = None
    if not args.no_longform:
        torch.cuda.empty_cache()
        if rank == 0 and torch.cuda.device_count() >= world:
            try:
                model._workspace = None
                torch.cuda.empty_cache()
                longform = longform_leg(model, cfg, world)
            except Exception as e:  # noqa: BLE001
                longform = {"error": str(e)}
        barrier()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    pk, pk_kind = peaks()
    layers = {}
    for p in prof:
        if p["ms"] <= 0:
            continue
        layers[p["name"]] = {
            "ms_per_step": p["ms"] / args.steps, "share": p["ms"] / ms,
            "tflops": p["flops"] / (p["ms"] * 1e-3) / 1e12, "hbm_gbs": p["bytes"] / (p["ms"] * 1e-3) / 1e9,
            "tensor_frac": p["flops"] / (p["ms"] * 1e-3) / 1e12 / pk["bf16_tflops_sustained"],
            "hbm_frac": p["bytes"] / (p["ms"] * 1e-3) / 1e9 / pk["hbm_gbs"],
        }
    dom = max((p for p in prof if p["name"].startswith("mrf")), key=lambda p: p["ms"])
    dom_l = layers[dom["name"]]
    spans = max(1, dom["spans"])
    if args.mode == "fp32":
        roof = {"bound": "tensor", "kernel": f"conv_f32_kernel ({dom['name']})", "achieved": dom_l["tflops"], "peak": pk["bf16_tflops_sustained"],
                "unit": "TFLOP/s", "frac": dom_l["tensor_frac"], "traffic": None, "peak_kind": pk_kind + " bf16 sustained (fp32 FFMA kernels cannot reach it)"}
    else:
        roof = {"bound": "tensor", "kernel": f"conv_mma_kernel ({dom['name']}: the stage's 18 ResBlock convs)", "achieved": dom_l["tflops"],
                "peak": pk["bf16_tflops_sustained"], "unit": "TFLOP/s", "frac": dom_l["tensor_frac"], "traffic": None,
                "peak_kind": pk_kind + " bf16 sustained (kernel timed inside a long step)",
                "avg_launch_ms": dom["ms"] / spans, "flops_per_launch": dom["flops"] / spans,
                "hbm_gbs": dom_l["hbm_gbs"], "hbm_frac": dom_l["hbm_frac"]}
    if args.mode != "fp32":
        roof["traffic"], roof["traffic_note"] = measured_traffic(B, T, args.steps, spans)
        roof["traffic_source"] = "scaled from the committed ncu --set full capture (not measured in this run)"
        roof["algorithmic_bytes_per_launch"] = dom["bytes"] / spans
    total_tflops = 2 * MMAC_PER_FRAME * T * B * args.steps / (ms * 1e-3) / 1e12
    cpu = parity = synth = rmv = hub = None
    if world == 1 and not args.no_cpu:
        r = cpu_reference_runs(1, 1, full=False)
        cpu = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
        parity = parity_check(model, cfg, dev)
        try:
            model._workspace = None
            torch.cuda.empty_cache()
            synth = synthesizer_leg(dev)
        except Exception as e:  # noqa: BLE001
            synth = {"error": str(e)}
        try:
            rmv = rmvpe_leg(dev)
        except Exception as e:  # noqa: BLE001
            rmv = {"error": str(e)}
        try:
            hub = hubert_leg(dev)
        except Exception as e:  # noqa: BLE001
            hub = {"error": str(e)}
    line = {
        "metric": METRIC, "value": value, "unit": "audio-s/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32" if args.mode == "fp32" else "bf16/f16 operands, f32 accumulate", "data": "synthetic",
        "config": workload_config(B, T, args.mode),
        "clocks": clocks, "gpu_launches": launches,
        "e2e": {"value": e2e_value, "unit": "audio-s/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "roofline": roof, "peaks": {"source": pk_kind, "hbm_gbs": pk["hbm_gbs"], "bf16_tflops_sustained": pk["bf16_tflops_sustained"],
                                    "bf16_tflops_burst": pk.get("bf16_tflops"),
                                    "note": "per-layer tensor_frac is against the sustained figure; a sub-millisecond launch (conv_pre) can exceed it, the burst figure bounds it"},
        "whole_decoder_tflops": total_tflops, "whole_decoder_tensor_frac": total_tflops / pk["bf16_tflops_sustained"],
        "layers": layers,
    }
    if cpu:
        line["cpu_baseline"] = cpu
    if parity:
        line["parity"] = parity
    if longform:
        line["longform"] = longform
    if synth:
        line["synthesizer"] = synth
    if rmv:
        line["rmvpe"] = rmv
    if hub:
        line["hubert"] = hub
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="fast", choices=["fast", "fp32", "fast_f16", "fast_bf16"])
    ap.add_argument("--chunks", type=int, default=CHUNKS_PER_GPU)
    ap.add_argument("--frames", type=int, default=T_FRAMES)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-longform", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
