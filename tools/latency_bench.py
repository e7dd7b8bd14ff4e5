"""Single-utterance latency (BASELINE configs[2] shape: B = 1, T = 1350 = 13.5 s; and a 30 s chunk) with and without
CUDA-graph replay of the launch sequence.  Usage on the GPU box: python tools/latency_bench.py > gpurun_out/latency.json

  device : rvcdec_forward through the module, inputs resident, CUDA events around 50 calls
  host   : rvcdec_forward_host (pinned host buffers in and out), wall clock around 50 calls - the reference's per-chunk
           call (rvc/infer/pipeline.py:366-371)
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from bench import build_model, synth_batch  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    model, cfg = build_model(dev)
    model = model.to(dev)
    res = []
    for T in (1350, 3000, 500):
        x_h, f0_h, g_h = synth_batch(cfg, 1, T, seed=5, pin=True)
        x, f0, g = x_h.to(dev), f0_h.to(dev), g_h.to(dev)
        out_h = torch.empty(1, 1, T * cfg.upp, dtype=torch.float32).pin_memory()
        for graphs in (0, 1):
            with torch.no_grad():
                model(x, f0, g, mode="fast")
                model._handle.set_option("graphs", graphs)
                for _ in range(5):
                    model(x, f0, g, mode="fast")
                    model.decode_host(x_h, f0_h, g_h, mode="fast", out=out_h)
                torch.cuda.synchronize()
                n = 50
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(n):
                    model(x, f0, g, mode="fast")
                e1.record()
                torch.cuda.synchronize()
                dev_ms = e0.elapsed_time(e1) / n
                t0 = time.perf_counter()
                for _ in range(n):
                    model.decode_host(x_h, f0_h, g_h, mode="fast", out=out_h)
                host_ms = (time.perf_counter() - t0) / n * 1e3
                hs = {k: model._handle.stat(k) for k in ("graph_hits", "graph_captures", "graphs_cached", "host_setup_s", "host_enqueue_s", "host_wait_s")}
                # the same call straight through ctypes (no module bookkeeping)
                hd = model._handle
                t0 = time.perf_counter()
                for _ in range(n):
                    hd.forward_host(x_h.data_ptr(), f0_h.data_ptr(), g_h.data_ptr(), None, 1, out_h.data_ptr(), 1, T, 1, 0)
                raw_ms = (time.perf_counter() - t0) / n * 1e3
            res.append({"T": T, "audio_s": T / 100.0, "graphs": graphs, "launches": model.last_launch_count, "device_ms": dev_ms,
                        "host_ms": host_ms, "host_raw_ctypes_ms": raw_ms, "stats": hs, "device_audio_s_per_s": T / 100.0 / dev_ms * 1e3, "host_audio_s_per_s": T / 100.0 / host_ms * 1e3})
    print(json.dumps({"what": "single-utterance latency, fast mode, in-kernel noise, B=1", "runs": res}, indent=1))


if __name__ == "__main__":
    main()
