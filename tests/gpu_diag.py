"""Bring-up diagnostic (not a pytest file): per-layer max-abs error of the CUDA path vs the oracle for
both back-ends, written to gpurun_out/diag.txt.  Usage on the GPU box: python tests/gpu_diag.py"""
import os
import sys
import traceback

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import torch  # noqa: E402

from test_gpu import layer_errors  # noqa: E402

if __name__ == "__main__":
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    dev = torch.device("cuda", 0)
    with open(os.path.join(ROOT, "gpurun_out", "diag.txt"), "w") as f:
        for name, mode in [("tiny_24k_B2_T40", "fp32"), ("v2_48k_B2_T24", "fp32"), ("v2_48k_B2_T24", "fast_f16"),
                           ("v2_48k_B2_T24", "fast"), ("v2_40k_B1_T32", "fast")]:
            try:
                errs = layer_errors(name, mode, dev)
                for k, (e, s) in errs.items():
                    line = f"{name:18s} {mode:9s} {k:12s} max-abs {e:.3e} (ref max {s:.3e})"
                    print(line); f.write(line + "\n")
            except Exception:
                msg = f"{name} {mode} FAILED:\n{traceback.format_exc()}"
                print(msg); f.write(msg + "\n")
                f.flush()
                if "CUDA" in msg or "cuda" in msg:
                    break
