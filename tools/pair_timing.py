"""Where does the fused-pair kernel (mrf_pair.cu) wait?  One decode with the instrumented twin; prints the mean
clock64 spans of the MMA warp and of one epilogue warp for each of the 9 pair launches.
    python tools/pair_timing.py [chunks] [frames]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bench import build_model, synth_batch

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
T = int(sys.argv[2]) if len(sys.argv) > 2 else 3000
dev = torch.device("cuda", 0)
m, cfg = build_model(dev)
x, f0, g = (t.to(dev) for t in synth_batch(cfg, B, T, 1234, False))
with torch.no_grad():
    m(x, f0, g, mode="fast")
    m._handle.set_option("fused_timing", 1)
    m(x, f0, g, mode="fast")
    t = m._handle.fused_timing()
names = ["mma:wait_a0", "mma:wait_a1", "mma:wait_w", "mma:tap_loops", "", "mma:total", "fin:tmem_ld", "fin:store_x",
         "epi:wait_acc1", "epi:conv1", "epi:wait_x", "epi:final", "epi:first_x", "epi:total", "fin:stg", ""]
for i in range(min(9, t.shape[0])):
    a = t[i].astype(float).mean(axis=0)
    print(f"pair {i}: " + "  ".join(f"{n}={a[j]/1e3:.0f}k" for j, n in enumerate(names) if n))
