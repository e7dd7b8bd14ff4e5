"""Where does the fused ResBlock kernel wait?  Runs one decode with the instrumented kernel twin and prints,
per fused launch, the mean clock64 spans of the MMA warp and of one epilogue warp.
    python tools/fused_timing.py [chunks] [frames] [fused_mrf]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bench import build_model, synth_batch

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
T = int(sys.argv[2]) if len(sys.argv) > 2 else 3000
fm = int(sys.argv[3]) if len(sys.argv) > 3 else 1
dev = torch.device("cuda", 0)
m, cfg = build_model(dev)
m.fused_mrf = fm
if len(sys.argv) > 4:
    m._ensure_handle(dev).set_option("fused_epi_warps", int(sys.argv[4]))
x, f0, g = (t.to(dev) for t in synth_batch(cfg, B, T, 1234, False))
with torch.no_grad():
    m(x, f0, g, mode="fast")
    m._handle.set_option("fused_timing", 1)
    m(x, f0, g, mode="fast")
    t = m._handle.fused_timing()
names = ["mma:wait_rows", "fin:out(sum+math+stg)", "mma:wait_w", "mma:total", "rows:c1b0", "rows:c1b1+", "mma:tap_loops", "rows:c2",
         "epi:wait_acc1", "epi:wait_x", "epi:xload", "epi:total", "epi1:tmem_ld", "epi1:math+sts", "epi1:fence+arrive", "fin:next_x(ldg wait+sttm+sts+publish)"]
for i in range(t.shape[0]):
    a = t[i].astype(float).mean(axis=0)
    print(f"launch {i}: " + "  ".join(f"{n}={a[j]/1e3:.0f}k" for j, n in enumerate(names) if n))
