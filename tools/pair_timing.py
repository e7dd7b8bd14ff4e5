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
K = int(sys.argv[3]) if len(sys.argv) > 3 else 1  # 1 = mrf_pair.cu, 2 = mrf_pair2.cu (CTA pairs)
dev = torch.device("cuda", 0)
m, cfg = build_model(dev)
h = m._ensure_handle(dev)
h.set_option("pair_kernel", K)
h.set_option("pair_mrf", 1)
x, f0, g = (t.to(dev) for t in synth_batch(cfg, B, T, 1234, False))
with torch.no_grad():
    m(x, f0, g, mode="fast")
    m._handle.set_option("fused_timing", 1)
    m(x, f0, g, mode="fast")
    t = m._handle.fused_timing()
names = ["mma:wait_a0", "mma:wait_a1", "mma:wait_w", "mma:tap_loops", "", "mma:total", "fin:tmem_ld", "fin:store_x",
         "epi:wait_acc1", "epi:conv1", "epi:wait_x", "epi:final", "epi:first_x", "epi:total", "fin:stg", ""]
if K == 2:
    names2 = ["mma:wait_in", "mma:wait_mid", "mma:wait_w", "F:load_issue", "M:tmem_ld", "mma:total", "F:L", "prod:wait_empty",
              "epi:wait_acc1", "epi:M", "epi:wait_x", "epi:F+L", "epi:prologue", "epi:total", "M:fence+arrive", "F:out"]
    for i in range(min(9, t.shape[0])):
        lead = t[i][0::2].astype(float).mean(axis=0)
        peer = t[i][1::2].astype(float).mean(axis=0)
        print(f"pair {i} leader: " + "  ".join(f"{n}={lead[j]/1e3:.0f}k" for j, n in enumerate(names2) if n))
        print(f"pair {i} peer  : relay:wait_w={peer[0]/1e3:.0f}k  " + "  ".join(f"{n}={peer[j]/1e3:.0f}k" for j, n in enumerate(names2) if n and j >= 3 and j != 5))
    sys.exit(0)
for i in range(min(9, t.shape[0])):
    a = t[i].astype(float).mean(axis=0)
    print(f"pair {i}: " + "  ".join(f"{n}={a[j]/1e3:.0f}k" for j, n in enumerate(names) if n))
