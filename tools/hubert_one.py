"""One forward of the HuBERT embedder (for an ncu launch list): python tools/hubert_one.py [seconds] [batch]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import rvc_b200  # noqa: E402

secs = float(sys.argv[1]) if len(sys.argv) > 1 else 30.0
B = int(sys.argv[2]) if len(sys.argv) > 2 else 1
torch.manual_seed(1)
m = rvc_b200.HubertModelWithFinalProj().eval()
a = (0.3 * torch.randn(B, int(secs * 16000))).cuda()
m(a)
torch.cuda.synchronize()
y = m(a)["last_hidden_state"]
torch.cuda.synchronize()
print(tuple(y.shape), float(y.abs().mean()))
