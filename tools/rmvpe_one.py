import sys, torch, numpy as np
sys.path.insert(0, '/root/repo')
import rvc_b200
dev = torch.device("cuda", 0)
torch.manual_seed(1)
p = rvc_b200.RMVPE0Predictor(None, dev); p.model.to(dev)
T = int(sys.argv[1])
mel = torch.randn(1, 128, T, device=dev)
p.model(mel); torch.cuda.synchronize()
p.model(mel); torch.cuda.synchronize()
