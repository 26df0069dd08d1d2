import sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/3dgs-mcmc_b200")
import torch, gs_synth
from simple_knn._C import distCUDA2
pts = torch.from_numpy(gs_synth.scene_garden(3_000_000, 0)["means3D"]).cuda()
distCUDA2(pts); torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); distCUDA2(pts); e1.record(); torch.cuda.synchronize()
print("ms", e0.elapsed_time(e1))
import numpy as np
p = pts.cpu().numpy()
print("extent", p.min(0), p.max(0))
