"""One device-resident bench step (profiling target): python tools/one_step.py [batch] [steps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from my_orb_slam2_b200 import orbfe

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
dev = torch.device("cuda", 0)
lefts, rights = bench.make_pairs(bench.N_UNIQUE, 0)
idx = np.arange(B) % bench.N_UNIQUE
d_left = torch.from_numpy(lefts[idx]).to(dev)
d_right = torch.from_numpy(rights[idx]).to(dev)
L = orbfe.Extractor(*bench.PARAMS, bench.W, bench.H, max_batch=B)
R = orbfe.Extractor(*bench.PARAMS, bench.W, bench.H, max_batch=B)
R.share_stream(L)
lib = orbfe.lib()
for _ in range(steps):
    L.upload(d_left.data_ptr(), B, on_device=True)
    R.upload(d_right.data_ptr(), B, on_device=True)
    L.run(B)
    R.run(B)
    orbfe.check(lib.orbfe_stereo_match_run(L._h, R._h, B, bench.MBF, bench.MB))
torch.cuda.synchronize()
R.close()
L.close()
print("ok")
