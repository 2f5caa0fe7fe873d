"""Does running the left and right handle on separate streams (kernel overlap) beat one in-order stream?"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, bench
from my_orb_slam2_b200 import orbfe
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dev = torch.device("cuda", 0)
lefts, rights = bench.make_pairs(bench.N_UNIQUE, 0)
idx = np.arange(B) % bench.N_UNIQUE
d_left = torch.from_numpy(lefts[idx]).to(dev); d_right = torch.from_numpy(rights[idx]).to(dev)
lib = orbfe.lib()
for share in (True, False):
    L = orbfe.Extractor(*bench.PARAMS, bench.W, bench.H, max_batch=B); R = orbfe.Extractor(*bench.PARAMS, bench.W, bench.H, max_batch=B)
    if share: R.share_stream(L)
    def step():
        L.upload(d_left.data_ptr(), B, on_device=True); R.upload(d_right.data_ptr(), B, on_device=True)
        L.run(B); R.run(B)
        orbfe.check(lib.orbfe_stereo_match_run(L._h, R._h, B, bench.MBF, bench.MB))
    for _ in range(3): step()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(20): step()
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 20
    print("share_stream=%s: %.3f ms/step  %.0f frames/s" % (share, dt * 1e3, B / dt))
    R.close(); L.close()
