"""Throughput of one step (64 pairs) run as K concurrent sub-batches (K rigs x 2 handles x own streams)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, bench
from my_orb_slam2_b200 import orbfe
TOTAL = 64
dev = torch.device("cuda", 0)
lefts, rights = bench.make_pairs(bench.N_UNIQUE, 0)
lib = orbfe.lib()
for K in (1, 2, 4):
    B = TOTAL // K
    idx = np.arange(B) % bench.N_UNIQUE
    d_left = torch.from_numpy(lefts[idx]).to(dev); d_right = torch.from_numpy(rights[idx]).to(dev)
    rigs = [(orbfe.Extractor(*bench.PARAMS, bench.W, bench.H, max_batch=B), orbfe.Extractor(*bench.PARAMS, bench.W, bench.H, max_batch=B)) for _ in range(K)]
    def step():
        for L, R in rigs:
            L.upload(d_left.data_ptr(), B, on_device=True); R.upload(d_right.data_ptr(), B, on_device=True)
        for L, R in rigs:
            L.run(B); R.run(B)
        for L, R in rigs:
            orbfe.check(lib.orbfe_stereo_match_run(L._h, R._h, B, bench.MBF, bench.MB))
    for _ in range(3): step()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(20): step()
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 20
    print("K=%d sub-batches of %d: %.3f ms/step  %.0f frames/s" % (K, B, dt * 1e3, TOTAL / dt))
    for L, R in rigs:
        R.close(); L.close()
