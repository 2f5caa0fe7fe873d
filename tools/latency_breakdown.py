"""Per-stage CUDA-event times of ONE KITTI stereo pair (batch 1): python tools/latency_breakdown.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from my_orb_slam2_b200 import orbfe

lefts, rights = bench.make_pairs(2, 0)
hl = torch.from_numpy(lefts).pin_memory(); hr = torch.from_numpy(rights).pin_memory()
L = orbfe.Extractor(*bench.PARAMS, bench.W, bench.H, max_batch=1)
R = orbfe.Extractor(*bench.PARAMS, bench.W, bench.H, max_batch=1)
lib = orbfe.lib()
cap = L.capacity
kp = torch.empty((2, cap, 28), dtype=torch.uint8).pin_memory(); ds = torch.empty((2, cap, 32), dtype=torch.uint8).pin_memory()
cnt = torch.zeros(2, dtype=torch.int32).pin_memory(); ur = torch.empty((1, cap), dtype=torch.float32).pin_memory()
dp = torch.empty((1, cap), dtype=torch.float32).pin_memory(); nm = torch.zeros(1, dtype=torch.int32).pin_memory()

def timed(fn, n=200):
    for _ in range(20): fn()
    t = []
    for _ in range(n):
        t0 = time.perf_counter(); fn(); t.append(1e3 * (time.perf_counter() - t0))
    return float(np.median(t))

def up():
    L.upload(hl[0].data_ptr(), 1); R.upload(hr[0].data_ptr(), 1); L.sync(); R.sync()
def up_run():
    L.upload(hl[0].data_ptr(), 1); R.upload(hr[0].data_ptr(), 1); L.run(1); R.run(1); L.sync(); R.sync()
def up_run_st():
    L.upload(hl[0].data_ptr(), 1); R.upload(hr[0].data_ptr(), 1); L.run(1); R.run(1)
    orbfe.check(lib.orbfe_stereo_match_run(L._h, R._h, 1, bench.MBF, bench.MB)); L.sync(); R.sync()
def one_L():
    L.upload(hl[0].data_ptr(), 1); L.run(1); L.sync()
def full():
    orbfe.check(lib.orbfe_stereo_frame(L._h, R._h, hl[0].data_ptr(), hr[0].data_ptr(), bench.W * bench.H, bench.W, 1, bench.MBF, bench.MB,
        kp[0].data_ptr(), ds[0].data_ptr(), cnt[0:].data_ptr(), kp[1].data_ptr(), ds[1].data_ptr(), cnt[1:].data_ptr(), ur.data_ptr(), dp.data_ptr(), nm.data_ptr(), cap))
print("upload L+R + sync          %.4f ms" % timed(up))
print("upload L + run L + sync    %.4f ms" % timed(one_L))
print("upload + run L,R + sync    %.4f ms" % timed(up_run))
print("... + stereo run + sync    %.4f ms" % timed(up_run_st))
print("orbfe_stereo_frame (all)   %.4f ms" % timed(full))
L.set_profiling(True)
for _ in range(50):
    L.upload(hl[0].data_ptr(), 1); L.run(1); L.sync()
    R.upload(hr[0].data_ptr(), 1); R.run(1); R.sync()
    orbfe.check(lib.orbfe_stereo_match_run(L._h, R._h, 1, bench.MBF, bench.MB)); L.sync()
st, runs, sruns = L.stage_times()
print("per-stage events, batch 1, plain launches (not the graph):", {k: round(v / max(runs if 'stereo' not in k else sruns, 1), 4) for k, v in st.items()})
