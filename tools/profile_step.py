"""Where does a bench step go?  CUDA-event timing of the pieces on one handle / two handles (debug aid)."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from my_orb_slam2_b200 import orbfe, synth
import bench

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dev = torch.device("cuda", 0)
lefts, rights = bench.make_pairs(4, 0)
idx = np.arange(B) % 4
d_left = torch.from_numpy(lefts[idx]).to(dev); d_right = torch.from_numpy(rights[idx]).to(dev)
h_left = torch.from_numpy(lefts[idx]).pin_memory()
L = orbfe.Extractor(*bench.PARAMS, bench.W, bench.H, max_batch=B); R = orbfe.Extractor(*bench.PARAMS, bench.W, bench.H, max_batch=B)
sL = torch.cuda.ExternalStream(L.stream, device=dev); sR = torch.cuda.ExternalStream(R.stream, device=dev)
lib = orbfe.lib()

def timeit(name, fn, stream=sL, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record(stream)
    for _ in range(n): fn()
    e1.record(stream); t_issue = time.perf_counter() - t0
    torch.cuda.synchronize(); wall = time.perf_counter() - t0
    print("%-34s device %.3f ms/iter   host-issue %.3f ms/iter  wall %.3f" % (name, e0.elapsed_time(e1) / n, 1e3 * t_issue / n, 1e3 * wall / n))

timeit("upload D2D (L)", lambda: L.upload(d_left.data_ptr(), B, on_device=True))
timeit("upload H2D pinned (L)", lambda: L.upload(h_left.data_ptr(), B))
timeit("run (L only)", lambda: L.run(B))
L.set_profiling(True)
for _ in range(5): L.run(B)
print("stages L alone:", {k: round(v / 5, 4) for k, v in L.stage_times()[0].items()})
L.set_profiling(False)
def both():
    L.run(B); R.run(B)
    orbfe.check(lib.orbfe_stereo_match_run(L._h, R._h, B, bench.MBF, bench.MB))
R.upload(d_right.data_ptr(), B, on_device=True); R.run(B)
timeit("run L + run R + stereo", both)
def full():
    L.upload(d_left.data_ptr(), B, on_device=True); R.upload(d_right.data_ptr(), B, on_device=True); both()
timeit("full device step", full)
