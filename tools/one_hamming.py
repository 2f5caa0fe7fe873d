"""Brute-force Hamming 8192 x 1M on one GPU (BASELINE.json configs[4]; profiling / timing target):
python tools/one_hamming.py [reps]   (ORBFE_BF_KERNEL=1 selects the round-1 kernel)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from my_orb_slam2_b200 import orbfe, synth

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
dev = torch.device("cuda", 0)
nq, nm = 8192, 1 << 20
q, m = synth.bruteforce_case(nq, nm, 4096, seed=1)
d_q, d_m = torch.from_numpy(q).to(dev), torch.from_numpy(m).to(dev)
best = torch.empty(nq, dtype=torch.int64, device=dev)
second = torch.empty(nq, dtype=torch.int64, device=dev)
mt = orbfe.Matcher()
st = torch.cuda.ExternalStream(mt.stream, device=dev)
for _ in range(2):
    mt.hamming_bruteforce_device(d_q.data_ptr(), nq, d_m.data_ptr(), nm, 0, best.data_ptr(), second.data_ptr())
mt.sync()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
for _ in range(reps):
    mt.hamming_bruteforce_device(d_q.data_ptr(), nq, d_m.data_ptr(), nm, 0, best.data_ptr(), second.data_ptr())
e1.record(st)
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
b = best.cpu().numpy().view(np.uint64)
s = second.cpu().numpy().view(np.uint64)
import hashlib
print("kernel", os.environ.get("ORBFE_BF_KERNEL", "2"), "ms %.3f  G pairs/s %.1f  popc peak %.3e  sha %s" % (
    ms, nq * nm / ms / 1e6, mt.popc_peak(), hashlib.sha256(b.tobytes() + s.tobytes()).hexdigest()[:16]))
