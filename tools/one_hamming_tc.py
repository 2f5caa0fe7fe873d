"""Tensor-core brute force (orbfe_hamming_bruteforce_tc_device) vs the POPC path: keys must be bit-equal.
python tools/one_hamming_tc.py [nq] [nm] [reps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from my_orb_slam2_b200 import orbfe, synth

nq = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
nm = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
dev = torch.device("cuda", 0)
q, m = synth.bruteforce_case(nq, nm, min(nq, nm) // 2, seed=1)
if nm > 40:
    m[17] = m[3]; m[nm - 1] = m[3]; q[0] = m[3]      # duplicates: ties to the lowest index
d_q, d_m = torch.from_numpy(q).to(dev), torch.from_numpy(m).to(dev)
mt = orbfe.Matcher()
st = torch.cuda.ExternalStream(mt.stream, device=dev)


def run(fn):
    best = torch.empty(nq, dtype=torch.int64, device=dev)
    second = torch.empty(nq, dtype=torch.int64, device=dev)
    for _ in range(2):
        fn(d_q.data_ptr(), nq, d_m.data_ptr(), nm, 7, best.data_ptr(), second.data_ptr())
    mt.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(reps):
        fn(d_q.data_ptr(), nq, d_m.data_ptr(), nm, 7, best.data_ptr(), second.data_ptr())
    e1.record(st)
    mt.sync()
    return e0.elapsed_time(e1) / reps, best.cpu().numpy().view(np.uint64), second.cpu().numpy().view(np.uint64)


t_p, bp, sp = run(mt.hamming_bruteforce_device)
t_t, bt, s_t = run(mt.hamming_bruteforce_tc_device)
eq_b, eq_s = np.array_equal(bp, bt), np.array_equal(sp, s_t)
print("nq %d nm %d: popc %.3f ms (%.0f G pairs/s)  tensor %.3f ms (%.0f G pairs/s)  best equal %s  second equal %s" % (
    nq, nm, t_p, nq * nm / t_p / 1e6, t_t, nq * nm / t_t / 1e6, eq_b, eq_s))
if not (eq_b and eq_s):
    bad = np.nonzero((bp != bt) | (sp != s_t))[0]
    print("mismatches:", len(bad), "first:", bad[:5])
    for i in bad[:5]:
        print(i, "popc", bp[i] >> np.uint64(32), bp[i] & np.uint64(0xFFFFFFFF), sp[i] >> np.uint64(32), sp[i] & np.uint64(0xFFFFFFFF),
              "| tc", bt[i] >> np.uint64(32), bt[i] & np.uint64(0xFFFFFFFF), s_t[i] >> np.uint64(32), s_t[i] & np.uint64(0xFFFFFFFF))
    sys.exit(1)
