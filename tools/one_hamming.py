"""One brute-force Hamming call at configs[4] size (profiling target): python tools/one_hamming.py [nq] [nm]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from my_orb_slam2_b200 import orbfe, synth

nq = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
nm = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20
q, m = synth.bruteforce_case(nq, nm, 4096, seed=1)
M = orbfe.Matcher()
for _ in range(2):
    best, second = M.hamming_bruteforce(q, m)
print("ok", int(best[0]))
