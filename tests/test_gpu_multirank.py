"""N > 1 ranks on hardware (SURVEY.md section 4 level 4): spawns tests/multirank_worker.py
under torch.distributed.run with one process per GPU (NCCL) for N = 2 and for all visible GPUs, and requires every
rank's sharded results (frames of configs[2], brute-force top-2 of configs[4] through every exchange path of the
library) to equal the single-GPU results and the CPU oracle.  Skipped on a box with fewer than two GPUs — the same
comparison then runs inside `bench.py --gpus N` (its `sharded_equal_to_single` key)."""
import json
import os
import socket
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    from my_orb_slam2_b200 import orbfe
    return orbfe.device_count()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(n):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n),
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(ROOT, "tests", "multirank_worker.py")]
    env = dict(os.environ, ORBFE_COMM_TIMEOUT_S="20")
    p = subprocess.run(cmd, cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert p.returncode == 0, p.stdout[-4000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("MULTIRANK_RESULT ")]
    assert line, p.stdout[-4000:]
    return json.loads(line[-1][len("MULTIRANK_RESULT "):])


def test_two_ranks_equal_single_gpu_and_oracle():
    if _gpus() < 2:
        pytest.skip("needs >= 2 GPUs (bench.py --gpus N carries the same check as sharded_equal_to_single)")
    r = _run(2)
    assert r["world"] == 2 and r["frames_sharded_equal_single_and_oracle"] == 16
    assert "nccl" in r["bruteforce_exchanges_equal_single_and_oracle"]
    print(r)


def test_all_visible_ranks_equal_single_gpu_and_oracle():
    n = _gpus()
    if n < 3:
        pytest.skip("needs > 2 GPUs")
    r = _run(n)
    assert r["world"] == n
    print(r)
