"""world_size-2 (and 3) gloo tests of the multi-GPU host logic (my_orb_slam2_b200/sharding.py) on the CPU:
frame sharding, map sharding with global index bases, the all-gather exchange of the per-shard top-2 keys and the
tie rule (lowest global index).  The per-shard top-2 comes from the CPU oracle here and the merge is a numpy
statement of orbfe_hamming_merge_device; on the GPU box the same exchange code runs over NCCL with the CUDA kernels
(tests/test_gpu_matcher.py::test_bruteforce_sharded_single_process, bench.py --gpus N)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from my_orb_slam2_b200 import sharding, synth  # noqa: E402


def merge_top2_host(best_all, second_all):
    keys = np.concatenate([np.asarray(best_all, np.uint64), np.asarray(second_all, np.uint64)], axis=0)
    keys = np.sort(keys, axis=0)
    return keys[0].copy(), keys[1].copy()


def test_frame_indices_partition():
    for n, world in [(256, 8), (10, 4), (3, 8), (0, 2)]:
        owned = [sharding.frame_indices(n, r, world) for r in range(world)]
        flat = sorted(i for o in owned for i in o)
        assert flat == list(range(n))
        assert max(len(o) for o in owned) - min(len(o) for o in owned) <= 1


def test_map_shard_partition():
    for nm, world in [(1 << 20, 8), (1000, 3), (5, 8), (0, 2)]:
        pos = 0
        for r in range(world):
            start, cnt = sharding.map_shard(nm, r, world)
            assert start == pos and cnt >= 0
            pos += cnt
        assert pos == nm


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, nq, nm, out_dir):
    import torch
    import torch.distributed as dist
    from oracle import pyoracle
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    q, m = synth.bruteforce_case(nq, nm, min(nq, nm) // 2, seed=7)
    m[nm - 1] = m[3]  # an exact duplicate in the LAST shard: the tie must go to the lower global index (shard 0)
    q[0] = m[3]
    start, cnt = sharding.map_shard(nm, rank, world)
    base = 1 << 20  # a non-zero global index base
    if cnt > 0:
        b, s = pyoracle.hamming_bruteforce(q, m[start:start + cnt], base + start)
    else:
        b = np.full(nq, np.iinfo(np.uint64).max, np.uint64)
        s = b.copy()
    bl = torch.from_numpy(b.view(np.int64).copy())
    sl = torch.from_numpy(s.view(np.int64).copy())
    ba, sa = sharding.exchange_top2(bl, sl)
    assert ba.shape == (world, nq)
    assert np.array_equal(ba[rank].numpy().view(np.uint64), b)  # gather layout: row r is rank r's vector
    gb, gs = merge_top2_host(ba.numpy().view(np.uint64), sa.numpy().view(np.uint64))
    ob, os_ = pyoracle.hamming_bruteforce(q, m, base)  # the unsharded answer
    ok = np.array_equal(gb, ob) and np.array_equal(gs, os_)
    # frames sharded by rank: every rank reports its frame list, rank 0 checks the partition
    mine = torch.zeros(16, dtype=torch.int64)
    for i in sharding.frame_indices(16, rank, world):
        mine[i] = rank + 1
    dist.all_reduce(mine)
    ok = ok and bool((mine == torch.tensor([(i % world) + 1 for i in range(16)])).all())
    open(os.path.join(out_dir, "ok_%d" % rank), "w").write("1" if ok else "0")
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,nq,nm", [(2, 64, 1000), (3, 33, 500)])
def test_sharded_bruteforce_exchange_gloo(oracle, tmp_path, world, nq, nm):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(world, port, nq, nm, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert open(tmp_path / ("ok_%d" % r)).read() == "1"
