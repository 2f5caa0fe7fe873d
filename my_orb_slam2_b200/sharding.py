"""Multi-GPU partitioning of the hot path: one process per GPU, torch.distributed for the plumbing.

* Extraction / stereo matching: units are independent frames -> frame i belongs to rank i % world; NO data-path
  collective (`frame_indices`).
* Brute-force map matching (BASELINE.json config 5): the MAP descriptors are sharded contiguously (`map_shard`), the
  queries are replicated; every rank computes its per-query local top-2 keys `dist << 32 | global index` with the CUDA
  kernel, the keys are exchanged (world x nq x 2 x 8 bytes) and a CUDA kernel takes the global top-2.  The exchange
  lives in the library (csrc/comm.cu, `orbfe_hamming_bruteforce_sharded`): peer-memory stores over NVLink + flags, or
  one ncclAllGather.  Ties resolve to the lowest global index by construction of the key.

`exchange_top2` is the same exchange written against torch.distributed (all_gather_into_tensor on int64 views): it
runs over gloo in the CPU tests (tests/test_sharding_gloo.py), where the local top-2 and the merge are supplied by the
test, and is kept as the comparison leg `exchange="torch"` of ShardedBruteForce.
"""

def frame_indices(n_frames, rank, world):
    """Frames owned by `rank`: i % world == rank (left and right image of a pair stay together)."""
    return list(range(rank, n_frames, world))


def map_shard(n_map, rank, world):
    """Contiguous shard [start, start + count) of the map descriptors owned by `rank`."""
    per = (n_map + world - 1) // world
    start = min(rank * per, n_map)
    return start, min(per, n_map - start)


def exchange_top2(best_local, second_local, group=None):
    """All-gather the per-rank (best, second) key vectors.  Inputs: torch int64 tensors [nq] (bit patterns of the
    uint64 keys) on the backend's device; returns two tensors [world, nq]."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    nq = best_local.numel()
    both = torch.stack([best_local, second_local]).contiguous()           # [2, nq]
    out = torch.empty((world, 2, nq), dtype=both.dtype, device=both.device)
    dist.all_gather_into_tensor(out.view(world * 2, nq), both, group=group)
    return out[:, 0, :].contiguous(), out[:, 1, :].contiguous()


def broadcast_comm_id(group=None):
    """Rank 0 makes the library communicator's id (orbfe_comm_unique_id) and every rank receives its 128 bytes
    through torch.distributed (any backend: the id is a small object)."""
    import numpy as np
    import torch.distributed as dist
    from . import orbfe
    box = [orbfe.comm_unique_id().tobytes() if dist.get_rank(group) == 0 else None]
    dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    return np.frombuffer(box[0], np.uint8).copy()


class ShardedBruteForce:
    """Config 5 on `world` GPUs through the C ABI (orbfe_hamming_bruteforce_sharded): queries replicated, map sharded by
    `map_shard`; the exchange of the per-rank top-2 keys happens INSIDE the library — stores into the peers' mapped
    buffers over NVLink + flags (default) or one ncclAllGather (exchange=orbfe.EXCHANGE_NCCL).  torch.distributed only
    carries the 128-byte communicator id at set-up.  `exchange="torch"` keeps the round-1 route (per-shard kernel,
    torch all_gather_into_tensor, merge kernel) as the comparison the bench reports."""

    def __init__(self, matcher, device, group=None, max_queries=8192):
        import torch
        import torch.distributed as dist
        from . import orbfe
        self.m, self.dev, self.group = matcher, device, group
        self.stream = torch.cuda.ExternalStream(matcher.stream, device=device)
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.comm = None
        if self.world > 1:
            uid = broadcast_comm_id(group)
            index = device.index if hasattr(device, "index") and device.index is not None else int(device)
            self.comm = orbfe.Comm(uid, self.world, self.rank, index, max_queries)

    def run(self, d_query, d_map_shard, index_base, exchange=0, tensor=False):
        """d_query [nq,32] / d_map_shard [nm_local,32]: torch uint8 CUDA tensors.  Returns (best, second) int64 CUDA
        tensors [nq] holding the uint64 keys of the GLOBAL top-2 (valid on the current stream).  tensor=True: the
        shard's top-2 by the tensor-core kernel (tcgen05 int8 GEMM on the expanded bits) instead of the POPC kernel."""
        import torch
        nq, nm = d_query.shape[0], d_map_shard.shape[0]
        best = torch.empty(nq, dtype=torch.int64, device=self.dev)
        second = torch.empty(nq, dtype=torch.int64, device=self.dev)
        cur = torch.cuda.current_stream(self.dev)
        self.stream.wait_stream(cur)
        if self.comm is None:
            fn = self.m.hamming_bruteforce_tc_device if tensor and nm > 0 else self.m.hamming_bruteforce_device
            fn(d_query.data_ptr(), nq, d_map_shard.data_ptr(), nm, index_base, best.data_ptr(), second.data_ptr())
            cur.wait_stream(self.stream)
            return best, second
        if exchange == "torch":
            return self._run_torch(d_query, d_map_shard, index_base, best, second, cur)
        from . import orbfe
        self.comm.hamming_bruteforce_sharded(self.m, d_query.data_ptr(), nq, d_map_shard.data_ptr(), nm, index_base,
                                             best.data_ptr(), second.data_ptr(),
                                             exchange | (orbfe.KERNEL_TENSOR if tensor else 0))
        cur.wait_stream(self.stream)
        return best, second

    def _run_torch(self, d_query, d_map_shard, index_base, best, second, cur):
        import torch
        nq, nm = d_query.shape[0], d_map_shard.shape[0]
        self.m.hamming_bruteforce_device(d_query.data_ptr(), nq, d_map_shard.data_ptr(), nm, index_base,
                                         best.data_ptr(), second.data_ptr())
        cur.wait_stream(self.stream)
        best_all, second_all = exchange_top2(best, second, self.group)    # NCCL through torch, on the current stream
        self.stream.wait_stream(cur)
        gb = torch.empty_like(best)
        gs = torch.empty_like(second)
        self.m.hamming_merge_device(best_all.data_ptr(), second_all.data_ptr(), self.world, nq, gb.data_ptr(),
                                    gs.data_ptr())
        cur.wait_stream(self.stream)
        return gb, gs
