"""Multi-GPU partitioning of the hot path: one process per GPU, torch.distributed for the plumbing.

* Extraction / stereo matching: units are independent frames -> frame i belongs to rank i % world; NO data-path
  collective (`frame_indices`).
* Brute-force map matching (BASELINE.json config 5): the MAP descriptors are sharded contiguously (`map_shard`), the
  queries are replicated; every rank computes its per-query local top-2 keys `dist << 32 | global index` with the CUDA
  kernel, ONE all-gather exchanges them (world x nq x 2 x 8 bytes) and the CUDA merge kernel takes the global top-2.
  Ties resolve to the lowest global index by construction of the key.

The exchange step is written against the torch.distributed API only (all_gather_into_tensor on int64 views), so the
same code runs over NCCL on the GPUs and over gloo in the CPU tests (tests/test_sharding_gloo.py), where the local
top-2 and the merge are supplied by the test (the product path below always uses the CUDA kernels).
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


class ShardedBruteForce:
    """Config 5 on `world` GPUs: queries replicated, map sharded by `map_shard`, CUDA kernels + one NCCL all-gather."""

    def __init__(self, matcher, device, group=None):
        import torch
        self.m, self.dev, self.group = matcher, device, group
        self.stream = torch.cuda.ExternalStream(matcher.stream, device=device)

    def run(self, d_query, d_map_shard, index_base):
        """d_query [nq,32] / d_map_shard [nm_local,32]: torch uint8 CUDA tensors.  Returns (best, second) int64 CUDA
        tensors [nq] holding the uint64 keys of the GLOBAL top-2."""
        import torch
        import torch.distributed as dist
        nq, nm = d_query.shape[0], d_map_shard.shape[0]
        best = torch.empty(nq, dtype=torch.int64, device=self.dev)
        second = torch.empty(nq, dtype=torch.int64, device=self.dev)
        cur = torch.cuda.current_stream(self.dev)
        self.stream.wait_stream(cur)
        self.m.hamming_bruteforce_device(d_query.data_ptr(), nq, d_map_shard.data_ptr(), nm, index_base,
                                         best.data_ptr(), second.data_ptr())
        cur.wait_stream(self.stream)
        world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        if world == 1:
            return best, second
        best_all, second_all = exchange_top2(best, second, self.group)    # NCCL, on the current stream
        self.stream.wait_stream(cur)
        gb = torch.empty_like(best)
        gs = torch.empty_like(second)
        self.m.hamming_merge_device(best_all.data_ptr(), second_all.data_ptr(), world, nq, gb.data_ptr(), gs.data_ptr())
        cur.wait_stream(self.stream)
        return gb, gs
