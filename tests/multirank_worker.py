"""Worker of tests/test_gpu_multirank.py: one process per GPU under torch.distributed.run (NCCL).

SURVEY.md section 4 test level (4): the same inputs sharded N ways must give the outputs of the single-GPU run.
 * configs[2] (stereo batch sharded by frame, no data-path collective): every rank extracts + stereo-matches its frames
   (frame i -> rank i % world), the results are gathered on rank 0 and compared, frame by frame, with ONE single-GPU
   batched call over all frames and with the CPU oracle.
 * configs[4] (brute force, map sharded, exchange of the top-2 keys inside the library): every exchange path
   (peer memory over NVLink, ncclAllGather in the library, torch all-gather) must leave the SAME global top-2 on EVERY
   rank, equal to the single-GPU result over the whole map and to the oracle's double loop (all 8192 queries).
Rank 0 prints one line `MULTIRANK_RESULT {json}`; any mismatch raises."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    from my_orb_slam2_b200 import orbfe, sharding, synth
    n_pairs = int(os.environ.get("MR_PAIRS", "16"))
    n_map = int(os.environ.get("MR_MAP", str(1 << 17)))
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    report = {"world": world}

    # ---------------- configs[2]: frames sharded over the ranks ----------------
    s = synth.SETTINGS["euroc"]
    w, h = synth.SHAPES["euroc"]
    params = (s["nfeatures"], s["scale_factor"], s["nlevels"], s["ini_th"], s["min_th"])
    mbf = np.float32(s["bf"])
    mb = mbf / np.float32(s["fx"])
    pairs = [synth.synth_pair(w, h, seed=100 + i) for i in range(n_pairs)]   # every rank renders the same set
    mine = sharding.frame_indices(n_pairs, rank, world)
    L = orbfe.Extractor(*params, w, h, max_batch=max(len(mine), 1), device=local)
    R = orbfe.Extractor(*params, w, h, max_batch=max(len(mine), 1), device=local)
    res = {}
    if mine:
        out = orbfe.stereo_frame(L, R, np.stack([pairs[i][0] for i in mine]), np.stack([pairs[i][1] for i in mine]), mbf, mb)
        kl, dl, nl, kr, dr, nr, ur, dp, nm = out
        for j, i in enumerate(mine):
            res[i] = (kl[j][:nl[j]].copy(), dl[j][:nl[j]].copy(), kr[j][:nr[j]].copy(), dr[j][:nr[j]].copy(),
                      ur[j][:nl[j]].copy(), dp[j][:nl[j]].copy(), int(nm[j]))
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(res, gathered, dst=0)
    if rank == 0:
        from oracle import pyoracle
        pyoracle.build()
        merged = {}
        for g in gathered:
            merged.update(g)
        assert sorted(merged) == list(range(n_pairs))
        L1 = orbfe.Extractor(*params, w, h, max_batch=n_pairs, device=local)
        R1 = orbfe.Extractor(*params, w, h, max_batch=n_pairs, device=local)
        kl, dl, nl, kr, dr, nr, ur, dp, nm = orbfe.stereo_frame(
            L1, R1, np.stack([p[0] for p in pairs]), np.stack([p[1] for p in pairs]), mbf, mb)
        oL, oR = pyoracle.Extractor(*params), pyoracle.Extractor(*params)
        for i in range(n_pairs):
            a = merged[i]
            single = (kl[i][:nl[i]], dl[i][:nl[i]], kr[i][:nr[i]], dr[i][:nr[i]], ur[i][:nl[i]], dp[i][:nl[i]], int(nm[i]))
            for x, y in zip(a, single):
                assert np.array_equal(x, y), "frame %d: sharded result differs from the single-GPU batch" % i
            okl, odl = oL.extract(pairs[i][0])
            okr, odr = oR.extract(pairs[i][1])
            our, odp, on = pyoracle.stereo_matches(oL, oR, okl, odl, okr, odr, mbf, mb)
            for x, y in zip(a, (okl, odl, okr, odr, our, odp, on)):
                assert np.array_equal(x, y), "frame %d: sharded result differs from the oracle" % i
        report["frames_sharded_equal_single_and_oracle"] = n_pairs
    dist.barrier()

    # ---------------- configs[4]: map sharded, exchange inside the library ----------------
    nq = 8192
    q, m = synth.bruteforce_case(nq, n_map, 4096, seed=1)
    m[n_map - 7] = m[11]   # exact duplicates in the first and the last shard: the lower index must win everywhere
    q[3] = m[11]
    start, cnt = sharding.map_shard(n_map, rank, world)
    mt = orbfe.Matcher(device=local)
    d_q = torch.from_numpy(q).to(dev)
    d_m = torch.from_numpy(m[start:start + cnt].copy()).to(dev)
    sb = sharding.ShardedBruteForce(mt, dev, max_queries=nq)
    info = sb.comm.info()
    report["comm"] = info
    modes = [("nccl", orbfe.EXCHANGE_NCCL), ("torch", "torch")]
    if info["peer_memory"]:
        modes.insert(0, ("peer", orbfe.EXCHANGE_PEER))
    modes.append(("tensor_core", orbfe.EXCHANGE_AUTO))   # the shard's top-2 by the tcgen05 kernel, default exchange
    results = {}
    for name, mode in modes:
        for rep in range(3):   # three calls in a row: both parities of the exchange buffer and its reuse
            best, second = sb.run(d_q, d_m, start, exchange=mode, tensor=(name == "tensor_core"))
        torch.cuda.synchronize()
        sb.comm.status()
        results[name] = (best.cpu().numpy().view(np.uint64), second.cpu().numpy().view(np.uint64))
    # an unbalanced call: nq smaller than max_queries, one rank with an EMPTY shard
    nq2 = 1000
    cnt2 = 0 if rank == world - 1 else cnt
    b2 = torch.empty(nq2, dtype=torch.int64, device=dev)
    s2 = torch.empty(nq2, dtype=torch.int64, device=dev)
    sb.comm.hamming_bruteforce_sharded(mt, d_q.data_ptr(), nq2, d_m.data_ptr(), cnt2, start, b2.data_ptr(), s2.data_ptr())
    mt.sync()
    sb.comm.status()
    results["ragged"] = (b2.cpu().numpy().view(np.uint64), s2.cpu().numpy().view(np.uint64))
    allres = [None] * world if rank == 0 else None
    dist.gather_object(results, allres, dst=0)
    if rank == 0:
        from oracle import pyoracle
        sbest, ssecond = orbfe.Matcher(device=local).hamming_bruteforce(q, m, 0)      # single GPU, whole map
        ob, os_ = pyoracle.hamming_bruteforce(q, m, 0)                                # CPU oracle, all 8192 queries
        assert np.array_equal(sbest, ob) and np.array_equal(ssecond, os_), "single-GPU brute force vs oracle"
        last = sharding.map_shard(n_map, world - 1, world)[0]
        rb, rs = pyoracle.hamming_bruteforce(np.ascontiguousarray(q[:nq2]), np.ascontiguousarray(m[:last]), 0)
        for r, rr in enumerate(allres):
            for name, (b, sc) in rr.items():
                wb, ws = (rb, rs) if name == "ragged" else (ob, os_)
                assert np.array_equal(b, wb) and np.array_equal(sc, ws), "rank %d, exchange %s" % (r, name)
        assert int(ob[3]) == 11 and int(os_[3]) == n_map - 7
        report["bruteforce_exchanges_equal_single_and_oracle"] = sorted(allres[0].keys())
        report["bruteforce_queries"] = nq
        report["bruteforce_map"] = n_map
        print("MULTIRANK_RESULT " + json.dumps(report), flush=True)
    dist.barrier()
    sb.comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
