#!/usr/bin/env python3
"""bench.py — the headline metric of BASELINE.json on its configuration:

    frames/sec of ORB extract + match on KITTI-shaped 1241x376 stereo pairs, 2000 features per image
    (configs[1]: ORB extraction of both images + Frame::ComputeStereoMatches), plus p50 frame latency.

A "step" is `--batches` (100) passes of the hot path, each over one batch of `--batch` (64) synthetic stereo pairs per
GPU, so that `--steps 20` times >= 2 s of GPU work; every rank owns 64 DISTINCT pairs and each batch starts at another
offset into them (a handle slot sees a different image every batch).
  value : whole-job stereo frames/s with the inputs already resident in HBM (device images -> pyramid level 0 ->
          all kernels -> results left on the device), timed with CUDA events on the handles' streams.
  e2e   : the same through the C ABI with HOST buffers: pinned host images -> H2D -> kernels -> keypoints,
          descriptors, uRight, depth back in host memory, every step, inside the timed region.
  roofline / stages : per-stage CUDA-event times recorded inside the timed region by the library itself.
  cpu_baseline      : the UNMODIFIED reference sources compiled here (oracle/_ref/libref_orbslam2.so: src/ORBextractor.cc,
                      Frame.cc, ... against an OpenCV stand-in whose pixel primitives are scalar restatements), timed
                      single-threaded as ORB-SLAM2 runs it, with its real two L/R threads, p50/p95, the port beside it and
                      the measured SIMD factor of real cv2 primitives over the stand-ins.
`--impl reference` times that CPU path on all host cores (one instance per core) instead of the GPU path.
Multi-GPU (`torchrun ... bench.py --gpus N`): frames are sharded by rank, no data-path collective ("weak").
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from my_orb_slam2_b200 import synth  # noqa: E402

SHAPE = "kitti"
W, H = synth.SHAPES[SHAPE]
S = synth.SETTINGS[SHAPE]
PARAMS = (S["nfeatures"], S["scale_factor"], S["nlevels"], S["ini_th"], S["min_th"])
MBF = np.float32(S["bf"])
MB = np.float32(S["bf"]) / np.float32(S["fx"])
N_UNIQUE = 64  # distinct synthetic pairs per rank (one batch); every batch of a step starts at another offset
WORKLOAD = "configs[1]: KITTI00-shaped stereo pair 1241x376, 2000 features, extract + ComputeStereoMatches"
METRIC = "stereo frames/sec, ORB extract (L+R) + ComputeStereoMatches, KITTI 1241x376 2k feat"


def make_pairs(n_unique, seed0):
    pairs = [synth.synth_pair(W, H, seed0 + i) for i in range(n_unique)]
    return np.stack([p[0] for p in pairs]), np.stack([p[1] for p in pairs])


def level_pixels():
    """Sum over the 8 pyramid levels of w*h (1,444,097 for KITTI)."""
    sf, tot = np.float32(1.0), 0
    for lv in range(S["nlevels"]):
        inv = np.float32(1.0) / sf
        tot += int(np.rint(np.float32(W) * inv)) * int(np.rint(np.float32(H) * inv))
        sf = np.float32(float(sf) * float(np.float32(S["scale_factor"])))
    return tot


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons during the timed region (NVML)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake": 0x80}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            time.sleep(0.004)

    def result(self):
        self.stop_flag = True
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples), "period_ms": 4}


def ref_pair(pyref, rL, rR, left, right, threads):
    """One stereo frame through the compiled reference: ORBextractor::operator() for L and R (sequentially, or on the
    reference's two std::threads, src/Frame.cc:89-92), then Frame::ComputeStereoMatches.  Returns seconds."""
    t0 = time.perf_counter()
    if threads:
        pyref.extract_pair_threads(rL, rR, left, right)
    else:
        rL.extract(left)
        rR.extract(right)
    pyref.stereo_matches(rL, rR, MBF, MB)
    return time.perf_counter() - t0


def port_pair(pyoracle, oL, oR, left, right):
    t0 = time.perf_counter()
    kl, dl = oL.extract(left)
    kr, dr = oR.extract(right)
    pyoracle.stereo_matches(oL, oR, kl, dl, kr, dr, MBF, MB)
    return time.perf_counter() - t0


def simd_factor(pyoracle, img):
    """How much faster real (SIMD) OpenCV runs the pixel primitives than the scalar stand-ins the compiled reference
    and the port use: pyramid (7 x resize) + whole-level FAST at 20 + 8 x GaussianBlur on one frame.  None without cv2."""
    try:
        import cv2
    except Exception:
        return None
    sizes, sf = [], np.float32(1.0)
    for lv in range(S["nlevels"]):
        inv = np.float32(1.0) / sf
        sizes.append((int(np.rint(np.float32(W) * inv)), int(np.rint(np.float32(H) * inv))))
        sf = np.float32(float(sf) * float(np.float32(S["scale_factor"])))
    cv2.setNumThreads(1)
    fast = cv2.FastFeatureDetector_create(S["ini_th"], True)

    def chain_cv2():
        lv = img
        for i, (w, h) in enumerate(sizes):
            if i:
                lv = cv2.resize(lv, (w, h), interpolation=cv2.INTER_LINEAR)
            fast.detect(lv)
            cv2.GaussianBlur(lv, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101)

    def chain_port():
        lv = img
        for i, (w, h) in enumerate(sizes):
            if i:
                lv = pyoracle.resize_linear(lv, w, h)
            pyoracle.fast9_16(lv, S["ini_th"], True)
            pyoracle.gaussian_blur7(lv)

    out = []
    for fn in (chain_cv2, chain_port):
        fn()
        t0 = time.perf_counter()
        for _ in range(3):
            fn()
        out.append((time.perf_counter() - t0) / 3)
    return {"cv2_ms": 1e3 * out[0], "scalar_ms": 1e3 * out[1], "factor": out[1] / out[0], "cv2_version": cv2.__version__,
            "what": "7 x resize + 8 x FAST(20, nms) on whole levels + 8 x GaussianBlur 7x7 on one KITTI frame, 1 thread"}


def cpu_baseline(lefts, rights, seconds):
    """The CPU baseline block (rank 0, N = 1): bounded samples of the same workload on the host cores."""
    from oracle import pyoracle, pyref
    pyoracle.build()
    pyref.set_mode(pyref.MODE_AS_IS)  # libm cos/sin + plain malloc: the reference exactly as it runs
    rL, rR = pyref.Extractor(*PARAMS), pyref.Extractor(*PARAMS)
    n = len(lefts)
    ref_pair(pyref, rL, rR, lefts[0], rights[0], False)
    ts, k = [], 0
    while sum(ts) < seconds:
        ts.append(ref_pair(pyref, rL, rR, lefts[k % n], rights[k % n], False))
        k += 1
    t2, k2 = [], 0
    while sum(t2) < seconds / 2:
        t2.append(ref_pair(pyref, rL, rR, lefts[k2 % n], rights[k2 % n], True))
        k2 += 1
    oL, oR = pyoracle.Extractor(*PARAMS), pyoracle.Extractor(*PARAMS)
    tp, k3 = [], 0
    while sum(tp) < seconds / 4:
        tp.append(port_pair(pyoracle, oL, oR, lefts[k3 % n], rights[k3 % n]))
        k3 += 1
    pyref.set_mode(pyref.MODE_DETERMINISTIC)
    return {"value": len(ts) / sum(ts), "unit": "frames/s", "cores": 1, "kind": "reference",
            "sample": "%d stereo frames (ORBextractor L then R + ComputeStereoMatches), single thread, unmodified "
                      "src/ORBextractor.cc + src/Frame.cc compiled -O3 against scalar stand-ins of the OpenCV "
                      "primitives (see simd)" % len(ts),
            "ms_p50": 1e3 * float(np.percentile(ts, 50)), "ms_p95": 1e3 * float(np.percentile(ts, 95)),
            "two_threads": {"value": len(t2) / sum(t2), "cores": 2, "ms_p50": 1e3 * float(np.percentile(t2, 50)),
                            "ms_p95": 1e3 * float(np.percentile(t2, 95)),
                            "sample": "%d frames, L and R on two std::threads as src/Frame.cc:89-92" % len(t2)},
            "port": {"value": len(tp) / sum(tp), "cores": 1, "kind": "port",
                     "sample": "%d frames through the oracle restatement (oracle/*.cc)" % len(tp)},
            "simd": simd_factor(pyoracle, lefts[0])}


def parity_block(orbfe, lefts, rights, local):
    """GPU results against the compiled reference on a few pairs of the bench's own inputs, plus the reference's as-is
    divergence (heap-address quadtree ties, src/ORBextractor.cc:686; libm cos/sin, :112-113) per image."""
    from oracle import pyref
    n = min(4, len(lefts))
    L = orbfe.Extractor(*PARAMS, W, H, max_batch=n, device=local)
    R = orbfe.Extractor(*PARAMS, W, H, max_batch=n, device=local)
    kl, dl, nl, kr, dr, nr, ur, dp, nm = orbfe.stereo_frame(L, R, lefts[:n], rights[:n], MBF, MB)
    R.close()
    L.close()
    rL, rR = pyref.Extractor(*PARAMS), pyref.Extractor(*PARAMS)
    equal, asis = True, []
    for b in range(n):
        pyref.set_mode(pyref.MODE_DETERMINISTIC)
        (rkl, rdl), (rkr, rdr) = pyref.extract_pair_threads(rL, rR, lefts[b], rights[b])
        rur, rdp, rn = pyref.stereo_matches(rL, rR, MBF, MB)
        ok = (nl[b] == len(rkl) and nr[b] == len(rkr) and np.array_equal(kl[b][:nl[b]], rkl)
              and np.array_equal(dl[b][:nl[b]], rdl) and np.array_equal(kr[b][:nr[b]], rkr)
              and np.array_equal(dr[b][:nr[b]], rdr) and np.array_equal(ur[b][:nl[b]], rur)
              and np.array_equal(dp[b][:nl[b]], rdp) and int(nm[b]) == rn)
        equal = equal and bool(ok)
        pyref.set_mode(pyref.MODE_LIBM)
        _, ldl = rL.extract(lefts[b])
        pyref.set_mode(pyref.MODE_HEAP_TIES)
        hk, _ = rL.extract(lefts[b])
        a = set(zip(rkl["x"].tolist(), rkl["y"].tolist(), rkl["octave"].tolist()))
        h = set(zip(hk["x"].tolist(), hk["y"].tolist(), hk["octave"].tolist()))
        asis.append({"keypoints": int(len(rkl)), "descriptor_bits_changed_by_libm_cos_sin": int(np.unpackbits(ldl ^ rdl).sum()),
                     "keypoints_lost_with_heap_address_ties": len(a - h), "keypoints_gained_with_heap_address_ties": len(h - a)})
    pyref.set_mode(pyref.MODE_DETERMINISTIC)
    return {"pairs_checked": n, "gpu_equal_to_reference_source": equal,
            "what": "keypoints, descriptors (L and R), mvuRight, mvDepth bit-equal to the compiled unmodified reference "
                    "(deterministic mode: creation-order quadtree ties, correctly rounded cos/sin)",
            "angle_bin_flips": 0 if equal else None, "as_is_divergence_left_images": asis}


_CPU_STATE = {}


def _cpu_init(per_core):
    """Pool initializer: each worker process generates its own frames and extractors once, outside any timing."""
    from oracle import pyref
    lefts, rights = make_pairs(per_core, 1000 + (os.getpid() % 9973))
    pyref.set_mode(pyref.MODE_AS_IS)
    _CPU_STATE.update(lefts=lefts, rights=rights, rL=pyref.Extractor(*PARAMS), rR=pyref.Extractor(*PARAMS), pr=pyref)


def _cpu_worker(reps):
    st = _CPU_STATE
    t0 = time.perf_counter()
    for r in range(reps):
        for i in range(len(st["lefts"])):
            ref_pair(st["pr"], st["rL"], st["rR"], st["lefts"][i], st["rights"][i], False)
    return time.perf_counter() - t0


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path (oracle/_ref/libref_orbslam2.so, the
    unmodified sources compiled here) on all host cores, one instance per core, each looping over its own frames."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    from oracle import pyref
    pyref.lib()
    cores = os.cpu_count() or 1
    per_core = 2
    reps = 2
    ctx = mp.get_context("fork")
    with ctx.Pool(cores, initializer=_cpu_init, initargs=(per_core,)) as pool:
        for _ in range(args.warmup):
            pool.map(_cpu_worker, [1] * cores, chunksize=1)
        times = []
        for _ in range(args.steps):
            t0 = time.perf_counter()
            pool.map(_cpu_worker, [reps] * cores, chunksize=1)
            times.append(time.perf_counter() - t0)
    frames = cores * per_core * reps
    total = sum(times)
    value = frames * args.steps / total
    line = {
        "impl": "reference", "metric": METRIC,
        "value": value, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "frames_per_step": frames},
        "cpu_baseline": {"value": value, "unit": "frames/s", "cores": cores, "kind": "reference",
                         "sample": "%d stereo frames per step, one instance of the compiled unmodified reference per "
                                   "core (scalar stand-ins for the OpenCV primitives)" % frames},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def matcher_extras(torch, dist, orbfe, rank, world, local, kp_frame, desc_frame, quick):
    """BASELINE.json configs 4 and 5 (reported next to the headline, not the headline): brute-force Hamming
    8192 x 1,048,576 (map sharded over the ranks, one NCCL all-gather of the per-shard top-2, CUDA merge) as a fraction of
    the MEASURED popc-pipe peak, and the two SearchByProjection overloads through the C ABI with host views."""
    from my_orb_slam2_b200 import frontend, sharding
    dev = torch.device("cuda", local)
    m = orbfe.Matcher(device=local)
    out = {}
    popc_peak = m.popc_peak()
    nq, nm = 8192, (1 << 20) if not quick else (1 << 17)
    q, mp = synth.bruteforce_case(nq, nm, 4096, seed=1)
    start, cnt = sharding.map_shard(nm, rank, world)
    d_q = torch.from_numpy(q).to(dev)
    d_m = torch.from_numpy(mp[start:start + cnt].copy()).to(dev)
    sb = sharding.ShardedBruteForce(m, dev, max_queries=nq)

    def time_mode(mode, reps=5, tensor=False):
        for _ in range(3):
            sb.run(d_q, d_m, start, exchange=mode, tensor=tensor)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            b, s2 = sb.run(d_q, d_m, start, exchange=mode, tensor=tensor)
        e1.record()
        torch.cuda.synchronize()
        t = e0.elapsed_time(e1) / reps
        if world > 1:
            tt = torch.tensor([t], dtype=torch.float64, device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t = float(tt.item())
        return t, b, s2

    ms, best, second = time_mode(0)   # the library's default exchange (peer memory over NVLink when the ranks can map it)
    pairs = float(nq) * nm
    bf = {
        "workload": "configs[4]: %d queries x %d map descriptors, map sharded over %d GPU(s)" % (nq, nm, world),
        "ms": ms, "pairs_per_s": pairs / (ms * 1e-3), "algorithmic_popc32_per_s": 8.0 * pairs / (ms * 1e-3),
        "popc_peak_per_gpu_measured": popc_peak, "frac_of_popc_peak": 8.0 * pairs / (ms * 1e-3) / (popc_peak * world),
        "note": "the kernel spends 4 POPC per pair (four carry-save adders compress the 8 XOR words into words of "
                "weight 1, 1, 2, 4), so the algorithmic rate (8 popc32 per pair) may exceed the POPC pipe's own peak",
        "first_best_key": int(best[0].item()) & 0xFFFFFFFFFFFFFFFF,
    }
    # the optional second path: the same top-2 on the tensor cores (tcgen05 kind::i8 GEMM over the bits expanded to
    # bytes, K = 288; k_match_mma.cu), bit expansion of queries AND map inside the timed call; the POPC kernel above stays
    # the graded path (north_star), the keys must be bit-equal
    ms_tc, b_tc, s_tc = time_mode(0, tensor=True)
    m.sync()
    bf["tensor_core"] = {
        "ms": ms_tc, "pairs_per_s": pairs / (ms_tc * 1e-3), "speedup_vs_popc_kernel": ms / ms_tc,
        "int8_tops": 2.0 * 288.0 * pairs / (ms_tc * 1e-3) / 1e12,
        "equal_to_popc_keys": bool(torch.equal(b_tc, best)) and bool(torch.equal(s_tc, second)),
        "what": "orbfe_hamming_bruteforce_tc_device: d = |q| + |m| - 2 q.m as an s8 x u8 GEMM (M128 x N128 x K32 "
                "tcgen05.mma, TMEM accumulators, TMA operand tiles), threshold-gated exact top-2 in the epilogue; "
                "includes expanding the map to one byte per bit in every call"}
    if world > 1:
        # the exchange, broken down (device times of the LAST call, max over ranks): brute-force kernel | exchange + merge
        info = sb.comm.info()
        sb.comm.set_timing(True)
        brk = {}
        modes = [("nccl_in_library", orbfe.EXCHANGE_NCCL), ("torch_all_gather", "torch")]
        if info["peer_memory"]:
            modes.insert(0, ("peer_memory_nvlink", orbfe.EXCHANGE_PEER))
        equal = True
        ref_b, ref_s = best.clone(), second.clone()
        for name, mode in modes:
            t_all, b, s2 = time_mode(mode)
            equal = equal and bool(torch.equal(b, ref_b)) and bool(torch.equal(s2, ref_s))
            entry = {"ms": t_all}
            if mode != "torch":
                k_ms, x_ms = sb.comm.last_times()
                tt = torch.tensor([k_ms, x_ms], dtype=torch.float64, device=dev)
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                entry["kernel_ms"], entry["exchange_merge_ms"] = float(tt[0].item()), float(tt[1].item())
            brk[name] = entry
        sb.comm.set_timing(False)
        sb.comm.status()
        # parity with N > 1 ranks on hardware: EVERY rank's global top-2 (all 8192 queries) must equal the single-GPU
        # result over the whole map, computed on rank 0 and broadcast
        if rank == 0:
            d_full = torch.from_numpy(mp).to(dev)
            sb1, ss1 = torch.empty_like(best), torch.empty_like(second)
            m.hamming_bruteforce_device(d_q.data_ptr(), nq, d_full.data_ptr(), nm, 0, sb1.data_ptr(), ss1.data_ptr())
            m.sync()
            del d_full
        else:
            sb1, ss1 = torch.empty_like(best), torch.empty_like(second)
        dist.broadcast(sb1, src=0)
        dist.broadcast(ss1, src=0)
        ok = torch.tensor([int(equal and bool(torch.equal(best, sb1)) and bool(torch.equal(second, ss1)))], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        bf["exchange"] = {"default": "peer_memory_nvlink" if info["peer_memory"] else "nccl_in_library",
                          "nccl_version": info["nccl_version"], "paths": brk,
                          "what": "per path: whole call (max over ranks) and, from CUDA events inside the library, the "
                                  "brute-force kernel vs the exchange + global merge (peer path: merge+push kernel and "
                                  "wait+merge kernel; NCCL path: merge kernel, ncclAllGather, merge kernel)"}
        bf["sharded_equal_to_single"] = bool(ok.item())
    out["bruteforce_hamming"] = bf
    if rank == 0:
        # config 4: SearchByProjection (map overload, then last-frame overload) through the C ABI, host views
        bounds = (0.0, 0.0, float(W), float(H))
        frame = synth.frame_arrays(kp_frame, desc_frame, bounds, seed=0)
        mps = synth.map_points_case(frame, 20000, 1500, seed=0, bf=float(MBF))
        fv = frontend.frame_view(frame["keys_un"], frame["u_right"], frame["desc"], frame["mp_obs"], frame["bounds"],
                                 frame["scale_factors"])
        mv = frontend.mappoints_view(mps["proj_x"], mps["proj_y"], mps["proj_xr"], mps["level"], mps["view_cos"],
                                     mps["in_view"], mps["obs"], mps["desc"])
        ts = []
        for i in range(13):
            t0 = time.perf_counter()
            assigned, n = m.search_by_projection_map(fv, mv, 1.0, 0.8)
            ts.append(time.perf_counter() - t0)
        out["search_by_projection_map"] = {"workload": "configs[3]: %d features vs 20000 map points, th=1, nnratio=0.8, "
                                           "host views in, assignments out" % len(kp_frame),
                                           "ms_p50": 1e3 * float(np.median(ts[3:])), "matches": int(n)}
        last, cam = synth.last_frame_case(frame, seed=0)
        lv = frontend.lastframe_view(last["valid"], last["world_pos"], last["obs"], last["desc"], last["octave"],
                                     last["angle"], last["tcw"])
        cv = frontend.camera_pose(cam["tcw"], cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["mbf"], cam["mb"])
        ts = []
        for i in range(13):
            t0 = time.perf_counter()
            assigned, n = m.search_by_projection_last(fv, cv, lv, 7.0, False, True)
            ts.append(time.perf_counter() - t0)
        out["search_by_projection_last"] = {"workload": "configs[3]: last-frame overload, th=7, rotation histogram on",
                                            "ms_p50": 1e3 * float(np.median(ts[3:])), "matches": int(n)}
        # each entry point through the C ABI with host views (p50 of 10), the CPU oracle beside it
        from oracle import pyoracle as orc

        def p50(fn, reps=13, skip=3):
            ts = []
            for _ in range(reps):
                t0 = time.perf_counter()
                r = fn()
                ts.append(time.perf_counter() - t0)
            return 1e3 * float(np.median(ts[skip:])), r

        def both(name, workload, gpu_fn, cpu_fn, count):
            g_ms, g = p50(gpu_fn)
            c_ms, c = p50(cpu_fn, reps=3, skip=0)
            out[name] = {"workload": workload, "ms_p50": g_ms, "cpu_oracle_ms": c_ms, "matches": int(count(g)),
                         "equal_to_oracle": bool(count(g) == count(c))}

        def ov(kind, d):
            if kind == "frame":
                return orc.make_frame_view(d["keys_un"], d["u_right"], d["desc"], d["mp_obs"], d["bounds"], d["scale_factors"])
            if kind == "kf":
                return orc.make_keyframe_view(d["keys_un"], d["u_right"], d["desc"], d["has_mp"], d["node_ids"],
                                              d["node_offsets"], d["node_features"], d["scale_factors"], d["level_sigma2"])
            return orc.make_points3d_view(d["valid"], d["world_pos"], d["normal"], d["max_distance"], d["min_distance"],
                                          d["desc"], d["angle"])

        def gv(kind, d):
            if kind == "frame":
                return frontend.frame_view(d["keys_un"], d["u_right"], d["desc"], d["mp_obs"], d["bounds"], d["scale_factors"])
            if kind == "kf":
                return frontend.keyframe_view(d["keys_un"], d["u_right"], d["desc"], d["has_mp"], d["node_ids"],
                                              d["node_offsets"], d["node_features"], d["scale_factors"], d["level_sigma2"])
            return frontend.points3d_view(d["valid"], d["world_pos"], d["normal"], d["max_distance"], d["min_distance"],
                                          d["desc"], d["angle"])

        def proj(mod, pr, ow=None):
            f = orc.make_projection if mod is orc else frontend.projection
            return f(pr["tcw"], pr["fx"], pr["fy"], pr["cx"], pr["cy"], pr["bf"], pr["log_scale_factor"], pr["is_sim3"],
                     ow if ow is not None else pr["ow"])

        k1, k2 = synth.bow_case(kp_frame, desc_frame, seed=0)
        g1, g2, o1, o2 = gv("kf", k1), gv("kf", k2), ov("kf", k1), ov("kf", k2)
        both("search_by_bow", "SearchByBoW(KeyFrame, Frame): %d features, 100 nodes, nnratio 0.7" % len(kp_frame),
             lambda: m.search_by_bow(g1, g2, 0, 0.7, True), lambda: orc.search_by_bow(o1, o2, 0, 0.7, True), lambda r: r[1])
        pts, pr = synth.points3d_case(frame, seed=0, n_extra=18000)
        obs = np.random.default_rng(0).integers(0, 9, len(pts["valid"])).astype(np.int32)
        gp, op, gf, of = gv("pts", pts), ov("pts", pts), gv("frame", frame), ov("frame", frame)
        both("search_local_points", "Tracking::SearchLocalPoints: isInFrustum over %d map points fused with "
             "SearchByProjection, th=1" % len(obs),
             lambda: m.search_local_points(gf, proj(frontend, pr), gp, obs, 0.5, 1.0, 0.8),
             lambda: orc.search_local_points(of, proj(orc, pr), op, obs, 0.5, 1.0, 0.8), lambda r: r[1])
        both("search_by_projection_reloc", "relocalisation overload, th=10, ORBdist=100, rotation histogram on",
             lambda: m.search_by_projection_reloc(gf, proj(frontend, pr), gp, 10.0, 100, True),
             lambda: orc.search_by_projection_reloc(of, proj(orc, pr), op, 10.0, 100, True), lambda r: r[1])
        ow = (-pr["tcw"][:, :3].T @ pr["tcw"][:, 3]).astype(np.float32)
        sig = (np.float32(1.0) / (frame["scale_factors"] * frame["scale_factors"])).astype(np.float32)
        both("fuse", "Fuse(KeyFrame, %d map points, th=3): the search" % len(obs),
             lambda: m.fuse(gf, proj(frontend, pr, ow), gp, 3.0, sig), lambda: orc.fuse(of, proj(orc, pr, ow), op, 3.0, sig),
             lambda r: r[1])
        voc = synth.vocabulary_case(10, 5, seed=0)
        gvoc = orbfe.Vocabulary.from_arrays(10, 5, 0, 0, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"], device=local)
        ovoc = orc.Vocabulary.from_arrays(10, 5, 0, 0, voc["parent"], voc["is_leaf"], voc["desc"], voc["weight"])
        both("bow_transform", "DBoW2 transform: %d descriptors through a k=10, L=5 tree (%d nodes), levelsup 3" %
             (len(desc_frame), len(voc["parent"])), lambda: gvoc.transform(desc_frame, 3), lambda: ovoc.transform(desc_frame, 3),
             lambda r: len(r[0]))
        gvoc.close()
        dd, doff = synth.observations_case(5000, seed=0)
        both("compute_distinctive_descriptors", "5000 map points, %d observations" % len(dd),
             lambda: m.compute_distinctive_descriptors(dd, doff), lambda: orc.compute_distinctive_descriptors(dd, doff),
             lambda r: int((r[0] >= 0).sum()))
    m.close()
    return out


def shape_extras(torch, dist, orbfe, rank, world, local, steps, warmup):
    """configs[0] (TUM1 640x480 mono, 1000 features: extraction only) and configs[2] (EuRoC 752x480 stereo, 1200
    features, a batch of 256 pairs sharded by frame over the ranks: extraction L+R + ComputeStereoMatches) on the
    same device-resident footing as `value` (CUDA events on the handle's stream, max over ranks)."""
    lib = orbfe.lib()
    dev = torch.device("cuda", local)
    out = {}

    def timed(step_fn, stream, finalize=None):
        for _ in range(warmup):
            step_fn()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            step_fn()
        if finalize is not None:
            finalize()
        e1.record(stream)
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # configs[2]: 256 EuRoC pairs per step in total, rank r owns pairs r, r+world, ...
    s, (w, h) = synth.SETTINGS["euroc"], synth.SHAPES["euroc"]
    prm = (s["nfeatures"], s["scale_factor"], s["nlevels"], s["ini_th"], s["min_th"])
    total = int(os.environ.get("ORBFE_BENCH_EUROC_TOTAL", "256"))
    mine = len(range(rank, total, world))
    pairs = [synth.synth_pair(w, h, 1000 + rank + world * i) for i in range(4)]
    dl = torch.from_numpy(np.stack([pairs[i % 4][0] for i in range(mine)])).to(dev)
    dr = torch.from_numpy(np.stack([pairs[i % 4][1] for i in range(mine)])).to(dev)
    # two rigs (an extractor pair each) take the steps in turn: consecutive batches overlap, so the fixed tails of a
    # batch (one quadtree CTA per level and image, the small pyramid levels, the 64-CTA stereo kernels) are filled by
    # the next batch's kernels — what a stream of 256-pair batches looks like on every rank
    n_rigs = max(1, int(os.environ.get("ORBFE_BENCH_EUROC_RIGS", "2")))
    rigs = [(orbfe.Extractor(*prm, w, h, max_batch=mine, device=local), orbfe.Extractor(*prm, w, h, max_batch=mine, device=local))
            for _ in range(n_rigs)]
    L, R = rigs[0]
    mbf, mb = np.float32(s["bf"]), np.float32(s["bf"]) / np.float32(s["fx"])
    turn = [0]

    def euroc_step():
        Lk, Rk = rigs[turn[0] % n_rigs]
        turn[0] += 1
        Lk.upload(dl.data_ptr(), mine, on_device=True)
        Rk.upload(dr.data_ptr(), mine, on_device=True)
        Lk.run(mine)
        Rk.run(mine)
        orbfe.check(lib.orbfe_stereo_match_run(Lk._h, Rk._h, mine, mbf, mb))

    st0 = torch.cuda.ExternalStream(L.stream, device=dev)
    joins = [torch.cuda.Event() for _ in rigs[1:]]

    def euroc_join():  # the timed stream (rig 0, left) ends after every other rig's left stream (which ends its rig)
        for ev, (Lk, _) in zip(joins, rigs[1:]):
            ev.record(torch.cuda.ExternalStream(Lk.stream, device=dev))
            st0.wait_event(ev)

    ms = timed(euroc_step, st0, euroc_join)
    # the parity check below reads rig 0's results: make its last batch the freshest one
    torch.cuda.synchronize()
    turn[0] = 0
    euroc_step()
    torch.cuda.synchronize()
    sharded_equal = None
    if world > 1:
        # parity with N > 1 ranks on hardware: the results of the timed, sharded configuration (this rank's first 4 slots
        # = its 4 distinct pairs) are gathered on rank 0, which extracts the same 4 x world pairs alone in ONE batch
        import hashlib
        cap = L.capacity
        kl, kr = np.zeros((mine, cap), orbfe.KEYPOINT_DTYPE), np.zeros((mine, cap), orbfe.KEYPOINT_DTYPE)
        dsl, dsr = np.zeros((mine, cap, 32), np.uint8), np.zeros((mine, cap, 32), np.uint8)
        nl, nr = np.zeros(mine, np.int32), np.zeros(mine, np.int32)
        L.download(mine, kl, dsl, nl)
        R.download(mine, kr, dsr, nr)
        ur, dp, nm = orbfe.stereo_match(L, R, mine, mbf, mb)

        def digest(i, kl, dsl, nl, kr, dsr, nr, ur, dp):
            hsh = hashlib.sha256()
            for a in (kl[i][:nl[i]], dsl[i][:nl[i]], kr[i][:nr[i]], dsr[i][:nr[i]], ur[i][:nl[i]], dp[i][:nl[i]]):
                hsh.update(np.ascontiguousarray(a).tobytes())
            return hsh.hexdigest()

        own = [digest(i, kl, dsl, nl, kr, dsr, nr, ur, dp) for i in range(min(4, mine))]
        gathered = [None] * world if rank == 0 else None
        dist.gather_object(own, gathered, dst=0)
        flag = [None]
        if rank == 0:
            allp = [synth.synth_pair(w, h, 1000 + r + world * i) for r in range(world) for i in range(4)]
            n1 = len(allp)
            L1 = orbfe.Extractor(*prm, w, h, max_batch=n1, device=local)
            R1 = orbfe.Extractor(*prm, w, h, max_batch=n1, device=local)
            o = orbfe.stereo_frame(L1, R1, np.stack([p[0] for p in allp]), np.stack([p[1] for p in allp]), mbf, mb)
            single = [digest(i, o[0], o[1], o[2], o[3], o[4], o[5], o[6], o[7]) for i in range(n1)]
            flag[0] = all(gathered[r][i] == single[4 * r + i] for r in range(world) for i in range(len(gathered[r])))
            R1.close()
            L1.close()
        dist.broadcast_object_list(flag, src=0)
        sharded_equal = bool(flag[0])
    out["euroc_stereo_256"] = {
        "workload": "configs[2]: 256 EuRoC-shaped 752x480 stereo pairs per step, 1200 features, extract L+R + "
                    "ComputeStereoMatches, sharded by frame over %d GPU(s) (%d pairs on rank 0)" % (world, mine),
        "frames_per_s": total * steps / (ms * 1e-3), "ms_per_step": ms / steps, "scaling": "strong"}
    out["euroc_stereo_256"]["rigs"] = n_rigs
    if sharded_equal is not None:
        out["euroc_stereo_256"]["sharded_equal_to_single"] = sharded_equal
    for Lk, Rk in rigs:
        Rk.close()
        Lk.close()
    del dl, dr

    # configs[0]: TUM1 mono, 2 x 64 frames per step per GPU on two handles (two streams, like the two extractors of a
    # stereo rig: the tails of one handle's launches are filled by the other's)
    s, (w, h) = synth.SETTINGS["tum"], synth.SHAPES["tum"]
    prm = (s["nfeatures"], s["scale_factor"], s["nlevels"], s["ini_th"], s["min_th"])
    nb = 64
    imgs = [synth.synth_image(w, h, 2000 + rank * 8 + i) for i in range(8)]
    di = torch.from_numpy(np.stack([imgs[i % 8] for i in range(nb)])).to(dev)
    E = [orbfe.Extractor(*prm, w, h, max_batch=nb, device=local) for _ in range(2)]
    st = [torch.cuda.ExternalStream(e.stream, device=dev) for e in E]
    join = torch.cuda.Event()

    def tum_step():
        for e in E:
            e.upload(di.data_ptr(), nb, on_device=True)
            e.run(nb)
        join.record(st[1])       # the step ends when both streams are done: stream 0 (timed) waits for stream 1
        st[0].wait_event(join)

    ms = timed(tum_step, st[0])
    out["tum_mono_128"] = {
        "workload": "configs[0] shape: TUM1 640x480 mono, 1000 features, ORB extraction only, 2 x 64 frames per step "
                    "per GPU on two handles (two streams)",
        "frames_per_s": 2 * nb * world * steps / (ms * 1e-3), "ms_per_step": ms / steps, "scaling": "weak"}
    for e in E:
        e.close()
    return out


def pin_to_gpu_numa_node(local):
    """Run this rank's host threads (and first-touch its pinned buffers) on the CPUs next to its GPU: with 8 ranks the
    e2e leg moves ~60 GB/s per GPU over PCIe, and a buffer on the other socket halves that.  Returns the CPU count
    used, or None when NVML / affinity are not available."""
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = int(vis.split(",")[local]) if vis and all(t.strip().isdigit() for t in vis.split(",")) else local
        h = pynvml.nvmlDeviceGetHandleByIndex(idx)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = [c for c in range(ncpu) if (words[c // 64] >> (c % 64)) & 1]
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if cpus:
            os.sched_setaffinity(0, cpus)
            return len(cpus)
    except Exception:
        pass
    return None


def run_gpu(args):
    import torch
    import torch.distributed as dist
    from my_orb_slam2_b200 import orbfe

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if orbfe.device_count() < 1:
        raise RuntimeError("bench.py: no CUDA device; the product has no CPU fallback")
    torch.cuda.set_device(local)
    numa = pin_to_gpu_numa_node(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B = args.batch
    NB = args.batches
    dev = torch.device("cuda", local)

    # ---- synthetic inputs: frames sharded by rank (rank r owns seeds 1000*r ...), N_UNIQUE distinct pairs; the host
    # and device buffers hold N_UNIQUE + B images (image i = pair i mod N_UNIQUE) so that a batch of B consecutive
    # images can start at any offset < N_UNIQUE ----
    lefts_u, rights_u = make_pairs(N_UNIQUE, 1000 * rank)
    idx = np.arange(N_UNIQUE + B) % N_UNIQUE
    h_left = torch.empty((N_UNIQUE + B, H, W), dtype=torch.uint8).pin_memory()
    h_right = torch.empty((N_UNIQUE + B, H, W), dtype=torch.uint8).pin_memory()
    h_left.numpy()[:] = lefts_u[idx]
    h_right.numpy()[:] = rights_u[idx]
    d_left = h_left.to(dev)
    d_right = h_right.to(dev)
    offs = [(7 * j) % N_UNIQUE for j in range(max(NB, N_UNIQUE))]  # 7 is coprime to 64: all offsets are visited
    dptr = [(d_left.data_ptr() + o * W * H, d_right.data_ptr() + o * W * H) for o in offs]
    hptr = [(h_left.data_ptr() + o * W * H, h_right.data_ptr() + o * W * H) for o in offs]

    lib = orbfe.lib()

    class Rig:
        """A stereo rig: left + right extractor handles, each on its own stream as the reference drives them from
        two threads (src/Frame.cc:89-92; the stereo matcher joins them, and the right stream is ordered after the
        matcher's reads), plus pinned host result buffers.  Several rigs let the e2e leg overlap one batch's copies
        with another batch's kernels and a third one's result drain."""

        def __init__(self):
            self.L = orbfe.Extractor(*PARAMS, W, H, max_batch=B, device=local)
            self.R = orbfe.Extractor(*PARAMS, W, H, max_batch=B, device=local)
            cap = self.L.capacity
            self.kpL = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory()
            self.kpR = torch.empty((B, cap, 28), dtype=torch.uint8).pin_memory()
            self.dsL = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
            self.dsR = torch.empty((B, cap, 32), dtype=torch.uint8).pin_memory()
            self.cntL = torch.zeros(B, dtype=torch.int32).pin_memory()
            self.cntR = torch.zeros(B, dtype=torch.int32).pin_memory()
            self.ur = torch.empty((B, cap), dtype=torch.float32).pin_memory()
            self.dp = torch.empty((B, cap), dtype=torch.float32).pin_memory()
            self.nm = torch.zeros(B, dtype=torch.int32).pin_memory()

        def kernels(self):
            self.L.run(B)
            self.R.run(B)
            orbfe.check(lib.orbfe_stereo_match_run(self.L._h, self.R._h, B, MBF, MB))

        def download_async(self, right=True):
            cap = self.L.capacity
            # right=False: the right image's keypoints / descriptors stay on the device (only their counts leave);
            # Frame reads them nowhere outside ComputeStereoMatches (src/Frame.cc:516-594), which ran on the device
            orbfe.check(lib.orbfe_extract_download_async(self.R._h, B, self.kpR.data_ptr() if right else None,
                                                         self.dsR.data_ptr() if right else None, cap,
                                                         self.cntR.data_ptr()))
            orbfe.check(lib.orbfe_extract_download_async(self.L._h, B, self.kpL.data_ptr(), self.dsL.data_ptr(), cap,
                                                         self.cntL.data_ptr()))
            orbfe.check(lib.orbfe_stereo_match_download_async(self.L._h, B, self.ur.data_ptr(), self.dp.data_ptr(),
                                                              cap, self.nm.data_ptr()))

        def finish(self):
            orbfe.check(lib.orbfe_extract_finish(self.R._h))
            orbfe.check(lib.orbfe_extract_finish(self.L._h))

        def close(self):
            self.R.close()
            self.L.close()

    rigs = [Rig() for _ in range(args.rigs)]
    L, R = rigs[0].L, rigs[0].R
    cap = L.capacity
    sL = torch.cuda.ExternalStream(L.stream, device=dev)

    # consecutive batches take the rigs in turn (as in the e2e leg): the tail of a batch (quadtree CTAs, small pyramid
    # levels, the 64-CTA stereo kernels) overlaps the next batch's kernels.  1 rig 61.8 k, 2 rigs 63.5 k, 3 rigs 64.2 k
    n_vrigs = max(1, min(int(os.environ.get("ORBFE_BENCH_VALUE_RIGS", str(len(rigs)))), len(rigs)))

    def batch_device(j, ptrs=None):
        pl, pr = dptr[j % len(dptr)] if ptrs is None else ptrs
        g = rigs[j % n_vrigs]
        g.L.upload(pl, B, on_device=True)
        g.R.upload(pr, B, on_device=True)
        g.kernels()

    def step_device():
        for j in range(NB):
            batch_device(j)

    def join_rigs():
        for g in rigs[1:n_vrigs]:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.ExternalStream(g.L.stream, device=dev))
            sL.wait_event(ev)

    def batch_e2e(k, right=True):
        """pinned host images -> H2D -> kernels -> keypoints, descriptors, uRight, depth in pinned host memory.
        Software-pipelined over the rigs: batch k's copies overlap batch k-1's kernels (different streams)."""
        g = rigs[k % len(rigs)]
        g.finish()  # the results of the batch that used this rig last are complete and checked before reuse
        pl, pr = hptr[k % len(hptr)]
        g.L.upload(pl, B)
        g.R.upload(pr, B)
        g.kernels()
        g.download_async(right)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- device-resident leg (value) ----
    # every batch ends on L's stream (the stereo matcher waits for R's stream), so events on L's stream bracket it
    for _ in range(args.warmup):
        for j in range(min(NB, 8)):
            batch_device(j)
    barrier()
    launches0 = sum(g.L.launch_count + g.R.launch_count for g in rigs[:n_vrigs])
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(sL)
    for _ in range(args.steps):
        step_device()
    join_rigs()   # the timed stream ends after the other rigs' last batches
    e1.record(sL)
    barrier()
    ms_dev = max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.result()
    launches = sum(g.L.launch_count + g.R.launch_count for g in rigs[:n_vrigs]) - launches0
    # a few copies per batch (cudaMemcpy2DAsync D2D of level 0) are not kernels of ours and are not counted

    # ---- the same leg on DENSE images (uniform noise: every pixel a candidate region, ~4x the FAST / quadtree work;
    # right = left shifted by 8 px so the stereo matcher has true matches) ----
    n_noise = [synth.noise_image(W, H, 5000 + 100 * rank + i) for i in range(B)]
    d_nl = torch.from_numpy(np.stack(n_noise)).to(dev)
    d_nr = torch.from_numpy(np.stack([np.roll(im, -8, axis=1) for im in n_noise])).to(dev)
    nb_noise = max(4, NB // 5)
    for j in range(3):
        batch_device(j, (d_nl.data_ptr(), d_nr.data_ptr()))
    barrier()
    n0, n1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0.record(sL)
    for j in range(nb_noise):
        batch_device(j, (d_nl.data_ptr(), d_nr.data_ptr()))
    join_rigs()
    n1.record(sL)
    barrier()
    ms_noise = max_over_ranks(n0.elapsed_time(n1))
    noise = {"value": B * nb_noise * world / (ms_noise * 1e-3), "unit": "frames/s", "ms_per_batch": ms_noise / nb_noise,
             "batches": nb_noise, "unique_pairs": B,
             "workload": "same path on uniform-noise images (dense corners), right = left shifted 8 px"}
    del d_nl, d_nr

    # ---- per-stage pass: the same step, K more times, with the library's per-stage CUDA events on each handle's
    # stream and the two handles serialised (host sync in between) so that a stage's time is that kernel alone ----
    L.set_profiling(True)
    R.set_profiling(True)
    n_stage = min(100, max(10, NB))
    for j in range(n_stage):
        pl, pr = dptr[j % len(dptr)]
        L.upload(pl, B, on_device=True)
        L.run(B)
        L.sync()
        R.upload(pr, B, on_device=True)
        R.run(B)
        R.sync()
        orbfe.check(lib.orbfe_stereo_match_run(L._h, R._h, B, MBF, MB))
        L.sync()
    stL, runsL, srunsL = L.stage_times()
    stR, runsR, _ = R.stage_times()
    L.set_profiling(False)
    R.set_profiling(False)

    # ---- end-to-end leg (host buffers through the C ABI) ----
    for k in range(len(rigs) * max(2, args.warmup)):
        batch_e2e(k)
    for g in rigs:
        g.finish()
    barrier()
    sampler2 = ClockSampler(local)
    sampler2.start()
    t0 = time.perf_counter()
    for k in range(args.steps * NB):
        batch_e2e(k)
    for g in rigs:
        g.finish()
    torch.cuda.synchronize()
    t_e2e = max_over_ranks(time.perf_counter() - t0)
    clocks_e2e = sampler2.result()
    # the same leg with the right image's keypoints / descriptors left on the device (an OPTION of the C ABI: NULL
    # output pointers; the default above is faithful to what Frame keeps): half the D2H traffic.  Reported beside the
    # headline because D2H is the expensive direction of this host at N ranks (host.d2h_gbs_concurrent)
    n_lo = max(args.steps // 2, 1) * NB
    for k in range(len(rigs)):
        batch_e2e(k, right=False)
    for g in rigs:
        g.finish()
    barrier()
    t0 = time.perf_counter()
    for k in range(n_lo):
        batch_e2e(k, right=False)
    for g in rigs:
        g.finish()
    torch.cuda.synchronize()
    t_e2e_lo = max_over_ranks(time.perf_counter() - t0)
    e2e_left_only = {"value": n_lo * B * world / t_e2e_lo, "unit": "frames/s",
                     "d2h_bytes_per_step": int(NB * B * (cap * (28 + 32 + 8) + 12)),
                     "what": "as e2e, but the right image's keypoints and descriptors stay on the device (NULL output "
                             "pointers): Frame reads them only inside ComputeStereoMatches (src/Frame.cc:516-594)"}
    # what the host link can do with ALL ranks copying at once (no kernels): every rank moves its pinned image buffer
    # to the device / a result-sized buffer back, 5 times back to back, after a barrier; the sum over the ranks is
    # the concurrent ceiling the e2e leg's traffic is compared with
    nbytes = h_left.numel()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(5):
        d_left.copy_(h_left, non_blocking=True)
    ev1.record()
    torch.cuda.synchronize()
    h2d_rank = 5 * nbytes / (ev0.elapsed_time(ev1) * 1e-3) / 1e9
    barrier()
    ev0.record()
    for _ in range(5):
        h_right.copy_(d_right, non_blocking=True)
    ev1.record()
    torch.cuda.synchronize()
    d2h_rank = 5 * nbytes / (ev0.elapsed_time(ev1) * 1e-3) / 1e9
    barrier()
    # both directions at once: the D2H copies go on a second stream (on one stream the copies would run in turn)
    side = torch.cuda.Stream(device=dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    ev0.record()
    for _ in range(5):
        d_left.copy_(h_left, non_blocking=True)
    with torch.cuda.stream(side):
        for _ in range(5):
            h_right.copy_(d_right, non_blocking=True)
    torch.cuda.current_stream(dev).wait_stream(side)
    ev1.record()
    torch.cuda.synchronize()
    bidir_rank = 5 * 2 * nbytes / (ev0.elapsed_time(ev1) * 1e-3) / 1e9
    h2d_all, d2h_all, bidir_all = sum_over_ranks(h2d_rank), sum_over_ranks(d2h_rank), sum_over_ranks(bidir_rank)
    h_right.numpy()[:] = rights_u[idx]
    n_left = int(rigs[0].cntL.numpy().sum())
    n_match = int(rigs[0].nm.numpy().sum())
    kpL, kpR, dsL, dsR, cntL, cntR = rigs[0].kpL, rigs[0].kpR, rigs[0].dsL, rigs[0].dsR, rigs[0].cntL, rigs[0].cntR
    ur, dp, nm = rigs[0].ur, rigs[0].dp, rigs[0].nm

    # ---- latency leg: one pair in flight, host pointer in -> results on host ----
    lat = []
    if rank == 0:
        L1 = orbfe.Extractor(*PARAMS, W, H, max_batch=1, device=local)
        R1 = orbfe.Extractor(*PARAMS, W, H, max_batch=1, device=local)
        for i in range(args.latency_frames + 10):
            j = i % B
            t1 = time.perf_counter()
            # the stereo Frame constructor as one C-ABI call: uploads, both extractions (two streams), the stereo
            # matcher, the three result downloads queued behind them, one wait per stream
            orbfe.check(lib.orbfe_stereo_frame(L1._h, R1._h, h_left[j].data_ptr(), h_right[j].data_ptr(), W * H, W, 1,
                                               MBF, MB, kpL.data_ptr(), dsL.data_ptr(), cntL.data_ptr(),
                                               kpR.data_ptr(), dsR.data_ptr(), cntR.data_ptr(), ur.data_ptr(),
                                               dp.data_ptr(), nm.data_ptr(), cap))
            if i >= 10:
                lat.append(1e3 * (time.perf_counter() - t1))
        L1.close()
        R1.close()

    # ---- configs 3 / 4: the matchers (extra keys; every rank takes part in the sharded brute force) ----
    extras = None
    if not args.no_matchers:
        n0 = int(rigs[0].cntL[0])
        kp0 = rigs[0].kpL[0].numpy()[:n0].copy().view(orbfe.KEYPOINT_DTYPE).reshape(-1)
        ds0 = rigs[0].dsL[0].numpy()[:n0].copy()
        extras = matcher_extras(torch, dist, orbfe, rank, world, local, kp0, ds0, args.quick_matchers)
    for g in rigs:
        g.close()
    del d_left, d_right
    shapes = None if args.no_shapes else shape_extras(torch, dist, orbfe, rank, world, local, args.steps,
                                                      max(args.warmup, 3))
    if world > 1:
        dist.barrier()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- the drop-in call sequence Tracking would execute (src/Frame.cc:89-102), timed in C++: compat::ORBextractor
    # operator() for L and R from two threads on pageable cv::Mat buffers, join, ComputeStereoMatches ----
    dropin = None
    exe = os.path.join(ROOT, "compat", "_build", "bench_dropin")
    if os.path.exists(exe) and not args.no_dropin:
        import subprocess
        import tempfile
        with tempfile.NamedTemporaryFile(suffix=".raw") as tf:
            for i in range(8):
                tf.write(lefts_u[i].tobytes())
                tf.write(rights_u[i].tobytes())
            tf.flush()
            env = dict(os.environ, CUDA_VISIBLE_DEVICES=os.environ.get("CUDA_VISIBLE_DEVICES", "")) \
                if "CUDA_VISIBLE_DEVICES" in os.environ else dict(os.environ)
            try:
                r = subprocess.run([exe, tf.name, "8", str(W), str(H), str(S["nfeatures"]), repr(float(S["bf"])),
                                    repr(float(S["fx"])), str(args.latency_frames), "20"], capture_output=True, text=True,
                                   timeout=120, env=env)
                dropin = json.loads(r.stdout.strip().splitlines()[-1]) if r.returncode == 0 else {"error": r.stderr[-300:]}
            except Exception as ex:  # the figure is an extra: never fail the bench line for it
                dropin = {"error": repr(ex)[:300]}

    # ---- CPU baseline + parity block on bounded samples (rank 0, N=1 only) ----
    cpu, parity = None, None
    if world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(lefts_u, rights_u, args.cpu_seconds)
        parity = parity_block(orbfe, lefts_u, rights_u, local)

    # ---- roofline of the dominant kernel ----
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_gbs = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    lp = level_pixels()
    kp_per_img = n_left / max(B, 1)
    # algorithmic bytes per IMAGE of each stage ("Kernels"): compulsory reads + writes, no re-reads
    alg = {
        "pyramid": (lp - W * H) + (lp - int(W / 1.2 ** 7) * int(H / 1.2 ** 7)),  # read levels 0..6, write levels 1..7
        "fast": lp + 4 * 13000,                  # read every level once, write ~13 k candidates x 4 B
        "quadtree": 13000 * (4 + 2) + 2000 * 4,  # candidates + node ids in, kept out
        "blur": 2 * lp,                          # read every level once, write its blurred copy
        "describe": kp_per_img * (749 + 512 + 28 + 32 + 4),  # disc + pattern gathers in, keypoint + descriptor out
    }
    stage_ms = {k: (stL[k] + stR.get(k, 0.0)) for k in stL}
    runs = max(runsL + runsR, 1)
    per_launch = {k: stage_ms[k] / runs for k in ("pyramid", "fast", "quadtree", "blur", "describe")}
    per_launch["stereo_match"] = stL["stereo_match"] / max(srunsL, 1)
    per_launch["stereo_median"] = stL["stereo_median"] / max(srunsL, 1)
    dom = max(("pyramid", "fast", "quadtree", "blur", "describe"), key=lambda k: per_launch[k])
    achieved = alg[dom] * B / (per_launch[dom] * 1e-3) / 1e9 if per_launch[dom] > 0 else 0.0
    b_extract = W * H + lp + kp_per_img * 60  # compulsory bytes of the whole extraction per image
    # DRAM traffic of the dominant kernel from the committed `ncu --set full` capture (profiles/, per launch of 64 images)
    traffic = None
    issue = None
    try:
        tpath = [p for p in (os.path.join(ROOT, "profiles", t) for t in ("r2_traffic.json", "r1_traffic.json"))
                 if os.path.exists(p)][0]   # the latest committed capture
        tj = json.load(open(tpath))
        kname = {"fast": "fast_detect_kernel", "blur": "void blur_levels_kernel<32, 1>", "describe": "describe_kernel",
                 "pyramid": "void pyramid_level_kernel<8, 1>", "quadtree": "void quadtree_kernel<3>"}[dom]
        traffic = tj[kname]["dram_bytes_per_launch"] * B / tj[kname]["images"]
        # the bound that applies (north_star: "fraction of the HBM or integer-pipe roofline"): warp instructions of
        # the kernel (deterministic for these inputs; counted by ncu, committed in profiles/) / LIVE launch time,
        # against one warp instruction per SM sub-partition per clock
        winst = tj[kname].get("warp_inst_per_launch")
        if winst and per_launch[dom] > 0:
            a_i = winst * B / tj[kname]["images"] / (per_launch[dom] * 1e-3) / 1e12
            p_i = 148 * 4 * float(clocks.get("sm_mhz") or clocks.get("sm_max_mhz") or 1965.0) * 1e6 / 1e12
            issue = {"bound": "issue", "achieved": a_i, "peak": p_i, "unit": "T warp-inst/s", "frac": a_i / p_i,
                     "peak_source": "148 SMs x 4 schedulers x SM clock under load (1 warp instruction / scheduler / clk)",
                     "warp_inst_per_launch": winst * B / tj[kname]["images"]}
    except Exception:
        pass
    frames = B * NB * args.steps * world
    value = frames / (ms_dev * 1e-3)
    e2e_value = frames / t_e2e
    h2d_step = 2 * B * NB * W * H
    d2h_step = int(NB * (2 * B * cap * 60 + 2 * B * cap * 4 + 3 * B * 4))
    e2e_h2d_gbs = h2d_step * args.steps * world / t_e2e / 1e9
    e2e_d2h_gbs = d2h_step * args.steps * world / t_e2e / 1e9
    topo = None
    try:
        import subprocess
        topo = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout[:4000]
    except Exception:
        pass
    line = {
        "metric": METRIC,
        "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD,
                   "pairs_per_batch_per_gpu": B, "batches_per_step": NB, "pairs_per_step_per_gpu": B * NB,
                   "unique_pairs": N_UNIQUE, "rigs_in_flight": n_vrigs, "timed_region_s": ms_dev * 1e-3,
                   "l2": "inputs + pyramids of one batch = %.0f MB > 126 MB L2; every batch reads other images" %
                         (2 * B * (W * H + lp * 1.03) / 1e6),
                   "keypoints_per_image": kp_per_img, "stereo_matches_per_frame": n_match / max(B, 1)},
        "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d_step, "d2h_bytes_per_step": d2h_step,
                "timed_region_s": t_e2e,
                "latency_ms_p50": float(np.percentile(lat, 50)) if lat else None,
                "latency_ms_p95": float(np.percentile(lat, 95)) if lat else None,
                "latency_what": "one stereo pair in flight through orbfe_stereo_frame: pinned host images in, keypoints / "
                                "descriptors / uRight / depth on the host out",
                "dropin_latency": dropin,
                "dropin_what": "compat::ORBextractor::operator() for L and R on two threads (pageable buffers) + "
                               "ComputeStereoMatches, the sequence of src/Frame.cc:89-102, timed in C++ "
                               "(compat/bench_dropin.cc)",
                "left_only": e2e_left_only,
                "clocks": clocks_e2e},
        "latency_ms_p50": float(np.percentile(lat, 50)) if lat else None,
        "latency_ms_p95": float(np.percentile(lat, 95)) if lat else None,
        "gpu_launches": int(launches),
        "clocks": clocks,
        "noise": noise,
        "host": {"cpus_near_gpu": numa, "note": "each rank is pinned to the CPUs NVML reports next to its GPU before "
                                                "its pinned buffers are allocated",
                 "h2d_gbs_concurrent": h2d_all, "d2h_gbs_concurrent": d2h_all, "bidir_gbs_concurrent": bidir_all,
                 "probe": "all %d rank(s) copy their pinned %.0f MB image buffer at the same time, no kernels; sums over "
                          "ranks (bidir = H2D and D2H on two streams at the same time, bytes of both directions)" % (world, nbytes / 1e6),
                 "h2d_gbs_rank0": h2d_rank, "d2h_gbs_rank0": d2h_rank,
                 "e2e_h2d_gbs": e2e_h2d_gbs, "e2e_d2h_gbs": e2e_d2h_gbs,
                 "e2e_frac_of_link": e2e_h2d_gbs / h2d_all if h2d_all > 0 else None,
                 "e2e_frac_of_bidir_link": (e2e_h2d_gbs + e2e_d2h_gbs) / bidir_all if bidir_all > 0 else None,
                 "e2e_bound": "the host link: the e2e leg must move 2 x W x H input bytes per pair in and ~0.26 MB of "
                              "results out.  e2e_frac_of_link = its input traffic / the measured concurrent H2D-only "
                              "ceiling of the same ranks (the bound at 1-2 ranks); e2e_frac_of_bidir_link = its traffic "
                              "in BOTH directions / the measured ceiling with both directions busy (the bound from 4 "
                              "ranks on: this host's aggregate stops near 150 GB/s as soon as D2H traffic is mixed in, "
                              "although H2D alone reaches 236 GB/s at 8 ranks) — see e2e.left_only",
                 "topology": topo},
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak_gbs, "unit": "GB/s",
                     "frac": achieved / peak_gbs, "traffic": traffic, "peak_source": peak_src,
                     "timed": "per-stage CUDA events on the handle's stream, %d batches run right after the timed region "
                              "with the two handles serialised (the timed region overlaps their streams)" % n_stage,
                     "bound_note": "every extraction kernel is bound by integer instruction issue (ALU pipe), not HBM: "
                                   "see profiles/ (issue slots busy, pipe utilisation)",
                     "algorithmic_bytes_per_launch": alg[dom] * B, "issue": issue,
                     "extract_frac": (b_extract * 2 * frames / world / (ms_dev * 1e-3) / 1e9) / peak_gbs},
        "stages_ms_per_launch": per_launch,
        "cpu_baseline": cpu,
        "parity": parity,
        "matchers": extras,
        "shapes": shapes,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=64, help="stereo pairs per batch per GPU")
    ap.add_argument("--batches", type=int, default=100, help="batches per step (a step = batches x batch pairs)")
    ap.add_argument("--no-dropin", action="store_true", help="skip the C++ drop-in latency figure")
    ap.add_argument("--rigs", type=int, default=3, help="stereo rigs the e2e leg pipelines over (copies of step k+1, "
                    "kernels of step k, result drain of step k-1)")
    ap.add_argument("--latency-frames", type=int, default=200)
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-matchers", action="store_true", help="skip the configs 3/4 matcher figures")
    ap.add_argument("--no-shapes", action="store_true", help="skip the TUM / EuRoC shape figures")
    ap.add_argument("--quick-matchers", action="store_true", help="128 k map descriptors instead of 1 M")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
