"""GPU parity of the ORBmatcher path against the CPU oracle: DescriptorDistance, brute-force top-2,
SearchByProjection (both overloads, incl. the sequential claim and the rotation histogram) and
SearchForTriangulation.  All outputs are indices / integers: bit-exact."""
import numpy as np
import pytest

from cases import extract_with_oracle, views
from my_orb_slam2_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fe():
    from my_orb_slam2_b200 import frontend
    return frontend


@pytest.fixture(scope="module")
def kitti_features(oracle):
    params, img, ex, (kp, desc) = extract_with_oracle(oracle, "kitti", 0)
    return kp, desc


def test_descriptor_distance(fe, oracle):
    from my_orb_slam2_b200 import orbfe
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (1000, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (1000, 32), dtype=np.uint8)
    a[0] = 0; b[0] = 0; a[1] = 0; b[1] = 255
    got = orbfe.Matcher().descriptor_distance(a, b)
    want = np.unpackbits(a ^ b, axis=1).sum(1)
    assert np.array_equal(got, want) and got[0] == 0 and got[1] == 256
    assert fe.ORBmatcher.DescriptorDistance(a[5], b[5]) == oracle.descriptor_distance(a[5], b[5])


@pytest.mark.parametrize("nq,nm,base", [(1, 1, 0), (7, 33, 5), (600, 5000, (1 << 32) - 5000), (8192, 20000, 0)])
def test_bruteforce_vs_oracle(oracle, nq, nm, base):
    from my_orb_slam2_b200 import orbfe
    q, m = synth.bruteforce_case(nq, nm, min(nq, nm) // 2, seed=nq)
    if nm > 40:
        m[17] = m[3]  # exact duplicates: ties must go to the lowest index
        m[29] = m[3]
        q[0] = m[3]
    best, second = orbfe.Matcher().hamming_bruteforce(q, m, base)
    ob, os_ = oracle.hamming_bruteforce(q, m, base)
    assert np.array_equal(best, ob) and np.array_equal(second, os_)


@pytest.mark.parametrize("nq,nm,base", [(1, 1, 0), (7, 33, 5), (300, 5000, (1 << 32) - 5000), (1000, 100000, 11),
                                        (8192, 20000, 0)])
def test_bruteforce_tensor_core_path_vs_oracle(oracle, nq, nm, base):
    """orbfe_hamming_bruteforce_tc_device (tcgen05 kind::i8 GEMM on the expanded bits, k_match_mma.cu): keys bit-equal
    to the oracle's double loop over DescriptorDistance, ragged sizes, duplicates (ties to the lowest index), twice in a
    row on the same matcher (scratch and TMEM reuse)."""
    import torch
    from my_orb_slam2_b200 import orbfe
    q, m = synth.bruteforce_case(nq, nm, min(nq, nm) // 2, seed=nq + 1)
    if nm > 40:
        m[17] = m[3]
        m[nm - 1] = m[3]
        q[0] = m[3]
    ob, os_ = oracle.hamming_bruteforce(q, m, base)
    dev = torch.device("cuda", 0)
    d_q, d_m = torch.from_numpy(q).to(dev), torch.from_numpy(m).to(dev)
    mt = orbfe.Matcher()
    for _ in range(2):
        best = torch.zeros(nq, dtype=torch.int64, device=dev)
        second = torch.zeros(nq, dtype=torch.int64, device=dev)
        torch.cuda.synchronize()
        mt.hamming_bruteforce_tc_device(d_q.data_ptr(), nq, d_m.data_ptr(), nm, base, best.data_ptr(), second.data_ptr())
        mt.sync()
        assert np.array_equal(best.cpu().numpy().view(np.uint64), ob)
        assert np.array_equal(second.cpu().numpy().view(np.uint64), os_)


def test_bruteforce_tensor_core_equals_popc_on_random_shapes():
    """Seeded random (nq, nm, index_base) incl. sizes around the 128 / 256 tile edges: both device paths, same keys."""
    import torch
    from my_orb_slam2_b200 import orbfe
    rng = np.random.default_rng(123)
    dev = torch.device("cuda", 0)
    mt = orbfe.Matcher()
    shapes = [(127, 127), (128, 128), (129, 129), (255, 4095), (256, 4096), (257, 4097), (513, 16385)]
    shapes += [(int(rng.integers(1, 700)), int(rng.integers(1, 70000))) for _ in range(8)]
    for nq, nm in shapes:
        base = int(rng.integers(0, (1 << 32) - nm))
        q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
        m = rng.integers(0, 256, (nm, 32), dtype=np.uint8)
        m[rng.integers(0, nm, max(nm // 50, 1))] = q[rng.integers(0, nq, max(nm // 50, 1))]   # exact hits and ties
        d_q, d_m = torch.from_numpy(q).to(dev), torch.from_numpy(m).to(dev)
        out = [torch.zeros(nq, dtype=torch.int64, device=dev) for _ in range(4)]
        mt.hamming_bruteforce_device(d_q.data_ptr(), nq, d_m.data_ptr(), nm, base, out[0].data_ptr(), out[1].data_ptr())
        mt.hamming_bruteforce_tc_device(d_q.data_ptr(), nq, d_m.data_ptr(), nm, base, out[2].data_ptr(), out[3].data_ptr())
        mt.sync()
        assert torch.equal(out[0], out[2]) and torch.equal(out[1], out[3]), (nq, nm, base)


def test_bruteforce_full_size_properties():
    """Config 5 per-GPU shard (8192 x 131072) and a 1M map: checked through properties the oracle-free way:
    planted duplicates are found, sharded top-2 merged == unsharded top-2, best <= second."""
    from my_orb_slam2_b200 import orbfe
    q, m = synth.bruteforce_case(8192, 1 << 20, 4096, seed=1)
    mt = orbfe.Matcher()
    best, second = mt.hamming_bruteforce(q, m, 0)
    assert np.all(best <= second)
    # exact check on a sample of queries with numpy
    for qi in [0, 1, 777, 8191]:
        d = np.unpackbits(q[qi][None] ^ m, axis=1).sum(1).astype(np.uint64)
        keys = np.sort((d << np.uint64(32)) | np.arange(len(m), dtype=np.uint64))
        assert best[qi] == keys[0] and second[qi] == keys[1]
    assert (best >> np.uint64(32) <= 20).sum() >= 4096 - 8  # planted near-duplicates found
    # 8 shards merged on the host == unsharded
    shard = (1 << 20) // 8
    parts = [mt.hamming_bruteforce(q, m[i * shard:(i + 1) * shard], i * shard) for i in range(8)]
    allk = np.sort(np.concatenate([np.stack(p, 1) for p in parts], 1), axis=1)
    assert np.array_equal(allk[:, 0], best) and np.array_equal(allk[:, 1], second)


@pytest.mark.parametrize("th,seed", [(1.0, 0), (3.0, 1), (5.0, 2)])
def test_search_by_projection_map(fe, oracle, kitti_features, th, seed):
    kp, desc = kitti_features
    frame = synth.frame_arrays(kp, desc, (0.0, 0.0, 1241.0, 376.0), seed=seed)
    mps = synth.map_points_case(frame, 20000, 1500, seed=seed)
    ofv, omv = views(oracle, frame=frame, mps=mps)
    want, wn = oracle.search_by_projection_map(ofv, omv, th, 0.8)
    fv, mv = views(fe, frame=frame, mps=mps)
    got, gn = fe.ORBmatcher(0.8).SearchByProjection(fv, mv, th)
    assert gn == wn and np.array_equal(got, want)
    assert wn > 300


def test_search_by_projection_map_heavy_conflicts(fe, oracle, kitti_features):
    # every map point is a "true" one drawn with replacement from 200 features: long claim chains
    kp, desc = kitti_features
    frame = synth.frame_arrays(kp[:200], desc[:200], (0.0, 0.0, 1241.0, 376.0), seed=4, taken_frac=0.0)
    mps = synth.map_points_case(frame, 3000, 3000, seed=4)
    ofv, omv = views(oracle, frame=frame, mps=mps)
    want, wn = oracle.search_by_projection_map(ofv, omv, 5.0, 0.8)
    fv, mv = views(fe, frame=frame, mps=mps)
    got, gn = fe.ORBmatcher(0.8).SearchByProjection(fv, mv, 5.0)
    assert gn == wn and np.array_equal(got, want)


@pytest.mark.parametrize("tz,mono,check_ori,th", [(0.0, False, True, 7.0), (2.0, False, True, 7.0),
                                                    (-2.0, False, True, 15.0), (2.0, True, False, 15.0)])
def test_search_by_projection_last(fe, oracle, kitti_features, tz, mono, check_ori, th):
    kp, desc = kitti_features
    frame = synth.frame_arrays(kp, desc, (0.0, 0.0, 1241.0, 376.0), seed=7)
    last, cam = synth.last_frame_case(frame, seed=7, tz=tz)
    ofv, olv, ocv = views(oracle, frame=frame, last=last, cam=cam)
    want, wn = oracle.search_by_projection_last(ofv, ocv, olv, th, mono, check_ori)
    fv, lv, cv = views(fe, frame=frame, last=last, cam=cam)
    got, gn = fe.ORBmatcher(0.9, check_ori).SearchByProjection(fv, lv, th, bMono=mono, cam=cv)
    assert gn == wn and np.array_equal(got, want)
    assert wn > 300


@pytest.mark.parametrize("only_stereo,check_ori", [(False, False), (False, True), (True, True)])
def test_search_for_triangulation(fe, oracle, kitti_features, only_stereo, check_ori):
    kp, desc = kitti_features
    k1, k2, F12, ep = synth.keyframe_pair_case(kp, desc, seed=3)
    want = oracle.search_for_triangulation(views(oracle, kf=k1), views(oracle, kf=k2), F12, float(ep[0]), float(ep[1]),
                                           only_stereo, check_ori)
    got = fe.ORBmatcher(0.6, check_ori).SearchForTriangulation(views(fe, kf=k1), views(fe, kf=k2), F12, ep, only_stereo)
    assert np.array_equal(got, want)
    assert len(want) > (50 if only_stereo else 300)


def test_empty_inputs(fe):
    from my_orb_slam2_b200 import orbfe
    mt = orbfe.Matcher()
    kp = np.zeros(0, orbfe.KEYPOINT_DTYPE)
    frame = synth.frame_arrays(kp, np.zeros((0, 32), np.uint8), (0.0, 0.0, 100.0, 100.0))
    mps = synth.map_points_case(synth.frame_arrays(np.zeros(1, orbfe.KEYPOINT_DTYPE), np.zeros((1, 32), np.uint8),
                                                   (0.0, 0.0, 100.0, 100.0)), 10, 5)
    fv, mv = views(fe, frame=frame, mps=mps)
    got, n = mt.search_by_projection_map(fv, mv, 1.0, 0.8)
    assert n == 0 and len(got) == 0


def test_bruteforce_sharded_single_process(oracle):
    """The multi-GPU path of config 5 on one GPU: per-shard CUDA top-2 with global index bases + the CUDA merge kernel
    (the NCCL all-gather in between is exercised by bench.py --gpus N; its layout by tests/test_sharding_gloo.py)."""
    import torch
    from my_orb_slam2_b200 import orbfe, sharding
    nq, nm, world = 300, 7000, 3
    q, m = synth.bruteforce_case(nq, nm, 150, seed=5)
    m[nm - 1] = m[2]
    q[1] = m[2]  # a tie across shards: the lower global index must win
    mt = orbfe.Matcher()
    dev = torch.device("cuda", 0)
    d_q = torch.from_numpy(q).to(dev)
    base = 12345
    bests, seconds = [], []
    sb = sharding.ShardedBruteForce(mt, dev)
    for r in range(world):
        start, cnt = sharding.map_shard(nm, r, world)
        d_m = torch.from_numpy(m[start:start + cnt].copy()).to(dev)
        b, s = sb.run(d_q, d_m, base + start)  # world == 1 (no process group): the local top-2
        bests.append(b)
        seconds.append(s)
    ba, sa = torch.stack(bests).contiguous(), torch.stack(seconds).contiguous()
    gb, gs = torch.empty_like(bests[0]), torch.empty_like(bests[0])
    torch.cuda.synchronize()
    mt.hamming_merge_device(ba.data_ptr(), sa.data_ptr(), world, nq, gb.data_ptr(), gs.data_ptr())
    mt.sync()
    ob, os_ = oracle.hamming_bruteforce(q, m, base)
    assert np.array_equal(gb.cpu().numpy().view(np.uint64), ob)
    assert np.array_equal(gs.cpu().numpy().view(np.uint64), os_)
    assert mt.popc_peak() > 1e11
