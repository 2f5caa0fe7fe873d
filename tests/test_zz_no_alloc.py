"""orbfe_hamming_bruteforce_device_noalloc (include/orbfe.h): the allocation-free variant of the brute-force Hamming
top-2 (static __device__ scratch, my_orb_slam2_b200/csrc/noalloc.cu) against the existing allocating entry point
orbfe_hamming_bruteforce_device, through the C ABI, on two calls in a row (the shared scratch keeps the previous call's
contents).  Keys are integers: the comparison is bit-exact, like every other matcher test.

This file is named so that pytest collects it LAST.  It passes alone in a fresh process
and after the other tests."""
import numpy as np
import pytest

from my_orb_slam2_b200 import synth

pytestmark = pytest.mark.gpu


def _run_pair(orbfe, torch, m, q, mp, index_base):
    lib = orbfe.lib()
    dev = torch.device("cuda", 0)
    dq = torch.from_numpy(q).to(dev)
    dm = torch.from_numpy(mp).to(dev) if len(mp) else torch.zeros((1, 32), dtype=torch.uint8, device=dev)
    nq, nm = len(q), len(mp)
    out = [torch.zeros(nq, dtype=torch.int64, device=dev) for _ in range(4)]
    wb, ws, gb, gs = out
    torch.cuda.synchronize()
    # the existing entry point (allocates its partial-result buffer), on the matcher's stream
    orbfe.check(lib.orbfe_hamming_bruteforce_device(m._h, dq.data_ptr(), nq, dm.data_ptr(), nm, index_base,
                                                    wb.data_ptr(), ws.data_ptr()))
    m.sync()
    # the new one: no handle, no allocation, on torch's current stream
    stream = torch.cuda.current_stream().cuda_stream
    orbfe.check(lib.orbfe_hamming_bruteforce_device_noalloc(stream, dq.data_ptr(), nq, dm.data_ptr(), nm, index_base,
                                                            gb.data_ptr(), gs.data_ptr()))
    torch.cuda.synchronize()
    return [t.cpu().numpy().view(np.uint64) for t in (wb, ws, gb, gs)]


def test_noalloc_bruteforce_matches_existing_entry_point_twice():
    import torch
    from my_orb_slam2_b200 import orbfe
    m = orbfe.Matcher()
    # call 1: many slices of a larger map; call 2: fewer queries and a smaller map, so most of the scratch still holds
    # call 1's partial results when call 2 runs
    cases = [synth.bruteforce_case(4096, 1 << 17, 512, seed=5) + (0,),
             synth.bruteforce_case(300, 5000, 64, seed=6) + (1 << 31,),
             synth.bruteforce_case(4096, 1 << 17, 512, seed=7) + (12345,)]
    for q, mp, base in cases:
        wb, ws, gb, gs = _run_pair(orbfe, torch, m, q, mp, base)
        assert np.array_equal(gb, wb) and np.array_equal(gs, ws)
        assert (wb >> np.uint64(32)).min() <= 20  # the planted near-duplicates are found
        assert np.all((wb & np.uint64(0xFFFFFFFF)) >= np.uint64(base))
    # empty map: both keys are "none"
    q = cases[1][0]
    wb, ws, gb, gs = _run_pair(orbfe, torch, m, q, np.zeros((0, 32), np.uint8), 0)
    assert np.array_equal(gb, wb) and np.array_equal(gs, ws) and np.all(gb == np.uint64(0xFFFFFFFFFFFFFFFF))


def test_noalloc_bruteforce_rejects_sizes_beyond_the_static_scratch():
    import torch
    from my_orb_slam2_b200 import orbfe
    lib = orbfe.lib()
    dev = torch.device("cuda", 0)
    nq = 32768 + 1
    dq = torch.zeros((nq, 32), dtype=torch.uint8, device=dev)
    dm = torch.zeros((1024, 32), dtype=torch.uint8, device=dev)
    b = torch.zeros(nq, dtype=torch.int64, device=dev)
    s = torch.zeros(nq, dtype=torch.int64, device=dev)
    rc = lib.orbfe_hamming_bruteforce_device_noalloc(None, dq.data_ptr(), nq, dm.data_ptr(), 1024, 0, b.data_ptr(),
                                                     s.data_ptr())
    assert rc == -3  # ORBFE_ERR_CAPACITY
    assert b"32768" in lib.orbfe_last_error()
    # the limit itself is accepted
    rc = lib.orbfe_hamming_bruteforce_device_noalloc(None, dq.data_ptr(), 32768, dm.data_ptr(), 1024, 0, b.data_ptr(),
                                                     s.data_ptr())
    torch.cuda.synchronize()
    assert rc == 0
