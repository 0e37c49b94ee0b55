"""orbm_knn2_static — the allocation-free 2-NN entry point (include/orbx.h; no counterpart in the reference).
Called through the C ABI with device pointers, twice in a row (the calls share one
static device scratch array), and compared bit for bit with the existing entry point orbm_knn2 and the CPU oracle.
Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

from oracle_lib import OraclePrims
from orbslam2_b200.synth import synth_descriptors

pytestmark = pytest.mark.gpu


def test_knn2_static_matches_knn2_on_two_calls_in_a_row():
    import torch
    import orbslam2_b200 as ob
    from orbslam2_b200 import _lib
    L = _lib.lib()
    t, q = synth_descriptors(70000, 6000, seed=31, dup=400)
    t[0] = t[69999]; q[0] = t[0]                       # a tie across the first and the last train split
    dev = torch.device("cuda", 0)
    dt = torch.from_numpy(t).to(dev)
    m = ob.ORBmatcher(device=0)
    m.set_knn_mode(1)                                  # the existing entry point on the same (POPC) path ...
    m.set_train(dt)
    stream = torch.cuda.current_stream(dev)
    P = OraclePrims()
    # two calls in a row with different query sets and sizes: the second must not see anything of the first
    for call, (lo, hi) in enumerate(((0, 6000), (1500, 1900))):
        qq = np.ascontiguousarray(q[lo:hi]); n = hi - lo
        dq = torch.from_numpy(qq).to(dev)
        i1 = torch.full((n,), -7, dtype=torch.int32, device=dev)
        d1 = torch.full((n,), -7, dtype=torch.int16, device=dev)
        d2 = torch.full((n,), -7, dtype=torch.int16, device=dev)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info(dev)[0]
        rc = L.orbm_knn2_static(0, C.c_void_p(stream.cuda_stream), dq.data_ptr(), n, dt.data_ptr(), len(t), i1.data_ptr(),
                                d1.data_ptr(), d2.data_ptr())
        assert rc == 0, _lib.check(rc)
        stream.synchronize()
        if call:                                       # (the very first call may load the module lazily)
            assert torch.cuda.mem_get_info(dev)[0] == free0, "the allocation-free entry point changed free device memory"
        r1, e1, e2 = m.knn2(dq)
        assert m.last_knn_path() == 1
        assert torch.equal(i1, r1) and torch.equal(d1, e1) and torch.equal(d2, e2)
        o1, f1, f2 = P.knn2(qq[:300], t, popcnt=True)
        assert np.array_equal(i1.cpu().numpy()[:300], o1)
        assert np.array_equal(d1.cpu().numpy().view(np.uint16)[:300], f1) and np.array_equal(d2.cpu().numpy().view(np.uint16)[:300], f2)
    # ... and the tensor-core path of the existing entry point gives the same answer too
    m.set_knn_mode(2)
    r1, e1, e2 = m.knn2(dq)
    assert torch.equal(i1, r1) and torch.equal(d1, e1) and torch.equal(d2, e2)
    m.close()


def test_knn2_static_rejects_sizes_beyond_the_scratch():
    import torch
    from orbslam2_b200 import _lib
    L = _lib.lib()
    L.orbx_last_error_string.restype = C.c_char_p
    buf = torch.zeros(1024, dtype=torch.uint8, device="cuda:0")
    out = torch.zeros(64, dtype=torch.int32, device="cuda:0")
    rc = L.orbm_knn2_static(0, None, buf.data_ptr(), 65537, buf.data_ptr(), 32, out.data_ptr(), out.data_ptr(), out.data_ptr())
    assert rc == _lib.ORBX_EINVAL
    assert b"65536" in L.orbx_last_error_string()
    # the limit itself is the query count only; misaligned pointers and empty train sets are refused as well
    assert L.orbm_knn2_static(0, None, buf.data_ptr() + 1, 4, buf.data_ptr(), 32, out.data_ptr(), out.data_ptr(), out.data_ptr()) == _lib.ORBX_EINVAL
    assert L.orbm_knn2_static(0, None, buf.data_ptr(), 4, buf.data_ptr(), 0, out.data_ptr(), out.data_ptr(), out.data_ptr()) == _lib.ORBX_EINVAL
