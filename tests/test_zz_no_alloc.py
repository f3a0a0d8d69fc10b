"""orbx_hamming_best2_noalloc (csrc/no_alloc.cu): the allocation-free variant of the raw all-pairs Hamming entry point,
called through the C ABI with device buffers, against orbx_hamming_best2 on the same inputs (integer results: equal).
Two calls in a row share the library's fixed scratch table: the second (fewer queries, other data) must not see the
first one's rows.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _noalloc(torch, capi, E, a, b, stream=None, synchronize=1):
    dev = torch.device("cuda", 0)
    da = torch.from_numpy(a).to(dev)
    db = torch.from_numpy(b).to(dev)
    out = [torch.full((len(a),), -7, dtype=torch.int32, device=dev) for _ in range(3)]
    torch.cuda.synchronize()
    rc = capi.lib().orbx_hamming_best2_noalloc(E._h, da.data_ptr(), len(a), db.data_ptr(), len(b), out[0].data_ptr(),
                                               out[1].data_ptr(), out[2].data_ptr(), stream, synchronize)
    capi.check(rc, E._h)
    torch.cuda.synchronize()
    return [o.cpu().numpy() for o in out]


def test_noalloc_matches_existing_entry_point_twice():
    import torch
    from orb_slam_tracking_b200 import capi, frontend
    rng = np.random.default_rng(11)
    E = frontend.ORBextractor(100, 1.2, 2, 20, 7, 640, 480)
    a1 = rng.integers(0, 256, (3000, 32), dtype=np.uint8)
    b1 = rng.integers(0, 256, (2500, 32), dtype=np.uint8)
    b1[100:140] = a1[:40]   # exact matches, twice: the first in index order wins, second distance 0
    b1[300:340] = a1[:40]
    a2 = rng.integers(0, 256, (700, 32), dtype=np.uint8)   # fewer queries than the first call: stale rows stay behind them
    b2 = rng.integers(0, 256, (1, 32), dtype=np.uint8)     # one candidate: second distance -1
    for a, b in ((a1, b1), (a2, b2)):
        bi, bd, sd, _ = E.hamming_best2(a, b)
        gi, gd, gs = _noalloc(torch, capi, E, a, b)
        assert np.array_equal(gi, bi) and np.array_equal(gd, bd) and np.array_equal(gs, sd)
    # on the caller's stream, enqueue only
    bi, bd, sd, _ = E.hamming_best2(a1, b1)
    gi, gd, gs = _noalloc(torch, capi, E, a1, b1, stream=C.c_void_p(torch.cuda.current_stream().cuda_stream), synchronize=0)
    assert np.array_equal(gi, bi) and np.array_equal(gd, bd) and np.array_equal(gs, sd)
    # NULL outputs are allowed
    dev = torch.device("cuda", 0)
    da, db = torch.from_numpy(a2).to(dev), torch.from_numpy(b2).to(dev)
    capi.check(capi.lib().orbx_hamming_best2_noalloc(E._h, da.data_ptr(), len(a2), db.data_ptr(), 1, None, None, None, None, 1), E._h)
    E.close()


def test_noalloc_rejects_sizes_beyond_the_fixed_table():
    import torch
    from orb_slam_tracking_b200 import capi, frontend
    E = frontend.ORBextractor(100, 1.2, 2, 20, 7, 640, 480)
    dev = torch.device("cuda", 0)
    n = 16384 + 1
    da = torch.zeros((n, 32), dtype=torch.uint8, device=dev)
    out = torch.zeros(n, dtype=torch.int32, device=dev)
    L = capi.lib()
    rc = L.orbx_hamming_best2_noalloc(E._h, da.data_ptr(), n, da.data_ptr(), 8, out.data_ptr(), None, None, None, 1)
    assert rc == -5  # ORBX_ERR_CAPACITY
    with pytest.raises(capi.OrbxError) as ei:
        capi.check(rc, E._h)
    assert "16384" in str(ei.value)
    assert L.orbx_hamming_best2_noalloc(E._h, da.data_ptr(), 16384, da.data_ptr(), 8, out.data_ptr(), None, None, None, 1) == 0
    assert L.orbx_hamming_best2_noalloc(E._h, None, 4, da.data_ptr(), 8, None, None, None, None, 1) == -1
    E.close()
