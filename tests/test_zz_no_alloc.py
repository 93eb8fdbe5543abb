"""mtd_occupancy_noalloc (include/mtd_b200.h): the occupancy count with its scratch in a static __device__ array.

No counterpart in the reference.  Collected last on purpose:
a failure here must not hide another test.  The entry point is called through the C ABI, twice
in a row (the static scratch keeps the first call's contents), and compared exactly -- integer counts -- with the
handle's own mtd_occupancy on the same states.
"""
import ctypes as C

import numpy as np
import pytest

from mtdcovabm_b200 import capi


def _noalloc(lib, seir_dev, n, out_dev, stream=None):
    err = C.create_string_buffer(256)
    rc = lib.mtd_occupancy_noalloc(C.c_int32(0), C.c_void_p(seir_dev.data_ptr()), C.c_int64(n),
                                   C.c_void_p(out_dev.data_ptr()), C.c_void_p(stream), err, C.c_int64(len(err)))
    return rc, err.value.decode()


@pytest.mark.gpu
def test_noalloc_occupancy_matches_handle_path_twice_in_a_row():
    import torch
    lib = capi.load()
    rng = np.random.default_rng(20)
    # first call: 3.2M agents (49 blocks, ragged tail); second call: fewer agents and other states, so rows of the
    # static scratch that the first call filled are stale when the second one runs
    cases = [(3_200_013, [0.5, 0.1, 0.1, 0.1, 0.15, 0.05]), (1_000_003, [0.05, 0.3, 0.05, 0.2, 0.1, 0.3])]
    handles = [capi.Handle(n, 4, 64) for n, _ in cases]
    try:
        want, got = [], []
        for (n, p), h in zip(cases, handles):
            seir = rng.choice(6, size=n, p=p).astype(np.int8)
            h.set_state(seir=seir)
            want.append(np.asarray(h.occupancy(), np.int64))
            assert want[-1].sum() == n and np.array_equal(want[-1], np.bincount(seir, minlength=6))
            d_seir = torch.from_numpy(seir).cuda()
            d_out = torch.full((6,), -1, dtype=torch.int64, device="cuda")
            torch.cuda.synchronize()
            rc, msg = _noalloc(lib, d_seir, n, d_out, torch.cuda.current_stream().cuda_stream)
            assert rc == 0, msg
            torch.cuda.synchronize()
            got.append(d_out.cpu().numpy())
        for w, g in zip(want, got):
            assert np.array_equal(w, g), (w, g)
        # back-to-back on one stream without a host sync in between: the second call must not see the first one's ticket
        n, p = cases[0]
        seir_a = torch.from_numpy(rng.choice(6, size=n, p=p).astype(np.int8)).cuda()
        seir_b = torch.from_numpy(rng.choice(6, size=70_001).astype(np.int8)).cuda()
        out_a = torch.zeros(6, dtype=torch.int64, device="cuda")
        out_b = torch.zeros(6, dtype=torch.int64, device="cuda")
        s = torch.cuda.current_stream().cuda_stream
        assert _noalloc(lib, seir_a, n, out_a, s)[0] == 0
        assert _noalloc(lib, seir_b, 70_001, out_b, s)[0] == 0
        torch.cuda.synchronize()
        assert np.array_equal(out_a.cpu().numpy(), np.bincount(seir_a.cpu().numpy(), minlength=6))
        assert np.array_equal(out_b.cpu().numpy(), np.bincount(seir_b.cpu().numpy(), minlength=6))
    finally:
        for h in handles:
            h.close()


@pytest.mark.gpu
def test_noalloc_occupancy_rejects_sizes_beyond_the_static_scratch():
    import torch
    lib = capi.load()
    d = torch.zeros(16, dtype=torch.int8, device="cuda")
    out = torch.zeros(6, dtype=torch.int64, device="cuda")
    rc, msg = _noalloc(lib, d, capi.NOALLOC_MAX_AGENTS + 1, out)
    assert rc == 3 and "MTD_NOALLOC_MAX_AGENTS" in msg          # MTD_E_CAPACITY, nothing launched
    rc, msg = _noalloc(lib, d, 16, out)
    assert rc == 0, msg
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [16, 0, 0, 0, 0, 0]
    rc, msg = _noalloc(lib, d, 0, out)                          # empty population: six zeros
    assert rc == 0, msg
    torch.cuda.synchronize()
    assert out.cpu().tolist() == [0] * 6
