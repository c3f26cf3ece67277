"""-m gpu, collected last on purpose (the file name sorts after every other test file): if this
optional entry point breaks, a ``pytest -x`` run has already executed all the other tests.

``b200km_assign_static_f32`` labels rows without any caller-provided scratch: the centre pack and
the list of rows that need the float64 re-decision sit in a fixed 6 MiB array inside the library.
It is for callers that may not allocate device memory themselves (a plugin that only ever sees
the buffers its host hands it, a thread that must never block in cudaMalloc).  Here it is called
straight through the C ABI, twice in a row with different data and centres -- the second call
finds the first call's leftovers in the shared array -- and each result is compared with the
regular pack + assign + refine entry points on the same inputs.
"""
import ctypes

import numpy as np
import numpy.testing as npt
import pytest

from conftest import blobs

pytestmark = pytest.mark.gpu


def _regular(eng, Xd, C64d):
    pack = eng.pack(C64d)
    labels, mind2 = eng.assign(Xd, pack)
    return labels.cpu().numpy(), mind2.cpu().numpy()


def _static(eng, Xd, C64d):
    import torch

    n, d = Xd.shape
    k = C64d.shape[0]
    labels = torch.full((n,), -1, dtype=torch.int32, device=eng.device)
    mind2 = torch.full((n,), -1.0, dtype=torch.float32, device=eng.device)
    rc = eng.lib.b200km_assign_static_f32(Xd.data_ptr(), n, d, d, C64d.data_ptr(), k, labels.data_ptr(),
                                          mind2.data_ptr(), eng._stream())
    assert rc == 0, eng.lib.b200km_error_string(rc)
    torch.cuda.synchronize()
    return labels.cpu().numpy(), mind2.cpu().numpy()


def test_static_scratch_entry_point_matches_regular_path_twice():
    import torch
    from dask_ml_b200._engine import get_engine

    torch.cuda.set_device(0)
    eng = get_engine()
    # two different problems back to back: the second one sees the first one's pack and flag list
    cases = [(50000, 64, 1024, 0), (7001, 20, 150, 1)]
    for n, d, k, seed in cases:
        X, C = blobs(n, d, k, seed=seed, dtype=np.float32)
        C = C + 0.4 * np.random.RandomState(seed).standard_normal(C.shape)  # near ties -> non-empty flag list
        Xd = eng.to_device(X)
        Cd = eng.to_device(C, torch.float64)
        lab_s, d2_s = _static(eng, Xd, Cd)
        lab_r, d2_r = _regular(eng, Xd, Cd)
        npt.assert_array_equal(lab_s, lab_r)
        npt.assert_allclose(d2_s, d2_r, rtol=1e-5, atol=1e-5 * float(d2_r.max()))
        assert (lab_s >= 0).all() and (lab_s < k).all()


def test_static_scratch_entry_point_rejects_oversize():
    import torch
    from dask_ml_b200._engine import get_engine

    torch.cuda.set_device(0)
    eng = get_engine()
    lib = eng.lib
    x = torch.zeros(8, 128, dtype=torch.float32, device=eng.device)
    c = torch.zeros(8192, 128, dtype=torch.float64, device=eng.device)
    lab = torch.zeros(8, dtype=torch.int32, device=eng.device)
    # centre pack larger than the built-in 2 MiB
    rc = lib.b200km_assign_static_f32(x.data_ptr(), 8, 128, 128, c.data_ptr(), 8192, lab.data_ptr(), None, None)
    assert rc == -6 and b"B200KM_STATIC_MAX_PACK_BYTES" in lib.b200km_error_string(rc)
    # more rows than the built-in re-decision list
    rc = lib.b200km_assign_static_f32(x.data_ptr(), (1 << 20) + 1, 128, 128, c.data_ptr(), 8, lab.data_ptr(), None, None)
    assert rc == -5 and b"B200KM_STATIC_MAX_ROWS" in lib.b200km_error_string(rc)
