"""mmd_pairwise_accuracy_static (include/mmd_b200.h): the allocation-free variant of mmd_pairwise_accuracy whose scratch is
a file-scope `static __device__` array inside libmmd_b200.so.  Called through the C ABI and compared with the existing
entry point on the same inputs.  Tolerance: the accuracy tests of this suite (tests/test_gpu_eval_null.py) require exact
equality, and so does this one (same kernels, same order of operations).

The file name keeps this test last in pytest's collection order."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_ROWS, MAX_DIMS = 256, 1024          # MMD_STATIC_MAX_ROWS / MMD_STATIC_MAX_DIMS


def _static(lib, torch, lat, pred, std_pred, std_lat, device=0):
    """One call of the new entry point: device inputs and output, work enqueued on torch's current stream."""
    dev = torch.device("cuda", device)
    d_lat = torch.from_numpy(np.ascontiguousarray(lat, dtype=np.float64)).to(dev)
    d_pred = torch.from_numpy(np.ascontiguousarray(pred, dtype=np.float64)).to(dev)
    d_acc = torch.full((1,), -1.0, dtype=torch.float64, device=dev)
    err = C.create_string_buffer(512)
    stream = torch.cuda.current_stream(dev).cuda_stream
    rc = lib.mmd_pairwise_accuracy_static(device, stream, d_lat.data_ptr(), d_lat.shape[0], d_pred.data_ptr(), d_pred.shape[0],
                                          d_lat.shape[1], int(std_pred), int(std_lat), d_acc.data_ptr(), err, len(err))
    torch.cuda.synchronize(dev)
    # the inputs are borrowed read-only
    assert np.array_equal(d_lat.cpu().numpy(), lat) and np.array_equal(d_pred.cpu().numpy(), pred)
    return rc, float(d_acc.cpu()[0]), err.value.decode()


def _existing(lib, h, lat, pred, std_pred, std_lat):
    acc = C.c_double(-1.0)
    rc = lib.mmd_pairwise_accuracy(h, lat.ctypes.data, len(lat), pred.ctypes.data, len(pred), lat.shape[1], int(std_pred),
                                   int(std_lat), C.byref(acc))
    assert rc == 0, lib.mmd_last_error(h)
    return acc.value


def test_static_scratch_entry_point_matches_the_allocating_one():
    import torch
    from multimodal_decoding_b200 import _lib
    lib = _lib.load()
    h = C.c_void_p()
    assert lib.mmd_create(0, C.byref(h)) == 0
    try:
        rng = np.random.default_rng(20261105)
        # the searchlight's own shapes (140 test + 3 imagery rows, 1024 dims), the limits, ragged and tiny shapes
        cases = [(143, 143, 1024), (143, 3, 1024), (MAX_ROWS, MAX_ROWS, MAX_DIMS), (200, 180, 129), (70, 70, 33), (1, 1, 1),
                 (5, 1, 7)]
        for call in range(2):                  # two calls in a row per configuration: the second finds the first's scratch
            for n_lat, n_pred, d in cases:
                lat = rng.standard_normal((n_lat, d))
                pred = 0.3 * lat[rng.integers(0, n_lat, n_pred)] + rng.standard_normal((n_pred, d))
                for std_pred, std_lat in ((False, False), (True, False), (True, True)):
                    rc, got, msg = _static(lib, torch, lat, pred, std_pred, std_lat)
                    assert rc == 0, msg
                    want = _existing(lib, h, lat, pred, std_pred, std_lat)
                    if n_lat * n_pred == min(n_lat, n_pred):          # 0 / 0 comparisons: NaN from both
                        assert np.isnan(got) and np.isnan(want)
                    else:
                        assert got == want, (call, n_lat, n_pred, d, std_pred, std_lat, got, want)
        # the very same inputs twice, back to back, without anything in between
        lat = rng.standard_normal((143, 1024))
        pred = lat + 30.0 * rng.standard_normal((143, 1024))
        first = _static(lib, torch, lat, pred, True, False)
        second = _static(lib, torch, lat, pred, True, False)
        assert first[0] == 0 and second[0] == 0 and first[1] == second[1] == _existing(lib, h, lat, pred, True, False)
        assert 0.5 < first[1] < 1.0
    finally:
        lib.mmd_destroy(h)


def test_static_scratch_entry_point_rejects_what_does_not_fit():
    import torch
    from multimodal_decoding_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(1)
    for n_lat, n_pred, d in ((MAX_ROWS + 1, 10, 16), (MAX_ROWS + 1, MAX_ROWS + 1, 16), (20, 10, MAX_DIMS + 1)):
        rc, acc, msg = _static(lib, torch, rng.standard_normal((n_lat, d)), rng.standard_normal((n_pred, d)), True, False)
        assert rc != 0 and "exceed the static scratch" in msg and str(MAX_ROWS) in msg and str(MAX_DIMS) in msg
        assert acc == -1.0                                           # nothing was launched
    rc, acc, msg = _static(lib, torch, rng.standard_normal((3, 8)), rng.standard_normal((5, 8)), False, False)
    assert rc != 0 and "cannot be" in msg
    # host pointers are refused: this entry point stages nothing
    lat = rng.standard_normal((4, 8))
    d_acc = torch.zeros(1, dtype=torch.float64, device="cuda:0")
    err = C.create_string_buffer(256)
    rc = lib.mmd_pairwise_accuracy_static(0, None, lat.ctypes.data, 4, lat.ctypes.data, 4, 8, 0, 0, d_acc.data_ptr(), err, len(err))
    assert rc != 0 and b"device memory" in err.value
    # and a call after the refusals still works
    rc, acc, msg = _static(lib, torch, lat, lat, False, False)
    assert rc == 0 and acc == 1.0, msg
