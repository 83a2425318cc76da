"""magic_cosine_distance_noalloc: the allocation-free variant of cosine_distance (no counterpart in the reference; added
for callers that cannot call cudaMalloc).  Its scratch is a static device array
inside libmagic_b200.so, shared by every call and keeping the last call's contents -- so the check is two calls in a
row, on different data and different shapes, each against the existing magic_cosine_distance entry point.

Collected last on purpose (file name): a failure here must not hide other tests.
The test allocates its own inputs / outputs with torch; the entry point under test allocates nothing."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _reference(lib, check, a: np.ndarray, b: np.ndarray) -> np.ndarray:
    out = np.empty(a.shape[0], dtype=np.float32)
    check(lib.magic_cosine_distance(a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), a.shape[0], a.shape[1],
                                    out.ctypes.data_as(C.c_void_p)))
    return out


def test_noalloc_cosine_distance_matches_the_allocating_entry_point_twice_in_a_row():
    from magic_b200 import _lib
    lib = _lib.load()
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    stream = torch.cuda.current_stream(dev)
    rng = np.random.default_rng(0)
    # second call: fewer columns than the first, so stale sums of the first call sit right behind the part it clears;
    # third: the maximum width
    for B, L in ((300, 200), (64, 20), (17, 4096)):
        a = rng.normal(size=(B, L)).astype(np.float32)
        b = rng.normal(size=(B, L)).astype(np.float32)
        want = _reference(lib, _lib.check, a, b)
        da, db = torch.from_numpy(a).to(dev), torch.from_numpy(b).to(dev)
        out = torch.full((B,), float("nan"), dtype=torch.float32, device=dev)
        free_before = torch.cuda.mem_get_info(dev)[0]
        _lib.check(lib.magic_cosine_distance_noalloc(C.c_void_p(da.data_ptr()), C.c_void_p(db.data_ptr()), B, L,
                                                     C.c_void_p(out.data_ptr()), C.c_void_p(stream.cuda_stream)))
        stream.synchronize()
        assert torch.cuda.mem_get_info(dev)[0] >= free_before - (1 << 20)       # no device memory taken by the call
        got = out.cpu().numpy()
        # the tolerance of tests/test_tango_api.py::test_cosine_distance_is_a_real_call; the two entry points run the same
        # kernels, they differ only in the order of the atomic column sums
        np.testing.assert_allclose(got, want, rtol=2e-5, atol=1e-6)
        # against the definition as well (8c/vae.py:438-452: l2-normalise over the batch axis, sum over the latent axis)
        na = a / np.sqrt(np.maximum((a.astype(np.float64) ** 2).sum(0), 1e-12))
        nb = b / np.sqrt(np.maximum((b.astype(np.float64) ** 2).sum(0), 1e-12))
        np.testing.assert_allclose(got, (na * nb).sum(1), rtol=2e-5, atol=1e-6)


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    from magic_b200 import _lib
    lib = _lib.load()
    torch.cuda.set_device(0)
    x = torch.zeros(4, 4097, dtype=torch.float32, device="cuda:0")
    out = torch.zeros(4, dtype=torch.float32, device="cuda:0")
    with pytest.raises(_lib.MagicError, match="exceeds the static scratch"):
        _lib.check(lib.magic_cosine_distance_noalloc(C.c_void_p(x.data_ptr()), C.c_void_p(x.data_ptr()), 4, 4097,
                                                     C.c_void_p(out.data_ptr()), None))
    with pytest.raises(_lib.MagicError, match="null argument"):
        _lib.check(lib.magic_cosine_distance_noalloc(None, C.c_void_p(x.data_ptr()), 4, 16, C.c_void_p(out.data_ptr()), None))
