"""za_gpu_checksum_noalloc (include/za_gpu.h): the per-row checksum / energy reduction with its accumulators in the library's
file-scope static device array -- an entry point that allocates no device memory.  It has no counterpart in the reference,
so the check is against the existing entry point za_gpu_checksum on the same
device rows: the wrapping 64-bit sums bit for bit, the energies to the tolerance tests/test_hostpath_gpu.py uses (rtol 1e-9;
the double atomics commute only up to rounding).  Two calls in a row, the second on other data and fewer rows, so that sums
left in the shared array by the first call would show.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu(built_lib):
    import za_gpu
    if za_gpu.lib().za_gpu_device_count() < 1:
        pytest.fail("GPU test run without a CUDA device")
    return za_gpu


def both(gpu, x):
    """(sum, energy) of the rows of x [rows][frames] from the allocating and from the no-allocation entry point."""
    L = gpu.lib()
    rows, nf = x.shape
    out = []
    for fn in (L.za_gpu_checksum, L.za_gpu_checksum_noalloc):
        s, e = np.full(rows, 0xDEADBEEF, np.uint64), np.full(rows, -1.0, np.float64)
        rc = fn(C.c_void_p(x.data_ptr()), rows, nf, x.stride(0), C.c_void_p(s.ctypes.data), C.c_void_p(e.ctypes.data), x.device.index or 0, None)
        assert rc == 0, L.za_gpu_last_error().decode()
        out.append((s, e))
    return out


def test_noalloc_checksum_matches_the_allocating_entry_point_twice(gpu):
    torch = pytest.importorskip("torch")
    dev = torch.device("cuda", 0)
    x1 = torch.empty((1000, 1, 20011), dtype=torch.float32, device=dev)
    gpu.synth_into(x1, 1, 77, 5, 0)
    x2 = torch.empty((37, 1, 513), dtype=torch.float32, device=dev)       # second call: fewer rows, other data, short rows
    gpu.synth_into(x2, 0, 78, 900, 12345)
    torch.cuda.synchronize()
    mem_before = torch.cuda.memory_allocated(dev)
    for x in (x1[:, 0], x2[:, 0], x1[:, 0]):
        (s_ref, e_ref), (s, e) = both(gpu, x)
        assert np.array_equal(s, s_ref)
        assert np.allclose(e, e_ref, rtol=1e-9, atol=0.0)
        assert e_ref.min() > 0.0 and len(np.unique(s_ref)) == len(s_ref)   # the comparison is not between two empty results
    assert torch.cuda.memory_allocated(dev) == mem_before                 # and torch's allocator was not involved either
    # a strided view (rows further apart than they are long) and zero frames
    (s_ref, e_ref), (s, e) = both(gpu, x1[:, 0, :4096])
    assert np.array_equal(s, s_ref) and np.allclose(e, e_ref, rtol=1e-9, atol=0.0)
    L = gpu.lib()
    s, e = np.ones(5, np.uint64), np.ones(5, np.float64)
    assert L.za_gpu_checksum_noalloc(C.c_void_p(x1.data_ptr()), 5, 0, 0, C.c_void_p(s.ctypes.data), C.c_void_p(e.ctypes.data), 0, None) == 0
    assert not s.any() and not e.any()


def test_noalloc_checksum_refuses_more_rows_than_its_static_scratch(gpu):
    torch = pytest.importorskip("torch")
    L = gpu.lib()
    assert gpu.CHECKSUM_NOALLOC_MAX_ROWS == 16384
    rows = gpu.CHECKSUM_NOALLOC_MAX_ROWS + 1
    x = torch.zeros((rows, 8), dtype=torch.float32, device="cuda")
    s, e = np.zeros(rows, np.uint64), np.zeros(rows, np.float64)
    rc = L.za_gpu_checksum_noalloc(C.c_void_p(x.data_ptr()), rows, 8, 8, C.c_void_p(s.ctypes.data), C.c_void_p(e.ctypes.data), 0, None)
    assert rc == -1                                                         # ZA_ERR_ARG (include/za_gpu.h)
    msg = L.za_gpu_last_error().decode()
    assert "16384" in msg and "exceed" in msg, msg
    # exactly the maximum is served, and equals the allocating entry point
    (s_ref, e_ref), (s2, e2) = both(gpu, torch.rand((gpu.CHECKSUM_NOALLOC_MAX_ROWS, 64), dtype=torch.float32, device="cuda"))
    assert np.array_equal(s2, s_ref) and np.allclose(e2, e_ref, rtol=1e-9, atol=0.0)
