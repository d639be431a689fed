"""b200_mul_mat_noalloc (no counterpart in the reference): the batch <= 8
MUL_MAT whose q8_1 activation scratch is a __device__ array inside the library, so the call allocates nothing.

Called through the C ABI (ctypes) and checked against the existing entry point b200_mul_mat on the same inputs -- the two
run the same kernels, so the results must be bit-identical (and both meet the oracle bound) -- on two calls in a row with
different shapes, so that scratch contents left by the first call cannot leak into the second.
This file is named to be collected LAST.  It passes alone in a fresh process and after
the other tests."""
import ctypes as C

import numpy as np
import pytest

import qref

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

NMSE_TIGHT = 1e-9


@pytest.fixture(scope="module")
def qmm():
    import b200pkg
    q = b200pkg.load()
    assert torch.cuda.is_available(), "these tests need the B200"
    assert q.lib().b200_device_ok(torch.cuda.current_device()) == 1
    return q


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def noalloc(qmm, t, wd, xd, out, m, n, k):
    """the raw C-ABI call: pointers and sizes only, outputs allocated by the caller beforehand"""
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    return qmm.lib().b200_mul_mat_noalloc(t, wd.data_ptr(), 0, xd.data_ptr(), k, out.data_ptr(), m, m, n, k, s)


@pytest.mark.parametrize("name,t", list(qref.WEIGHT_TYPES.items()))
def test_noalloc_matches_mul_mat_on_two_calls_in_a_row(qmm, name, t):
    rng = np.random.default_rng(t)
    assert qmm.lib().b200_noalloc_scratch_size() == 384 * 1024
    # (m, k, n): a long row with many columns first, then a short one -- the second call reuses the front of the scratch
    cases = [(520, 4096, 7), (264, 1024, 1), (1000, 14336, 3)]
    prepared = []
    for m, k, n in cases:
        w = qref.random_blocks(t, m, k, rng)
        x = rng.uniform(-1, 1, (n, k)).astype(np.float32)
        prepared.append((m, k, n, w, x, dev(w), dev(x), torch.full((n, m), 7.0, device="cuda"),
                         torch.full((n, m), -7.0, device="cuda")))
    torch.cuda.synchronize()
    mem_before = torch.cuda.memory_allocated()
    outs = []
    for m, k, n, w, x, wd, xd, out, _ in prepared:          # back to back on one stream, nothing in between
        assert noalloc(qmm, t, wd, xd, out, m, n, k) == 0, qmm.lib().b200_last_cuda_error()
        outs.append(out)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == mem_before      # (torch's allocator was not touched either)
    for (m, k, n, w, x, wd, xd, out, ref), got in zip(prepared, outs):
        want = qmm.mul_mat(t, wd, xd, m, k, out=ref)        # the existing entry point, caller-owned workspace
        torch.cuda.synchronize()
        assert torch.equal(got, want)
        assert qref.nmse(qref.mul_mat(t, w, x, m, k, "q8_1"), got.cpu().numpy()) < NMSE_TIGHT
    # the same three calls again: same bits (the scratch now holds the last call's contents, not zeros)
    for (m, k, n, w, x, wd, xd, out, ref) in prepared:
        out.fill_(3.0)
        assert noalloc(qmm, t, wd, xd, out, m, n, k) == 0
        torch.cuda.synchronize()
        assert torch.equal(out, ref)


def test_noalloc_rejects_what_does_not_fit(qmm):
    L = qmm.lib()
    t, m = qref.Q4_K, 16
    for n, k in [(9, 4096), (1, 65536), (64, 4096)]:       # beyond the GEMV path / a row no GEMV CTA spans / n too large
        wd = torch.zeros(m * qref.row_size(t, k), dtype=torch.uint8, device="cuda")
        xd = torch.zeros((n, k), device="cuda")
        out = torch.full((n, m), 5.0, device="cuda")
        st = noalloc(qmm, t, wd, xd, out, m, n, k)
        assert st == 8, (n, k, st)                          # B200_ERR_TOO_LARGE
        assert b"built-in" in L.b200_last_cuda_error()
        assert b"exceeds" in L.b200_status_string(st)
        torch.cuda.synchronize()
        assert bool((out == 5.0).all())                     # nothing was launched
    # argument errors keep b200_mul_mat's codes
    xd = torch.zeros((1, 256), device="cuda")
    out = torch.zeros((1, m), device="cuda")
    wd = torch.zeros(m * 144, dtype=torch.uint8, device="cuda")
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert L.b200_mul_mat_noalloc(0, wd.data_ptr(), 0, xd.data_ptr(), 256, out.data_ptr(), m, m, 1, 256, s) == 1   # type
    assert L.b200_mul_mat_noalloc(t, wd.data_ptr(), 0, xd.data_ptr(), 100, out.data_ptr(), m, m, 1, 100, s) == 2   # shape
    assert L.b200_mul_mat_noalloc(t, wd.data_ptr(), 0, xd.data_ptr(), 256, out.data_ptr(), m, 0, 1, 256, s) == 0   # m = 0
