"""ctypes binding of the UNMODIFIED reference build under oracle/_ref (made by oracle/Makefile.ref from /root/reference).
TEST INFRASTRUCTURE ONLY: used to pin the oracle and to generate tests/golden/*.npz.  Optional at test time."""
import ctypes as C
import glob
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFDIR = os.path.join(ROOT, "oracle", "_ref")

F32, Q4_0, Q8_0, Q8_1, Q4_K, Q6_K, Q8_K, IQ4_XS = 0, 2, 8, 9, 12, 14, 15, 23
NAMES = {Q4_0: "q4_0", Q8_0: "q8_0", Q4_K: "q4_K", Q6_K: "q6_K", IQ4_XS: "iq4_xs"}


def available():
    return os.path.exists(os.path.join(REFDIR, "libggml-base.so")) and bool(
        glob.glob(os.path.join(REFDIR, "libggml-cpu-*.so")))


_base = None
_cpu = None


def base():
    global _base
    if _base is None:
        _base = C.CDLL(os.path.join(REFDIR, "libggml-base.so"), mode=C.RTLD_GLOBAL)

        class InitParams(C.Structure):  # ggml/include/ggml.h struct ggml_init_params
            _fields_ = [("mem_size", C.c_size_t), ("mem_buffer", C.c_void_p), ("no_alloc", C.c_bool)]

        # the fp16->fp32 lookup table ggml-base dequantizes through is filled by the first ggml_init (ggml/src/ggml.c)
        _base.ggml_init.restype = C.c_void_p
        _base.ggml_init.argtypes = [InitParams]
        _base.ggml_free.argtypes = [C.c_void_p]
        _base.ggml_free(_base.ggml_init(InitParams(1 << 16, None, False)))
        _base.ggml_row_size.restype = C.c_size_t
        _base.ggml_row_size.argtypes = [C.c_int, C.c_int64]
        _base.ggml_quantize_chunk.restype = C.c_size_t
        _base.ggml_quantize_chunk.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64,
                                              C.c_void_p]
        for n in ("quantize_row_q8_1_ref", "quantize_row_q8_0_ref", "quantize_row_q8_K_ref", "dequantize_row_q4_0",
                  "dequantize_row_q8_0", "dequantize_row_q4_K", "dequantize_row_q6_K", "dequantize_row_iq4_xs"):
            f = getattr(_base, n)
            f.restype = None
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
    return _base


def cpu():
    """the portable (haswell) CPU variant: the real ggml-cpu SIMD kernels"""
    global _cpu
    if _cpu is None:
        base()
        _cpu = C.CDLL(os.path.join(REFDIR, "libggml-cpu-haswell.so"))
        for n in ("quantize_row_q8_0", "quantize_row_q8_K"):
            f = getattr(_cpu, n)
            f.restype = None
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_int64]
        for n in ("q4_0_q8_0", "q8_0_q8_0", "q4_K_q8_K", "q6_K_q8_K", "iq4_xs_q8_K"):
            f = getattr(_cpu, "ggml_vec_dot_" + n)
            f.restype = None
            f.argtypes = [C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_int]
    return _cpu


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def row_size(t, k):
    return base().ggml_row_size(t, k)


def quantize_weights(t, x):
    """ggml_quantize_chunk (ggml/src/ggml.c:6386) without imatrix; x [nrows, k] f32 -> uint8 [nrows, row_size]"""
    x = np.ascontiguousarray(x, dtype=np.float32)
    nrows, k = x.shape
    out = np.zeros((nrows, row_size(t, k)), dtype=np.uint8)
    n = base().ggml_quantize_chunk(t, _p(x), _p(out), 0, nrows, k, None)
    assert n == out.size
    return out


def quantize_act(t, x):
    x = np.ascontiguousarray(x, dtype=np.float32)
    k = x.shape[-1]
    rows = x.reshape(-1, k)
    out = np.zeros((rows.shape[0], row_size(t, k)), dtype=np.uint8)
    fn = {Q8_1: base().quantize_row_q8_1_ref, Q8_0: base().quantize_row_q8_0_ref, Q8_K: base().quantize_row_q8_K_ref}[t]
    for r in range(rows.shape[0]):
        fn(_p(rows[r]), _p(out[r]), k)
    return out


def dequantize(t, w, k):
    w = np.ascontiguousarray(w, dtype=np.uint8)
    rs = row_size(t, k)
    nrows = w.size // rs
    w = w.reshape(nrows, rs)
    y = np.empty((nrows, k), dtype=np.float32)
    fn = getattr(base(), "dequantize_row_" + NAMES[t])
    for r in range(nrows):
        fn(_p(w[r]), _p(y[r]), k)
    return y


def mul_mat_cpu(t, w, x, m, k):
    """what ggml-cpu computes for MUL_MAT (ggml-cpu.c:1266): from_float(vec_dot_type) then the SIMD vec_dot"""
    lib = cpu()
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, k)
    n = x.shape[0]
    w = np.ascontiguousarray(w, dtype=np.uint8).reshape(m, -1)
    if t in (Q4_0, Q8_0):
        qt, qfn = Q8_0, lib.quantize_row_q8_0
    else:
        qt, qfn = Q8_K, lib.quantize_row_q8_K
    xq = np.zeros((n, row_size(qt, k)), dtype=np.uint8)
    for j in range(n):
        qfn(_p(x[j]), _p(xq[j]), k)
    dot = getattr(lib, "ggml_vec_dot_%s_%s" % (NAMES[t], "q8_0" if qt == Q8_0 else "q8_K"))
    dst = np.zeros((n, m), dtype=np.float32)
    s = C.c_float()
    for j in range(n):
        for i in range(m):
            dot(k, C.byref(s), 0, _p(w[i]), 0, _p(xq[j]), 0, 1)
            dst[j, i] = s.value
    return dst
