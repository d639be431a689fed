"""ctypes binding of the oracle (oracle/libggml_qref.so) -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "libggml_qref.so")

F32, Q4_0, Q8_0, Q8_1, Q4_K, Q6_K, Q8_K, IQ4_XS = 0, 2, 8, 9, 12, 14, 15, 23
WEIGHT_TYPES = {"q4_0": Q4_0, "q8_0": Q8_0, "q4_K": Q4_K, "q6_K": Q6_K, "iq4_xs": IQ4_XS}
BLCK = {Q4_0: 32, Q8_0: 32, Q8_1: 32, Q4_K: 256, Q6_K: 256, Q8_K: 256, IQ4_XS: 256}
TSIZE = {Q4_0: 18, Q8_0: 34, Q8_1: 36, Q4_K: 144, Q6_K: 210, Q8_K: 292, IQ4_XS: 136}


def build():
    src = os.path.join(ROOT, "oracle", "ggml_qref.c")
    if (not os.path.exists(LIB)) or os.path.getmtime(LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-f", "oracle/Makefile"], cwd=ROOT, stdout=subprocess.DEVNULL)
    return LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        vp, i64, sz, i32 = C.c_void_p, C.c_int64, C.c_size_t, C.c_int
        _lib.qref_row_size.restype = sz
        _lib.qref_row_size.argtypes = [i32, i64]
        _lib.qref_fp32_to_fp16.restype = C.c_uint16
        _lib.qref_fp32_to_fp16.argtypes = [C.c_float]
        _lib.qref_fp16_to_fp32.restype = C.c_float
        _lib.qref_fp16_to_fp32.argtypes = [C.c_uint16]
        for name in ("q8_1", "q8_0", "q8_K", "q4_0"):
            f = getattr(_lib, "qref_quantize_row_" + name)
            f.restype = None
            f.argtypes = [vp, vp, i64]
        _lib.qref_dequantize_rows.restype = None
        _lib.qref_dequantize_rows.argtypes = [i32, vp, vp, i64, i64]
        for name in ("qref_mul_mat", "qref_mul_mat_q8_1", "qref_mul_mat_f64"):
            f = getattr(_lib, name)
            f.restype = None
            f.argtypes = [i32, vp, sz, vp, vp, i64, i64, i64, i32]
        _lib.qref_mul_mat_id.restype = None
        _lib.qref_mul_mat_id.argtypes = [i32, vp, sz, vp, vp, i64, vp, i64, i64, i64, i64, i64, i64, i32, i32]
        _lib.qref_nmse.restype = C.c_double
        _lib.qref_nmse.argtypes = [vp, vp, sz]
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def row_size(t, k):
    return k // BLCK[t] * TSIZE[t]


def quantize(t, x):
    """x: [..., k] f32 -> uint8 [..., row_size]; t in {Q8_1, Q8_0, Q8_K, Q4_0}"""
    x = np.ascontiguousarray(x, dtype=np.float32)
    k = x.shape[-1]
    rows = x.reshape(-1, k)
    out = np.zeros((rows.shape[0], row_size(t, k)), dtype=np.uint8)
    fn = {Q8_1: lib().qref_quantize_row_q8_1, Q8_0: lib().qref_quantize_row_q8_0, Q8_K: lib().qref_quantize_row_q8_K,
          Q4_0: lib().qref_quantize_row_q4_0}[t]
    for r in range(rows.shape[0]):
        fn(_p(rows[r]), _p(out[r]), k)
    return out.reshape(x.shape[:-1] + (out.shape[1],))


def dequantize(t, w, k):
    """w: uint8 [nrows, row_size] -> f32 [nrows, k]"""
    w = np.ascontiguousarray(w, dtype=np.uint8)
    nrows = w.size // row_size(t, k)
    y = np.empty((nrows, k), dtype=np.float32)
    lib().qref_dequantize_rows(t, _p(w), _p(y), nrows, k)
    return y


def mul_mat(t, w, x, m, k, semantics="cpu", nthreads=0):
    """w: uint8 m rows; x: f32 [n, k] -> f32 [n, m]"""
    w = np.ascontiguousarray(w, dtype=np.uint8)
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1, k)
    n = x.shape[0]
    dst = np.zeros((n, m), dtype=np.float32)
    fn = {"cpu": lib().qref_mul_mat, "q8_1": lib().qref_mul_mat_q8_1, "f64": lib().qref_mul_mat_f64}[semantics]
    fn(t, _p(w), 0, _p(x), _p(dst), m, n, k, nthreads)
    return dst


def mul_mat_id(t, w, b, ids, m, k, n_expert, semantics="cpu", nthreads=0):
    """w: uint8 [n_expert * m rows]; b: f32 [n_tokens, ne11, k]; ids: i32 [n_tokens, n_used] -> [n_tokens, n_used, m]"""
    w = np.ascontiguousarray(w, dtype=np.uint8)
    b = np.ascontiguousarray(b, dtype=np.float32)
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    n_tokens, ne11, _ = b.shape
    n_used = ids.shape[1]
    dst = np.zeros((n_tokens, n_used, m), dtype=np.float32)
    sem = {"cpu": 0, "q8_1": 1, "f64": 2}[semantics]
    lib().qref_mul_mat_id(t, _p(w), 0, _p(b), _p(ids), n_used, _p(dst), m, k, n_expert, n_used, ne11, n_tokens, sem,
                          nthreads)
    return dst


def nmse(a, b):
    a = np.ascontiguousarray(a, dtype=np.float32).ravel()
    b = np.ascontiguousarray(b, dtype=np.float32).ravel()
    assert a.size == b.size
    return float(lib().qref_nmse(_p(a), _p(b), a.size))


def random_blocks(t, nrows, k, rng, scale=0.02):
    """Random VALID weight blocks written directly in the quantized format (every bit pattern of the packed fields is
    exercised; fp16 scales are finite and of sane magnitude).  Returns uint8 [nrows, row_size]."""
    nb = nrows * (k // BLCK[t])
    raw = rng.integers(0, 256, size=(nb, TSIZE[t]), dtype=np.uint8)

    def put_f16(col, vals):
        h = np.asarray(vals, dtype=np.float16).view(np.uint16)
        raw[:, col] = (h & 0xFF).astype(np.uint8)
        raw[:, col + 1] = (h >> 8).astype(np.uint8)

    sgn = rng.choice([-1.0, 1.0], size=nb)
    if t == Q4_0:
        put_f16(0, sgn * rng.uniform(0.1, 1.0, nb) * scale / 4)
    elif t == Q8_0:
        put_f16(0, sgn * rng.uniform(0.1, 1.0, nb) * scale / 64)
    elif t == Q4_K:
        put_f16(0, rng.uniform(0.1, 1.0, nb) * scale / 200)
        put_f16(2, rng.uniform(0.1, 1.0, nb) * scale / 60)
    elif t == Q6_K:
        put_f16(208, sgn * rng.uniform(0.1, 1.0, nb) * scale / 2000)
    elif t == IQ4_XS:
        put_f16(0, sgn * rng.uniform(0.1, 1.0, nb) * scale / 2000)
    else:
        raise ValueError(t)
    return raw.reshape(nrows, -1)


PLANE_KC = 256   # scale granularity of the tensor-core path; GPU tests set it from b200_mmq_chunk()


def mul_mat_plane(t, w, x, m, k, kc=None):
    """numpy restatement of the tensor-core path's arithmetic (llama.cpp.dsp_b200/csrc/mmq.cu): weights dequantized
    bit-exactly, re-expressed as int8 with one f32 scale per (row, chunk) -- Q8_0: value + residual plane; activations
    int8 with one scale per (column, chunk) -- against Q4_0 / Q8_0: value + residual plane; exact integer chunk dot
    products; f32 scale-accumulate over the segments W0.X0 | W0.X1 | W1.X0 in chunk order.  -> f32 [n, m]"""
    f32 = np.float32
    kc = kc or PLANE_KC
    x = np.ascontiguousarray(x, dtype=f32).reshape(-1, k)
    n = x.shape[0]
    kp = (k + kc - 1) // kc * kc
    Wf = np.zeros((m, kp), dtype=f32)
    Wf[:, :k] = dequantize(t, w, k)
    Xf = np.zeros((n, kp), dtype=f32)
    Xf[:, :k] = x

    def q8(a):
        c = a.reshape(a.shape[0], kp // kc, kc)
        s = (np.abs(c).max(-1) / f32(127.0)).astype(f32)
        inv = np.where(s > 0, f32(1.0) / np.where(s > 0, s, f32(1.0)), f32(0.0)).astype(f32)
        q = np.rint((c * inv[:, :, None]).astype(f32)).astype(np.int32)
        return q, s

    def planes(a, count):
        out = []
        for _ in range(count):
            q, s = q8(a)
            out.append((q, s))
            a = (a - (s[:, :, None] * q.astype(f32)).astype(f32).reshape(a.shape)).astype(f32)   # fsub(fmul), per op
        return out

    WP = planes(Wf, 2 if t == Q8_0 else 1)
    XP = planes(Xf, 2 if t in (Q4_0, Q8_0) else 1)
    acc = np.zeros((n, m), dtype=f32)
    for wi, xi in [(0, 0), (0, 1), (1, 0)]:
        if wi >= len(WP) or xi >= len(XP):
            continue
        (P, S), (Q, T) = WP[wi], XP[xi]
        for c in range(kp // kc):
            D = (Q[:, c, :].astype(np.int64) @ P[:, c, :].astype(np.int64).T).astype(f32)        # [n, m], exact
            st = (S[None, :, c] * T[:, None, c]).astype(f32)
            acc = (D.astype(np.float64) * st.astype(np.float64) + acc.astype(np.float64)).astype(f32)   # one fma rounding
    return acc
