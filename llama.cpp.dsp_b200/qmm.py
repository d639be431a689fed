"""ctypes binding of libb200qmm.so + thin torch-pointer wrappers named after the reference interface they replace.

Reference interface (paths relative to the reference tree):
  quantize_row_q8_1_cuda      ggml/src/ggml-cuda/quantize.cu:129-141      -> quantize_row_q8_1
  ggml_get_to_fp32_cuda(type) ggml/src/ggml-cuda/convert.cu:654           -> dequantize_row
  ggml_cuda_mul_mat           ggml/src/ggml-cuda/ggml-cuda.cu:1877-1938   -> mul_mat
  ggml_cuda_mul_mat_id        ggml/src/ggml-cuda/ggml-cuda.cu:1994-2130   -> mul_mat_id
Tensor conventions are ggml's: weights [m rows][k], activations [n][k] f32, dst [n][m] f32.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200_QMM_LIB") or os.path.join(HERE, "lib", "libb200qmm.so")   # override: A/B builds

GGML_TYPE_Q4_0, GGML_TYPE_Q8_0, GGML_TYPE_Q8_1, GGML_TYPE_Q4_K, GGML_TYPE_Q6_K, GGML_TYPE_IQ4_XS = 2, 8, 9, 12, 14, 23
TYPE_BY_NAME = {"q4_0": 2, "q8_0": 8, "q4_K": 12, "q6_K": 14, "iq4_xs": 23}
BLOCK = {2: 32, 8: 32, 9: 32, 12: 256, 14: 256, 23: 256}
TYPE_SIZE = {2: 18, 8: 34, 9: 36, 12: 144, 14: 210, 23: 136}

EXPORTS = [
    "b200_qmm_abi_version", "b200_status_string", "b200_last_cuda_error", "b200_device_ok", "b200_sm_count",
    "b200_launch_count", "b200_type_supported", "b200_block_size", "b200_type_size", "b200_row_size",
    "b200_weight_alloc_size", "b200_quantize_q8_1", "b200_dequantize_f32", "b200_mul_mat_workspace_size",
    "b200_mul_mat", "b200_mul_mat_multi_workspace_size", "b200_mul_mat_multi", "b200_mmq_plane_size", "b200_mmq_build_plane", "b200_mul_mat_mmq_workspace_size",
    "b200_mul_mat_mmq", "b200_mul_mat_mmq_multi", "b200_mmq_min_batch", "b200_mmq_use_cached_plane", "b200_mmq_chunk", "b200_mul_mat_host_workspace_size", "b200_mul_mat_host",
    "b200_mul_mat_multi_host_workspace_size", "b200_mul_mat_multi_host", "b200_mul_mat_id_workspace_size", "b200_mul_mat_id", "b200_row_split",
    "b200_mul_mat_rowsplit_workspace_size", "b200_mul_mat_rowsplit_fused", "b200_rowsplit_wait", "b200_rowsplit_timeout_count", "b200_copy_async", "b200_copy_mapped_async", "b200_mul_mat_mmq_multi_rowsplit", "b200_rowsplit_push", "b200_rowsplit_push_join",
    "b200_ipc_get_handle", "b200_ipc_open_handle", "b200_ipc_close_handle", "b200_reload_env", "b200_debug_gemv_trace",
    "b200_noalloc_scratch_size", "b200_mul_mat_noalloc",
    "b200_sequence_plan_size", "b200_sequence_plan_build", "b200_sequence_run", "b200_sequence_status",
    "b200_debug_sequence_trace", "b200_sequence_plan_build_rowsplit",
    "b200_mul_mat_id_prefers_mmq", "b200_mul_mat_id_mmq_workspace_size", "b200_mul_mat_id_mmq",
]


class SeqGroup(C.Structure):   # b200_seq_group
    _fields_ = [("x", C.c_void_p), ("k", C.c_int64), ("nmat", C.c_int), ("dependent", C.c_int), ("type", C.c_int * 4),
                ("W", C.c_void_p * 4), ("dst", C.c_void_p * 4), ("m", C.c_int64 * 4)]


class B200Error(RuntimeError):
    pass


_lib = None


def lib():
    """Load the C-ABI library.  Fails loudly when it has not been built: there is no fallback path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200Error("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(make -f llama.cpp.dsp_b200/Makefile). There is no CPU fallback." % LIB_PATH)
    L = C.CDLL(LIB_PATH)
    vp, i64, sz, i32 = C.c_void_p, C.c_int64, C.c_size_t, C.c_int
    L.b200_qmm_abi_version.restype = i32
    L.b200_status_string.restype = C.c_char_p
    L.b200_status_string.argtypes = [i32]
    L.b200_last_cuda_error.restype = C.c_char_p
    L.b200_device_ok.argtypes = [i32]
    L.b200_sm_count.argtypes = [i32]
    L.b200_launch_count.restype = i64
    L.b200_type_supported.argtypes = [i32]
    L.b200_block_size.argtypes = [i32]
    L.b200_type_size.restype = sz
    L.b200_type_size.argtypes = [i32]
    L.b200_row_size.restype = sz
    L.b200_row_size.argtypes = [i32, i64]
    L.b200_weight_alloc_size.restype = sz
    L.b200_weight_alloc_size.argtypes = [i32, i64, i64]
    L.b200_quantize_q8_1.argtypes = [vp, i64, vp, i64, i64, vp]
    L.b200_dequantize_f32.argtypes = [i32, vp, i64, vp, i64, i64, vp]
    L.b200_mul_mat_workspace_size.restype = sz
    L.b200_mul_mat_workspace_size.argtypes = [i32, i64, i64, i64]
    L.b200_mul_mat.argtypes = [i32, vp, i64, vp, i64, vp, i64, i64, i64, i64, vp, sz, vp]
    L.b200_mmq_plane_size.restype = sz
    L.b200_mmq_plane_size.argtypes = [i32, i64, i64]
    L.b200_mmq_build_plane.argtypes = [i32, vp, i64, vp, sz, i64, i64, vp]
    L.b200_mul_mat_mmq_workspace_size.restype = sz
    L.b200_mul_mat_mmq_workspace_size.argtypes = [i64, i64, i64]
    L.b200_mul_mat_mmq.argtypes = [i32, vp, vp, i64, vp, i64, i64, i64, i64, vp, sz, vp]
    L.b200_mul_mat_host_workspace_size.restype = sz
    L.b200_mul_mat_host_workspace_size.argtypes = [i32, i64, i64, i64]
    L.b200_mul_mat_host.argtypes = [i32, vp, i64, vp, vp, i64, i64, i64, vp, sz, vp]
    L.b200_mul_mat_id_workspace_size.restype = sz
    L.b200_mul_mat_id_workspace_size.argtypes = [i32, i64, i64, i64, i64, i64]
    L.b200_mul_mat_id.argtypes = [i32, vp, i64, vp, i64, i64, i64, vp, i64, vp, i64, i64, i64, i64, i64, i64, i64, vp,
                                  sz, vp]
    L.b200_mul_mat_multi_workspace_size.restype = sz
    L.b200_mul_mat_multi_workspace_size.argtypes = [i32, vp, vp, i64, i64]
    L.b200_mul_mat_multi.argtypes = [i32, vp, vp, vp, i64, vp, vp, vp, i64, i64, vp, sz, vp]
    L.b200_mul_mat_mmq_multi.argtypes = [i32, vp, vp, vp, i64, vp, vp, vp, i64, i64, vp, sz, vp]
    L.b200_mul_mat_multi_host_workspace_size.restype = sz
    L.b200_mul_mat_multi_host_workspace_size.argtypes = [i32, vp, vp, i64, i64]
    L.b200_mul_mat_multi_host.argtypes = [i32, vp, vp, vp, vp, vp, i64, i64, vp, sz, vp]
    L.b200_row_split.argtypes = [i64, i32, i32, vp, C.POINTER(i64), C.POINTER(i64)]
    L.b200_mul_mat_rowsplit_workspace_size.restype = sz
    L.b200_mul_mat_rowsplit_workspace_size.argtypes = [i32, i64, i64]
    L.b200_mul_mat_rowsplit_fused.argtypes = [i32, vp, vp, i64, vp, vp, i64, i64, i64, i64, i64, i64, i32, i32,
                                              vp, i32, vp, sz, vp]
    L.b200_rowsplit_wait.argtypes = [vp, i32, vp]
    L.b200_rowsplit_timeout_count.argtypes = [C.POINTER(C.c_int), vp]
    L.b200_copy_async.argtypes = [vp, vp, sz, vp]
    L.b200_mmq_use_cached_plane.argtypes = [i32, i64]
    L.b200_rowsplit_push.argtypes = [i32, vp, vp, vp, i64, i32, i32, vp, vp, vp]
    L.b200_rowsplit_push_join.argtypes = [vp]
    L.b200_mul_mat_mmq_multi_rowsplit.argtypes = [i32, vp, vp, vp, i64, vp, vp, vp, i64, i64, i32, i32, vp, vp, vp, sz, vp]
    L.b200_copy_mapped_async.argtypes = [vp, vp, sz, vp]
    L.b200_ipc_get_handle.argtypes = [vp, vp, C.POINTER(i64)]
    L.b200_ipc_open_handle.argtypes = [vp, i64, C.POINTER(vp)]
    L.b200_ipc_close_handle.argtypes = [vp, i64]
    L.b200_noalloc_scratch_size.restype = sz
    L.b200_mul_mat_noalloc.argtypes = [i32, vp, i64, vp, i64, vp, i64, i64, i64, i64, vp]
    L.b200_sequence_plan_size.restype = sz
    L.b200_sequence_plan_size.argtypes = [i32]
    L.b200_sequence_plan_build.argtypes = [vp, i32, vp, sz, vp]
    L.b200_sequence_run.argtypes = [vp, vp]
    L.b200_sequence_plan_build_rowsplit.argtypes = [vp, i32, vp, sz, i32, i32, vp, sz, vp, vp, vp]
    L.b200_sequence_status.argtypes = [vp, C.POINTER(i32), vp]
    L.b200_debug_sequence_trace.argtypes = [vp, vp, vp]
    L.b200_mul_mat_id_prefers_mmq.argtypes = [i64, i64, i64, i64]
    L.b200_mul_mat_id_mmq_workspace_size.restype = sz
    L.b200_mul_mat_id_mmq_workspace_size.argtypes = [i32, i64, i64, i64, i64, i64]
    L.b200_mul_mat_id_mmq.argtypes = [i32, vp, vp, i64, i64, i64, vp, i64, vp, i64, i64, i64, i64, i64, i64, i64, vp, sz, vp]
    L.b200_reload_env.restype = None
    L.b200_debug_gemv_trace.restype = None
    L.b200_debug_gemv_trace.argtypes = [vp, i64]
    _lib = L
    return L


def check(status):
    if status != 0:
        L = lib()
        msg = L.b200_status_string(status).decode()
        if status in (5, 6, 8):
            msg += ": " + L.b200_last_cuda_error().decode()
        raise B200Error("b200_qmm: %s (status %d)" % (msg, status))


def row_size(t, k):
    return k // BLOCK[t] * TYPE_SIZE[t]


def reload_env():
    """B200_* knobs are cached by the library at first use; call this after changing os.environ"""
    lib().b200_reload_env()


def setenv(**knobs):
    """set (value) or remove (None) B200_* knobs and have the library read them again"""
    for k, v in knobs.items():
        if v is None:
            os.environ.pop(k, None)
        else:
            os.environ[k] = str(v)
    reload_env()


def launch_count():
    return int(lib().b200_launch_count())


def row_split(m, world, rank, tensor_split=None):
    lo, hi = C.c_int64(), C.c_int64()
    ts = None
    if tensor_split is not None:
        ts = (C.c_float * world)(*tensor_split)
    check(lib().b200_row_split(m, world, rank, ts, C.byref(lo), C.byref(hi)))
    return lo.value, hi.value


# ---- torch-pointer wrappers (torch is device-memory plumbing only) -------------------------------------------------

def _stream_ptr(stream=None):
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)


def _require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise B200Error("b200_qmm operates on CUDA tensors only (no CPU fallback)")


def quantize_row_q8_1(x, stream=None):
    """x: [nrows, k] f32 cuda (last dim contiguous) -> uint8 [nrows, k/32*36] ggml block_q8_1 bytes"""
    import torch
    _require_cuda(x)
    assert x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1
    nrows, k = x.shape
    y = torch.empty((nrows, k // 32 * 36), dtype=torch.uint8, device=x.device)
    check(lib().b200_quantize_q8_1(x.data_ptr(), x.stride(0) if nrows > 1 else k, y.data_ptr(), k, nrows,
                                   _stream_ptr(stream)))
    return y


def dequantize_row(t, w, k, stream=None):
    """w: uint8 [nrows, row_size] cuda -> f32 [nrows, k]"""
    import torch
    _require_cuda(w)
    assert w.dtype == torch.uint8 and w.dim() == 2 and w.stride(1) == 1
    nrows = w.shape[0]
    y = torch.empty((nrows, k), dtype=torch.float32, device=w.device)
    check(lib().b200_dequantize_f32(t, w.data_ptr(), w.stride(0) if nrows > 1 else 0, y.data_ptr(), k, nrows,
                                    _stream_ptr(stream)))
    return y


class Workspace:
    """grow-only device scratch owned by the caller (the ggml plug-in keeps one per backend stream)"""

    def __init__(self, device):
        self.device = device
        self.buf = None

    def get(self, nbytes):
        import torch
        if nbytes == 0:
            return None, 0
        if self.buf is None or self.buf.numel() < nbytes:
            self.buf = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=self.device)
        return self.buf.data_ptr(), self.buf.numel()


_ws = {}


def _workspace(device, nbytes):
    key = str(device)
    if key not in _ws:
        _ws[key] = Workspace(device)
    return _ws[key].get(nbytes)


def mul_mat(t, w, x, m, k, out=None, stream=None):
    """w: uint8 cuda holding m rows (packed); x: [n, k] f32 cuda; returns dst [n, m] f32"""
    import torch
    _require_cuda(w, x)
    assert x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1 and x.shape[1] == k
    n = x.shape[0]
    dst = out if out is not None else torch.empty((n, m), dtype=torch.float32, device=x.device)
    ws, wsz = _workspace(x.device, lib().b200_mul_mat_workspace_size(t, m, n, k))
    check(lib().b200_mul_mat(t, w.data_ptr(), 0, x.data_ptr(), x.stride(0) if n > 1 else k, dst.data_ptr(),
                             dst.stride(0) if n > 1 else m, m, n, k, ws, wsz, _stream_ptr(stream)))
    return dst


class MultiMatMul:
    """Prepared argument block for b200_mul_mat_multi: several weight matrices multiplied with the same activations.
    Built once per group (pointer arrays are ctypes arrays), called every token."""

    def __init__(self, types, ws, ms, k, outs, n=1):
        import torch
        nmat = len(types)
        _require_cuda(*ws, *outs)
        self.nmat, self.k, self.n = nmat, k, n
        self.keep = (list(ws), list(outs))
        self.types = (C.c_int * nmat)(*types)
        self.W = (C.c_void_p * nmat)(*[w.data_ptr() for w in ws])
        self.dst = (C.c_void_p * nmat)(*[o.data_ptr() for o in outs])
        self.dstride = (C.c_int64 * nmat)(*[(o.stride(0) if n > 1 else m) for o, m in zip(outs, ms)])
        self.m = (C.c_int64 * nmat)(*ms)
        self.ws_bytes = lib().b200_mul_mat_multi_workspace_size(nmat, self.types, self.m, n, k)
        self.device = outs[0].device
        assert all(o.dtype == torch.float32 for o in outs)

    def __call__(self, x, stream=None):
        n = x.shape[0]
        assert n == self.n and x.shape[1] == self.k and x.stride(1) == 1
        ws, wsz = _workspace(self.device, self.ws_bytes)
        check(lib().b200_mul_mat_multi(self.nmat, self.types, self.W, x.data_ptr(), x.stride(0) if n > 1 else self.k,
                                       self.dst, self.dstride, self.m, n, self.k, ws, wsz, _stream_ptr(stream)))


class MultiMatMulHost:
    """b200_mul_mat_multi_host: device-resident weights, HOST activations and HOST results (adjacent results return in
    one copy).  outs_host: list of host tensors [n, m_i]."""

    def __init__(self, types, ws, ms, k, outs_host, n=1):
        nmat = len(types)
        _require_cuda(*ws)
        self.nmat, self.k, self.n = nmat, k, n
        self.keep = (list(ws), list(outs_host))
        self.types = (C.c_int * nmat)(*types)
        self.W = (C.c_void_p * nmat)(*[w.data_ptr() for w in ws])
        self.dst = (C.c_void_p * nmat)(*[o.data_ptr() for o in outs_host])
        self.m = (C.c_int64 * nmat)(*ms)
        self.ws_bytes = lib().b200_mul_mat_multi_host_workspace_size(nmat, self.types, self.m, n, k)
        self.device = ws[0].device

    def __call__(self, x_host, stream=None):
        assert not x_host.is_cuda and x_host.is_contiguous() and tuple(x_host.shape) == (self.n, self.k)
        ws, wsz = _workspace(self.device, self.ws_bytes)
        check(lib().b200_mul_mat_multi_host(self.nmat, self.types, self.W, x_host.data_ptr(), self.dst, self.m, self.n,
                                            self.k, ws, wsz, _stream_ptr(stream)))


class MultiMatMulMMQ:
    """Prepared argument block for b200_mul_mat_mmq_multi: several weight PLANES multiplied with the same activations
    on the tensor cores (prefill)."""

    def __init__(self, types, planes, ms, k, outs, n):
        nmat = len(types)
        _require_cuda(*planes, *outs)
        self.nmat, self.k, self.n = nmat, k, n
        self.keep = (list(planes), list(outs))
        self.types = (C.c_int * nmat)(*types)
        self.P = (C.c_void_p * nmat)(*[p.data_ptr() for p in planes])
        self.dst = (C.c_void_p * nmat)(*[o.data_ptr() for o in outs])
        self.dstride = (C.c_int64 * nmat)(*[o.stride(0) for o in outs])
        self.m = (C.c_int64 * nmat)(*ms)
        self.ws_bytes = lib().b200_mul_mat_mmq_workspace_size(max(ms), n, k)
        self.device = outs[0].device

    def __call__(self, x, stream=None):
        assert x.shape == (self.n, self.k) and x.stride(1) == 1
        ws, wsz = _workspace(self.device, self.ws_bytes)
        check(lib().b200_mul_mat_mmq_multi(self.nmat, self.types, self.P, x.data_ptr(), x.stride(0), self.dst,
                                           self.dstride, self.m, self.n, self.k, ws, wsz, _stream_ptr(stream)))


class Sequence:
    """b200_sequence_*: a list of MUL_MAT groups at batch 1 run as ONE persistent launch.
    groups: [(x [1, k] or [k] f32 cuda, k, [(type, w, m, out [m] or [1, m] f32 cuda), ...], dependent)]"""

    def __init__(self, groups, stream=None, split=None):
        """split = (world, rank, peer_base (ctypes array of `world` pointers), span_bytes, peer_sync (same), peer_ll (same, or
        None)): this rank's plan of a row split.  The groups then describe the WHOLE matrices (m, out = all rows) with w = this
        rank's rows (qmm.row_split); every `out` must lie inside this rank's symmetric allocation"""
        import torch
        n = len(groups)
        arr = (SeqGroup * n)()
        self.keep = []
        for g, (x, k, mats, dep) in zip(arr, groups):
            _require_cuda(x)
            assert x.dtype == torch.float32 and x.numel() == k and x.is_contiguous()
            g.x, g.k, g.nmat, g.dependent = x.data_ptr(), k, len(mats), 1 if dep else 0
            self.keep.append(x)
            for i, (t, w, m, out) in enumerate(mats):
                _require_cuda(w, out)
                assert out.dtype == torch.float32 and out.numel() == m and out.is_contiguous()
                g.type[i], g.W[i], g.dst[i], g.m[i] = t, w.data_ptr(), out.data_ptr(), m
                self.keep += [w, out]
        nbytes = lib().b200_sequence_plan_size(n)
        self.plan = torch.zeros(nbytes, dtype=torch.uint8, device=self.keep[0].device)
        if split is None:
            check(lib().b200_sequence_plan_build(arr, n, self.plan.data_ptr(), nbytes, _stream_ptr(stream)))
        else:
            world, rank, peer_base, span, peer_sync, peer_ll = split
            check(lib().b200_sequence_plan_build_rowsplit(arr, n, self.plan.data_ptr(), nbytes, world, rank, peer_base, span,
                                                          peer_sync, peer_ll, _stream_ptr(stream)))

    def __call__(self, stream=None):
        check(lib().b200_sequence_run(self.plan.data_ptr(), _stream_ptr(stream)))

    def status(self, stream=None):
        err = C.c_int(0)
        check(lib().b200_sequence_status(self.plan.data_ptr(), C.byref(err), _stream_ptr(stream)))
        return err.value


def mul_mat_multi(types, ws, x, ms, k, outs=None, stream=None):
    """[W_i x^T for i] with x quantized once and same-type matrices sharing launches; returns the list of dst [n, m_i]"""
    import torch
    n = x.shape[0]
    if outs is None:
        outs = [torch.empty((n, m), dtype=torch.float32, device=x.device) for m in ms]
    MultiMatMul(types, ws, ms, k, outs, n)(x, stream)
    return outs


def build_plane(t, w, m, k, stream=None):
    """int8 prefill plane of a weight tensor (built once per tensor, +1 byte per weight): uint8 cuda tensor"""
    import torch
    _require_cuda(w)
    nbytes = lib().b200_mmq_plane_size(t, m, k)
    plane = torch.empty(nbytes, dtype=torch.uint8, device=w.device)
    check(lib().b200_mmq_build_plane(t, w.data_ptr(), 0, plane.data_ptr(), nbytes, m, k, _stream_ptr(stream)))
    return plane


def mul_mat_mmq(t, plane, x, m, k, out=None, stream=None):
    """tensor-core batched matmul against a prepared plane; x: [n, k] f32 cuda -> dst [n, m] f32"""
    import torch
    _require_cuda(plane, x)
    assert x.dtype == torch.float32 and x.dim() == 2 and x.stride(1) == 1 and x.shape[1] == k
    n = x.shape[0]
    dst = out if out is not None else torch.empty((n, m), dtype=torch.float32, device=x.device)
    ws, wsz = _workspace(x.device, lib().b200_mul_mat_mmq_workspace_size(m, n, k))
    check(lib().b200_mul_mat_mmq(t, plane.data_ptr(), x.data_ptr(), x.stride(0) if n > 1 else k, dst.data_ptr(),
                                 dst.stride(0) if n > 1 else m, m, n, k, ws, wsz, _stream_ptr(stream)))
    return dst


def mul_mat_host(t, w, x_host, dst_host, m, k, stream=None):
    """w: uint8 cuda (device-resident weights); x_host [n, k] / dst_host [n, m]: f32 HOST tensors (pin them for truly
    asynchronous copies).  Stream-ordered H2D -> matmul -> D2H; synchronize the stream before reading dst_host."""
    _require_cuda(w)
    if x_host.is_cuda or dst_host.is_cuda:
        raise B200Error("mul_mat_host takes host activations / results")
    n = x_host.shape[0]
    assert x_host.is_contiguous() and dst_host.is_contiguous() and tuple(dst_host.shape) == (n, m)
    ws, wsz = _workspace(w.device, lib().b200_mul_mat_host_workspace_size(t, m, n, k))
    check(lib().b200_mul_mat_host(t, w.data_ptr(), 0, x_host.data_ptr(), dst_host.data_ptr(), m, n, k, ws, wsz,
                                  _stream_ptr(stream)))
    return dst_host


def mul_mat_id(t, w, b, ids, m, k, n_expert, out=None, stream=None):
    """w: uint8 cuda, n_expert*m rows; b: [n_tokens, ne11, k] f32; ids: [n_tokens, n_used] int32 (row stride free)
    returns dst [n_tokens, n_used, m]"""
    import torch
    _require_cuda(w, b, ids)
    assert b.dtype == torch.float32 and b.dim() == 3 and b.stride(2) == 1
    assert ids.dtype == torch.int32 and ids.dim() == 2 and ids.stride(1) == 1
    n_tokens, ne11, _ = b.shape
    n_used = ids.shape[1]
    dst = out if out is not None else torch.empty((n_tokens, n_used, m), dtype=torch.float32, device=b.device)
    ws, wsz = _workspace(b.device, lib().b200_mul_mat_id_workspace_size(t, m, k, n_expert, n_used, n_tokens))
    check(lib().b200_mul_mat_id(t, w.data_ptr(), 0, b.data_ptr(), b.stride(0), b.stride(1), ne11, ids.data_ptr(),
                                ids.stride(0), dst.data_ptr(), dst.stride(0), dst.stride(1), m, k, n_expert, n_used,
                                n_tokens, ws, wsz, _stream_ptr(stream)))
    return dst


def mul_mat_id_mmq(t, plane, b, ids, m, k, n_expert, out=None, stream=None):
    """MUL_MAT_ID on the tensor cores: plane = build_plane(t, w, n_expert * m, k) of the stacked experts; b, ids, dst as in
    mul_mat_id.  Rows of dst whose id is invalid are left untouched."""
    import torch
    _require_cuda(plane, b, ids)
    assert b.dtype == torch.float32 and b.dim() == 3 and b.stride(2) == 1
    assert ids.dtype == torch.int32 and ids.dim() == 2 and ids.stride(1) == 1
    n_tokens, ne11, _ = b.shape
    n_used = ids.shape[1]
    dst = out if out is not None else torch.empty((n_tokens, n_used, m), dtype=torch.float32, device=b.device)
    ws, wsz = _workspace(b.device, lib().b200_mul_mat_id_mmq_workspace_size(t, m, k, n_expert, n_used, n_tokens))
    check(lib().b200_mul_mat_id_mmq(t, plane.data_ptr(), b.data_ptr(), b.stride(0), b.stride(1), ne11, ids.data_ptr(),
                                    ids.stride(0), dst.data_ptr(), dst.stride(0), dst.stride(1), m, k, n_expert, n_used,
                                    n_tokens, ws, wsz, _stream_ptr(stream)))
    return dst
