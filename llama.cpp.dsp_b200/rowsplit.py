"""Row-split MUL_MAT across one-process-per-GPU ranks (split-mode row / tensor_split).

Host-side mirror of the reference's split choreography for this path: which output rows a rank owns
(get_row_split / get_row_rounding, ggml/src/ggml-cuda/ggml-cuda.cu:727-753 -- computed by the C ABI's b200_row_split so
that the C++ plug-in and this module cannot disagree) and how the per-rank f32 slices of dst come together again
(reference: per-chunk peer copies into the main device's dst + events, ggml-cuda.cu:1560-1700; here: ONE all-gather per
matmul over torch.distributed -- NCCL on the GPUs, gloo in the CPU tests -- or no collective at all when the GEMV stores
straight into the peers, b200_mul_mat_rowsplit_fused).

torch.distributed is plumbing only: the slices are produced by the CUDA kernels through qmm.mul_mat.
"""
from . import qmm


class RowSplit:
    """The row partition of an m-row weight matrix over `world` ranks."""

    def __init__(self, m, world, rank, tensor_split=None):
        if tensor_split is not None and len(tensor_split) != world:
            raise ValueError("tensor_split needs one fraction per rank")
        self.m, self.world, self.rank = int(m), int(world), int(rank)
        self.ranges = [qmm.row_split(m, world, r, tensor_split) for r in range(world)]
        # the reference's invariants: contiguous cover of [0, m), interior boundaries on multiples of 128 rows
        assert self.ranges[0][0] == 0 and self.ranges[-1][1] == m
        for (_, hi), (lo, _) in zip(self.ranges[:-1], self.ranges[1:]):
            assert hi == lo and lo % 128 == 0
        self.lo, self.hi = self.ranges[rank]
        self.rows = self.hi - self.lo
        self.max_rows = max(hi - lo for lo, hi in self.ranges)

    def slice_weights(self, w_rows):
        """w_rows: anything indexable by row (tensor / ndarray of m rows) -> this rank's rows"""
        return w_rows[self.lo:self.hi]

    def new_local(self, n, like):
        """padded result buffer [n, max_rows] this rank's kernel writes its slice into (columns [0, rows))"""
        import torch
        return torch.zeros((n, self.max_rows), dtype=torch.float32, device=like.device)

    def gather(self, local, out=None, staging=None, group=None):
        """local: [n, max_rows] f32 with this rank's slice in [:, :rows].  Returns the full dst [n, m] on every rank.
        Uneven slices (the last rank takes the remainder, ggml-cuda.cu:746-751) are handled by gathering the padded
        buffers and dropping the padding."""
        import torch
        import torch.distributed as dist
        n = local.shape[0]
        if staging is None:
            staging = torch.empty((self.world, n, self.max_rows), dtype=local.dtype, device=local.device)
        if self.world == 1:
            staging[0].copy_(local)
        else:
            dist.all_gather_into_tensor(staging.view(self.world * n, self.max_rows), local.contiguous(), group=group)
        if out is None:
            out = torch.empty((n, self.m), dtype=local.dtype, device=local.device)
        for r, (lo, hi) in enumerate(self.ranges):
            if hi > lo:
                out[:, lo:hi].copy_(staging[r, :, :hi - lo])
        return out


def mul_mat_row_split(t, w_local, x, split, k, group=None):
    """dst = W x^T with W's rows split over the ranks of `group`: this rank multiplies its slice (CUDA kernels) and the
    slices are all-gathered.  w_local: this rank's rows of the quantized weight (uint8 [rows, row_size]); x: [n, k]."""
    local = split.new_local(x.shape[0], x)
    if split.rows > 0:
        qmm.mul_mat(t, w_local, x, split.rows, k, out=local[:, :split.rows])
    return split.gather(local, group=group)


# ---------------------------------------------------------------------------------------------------------------------
# fused path: the GEMV stores its rows into every rank's dst over NVLink, no collective launch
class PeerBuffers:
    """One f32 dst buffer [n, m] and `nsync` sync blocks per rank, every rank's copies mapped into every process through
    CUDA IPC (b200_ipc_*).  `ptr_dst` / `ptr_sync[i]` are ctypes arrays of `world` device pointers in rank order, ready
    for b200_mul_mat_rowsplit_fused."""

    def __init__(self, n, m, nsync, world, rank, device, group=None):
        import ctypes as C
        import torch
        import torch.distributed as dist
        self.world, self.rank, self.n, self.m, self.nsync = world, rank, n, m, nsync
        self.dst = torch.zeros((n, m), dtype=torch.float32, device=device)
        self.sync = torch.zeros((nsync, 2), dtype=torch.int32, device=device)
        torch.cuda.synchronize(device)
        self._opened = []
        L = qmm.lib()

        def export(t):
            h, off = C.create_string_buffer(64), C.c_int64()
            qmm.check(L.b200_ipc_get_handle(C.c_void_p(t.data_ptr()), h, C.byref(off)))
            return bytes(h.raw), off.value

        mine = (export(self.dst), export(self.sync))
        everyone = [mine]
        if world > 1:
            everyone = [None] * world
            dist.all_gather_object(everyone, mine, group=group)
        dst_ptrs, sync_ptrs = [], []
        for r, (hd, hs) in enumerate(everyone):
            if r == rank:
                dst_ptrs.append(self.dst.data_ptr()), sync_ptrs.append(self.sync.data_ptr())
                continue
            for (h, off), out in ((hd, dst_ptrs), (hs, sync_ptrs)):
                p = C.c_void_p()
                qmm.check(L.b200_ipc_open_handle(h, off, C.byref(p)))
                self._opened.append((p.value, off))
                out.append(p.value)
        self.ptr_dst = (C.c_void_p * world)(*dst_ptrs)
        self.ptr_sync = [(C.c_void_p * world)(*[p + 8 * i for p in sync_ptrs]) for i in range(nsync)]

    def close(self):
        import ctypes as C
        for p, off in self._opened:
            qmm.lib().b200_ipc_close_handle(C.c_void_p(p), off)
        self._opened = []


def mul_mat_row_split_fused(t, w_local, x, split, k, ptr_dst, ptr_sync, dst_row_stride, workspace, stream=None,
                            wait_sync=None, wait_calls=1):
    """launch this rank's part: rows [split.lo, split.hi) of dst are written into EVERY rank's dst buffer.
    wait_sync: ptr_sync array of the fused call(s) x depends on (the wait is folded into this call's quantizer);
    wait_calls: how many fused calls signalled that block (independent matmuls of one group share a block)"""
    import ctypes as C
    n = x.shape[0]
    qmm.check(qmm.lib().b200_mul_mat_rowsplit_fused(
        t, C.c_void_p(w_local.data_ptr() if w_local is not None else workspace.data_ptr()), C.c_void_p(x.data_ptr()),
        x.stride(0) if n > 1 else k, ptr_dst, ptr_sync, dst_row_stride, split.m, split.lo, split.hi, n, k, split.world,
        split.rank, C.c_void_p(wait_sync[split.rank]) if wait_sync is not None else None, split.world * wait_calls,
        C.c_void_p(workspace.data_ptr()), workspace.numel() * workspace.element_size(),
        qmm._stream_ptr(stream)))


class MultiMatMulMMQRowSplit:
    """Prepared arguments of b200_mul_mat_mmq_multi_rowsplit: this rank's row slices (prefill planes) of up to four matrices
    that share the activations; the GEMM's epilogue stores the results into every rank's dst and counts one arrival.
    dsts_local[i]: this rank's [n, rows_i] view of its own full dst (row stride = the full m); peer_delta: bytes from this
    rank's symmetric allocation to rank r's."""

    def __init__(self, types, planes, m_slices, k, dsts_local, n, world, rank, peer_delta):
        import ctypes as C
        nmat = len(types)
        self.keep = (list(planes), list(dsts_local))
        self.nmat, self.k, self.n, self.world, self.rank = nmat, k, n, world, rank
        self.types = (C.c_int * nmat)(*types)
        self.P = (C.c_void_p * nmat)(*[p.data_ptr() if p is not None else 0 for p in planes])
        self.dst = (C.c_void_p * nmat)(*[d.data_ptr() if d is not None else 0 for d in dsts_local])
        self.dstride = (C.c_int64 * nmat)(*[d.stride(0) if d is not None else 0 for d in dsts_local])
        self.m = (C.c_int64 * nmat)(*m_slices)
        self.delta = (C.c_int64 * world)(*peer_delta)
        self.ws_bytes = qmm.lib().b200_mul_mat_mmq_workspace_size(max(max(m_slices), 1), n, k)
        self.device = next(d for d in dsts_local if d is not None).device if any(d is not None for d in dsts_local) else None

    def __call__(self, x, ptr_sync, stream=None):
        import ctypes as C
        ws, wsz = qmm._workspace(x.device, self.ws_bytes)
        qmm.check(qmm.lib().b200_mul_mat_mmq_multi_rowsplit(
            self.nmat, self.types, self.P, C.c_void_p(x.data_ptr()), x.stride(0), self.dst, self.dstride, self.m, self.n, self.k,
            self.world, self.rank, self.delta, ptr_sync, ws, wsz, qmm._stream_ptr(stream)))


class RowSplitPush:
    """Prepared arguments of b200_rowsplit_push: this rank's result slices (views [n, rows_i] of its own full dst) go to the
    same place in every peer's dst through the copy engines, one side stream per peer, then count this rank's arrival."""

    def __init__(self, dsts_local, n, world, rank, peer_delta):
        import ctypes as C
        d = [t for t in dsts_local if t is not None]
        self.keep = d
        self.nmat, self.n, self.world, self.rank = len(d), n, world, rank
        self.src = (C.c_void_p * max(1, len(d)))(*[t.data_ptr() for t in d])
        self.stride = (C.c_int64 * max(1, len(d)))(*[t.stride(0) for t in d])
        self.cols = (C.c_int64 * max(1, len(d)))(*[t.shape[1] for t in d])
        self.delta = (C.c_int64 * world)(*peer_delta)

    def __call__(self, ptr_sync, stream=None):
        qmm.check(qmm.lib().b200_rowsplit_push(self.nmat, self.src, self.stride, self.cols, self.n, self.world, self.rank,
                                               self.delta, ptr_sync, qmm._stream_ptr(stream)))


def row_split_push_join(stream=None):
    qmm.check(qmm.lib().b200_rowsplit_push_join(qmm._stream_ptr(stream)))


def row_split_wait(ptr_sync, rank, world, stream=None, calls=1):
    """stream-ordered wait until the slices of all `world` ranks (x `calls` fused calls sharing the block) have landed
    in this rank's dst; re-arms the counter"""
    import ctypes as C
    qmm.check(qmm.lib().b200_rowsplit_wait(C.c_void_p(ptr_sync[rank]), world * calls, qmm._stream_ptr(stream)))
