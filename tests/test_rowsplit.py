"""Multi-GPU row split.  The reference's tests do not pin split-mode row at all ("parity unpinned" there),
so this file does: the row partition must follow get_row_split (ggml/src/ggml-cuda/ggml-cuda.cu:727-753), and the
gathered result of the per-rank slices must equal the unsplit matmul BIT FOR BIT (rows are independent).

CPU part: two gloo ranks, each computing its slice with the oracle as a stand-in for the kernel (tests may use the
oracle), gathered through the product's host logic (llama.cpp.dsp_b200/rowsplit.py).  GPU part: the same with the
CUDA kernels, ranks emulated one after the other on one device."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import b200pkg
import qref

T_Q4_K, T_Q8_0, T_Q6_K = 12, 8, 14


def _ref_ranges(m, world, tensor_split=None):
    """get_row_split restated (ggml-cuda.cu:740-753): cumulative fractions, boundaries rounded down to 128 rows"""
    ts = np.asarray(tensor_split if tensor_split is not None else [1.0] * world, dtype=np.float32)
    cum = np.concatenate([[np.float32(0)], np.cumsum(ts, dtype=np.float32)[:-1]]) / np.float32(ts.sum(dtype=np.float32))
    out = []
    for r in range(world):
        lo = int(np.float32(m) * np.float32(cum[r])) if r else 0
        lo -= lo % 128
        if r == world - 1:
            hi = m
        else:
            hi = int(np.float32(m) * np.float32(cum[r + 1]))
            hi -= hi % 128
        out.append((lo, hi))
    return out


@pytest.mark.parametrize("m,world,ts", [(4096, 2, None), (4096, 8, None), (128256, 8, None), (300, 2, None),
                                        (100, 2, None), (28672, 4, [3, 1, 1, 3]), (1024, 3, [0.2, 0.5, 0.3]),
                                        (8192, 16, None), (0, 2, None)])
def test_partition_follows_reference_rule(m, world, ts):
    pkg = b200pkg.load()
    from llama_cpp_dsp_b200.rowsplit import RowSplit
    want = _ref_ranges(m, world, ts)
    for r in range(world):
        s = RowSplit(m, world, r, ts)
        assert s.ranges == want
        assert (s.lo, s.hi) == want[r]
    assert pkg is not None


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank, world, port, case, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        b200pkg.load()
        from llama_cpp_dsp_b200.rowsplit import RowSplit
        t, m, k, n, ts = case
        rng = np.random.default_rng(7)                      # same data on every rank
        w = qref.random_blocks(t, m, k, rng)
        x = rng.uniform(-1, 1, (n, k)).astype(np.float32)
        split = RowSplit(m, world, rank, ts)
        local = split.new_local(n, torch.zeros(1))
        if split.rows:
            mine = qref.mul_mat(t, split.slice_weights(w), x, split.rows, k, "q8_1")   # stand-in for the CUDA kernel
            local[:, :split.rows] = torch.from_numpy(mine)
        local[:, split.rows:] = float("nan")                # padding must never leak into the result
        full = split.gather(local)
        want = torch.from_numpy(qref.mul_mat(t, w, x, m, k, "q8_1"))
        ret[rank] = bool(torch.equal(full, want)) and not bool(torch.isnan(full).any())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("case", [(T_Q4_K, 512, 256, 1, None), (T_Q8_0, 300, 64, 3, None), (T_Q6_K, 100, 256, 2, None),
                                  (T_Q4_K, 640, 256, 1, [3, 1])])
def test_gloo_two_ranks_gather_equals_unsplit(case):
    world = 2
    mgr = mp.get_context("spawn").Manager()
    ret = mgr.dict()
    mp.spawn(_rank_main, args=(world, _free_port(), case, ret), nprocs=world, join=True)
    assert dict(ret) == {0: True, 1: True}


@pytest.mark.gpu
@pytest.mark.parametrize("t,m,k,world,ts", [(T_Q4_K, 4096, 4096, 2, None), (T_Q6_K, 1024, 4096, 8, None),
                                            (T_Q8_0, 1200, 2048, 4, [1, 2, 2, 1]), (T_Q4_K, 28672, 8192, 8, None)])
def test_cuda_slices_assemble_to_the_unsplit_result(t, m, k, world, ts):
    qmm = b200pkg.load()
    from llama_cpp_dsp_b200.rowsplit import RowSplit
    rng = np.random.default_rng(3)
    w = torch.from_numpy(qref.random_blocks(t, m, k, rng)).cuda()
    x = torch.from_numpy(rng.uniform(-1, 1, (1, k)).astype(np.float32)).cuda()
    want = qmm.mul_mat(t, w, x, m, k)
    got = torch.full((1, m), float("nan"), device="cuda")
    for r in range(world):
        s = RowSplit(m, world, r, ts)
        local = s.new_local(1, x)
        if s.rows:
            qmm.mul_mat(t, s.slice_weights(w).contiguous(), x, s.rows, k, out=local[:, :s.rows])
        got[:, s.lo:s.hi] = local[:, :s.rows]
    assert torch.equal(got, want)


@pytest.mark.gpu
@pytest.mark.parametrize("t,m,k,n,world,ts", [(T_Q4_K, 4096, 4096, 1, 2, None), (T_Q6_K, 1024, 8192, 2, 4, None),
                                              (T_Q8_0, 100, 2048, 1, 2, None), (T_Q4_K, 28672, 8192, 1, 8, None)])
def test_fused_peer_store_row_split_one_device(t, m, k, n, world, ts):
    """b200_mul_mat_rowsplit_fused + b200_rowsplit_wait with every "rank" on ONE device (peer pointers are local
    pointers): all ranks' dst buffers must equal the unsplit result bit for bit, twice in a row on alternating sync
    blocks (the counters re-arm).  The real multi-process path over CUDA IPC is exercised by bench.py --gpus N."""
    import ctypes as C
    qmm = b200pkg.load()
    from llama_cpp_dsp_b200.rowsplit import RowSplit, mul_mat_row_split_fused, row_split_wait
    rng = np.random.default_rng(5)
    w = torch.from_numpy(qref.random_blocks(t, m, k, rng)).cuda()
    x = torch.from_numpy(rng.uniform(-1, 1, (n, k)).astype(np.float32)).cuda()
    want = qmm.mul_mat(t, w, x, m, k)
    dsts = [torch.full((n, m), float("nan"), device="cuda") for _ in range(world)]
    syncs = [torch.zeros((2, 2), dtype=torch.int32, device="cuda") for _ in range(world)]
    ws = torch.empty(qmm.lib().b200_mul_mat_rowsplit_workspace_size(t, n, k) + 16, dtype=torch.uint8, device="cuda")
    ptr_dst = (C.c_void_p * world)(*[d.data_ptr() for d in dsts])
    for it in range(4):
        blk = it % 2
        ptr_sync = (C.c_void_p * world)(*[s.data_ptr() + 8 * blk for s in syncs])
        for d in dsts:
            d.fill_(float("nan"))
        for r in range(world):
            s = RowSplit(m, world, r, ts)
            wl = s.slice_weights(w).contiguous() if s.rows else None
            mul_mat_row_split_fused(t, wl, x, s, k, ptr_dst, ptr_sync, m, ws)
        for r in range(world):
            row_split_wait(ptr_sync, r, world)
        torch.cuda.synchronize()
        for r in range(world):
            assert torch.equal(dsts[r], want), (it, r)
            assert int(syncs[r][blk, 0]) == 0 and int(syncs[r][blk, 1]) == 0     # re-armed
    # chained form: call i+1 waits for call i inside its own activation quantizer, only the last one uses a stream wait
    prev = None
    for it in range(5):
        ptr_sync = (C.c_void_p * world)(*[s.data_ptr() + 8 * (it % 2) for s in syncs])
        for r in range(world):
            s = RowSplit(m, world, r, ts)
            wl = s.slice_weights(w).contiguous() if s.rows else None
            mul_mat_row_split_fused(t, wl, x, s, k, ptr_dst, ptr_sync, m, ws, wait_sync=prev)
        prev = ptr_sync
    for r in range(world):
        row_split_wait(prev, r, world)
    torch.cuda.synchronize()
    for r in range(world):
        assert torch.equal(dsts[r], want), r
        assert int(syncs[r].abs().sum()) == 0                                    # all re-armed
    gave_up = C.c_int(-1)
    qmm.check(qmm.lib().b200_rowsplit_timeout_count(C.byref(gave_up), None))
    assert gave_up.value == 0                                                    # no in-kernel wait timed out


@pytest.mark.gpu
def test_fused_row_split_groups_share_a_sync_block_one_device():
    """Independent matmuls of one dependency group (q/k/v, gate/up) signal ONE sync block; the first matmul of the next
    group waits for world * calls arrivals.  Different weights per call, same dst: the last call's result must win."""
    import ctypes as C
    qmm = b200pkg.load()
    from llama_cpp_dsp_b200.rowsplit import RowSplit, mul_mat_row_split_fused, row_split_wait
    t, m, k, world = T_Q4_K, 2048, 4096, 2
    rng = np.random.default_rng(11)
    ws_ = [torch.from_numpy(qref.random_blocks(t, m, k, rng)).cuda() for _ in range(3)]
    x = torch.from_numpy(rng.uniform(-1, 1, (1, k)).astype(np.float32)).cuda()
    wants = [qmm.mul_mat(t, w, x, m, k) for w in ws_]
    dsts = [torch.full((1, m), float("nan"), device="cuda") for _ in range(world)]
    syncs = [torch.zeros((4, 2), dtype=torch.int32, device="cuda") for _ in range(world)]
    scratch = torch.empty(qmm.lib().b200_mul_mat_rowsplit_workspace_size(t, 1, k) + 16, dtype=torch.uint8, device="cuda")
    ptr_dst = (C.c_void_p * world)(*[d.data_ptr() for d in dsts])
    blocks = [(C.c_void_p * world)(*[s.data_ptr() + 8 * b for s in syncs]) for b in range(4)]
    prev, prev_calls = None, 0
    for step in range(6):                       # groups of 3, 1, 2, 1, 3, 2 calls
        calls = [3, 1, 2, 1, 3, 2][step]
        blk = blocks[step % 4]
        for r in range(world):
            s = RowSplit(m, world, r)
            for j in range(calls):
                mul_mat_row_split_fused(t, s.slice_weights(ws_[j]).contiguous(), x, s, k, ptr_dst, blk, m, scratch,
                                        wait_sync=prev if j == 0 else None, wait_calls=prev_calls)
        prev, prev_calls = blk, calls
    for r in range(world):
        row_split_wait(prev, r, world, calls=prev_calls)
    torch.cuda.synchronize()
    for r in range(world):
        assert torch.equal(dsts[r], wants[1]), r     # the last group had 2 calls: weights index 1 wrote last
        assert int(syncs[r].abs().sum()) == 0


@pytest.mark.gpu
def test_fused_row_split_prefill_one_device():
    """b200_mul_mat_mmq_multi_rowsplit with ONE device playing every rank (peer_delta between two dst buffers in one
    allocation): each "rank" multiplies its row slices of q/k/v on the tensor cores, the epilogue stores them into both
    copies, one arrival per rank; both copies must equal the unsplit b200_mul_mat_mmq result bit for bit."""
    import ctypes as C
    qmm = b200pkg.load()
    from llama_cpp_dsp_b200.rowsplit import MultiMatMulMMQRowSplit, row_split_wait
    rng = np.random.default_rng(21)
    world, n, k = 2, 160, 1024
    mats = [(qref.Q4_K, 1280), (qref.Q6_K, 256), (qref.Q4_K, 384)]
    ws_ = [torch.from_numpy(qref.random_blocks(t, m, k, rng)).cuda() for t, m in mats]
    x = torch.from_numpy(rng.uniform(-1, 1, (n, k)).astype(np.float32)).cuda()
    # the row-split launch never splits k (one launch = one arrival), so the unsplit reference must not either: the two-half
    # split-K adds its halves with red.add -- the same value up to fp32 summation order, not the same bits
    qmm.setenv(B200_MMQ_KSPLIT=1)
    try:
        want = [qmm.mul_mat_mmq(t, qmm.build_plane(t, w, m, k), x, m, k) for (t, m), w in zip(mats, ws_)]
        torch.cuda.synchronize()
    finally:
        qmm.setenv(B200_MMQ_KSPLIT=None)
    total = sum(n * m for _, m in mats)
    sym = torch.full((world, total), float("nan"), device="cuda")           # rank r's "symmetric allocation" = row r
    syncs = torch.zeros((world, 2), dtype=torch.int32, device="cuda")
    ptr_sync = (C.c_void_p * world)(*[syncs[r].data_ptr() for r in range(world)])
    for rank in range(world):
        delta = [(r - rank) * total * 4 for r in range(world)]
        planes, msl, dl, off = [], [], [], 0
        for (t, m), w in zip(mats, ws_):
            lo, hi = qmm.row_split(m, world, rank)
            planes.append(qmm.build_plane(t, w[lo:hi].contiguous(), hi - lo, k) if hi > lo else None)
            msl.append(hi - lo)
            full = sym[rank, off:off + n * m].view(n, m)
            dl.append(full[:, lo:hi] if hi > lo else None)
            off += n * m
        MultiMatMulMMQRowSplit([t for t, _ in mats], planes, msl, k, dl, n, world, rank, delta)(x, ptr_sync)
    for rank in range(world):
        row_split_wait(ptr_sync, rank, world)
    torch.cuda.synchronize()
    assert int(syncs.abs().sum()) == 0
    off = 0
    for (t, m), w_ in zip(mats, want):
        for rank in range(world):
            assert torch.equal(sym[rank, off:off + n * m].view(n, m), w_), (m, rank)
        off += n * m


@pytest.mark.gpu
def test_copy_engine_push_one_device():
    """b200_rowsplit_push with ONE device playing both ranks: each "rank" multiplies its row slices locally
    (b200_mul_mat_mmq_multi into its own copy of dst), the push copies them into the other copy through side streams and
    counts the arrivals; both copies must hold the whole result, and the sync blocks must be re-armed by the waits."""
    import ctypes as C
    qmm = b200pkg.load()
    from llama_cpp_dsp_b200.rowsplit import RowSplitPush, row_split_push_join, row_split_wait
    rng = np.random.default_rng(22)
    world, n, k = 2, 96, 1024
    mats = [(qref.Q4_K, 768), (qref.IQ4_XS, 256)]
    ws_ = [torch.from_numpy(qref.random_blocks(t, m, k, rng)).cuda() for t, m in mats]
    x = torch.from_numpy(rng.uniform(-1, 1, (n, k)).astype(np.float32)).cuda()
    total = sum(n * m for _, m in mats)
    sym = torch.full((world, total), float("nan"), device="cuda")
    syncs = torch.zeros((world, 2), dtype=torch.int32, device="cuda")
    ptr_sync = (C.c_void_p * world)(*[syncs[r].data_ptr() for r in range(world)])
    want = []
    for rank in range(world):
        delta = [(r - rank) * total * 4 for r in range(world)]
        dl, off = [], 0
        for (t, m), w in zip(mats, ws_):
            lo, hi = qmm.row_split(m, world, rank)
            full = sym[rank, off:off + n * m].view(n, m)
            if hi > lo:
                plane = qmm.build_plane(t, w[lo:hi].contiguous(), hi - lo, k)
                qmm.mul_mat_mmq(t, plane, x, hi - lo, k, out=full[:, lo:hi])
                dl.append(full[:, lo:hi])
            off += n * m
        RowSplitPush(dl, n, world, rank, delta)(ptr_sync)
    for rank in range(world):
        row_split_wait(ptr_sync, rank, world)
    row_split_push_join()
    torch.cuda.synchronize()
    assert int(syncs.abs().sum()) == 0
    off = 0
    for (t, m), w in zip(mats, ws_):
        a, b = sym[0, off:off + n * m].view(n, m), sym[1, off:off + n * m].view(n, m)
        assert not torch.isnan(a).any() and torch.equal(a, b), m
        for rank in range(world):                            # every slice equals the matmul on that rank's rows alone
            lo, hi = qmm.row_split(m, world, rank)
            if hi > lo:
                ref = qmm.mul_mat_mmq(t, qmm.build_plane(t, w[lo:hi].contiguous(), hi - lo, k), x, hi - lo, k)
                assert torch.equal(a[:, lo:hi], ref), (m, rank)
        off += n * m
