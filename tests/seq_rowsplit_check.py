"""Row-split sequence kernel on N GPUs: a dependent chain of groups, each rank computing its row slices in ONE persistent
launch and mirroring them into every rank's result buffer (b200_sequence_plan_build_rowsplit).  Checked bit for bit
against the same matmuls computed whole on every rank with b200_mul_mat, over several runs of the same plan (the epoch
scheme of the arrival counters) including runs with a deliberately late rank.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/seq_rowsplit_check.py
Prints one line per rank, exits non-zero on any mismatch.  tests/test_gpu_parity.py runs it when two GPUs are visible.
"""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))   # qref: the random-block generator only (numpy)
import b200pkg  # noqa: E402

qmm = b200pkg.load().qmm
from llama_cpp_dsp_b200.rowsplit import PeerBuffers  # noqa: E402
import qref  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    rng = np.random.default_rng(7)            # the same weights on every rank
    # (k, [(type, m), ...]) per group; m = 128 leaves most ranks without rows of that matrix; the chain feeds group g's
    # first result (padded / truncated to the next k) into group g + 1
    shapes = [(1024, [(qref.Q4_K, 1536), (qref.Q6_K, 256), (qref.Q4_0, 128)]),
              (1536, [(qref.Q8_0, 1024)]),
              (1024, [(qref.IQ4_XS, 2048), (qref.Q4_K, 2048)]),
              (2048, [(qref.Q6_K, 1024)]),
              (1024, [(qref.Q4_K, 4096)]),
              (4096, [(qref.Q4_K, 1024)])] * 3
    full_w, offs, total = [], [], 0
    for k, mats in shapes:
        ws, os_ = [], []
        for t, m in mats:
            ws.append(torch.from_numpy(qref.random_blocks(t, m, k, rng)).to(dev))
            os_.append(total)
            total += (m + 63) // 64 * 64
        full_w.append(ws), offs.append(os_)
    sym = PeerBuffers(1, total, 2 * len(shapes) + 8, world, rank, dev)
    x0 = torch.from_numpy(rng.standard_normal(shapes[0][0]).astype(np.float32)).to(dev)

    # reference: every matmul whole, on this rank, group g + 1 reading group g's first result
    ref = torch.zeros(total, device=dev)
    x = x0
    for (k, mats), ws, os_ in zip(shapes, full_w, offs):
        assert x.numel() == k
        for (t, m), w, o in zip(mats, ws, os_):
            qmm.mul_mat(t, w, x.view(1, k), m, k, out=ref[o:o + m].view(1, m))
        x = ref[os_[0]:os_[0] + mats[0][1]]
    torch.cuda.synchronize()

    groups = []
    x = x0
    keep = []
    for (k, mats), ws, os_ in zip(shapes, full_w, offs):
        ms = []
        for (t, m), w, o in zip(mats, ws, os_):
            lo, hi = qmm.row_split(m, world, rank)
            wl = w[lo:hi].contiguous() if hi > lo else w[:1].contiguous()
            keep.append(wl)
            ms.append((t, wl, m, sym.dst[0, o:o + m]))         # the whole matrix; w = this rank's rows
        groups.append((x.view(1, k), k, ms, True))
        x = sym.dst[0, os_[0]:os_[0] + mats[0][1]]
    use_ll = os.environ.get("SEQ_SPLIT_LL", "1") != "0"           # exchange protocol: flag-in-data (default) or counters
    llb = PeerBuffers(1, 2 * total, 1, world, rank, dev) if use_ll else None
    seq = qmm.Sequence(groups, split=(world, rank, sym.ptr_dst, total * 4, sym.ptr_sync[0], llb.ptr_dst if use_ll else None))
    dist.barrier()
    bad = 0
    for it in range(6):
        sym.dst.zero_()
        torch.cuda.synchronize()
        dist.barrier()
        if it >= 3 and rank == it % world:
            time.sleep(0.05)                   # a late rank: the others spin on its counters
        seq()
        torch.cuda.synchronize()
        st = seq.status()
        same = torch.equal(sym.dst[0], ref)
        if not same or st:
            bad += 1
            d = (sym.dst[0] != ref).nonzero()
            print("rank %d run %d: status %d, %d of %d elements differ, first at %s" %
                  (rank, it, st, d.numel(), total, d[0].item() if d.numel() else None), flush=True)
        dist.barrier()
    # timing, graph replay
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        seq(s)
        s.synchronize()
        dist.barrier()
        with torch.cuda.graph(g, stream=s):
            seq(s)
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        e0.record(s)
        for _ in range(20):
            g.replay()
        e1.record(s)
    s.synchronize()
    same = torch.equal(sym.dst[0], ref) and seq.status() == 0
    bad += 0 if same else 1
    print("rank %d/%d: %d groups, %s, %s, %.1f us per run" % (rank, world, len(groups), "ll" if use_ll else "counters", "OK" if not bad else "MISMATCH",
                                                        e0.elapsed_time(e1) * 1e3 / 20), flush=True)
    flag = torch.tensor([bad], device=dev)
    dist.all_reduce(flag)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(1 if flag.item() else 0)


if __name__ == "__main__":
    main()
