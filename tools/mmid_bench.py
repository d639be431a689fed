"""Mixtral-8x7B-shaped MUL_MAT_ID, batched decode (BASELINE config 4): experts [4096 -> 14336] x 8 (gate/up, broadcast b)
and [14336 -> 4096] x 8 (down), Q4_K, 2 of 8 used, n_tokens = 1..64, ids = random permutation prefix per token.
Distinct weight copies are cycled so that every call streams from HBM.  Algorithmic bytes = distinct experts touched x
expert bytes + activations + results.   usage: python tools/mmid_bench.py [B200_MMID_GROUP values...]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200pkg  # noqa: E402
import bench  # noqa: E402

qmm = b200pkg.load()
dev = torch.device("cuda", 0)
Q4K = 12
n_expert, n_used = 8, 2
modes = sys.argv[1:] or ["0", "1"]
gen = torch.Generator(device=dev)
gen.manual_seed(7)
rng = np.random.default_rng(0)
for label, m, k, bcast in (("gate/up", 14336, 4096, True), ("down", 4096, 14336, False)):
    copies = [bench.gen_weights(torch, Q4K, n_expert * m, k, gen, dev) for _ in range(3)]   # 3 x 264 MB > L2
    ebytes = m * bench.row_bytes(Q4K, k)
    for n_tokens in (1, 2, 4, 8, 16, 32, 64):
        ne11 = 1 if bcast else n_used
        b = torch.rand((n_tokens, ne11, k), device=dev, generator=gen) * 2 - 1
        ids_np = np.stack([rng.permutation(n_expert)[:n_used] for _ in range(n_tokens)]).astype(np.int32)
        ids = torch.from_numpy(ids_np).to(dev)
        out = torch.empty((n_tokens, n_used, m), device=dev)
        distinct = len(set(ids_np.flatten().tolist()))
        algo = distinct * ebytes + b.numel() * 4 + out.numel() * 4
        row = "%-7s n_tokens=%2d distinct experts=%d :" % (label, n_tokens, distinct)
        ref = None
        for mode in modes:
            os.environ["B200_MMID_GROUP"] = mode
            qmm.reload_env()
            s = torch.cuda.Stream()
            with torch.cuda.stream(s):
                for w in copies:
                    qmm.mul_mat_id(Q4K, w, b, ids, m, k, n_expert, out=out)
                s.synchronize()
                res = out.clone()
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=s):
                    for w in copies:
                        qmm.mul_mat_id(Q4K, w, b, ids, m, k, n_expert, out=out)
                g.replay()
                torch.cuda.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(s)
                for _ in range(10):
                    g.replay()
                e1.record(s)
                torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / (10 * len(copies))
            if ref is None:
                ref = res
            err = float((res - ref).abs().max() / ref.abs().max())
            row += "  [group=%s] %7.1f us %6.0f GB/s (rel diff %.1e)" % (mode, us, algo / us / 1e3, err)
        print(row, flush=True)
    del copies
    torch.cuda.empty_cache()
