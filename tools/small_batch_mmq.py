"""n = 5..16: GEMV (two passes above 4 columns) against the tensor-core path with cached planes, weights cycled > L2.
usage: python tools/small_batch_mmq.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200pkg  # noqa: E402
import bench  # noqa: E402

qmm = b200pkg.load()
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev)
gen.manual_seed(3)
m, k, copies = 4096, 14336, 8
for t in (bench.Q4_K, bench.Q6_K, bench.Q8_0):
    ws = [bench.gen_weights(torch, t, m, k, gen, dev) for _ in range(copies)]
    planes = [qmm.build_plane(t, w, m, k) for w in ws]
    for n in (4, 5, 6, 8, 12, 16):
        x = torch.rand((n, k), device=dev, generator=gen) * 2 - 1
        out = torch.empty((n, m), device=dev)
        res = {}
        for path in ("gemv", "mmq"):
            if path == "gemv" and n > 8:
                continue

            def calls():
                for i in range(copies):
                    if path == "mmq":
                        qmm.mul_mat_mmq(t, planes[i], x, m, k, out=out)
                    else:
                        qmm.mul_mat(t, ws[i], x, m, k, out=out)
            st = torch.cuda.Stream(device=dev)
            g = torch.cuda.CUDAGraph()
            with torch.cuda.stream(st):
                calls()
                st.synchronize()
                with torch.cuda.graph(g, stream=st):
                    calls()
                for _ in range(2):
                    g.replay()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(st)
                for _ in range(10):
                    g.replay()
                e1.record(st)
                st.synchronize()
            res[path] = e0.elapsed_time(e1) * 1e3 / (10 * copies)
        print("%-6s n=%2d  gemv %s us   tensor cores (cached plane) %.1f us" % (
            bench.TNAME[t], n, ("%.1f" % res["gemv"]) if "gemv" in res else "  - ", res["mmq"]), flush=True)
    del ws, planes
