"""Micro-benchmark of the GEMV kernel on distinct (> L2) weight matrices; sweeps env knobs in-process.
usage: python tools/gemv_sweep.py [type ...]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import b200pkg  # noqa: E402
import qref  # noqa: E402

qmm = b200pkg.load()


def make_weights(t, m, k, copies, rng):
    rs = qmm.row_size(t, k)
    one = torch.from_numpy(qref.random_blocks(t, m, k, rng)).cuda()
    ws = [one]
    for _ in range(copies - 1):
        ws.append(one.clone())
    return ws, rs * m


def time_gemv(t, m, k, n, ws, iters=20):
    x = torch.rand((n, k), device="cuda") * 2 - 1
    outs = [torch.empty((n, m), device="cuda") for _ in ws]
    for w, o in zip(ws, outs):
        qmm.mul_mat(t, w, x, m, k, out=o)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        for w, o in zip(ws, outs):
            qmm.mul_mat(t, w, x, m, k, out=o)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (iters * len(ws))  # us per matmul


def main():
    names = sys.argv[1:] or ["q4_K"]
    rng = np.random.default_rng(0)
    shapes = [(4096, 14336), (14336, 4096), (4096, 4096), (1024, 4096)]
    for name in names:
        t = qref.WEIGHT_TYPES[name]
        for (m, k) in shapes:
            nbytes = qmm.row_size(t, k) * m
            copies = max(2, int(600e6 // nbytes))          # > 4x L2
            ws, _ = make_weights(t, m, k, copies, rng)
            for n in (1, 2, 4, 8):
                for warps, stages, ctas, pdl in [(8, 3, 1, 1), (8, 3, 1, 0), (8, 4, 1, 1), (16, 2, 1, 1), (16, 3, 1, 1),
                                                 (4, 4, 2, 1), (8, 2, 2, 1), (4, 6, 1, 1)]:
                    if n > 1 and (warps, stages, ctas, pdl) != (8, 3, 1, 1):
                        continue
                    os.environ["B200_GEMV_WARPS"] = str(warps)
                    os.environ["B200_GEMV_STAGES"] = str(stages)
                    os.environ["B200_GEMV_CTAS_PER_SM"] = str(ctas)
                    os.environ["B200_PDL"] = str(pdl)
                    try:
                        us = time_gemv(t, m, k, n, ws)
                    except Exception as e:  # noqa: BLE001
                        print(name, m, k, n, warps, stages, ctas, "ERR", e)
                        continue
                    gbs = (nbytes + n * k * 4 + n * m * 4) / us / 1e3
                    print("%-6s m=%5d k=%5d n=%d warps=%2d stages=%d ctas/sm=%d pdl=%d : %8.2f us  %7.1f GB/s" %
                          (name, m, k, n, warps, stages, ctas, pdl, us, gbs), flush=True)
            del ws
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
