"""Micro-benchmark of the GEMV kernel on distinct (> L2) weight matrices, replayed from a CUDA graph so that host
launch overhead is out of the picture; sweeps env knobs in-process.
usage: python tools/gemv_sweep.py [--quick] [type ...]"""
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
KNOBS = ["B200_GEMV_STAGES", "B200_GEMV_V", "B200_GEMV_MINB", "B200_PDL", "B200_FUSED_QUANT_MAX_K", "B200_GEMV_WAVES",
         "B200_GEMV_RESIDENT", "B200_GEMV_TEAMS"]


def make_weights(t, m, k, copies, rng):
    one = torch.from_numpy(qref.random_blocks(t, m, k, rng)).cuda()
    return [one] + [one.clone() for _ in range(copies - 1)]


def time_gemv(t, m, k, n, ws, iters=10, graph=True):
    x = torch.rand((n, k), device="cuda") * 2 - 1
    outs = [torch.empty((n, m), device="cuda") for _ in ws]

    def run():
        for w, o in zip(ws, outs):
            qmm.mul_mat(t, w, x, m, k, out=o)

    run()
    torch.cuda.synchronize()
    s = torch.cuda.Stream()
    if graph:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.stream(s):
            run()
            s.synchronize()
            with torch.cuda.graph(g, stream=s):
                run()
        fn = g.replay
    else:
        fn = run
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (iters * len(ws))  # us per matmul


def main():
    args = sys.argv[1:]
    quick = "--quick" in args
    names = [a for a in args if not a.startswith("--")] or ["q4_K"]
    rng = np.random.default_rng(0)
    shapes = [(4096, 14336), (14336, 4096), (4096, 4096), (1024, 4096)]
    ns = (1, 2, 4, 8)
    if "--big" in args:      # ring geometry on the shapes that carry the bytes of a decode token
        shapes = [(28672, 4096), (14336, 4096), (4096, 14336), (6144, 4096), (4096, 4096)]
        ns = (1,)
    if "--70b" in args:      # the row lengths of the Llama-3-70B-shaped token (bench.py headline)
        shapes = [(10240, 8192), (8192, 8192), (57344, 8192), (8192, 28672)]
        ns = (1,)
    configs = [{}] if quick else [
        {}, {"B200_GEMV_V": 4}, {"B200_GEMV_V": 2}, {"B200_GEMV_V": 4, "B200_GEMV_STAGES": 12}, {"B200_GEMV_V": 4, "B200_GEMV_STAGES": 16},
        {"B200_GEMV_V": 2, "B200_GEMV_STAGES": 16}, {"B200_GEMV_V": 2, "B200_GEMV_STAGES": 24}, {"B200_GEMV_V": 8, "B200_GEMV_TEAMS": 3},
        {"B200_GEMV_V": 4, "B200_GEMV_TEAMS": 3, "B200_GEMV_STAGES": 9}, {"B200_GEMV_V": 4, "B200_GEMV_TEAMS": 2, "B200_GEMV_STAGES": 8},
    ] if "--70b" in args else [
        {}, {"B200_GEMV_V": 4}, {"B200_GEMV_V": 4, "B200_GEMV_STAGES": 24}, {"B200_GEMV_V": 4, "B200_GEMV_STAGES": 32},
        {"B200_GEMV_V": 8, "B200_GEMV_TEAMS": 6}, {"B200_GEMV_V": 8, "B200_GEMV_TEAMS": 4},
        {"B200_GEMV_V": 8, "B200_GEMV_TEAMS": 4, "B200_GEMV_STAGES": 12}, {"B200_GEMV_V": 2, "B200_GEMV_STAGES": 32},
        {"B200_GEMV_V": 2, "B200_GEMV_STAGES": 64}, {"B200_GEMV_V": 4, "B200_GEMV_TEAMS": 4, "B200_GEMV_STAGES": 24},
    ] if "--big" in args else [
        {}, {"B200_PDL": 0}, {"B200_GEMV_RESIDENT": 2}, {"B200_GEMV_RESIDENT": 2, "B200_FUSED_QUANT_MAX_K": 0},
        {"B200_FUSED_QUANT_MAX_K": 0}, {"B200_GEMV_V": 2}, {"B200_GEMV_V": 4}, {"B200_GEMV_V": 8},
        {"B200_GEMV_RESIDENT": 2, "B200_GEMV_V": 4}, {"B200_GEMV_RESIDENT": 2, "B200_GEMV_V": 8},
        {"B200_GEMV_STAGES": 2}, {"B200_GEMV_MINB": 1}, {"B200_GEMV_WAVES": 2},
    ]
    for name in names:
        t = qref.WEIGHT_TYPES[name]
        for (m, k) in shapes:
            nbytes = qmm.row_size(t, k) * m
            copies = max(2, int(600e6 // nbytes))          # > 4x L2
            ws = make_weights(t, m, k, copies, rng)
            for n in ns:
                for cfg in configs:
                    if n > 1 and cfg:
                        continue
                    for kn in KNOBS:
                        os.environ.pop(kn, None)
                    for kn, v in cfg.items():
                        os.environ[kn] = str(v)
                    qmm.reload_env()
                    try:
                        us = time_gemv(t, m, k, n, ws)
                    except Exception as e:  # noqa: BLE001
                        print(name, m, k, n, cfg, "ERR", str(e)[:100])
                        continue
                    gbs = (nbytes + n * k * 4 + n * m * 4) / us / 1e3
                    print("%-6s m=%5d k=%5d n=%d %-55s: %8.2f us  %7.1f GB/s" % (name, m, k, n, cfg, us, gbs), flush=True)
            del ws
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
