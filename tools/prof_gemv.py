"""Small fixed workload for ncu / debugging: a handful of GEMV launches on distinct matrices.
usage: python tools/prof_gemv.py [type] [m] [k] [n] [launches]"""
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
name = sys.argv[1] if len(sys.argv) > 1 else "q4_K"
m = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
k = int(sys.argv[3]) if len(sys.argv) > 3 else 14336
n = int(sys.argv[4]) if len(sys.argv) > 4 else 1
launches = int(sys.argv[5]) if len(sys.argv) > 5 else 8
t = qref.WEIGHT_TYPES[name]
rng = np.random.default_rng(0)
one = torch.from_numpy(qref.random_blocks(t, m, k, rng)).cuda()
ws = [one] + [one.clone() for _ in range(launches - 1)]
x = torch.rand((n, k), device="cuda") * 2 - 1
outs = [torch.empty((n, m), device="cuda") for _ in ws]
for rep in range(2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for w, o in zip(ws, outs):
        qmm.mul_mat(t, w, x, m, k, out=o)
    e1.record()
    torch.cuda.synchronize()
    print("%s m=%d k=%d n=%d: %.2f us per launch (eager, incl. host launch)" % (name, m, k, n, e0.elapsed_time(e1) * 1e3 / launches))
