"""Fixed workload for ncu: a few tensor-core matmuls.  usage: python tools/prof_mmq.py [type] [m] [k] [n] [reps]"""
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
k = int(sys.argv[3]) if len(sys.argv) > 3 else 4096
n = int(sys.argv[4]) if len(sys.argv) > 4 else 512
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 4
t = qref.WEIGHT_TYPES[name]
rng = np.random.default_rng(0)
w = torch.from_numpy(qref.random_blocks(t, m, k, rng)).cuda()
plane = qmm.build_plane(t, w, m, k)
x = torch.rand((n, k), device="cuda") * 2 - 1
out = torch.empty((n, m), device="cuda")
qmm.mul_mat_mmq(t, plane, x, m, k, out=out)
torch.cuda.synchronize()
s = torch.cuda.Stream()
g = torch.cuda.CUDAGraph()
with torch.cuda.stream(s):
    with torch.cuda.graph(g, stream=s):
        for _ in range(reps):
            qmm.mul_mat_mmq(t, plane, x, m, k, out=out)
g.replay()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    g.replay()
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) * 1e3 / (5 * reps)
print("%s m=%d k=%d n=%d BN=%s: %.1f us per matmul (graph replay, incl. activation quantizer)  %.1f TOP/s" % (
    name, m, k, n, os.environ.get("B200_MMQ_BN", "auto"), us, 2.0 * m * n * k / us / 1e6))
