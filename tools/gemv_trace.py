"""Timeline of back-to-back GEMV launches from in-kernel globaltimer stamps (B200_GEMV_TRACE_PTR).
usage: python tools/gemv_trace.py [type] [m] [k] [launches]
stamps per CTA: 0 entry, 1 compute warps past griddepcontrol.wait, 2 activations in registers, 3 first weights landed,
4 team 0 done, 5 producer past the wait, 6 last result written"""
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
launches = int(sys.argv[4]) if len(sys.argv) > 4 else 12
t = qref.WEIGHT_TYPES[name]
rng = np.random.default_rng(0)
one = torch.from_numpy(qref.random_blocks(t, m, k, rng)).cuda()
ws = [one] + [one.clone() for _ in range(launches - 1)]
x = torch.rand((1, k), device="cuda") * 2 - 1
outs = [torch.empty((1, m), device="cuda") for _ in ws]


def run():
    for w, o in zip(ws, outs):
        qmm.mul_mat(t, w, x, m, k, out=o)


run()
torch.cuda.synchronize()
trace = torch.zeros((launches, 4096, 8), dtype=torch.int64, device="cuda")
qmm.lib().b200_debug_gemv_trace(trace.data_ptr(), launches)
s = torch.cuda.Stream()
g = torch.cuda.CUDAGraph()
with torch.cuda.stream(s):
    with torch.cuda.graph(g, stream=s):
        run()
qmm.lib().b200_debug_gemv_trace(None, 0)
for _ in range(3):
    g.replay()
torch.cuda.synchronize()
tr = trace.cpu().numpy().astype(np.int64)
t0 = None
print("%s m=%d k=%d: per launch, ns relative to the first CTA entry of launch 0 (min / median / max over CTAs)" % (name, m, k))
labels = ["entry", "wait_done", "acts_ready", "first_w", "team0_done", "prod_wait", "last_write"]
for i in range(launches):
    a = tr[i]
    used = a[:, 0] != 0
    if not used.any():
        continue
    a = a[used]
    if t0 is None:
        t0 = a[:, 0].min()
    row = ["L%-2d ctas=%3d" % (i, len(a))]
    for j, lab in enumerate(labels):
        v = a[:, j][a[:, j] != 0] - t0
        if len(v):
            row.append("%s %6d/%6d/%6d" % (lab, v.min(), int(np.median(v)), v.max()))
    print("  ".join(row))
