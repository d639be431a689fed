"""Timeline of the GEMV launches of a few decoder layers (8B shapes, grouped calls) from in-kernel globaltimer stamps.
usage: python tools/layer_trace.py [layers]
stamps per CTA: 0 entry, 1 compute warps past the dependency wait, 2 activations in registers, 3 first weights landed,
4 team 0 done, 5 producer past the wait, 6 last result written"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import b200pkg  # noqa: E402
import bench  # noqa: E402

qmm = b200pkg.load()
layers = int(sys.argv[1]) if len(sys.argv) > 1 else 3
dev = torch.device("cuda", 0)
Q4K, Q6K = 12, 14
gen = torch.Generator(device=dev)
gen.manual_seed(1)
calls, names = [], []
xs = {k: torch.rand((1, k), device=dev) * 2 - 1 for k in (4096, 14336)}
for L in range(layers + 8):   # 8 extra layers in front so that the traced ones run in steady state, > L2 apart
    vt = Q6K if L % 2 == 0 else Q4K
    for nm, spec, k in (("qkv", [(Q4K, 4096), (Q4K, 1024), (vt, 1024)], 4096), ("o", [(Q4K, 4096)], 4096),
                        ("gate_up", [(Q4K, 14336), (Q4K, 14336)], 4096), ("down", [(vt, 4096)], 14336)):
        ws = [bench.gen_weights(torch, t, m, k, gen, dev) for t, m in spec]
        outs = [torch.zeros((1, m), device=dev) for _, m in spec]
        calls.append((qmm.MultiMatMul([t for t, _ in spec], ws, [m for _, m in spec], k, outs, 1), xs[k]))
        names.append("%s.%d" % (nm, L))


def run(first, last):
    for c, x in calls[first:last]:
        c(x)


run(0, len(calls))
torch.cuda.synchronize()
nl = 8 * layers
trace = torch.zeros((nl, 4096, 8), dtype=torch.int64, device="cuda")
s = torch.cuda.Stream()
g = torch.cuda.CUDAGraph()
with torch.cuda.stream(s):
    with torch.cuda.graph(g, stream=s):
        run(0, 8 * 4)
        qmm.lib().b200_debug_gemv_trace(trace.data_ptr(), nl)
        before = qmm.launch_count()
        run(8 * 4, len(calls))
        traced = qmm.launch_count() - before
        qmm.lib().b200_debug_gemv_trace(None, 0)
for _ in range(3):
    g.replay()
torch.cuda.synchronize()
tr = trace.cpu().numpy().astype(np.int64)
print("kernels launched in the traced part: %d (GEMVs + quantizers)" % traced)
labels = ["entry", "wait_done", "acts", "first_w", "team0_done", "prod_wait", "last_write"]
t0, prev_end = None, None
for i in range(nl):
    a = tr[i]
    a = a[a[:, 0] != 0]
    if not len(a):
        continue
    if t0 is None:
        t0 = a[:, 0].min()
    row = ["L%-2d ctas=%3d" % (i, len(a))]
    for j, lab in enumerate(labels):
        v = a[:, j][a[:, j] != 0] - t0
        if len(v):
            row.append("%s %6d/%6d/%6d" % (lab, v.min(), int(np.median(v)), v.max()))
    end = int((a[:, 6][a[:, 6] != 0] - t0).max())
    if prev_end is not None:
        row.append("| period %6d" % (end - prev_end))
    prev_end = end
    print("  ".join(row))
