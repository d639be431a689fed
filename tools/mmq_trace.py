"""Where the roles of one k_mmq launch spend their clocks (needs the -DB200_MMQ_TRACE build, B200_QMM_LIB=...trace.so).
usage: python tools/mmq_trace.py [type] [m] [k] [n]
Per CTA: producer total / wait sempty / wait empty; MMA warp total / wait tempty / wait full;
epilogue warp 2: main loop / wait sfull / wait tfull / final store."""
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
t = qref.WEIGHT_TYPES[name]
rng = np.random.default_rng(0)
w = torch.from_numpy(qref.random_blocks(t, m, k, rng)).cuda()
plane = qmm.build_plane(t, w, m, k)
x = torch.rand((n, k), device="cuda") * 2 - 1
out = torch.empty((n, m), device="cuda")
for _ in range(3):
    qmm.mul_mat_mmq(t, plane, x, m, k, out=out)
torch.cuda.synchronize()
trace = torch.zeros((4096, 16), dtype=torch.int64, device="cuda")
os.environ["B200_MMQ_TRACE_PTR"] = hex(trace.data_ptr())
qmm.mul_mat_mmq(t, plane, x, m, k, out=out)
torch.cuda.synchronize()
tr = trace.cpu().numpy()
tr = tr[tr[:, 0] != 0]
nch = (k + 255) // 256
print("%s m=%d k=%d n=%d BN=%s: %d CTAs, %d chunks each" % (name, m, k, n, os.environ.get("B200_MMQ_BN", "auto"), len(tr), nch))


def col(i):
    v = tr[:, i]
    return "%7d/%7d/%7d" % (v.min(), int(np.median(v)), v.max())


print("clocks, min/median/max over CTAs")
print("producer : total %s  wait sempty %s  wait empty %s" % (col(0), col(1), col(2)))
print("MMA warp : total %s  wait tempty %s  wait full  %s" % (col(4), col(5), col(6)))
print("epilogue : loop  %s  wait sfull  %s  wait tfull %s  store %s" % (col(8), col(9), col(10), col(11)))
print("per chunk (median): producer %.0f  mma %.0f (waits %.0f + %.0f)  epilogue %.0f (waits %.0f + %.0f)" % (
    np.median(tr[:, 0]) / nch, np.median(tr[:, 4]) / nch, np.median(tr[:, 5]) / nch, np.median(tr[:, 6]) / nch,
    np.median(tr[:, 8]) / nch, np.median(tr[:, 9]) / nch, np.median(tr[:, 10]) / nch))
g = tr[:, 12]
print("CTA end times (globaltimer): spread %d ns" % (g.max() - g.min()))
