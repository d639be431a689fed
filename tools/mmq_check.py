"""Correctness + timing probe of the tensor-core path.  usage: python tools/mmq_check.py [quick]"""
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
qref.PLANE_KC = qmm.lib().b200_mmq_chunk()
rng = np.random.default_rng(0)
quick = len(sys.argv) > 1 and sys.argv[1] == "quick"
for name, m, k, n in [("q4_K", 128, 256, 16), ("q4_K", 128, 512, 128), ("q4_K", 256, 1024, 256), ("q8_0", 300, 1024, 33),
                      ("q6_K", 512, 4096, 512)]:
    t = qref.WEIGHT_TYPES[name]
    w = qref.random_blocks(t, m, k, rng)
    x = rng.uniform(-1, 1, (n, k)).astype(np.float32)
    plane = qmm.build_plane(t, torch.from_numpy(w).cuda(), m, k)
    got = qmm.mul_mat_mmq(t, plane, torch.from_numpy(x).cuda(), m, k)
    torch.cuda.synchronize()
    got = got.cpu().numpy()
    want = qref.mul_mat_plane(t, w, x, m, k)
    cpu = qref.mul_mat(t, w, x, m, k, "cpu")
    print("%-5s m=%4d k=%5d n=%4d  nmse vs plane-oracle %.3e  vs ggml-cpu %.3e  maxabs %.4g (ref max %.4g)" % (
        name, m, k, n, qref.nmse(want, got), qref.nmse(cpu, got), np.abs(want - got).max(), np.abs(want).max()), flush=True)
    if qref.nmse(want, got) > 1e-6:
        bad = np.argwhere(np.abs(want - got) > 1e-3 * np.abs(want).max())
        print("   mismatches: %d of %d; first (col j, row i): %s" % (len(bad), got.size, bad[:8].tolist()))
        print("   got[0,:8]", got[0, :8], "\n   want[0,:8]", want[0, :8])
if quick:
    sys.exit(0)
for name in ("q4_K", "q6_K"):
    t = qref.WEIGHT_TYPES[name]
    for m, k in [(4096, 4096), (14336, 4096), (4096, 14336), (1024, 4096)]:
        w = torch.from_numpy(qref.random_blocks(t, m, k, rng)).cuda()
        plane = qmm.build_plane(t, w, m, k)
        for n in (512, 128, 32):
            x = torch.rand((n, k), device="cuda") * 2 - 1
            out = torch.empty((n, m), device="cuda")
            for bn in ("128", "256"):
                os.environ["B200_MMQ_BN"] = bn
                qmm.reload_env()
                for _ in range(3):
                    qmm.mul_mat_mmq(t, plane, x, m, k, out=out)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(20):
                    qmm.mul_mat_mmq(t, plane, x, m, k, out=out)
                e1.record()
                torch.cuda.synchronize()
                us = e0.elapsed_time(e1) * 1e3 / 20
                print("%-5s m=%5d k=%5d n=%3d BN=%s: %8.1f us  %7.1f TOP/s" % (name, m, k, n, bn, us, 2.0 * m * n * k / us / 1e6),
                      flush=True)
            os.environ.pop("B200_MMQ_BN", None)
            qmm.reload_env()
