"""CPU study (numpy, no GPU): NMSE of the tensor-core path's arithmetic against ggml-cpu semantics for the activation
quantization schemes under consideration, on quantizer-produced weights of all five types and on activations with
outlier channels.  TEST INFRASTRUCTURE: uses the oracle (tests/qref.py) and the reference build (tests/refggml.py).

  python tools/nmse_study.py > profiles/r02/mmq_nmse_schemes_cpu_study.log
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import qref      # noqa: E402
import refggml   # noqa: E402

f32 = np.float32


def q8_chunks(a, kc):
    c = a.reshape(a.shape[0], -1, kc)
    s = (np.abs(c).max(-1) / f32(127.0)).astype(f32)
    inv = np.where(s > 0, f32(1.0) / np.where(s > 0, s, f32(1.0)), f32(0.0)).astype(f32)
    q = np.rint((c * inv[:, :, None]).astype(f32)).astype(np.int32)
    return q, s


def plane_matmul(Wf, Xf, kc, wplanes, xplanes):
    """sum over (weight plane, activation plane) pairs of per-chunk integer dots scaled by s_w * t_x; the residual x
    residual cross term is dropped when both sides have two planes"""
    m, kp = Wf.shape
    n = Xf.shape[0]
    P, S = [], []
    r = Wf
    for _ in range(wplanes):
        p, s = q8_chunks(r, kc)
        P.append(p), S.append(s)
        r = (r - (s[:, :, None] * p.astype(f32)).reshape(m, kp)).astype(f32)
    Q, T = [], []
    r = Xf
    for _ in range(xplanes):
        q, t = q8_chunks(r, kc)
        Q.append(q), T.append(t)
        r = (r - (t[:, :, None] * q.astype(f32)).reshape(n, kp)).astype(f32)
    out = np.zeros((n, m), dtype=np.float64)
    for a in range(wplanes):
        for b in range(xplanes):
            if a and b:
                continue
            for c in range(kp // kc):
                D = Q[b][:, c, :].astype(np.int64) @ P[a][:, c, :].astype(np.int64).T
                out += D * (S[a][None, :, c].astype(np.float64) * T[b][:, None, c])
    return out.astype(f32)


def main():
    rng = np.random.default_rng(11)
    m, k, n = 256, 4096, 24
    wf = rng.normal(0, 1, (m, k)).astype(f32)
    print("m=%d k=%d n=%d; weights N(0,1) through ggml_quantize_chunk; NMSE vs ggml-cpu semantics (oracle) | vs f64" % (m, k, n))
    for dist in ("normal", "outliers x30 (6 ch)", "outliers x100 (6 ch)", "outliers x300 (3 ch)", "heavy-tail t(3)"):
        x = rng.normal(0, 1, (n, k)).astype(f32)
        if dist.startswith("outliers"):
            mag = float(dist.split("x")[1].split()[0])
            nch = int(dist.split("(")[1].split()[0])
            x[:, rng.choice(k, nch, replace=False)] *= mag
        elif dist.startswith("heavy"):
            x = rng.standard_t(3, (n, k)).astype(f32)
        print("--- activations:", dist)
        for name, t in qref.WEIGHT_TYPES.items():
            w = refggml.quantize_weights(t, wf)
            Wd = qref.dequantize(t, w, k)
            cpu = qref.mul_mat(t, w, x, m, k, "cpu")
            q81 = qref.mul_mat(t, w, x, m, k, "q8_1")
            f64 = qref.mul_mat(t, w, x, m, k, "f64")
            row = ["%-7s cpu-vs-f64 %.2e  q8_1-vs-cpu %.2e |" % (name, qref.nmse(f64, cpu), qref.nmse(cpu, q81))]
            wp = 2 if t == qref.Q8_0 else 1
            for kc in (256, 512):
                for xp in (1, 2):
                    got = plane_matmul(Wd, x, kc, wp, xp)
                    row.append("kc%d/x%d %.2e|%.2e" % (kc, xp, qref.nmse(cpu, got), qref.nmse(f64, got)))
            print("  ".join(row), flush=True)


if __name__ == "__main__":
    main()
