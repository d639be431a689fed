"""Generates tests/golden/*.npz from the UNMODIFIED reference build (oracle/_ref, made by oracle/Makefile.ref from
/root/reference).  Run here (the container that has /root/reference); the .npz files are committed so that the GPU box,
which has no reference tree, can still check the oracle and the CUDA path against the reference's own outputs.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import refggml as R  # noqa: E402

SEED = 20261104


def act_cases(rng):
    """activation rows that exercise the quantize_q8_1 corner cases (k = 32 each block)"""
    rows = []
    rows.append(rng.uniform(-1, 1, 512))                       # the harness distribution (test-backend-ops.cpp:39-61)
    rows.append(rng.normal(0, 3, 512))
    rows.append(0.1 + 2 * np.cos(np.arange(512) + 0.0))        # test-quantize-fns.cpp:31-35 synthetic data
    z = np.zeros(512); rows.append(z)                          # all-zero blocks: d = 0, id = 0
    e = np.zeros(512); e[::32] = 1e-30; e[5::32] = -1e-38; rows.append(e)   # tiny / denormal d
    b = rng.uniform(-1, 1, 512) * 1e30; rows.append(b)         # huge: d overflows fp16 -> inf
    t = np.zeros(512)                                          # exact .5 ties after scaling: amax = 127 => id = 1
    t[0::32] = 127.0; t[1::32] = 0.5; t[2::32] = -0.5; t[3::32] = 1.5; t[4::32] = -2.5; t[5::32] = 126.5
    rows.append(t)
    n = -np.abs(rng.normal(0, 1, 512)); n[7] = -0.0; rows.append(n)   # all negative, a negative zero
    s = rng.normal(0, 1, 512); s[::2] = 0; rows.append(s)
    o = rng.normal(0, 0.01, 512); o[100] = 37.5; o[300] = -81.25; rows.append(o)   # outliers
    return np.asarray(rows, dtype=np.float32)


def main():
    assert R.available(), "build oracle/_ref first: make -f oracle/Makefile.ref"
    rng = np.random.default_rng(SEED)
    out = {}
    x = act_cases(rng)
    out["act_x"] = x
    out["act_q8_1"] = R.quantize_act(R.Q8_1, x)
    out["act_q8_0"] = R.quantize_act(R.Q8_0, x)
    out["act_q8_K"] = R.quantize_act(R.Q8_K, x)
    m, k, n = 24, 512, 3
    xs = rng.uniform(-1, 1, (n, k)).astype(np.float32)
    out["mm_x"] = xs
    for t, name in R.NAMES.items():
        wf = rng.normal(0, 0.05, (m, k)).astype(np.float32)
        wf[3] *= 40.0                     # a loud row
        wf[5, :256] = 0.0                 # a zero super-block
        w = R.quantize_weights(t, wf)
        out["w_" + name] = w
        out["deq_" + name] = R.dequantize(t, w, k)
        out["mm_cpu_" + name] = R.mul_mat_cpu(t, w, xs, m, k)
    np.savez_compressed(os.path.join(HERE, "ggml_ref_vectors.npz"), **out)
    sz = os.path.getsize(os.path.join(HERE, "ggml_ref_vectors.npz"))
    print("wrote ggml_ref_vectors.npz", sz, "bytes;", len(out), "arrays")


if __name__ == "__main__":
    main()
