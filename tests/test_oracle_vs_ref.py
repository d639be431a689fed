"""The oracle against the reference itself (oracle/_ref, built from /root/reference by oracle/Makefile.ref) on fresh
seeded inputs, including random bit patterns that a quantizer would never emit.  Skipped where the ref build is absent."""
import numpy as np
import pytest

import qref
import refggml as R

pytestmark = pytest.mark.skipif(not R.available(), reason="oracle/_ref not built (make -f oracle/Makefile.ref)")

TYPES = list(qref.WEIGHT_TYPES.items())


@pytest.mark.parametrize("seed", [0, 1])
def test_q8_1_q8_0_q8_K_bit_exact(seed):
    rng = np.random.default_rng(seed)
    x = (rng.normal(0, 1, (64, 1024)) * rng.choice([1e-3, 1.0, 50.0], (64, 1))).astype(np.float32)
    x[3, :256] = 0
    for qt in (qref.Q8_1, qref.Q8_0, qref.Q8_K):
        got, want = qref.quantize(qt, x), R.quantize_act(qt, x)
        if qt == qref.Q8_K:
            g, w = got.reshape(-1, 292), want.reshape(-1, 292)
            live = w[:, :4].view(np.float32).ravel() != 0
            assert np.array_equal(g[live], w[live])
        else:
            assert np.array_equal(got, want)


@pytest.mark.parametrize("name,t", TYPES)
def test_dequantize_random_bit_patterns(name, t):
    rng = np.random.default_rng(11)
    k = 1024
    w = qref.random_blocks(t, 32, k, rng)
    got, want = qref.dequantize(t, w, k), R.dequantize(t, w, k)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("name,t", TYPES)
def test_mul_mat_against_ggml_cpu_kernels(name, t):
    rng = np.random.default_rng(5)
    m, k, n = 12, 1024, 2
    w = qref.random_blocks(t, m, k, rng)
    x = rng.uniform(-1, 1, (n, k)).astype(np.float32)
    want = R.mul_mat_cpu(t, w, x, m, k)
    got = qref.mul_mat(t, w, x, m, k, "cpu")
    assert qref.nmse(want, got) < 1e-9


def test_q4_0_weight_quantizer_bit_exact():
    rng = np.random.default_rng(2)
    x = rng.normal(0, 0.05, (16, 512)).astype(np.float32)
    assert np.array_equal(qref.quantize(qref.Q4_0, x), R.quantize_weights(R.Q4_0, x))
