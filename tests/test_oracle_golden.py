"""The oracle (oracle/ggml_qref.c) against the committed golden vectors produced by the UNMODIFIED reference build
(tests/golden/make_golden.py).  Bit-exact for quantize / dequantize; float-summation-order tolerance for dot products."""
import numpy as np
import pytest

import qref

TYPES = list(qref.WEIGHT_TYPES.items())


@pytest.mark.parametrize("qt,key", [(qref.Q8_1, "act_q8_1"), (qref.Q8_0, "act_q8_0"), (qref.Q8_K, "act_q8_K")])
def test_activation_quantizers_bit_exact(golden, qt, key):
    x = golden["act_x"]
    got = qref.quantize(qt, x)
    want = golden[key]
    if qt == qref.Q8_K:
        # reference leaves bsums of all-zero super-blocks unwritten: compare d/qs always, bsums where d != 0
        g = got.reshape(-1, 292)
        w = want.reshape(-1, 292)
        live = w[:, :4].view(np.float32).ravel() != 0
        assert np.array_equal(g[:, :260], w[:, :260])
        assert np.array_equal(g[live], w[live])
    else:
        assert np.array_equal(got, want)


@pytest.mark.parametrize("name,t", TYPES)
def test_dequantize_bit_exact(golden, name, t):
    w = golden["w_" + name]
    want = golden["deq_" + name]
    got = qref.dequantize(t, w, want.shape[1])
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("name,t", TYPES)
def test_mul_mat_cpu_semantics(golden, name, t):
    w = golden["w_" + name]
    x = golden["mm_x"]
    want = golden["mm_cpu_" + name]
    m, k = w.shape[0], x.shape[1]
    got = qref.mul_mat(t, w, x, m, k, "cpu")
    # same integers, different float summation order (and SIMD from_float rounds ties to even) => tiny differences
    assert qref.nmse(want, got) < 1e-9
    # and both are within the harness tolerance of the exact product (tests/test-backend-ops.cpp:1990)
    exact = qref.mul_mat(t, w, x, m, k, "f64")
    assert qref.nmse(exact, got) < 5e-4
    assert qref.nmse(exact, qref.mul_mat(t, w, x, m, k, "q8_1")) < 5e-4


def test_fp16_roundtrip_all_halves():
    lib = qref.lib()
    h = np.arange(65536, dtype=np.uint16)
    f = h.view(np.float16).astype(np.float32)
    for v in list(range(0, 65536, 97)) + [0, 1, 0x3ff, 0x400, 0x7bff, 0x7c00, 0x8000, 0xfbff, 0xfc00]:
        got = lib.qref_fp16_to_fp32(v)
        if np.isnan(f[v]):
            assert np.isnan(got)
        else:
            assert got == f[v]
            assert lib.qref_fp32_to_fp16(float(f[v])) == v


def test_fp32_to_fp16_matches_numpy_rne():
    rng = np.random.default_rng(7)
    x = np.concatenate([rng.normal(0, 1, 4000), rng.normal(0, 1e-6, 2000), rng.uniform(-7e4, 7e4, 2000),
                        [65504.0, 65519.99, 65520.0, 2.0 ** -24, 2.0 ** -25, 2.0 ** -25 * 1.0000001, 5.96e-8, 0.0, -0.0]]
                       ).astype(np.float32)
    with np.errstate(over="ignore"):
        want = x.astype(np.float16).view(np.uint16)
    lib = qref.lib()
    got = np.array([lib.qref_fp32_to_fp16(float(v)) for v in x], dtype=np.uint16)
    assert np.array_equal(got, want)


def test_mul_mat_id_matches_loop_of_mul_mat():
    rng = np.random.default_rng(3)
    m, k, n_expert, n_used, n_tokens = 16, 256, 4, 2, 5
    t = qref.Q4_K
    w = qref.random_blocks(t, n_expert * m, k, rng)
    for ne11 in (1, n_used):
        b = rng.uniform(-1, 1, (n_tokens, ne11, k)).astype(np.float32)
        ids = np.stack([rng.permutation(n_expert)[:n_used] for _ in range(n_tokens)]).astype(np.int32)
        got = qref.mul_mat_id(t, w, b, ids, m, k, n_expert, "cpu")
        for tok in range(n_tokens):
            for s in range(n_used):
                e = ids[tok, s]
                want = qref.mul_mat(t, w[e * m:(e + 1) * m], b[tok, s % ne11][None], m, k, "cpu")[0]
                assert np.array_equal(got[tok, s], want)
