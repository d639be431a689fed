"""GPU parity tests proper: the CUDA path, called through the C ABI (ctypes), against the oracle on the same seeded
inputs, against the committed golden vectors, and -- at BASELINE sizes -- through size-independent properties.
Bars: bit-exact for quantize / dequantize; NMSE <= 5e-4 vs the ggml-cpu semantics (tests/test-backend-ops.cpp:1990)
plus a tight bound vs the q8_1-semantics oracle (catches bit-decoding bugs the loose bound would hide)."""
import numpy as np
import pytest

import qref

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

NMSE_HARNESS = 5e-4   # max_nmse_err for MUL_MAT / MUL_MAT_ID in the reference harness
NMSE_TIGHT = 1e-9     # fp32-vs-double accumulation noise only


@pytest.fixture(scope="module")
def qmm():
    import b200pkg
    q = b200pkg.load()
    assert torch.cuda.is_available(), "these tests need the B200"
    assert q.lib().b200_device_ok(torch.cuda.current_device()) == 1
    qref.PLANE_KC = q.lib().b200_mmq_chunk()
    return q


@pytest.fixture(autouse=True)
def _knobs_reread_after_each_test(qmm):
    """the library caches B200_* knobs; monkeypatch restores the environment at teardown, so have them read again"""
    yield
    qmm.reload_env()


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


TYPES = list(qref.WEIGHT_TYPES.items())


# ------------------------------------------------------------------------------------------------ quantize_q8_1
def test_quantize_q8_1_golden_bit_exact(qmm, golden):
    x = golden["act_x"]
    got = qmm.quantize_row_q8_1(dev(x)).cpu().numpy()
    assert np.array_equal(got, golden["act_q8_1"])


@pytest.mark.parametrize("k", [32, 64, 96, 4096, 4096 + 32, 14336])
def test_quantize_q8_1_vs_oracle(qmm, k):
    rng = np.random.default_rng(k)
    nrows = 5
    x = (rng.normal(0, 1, (nrows, k)) * rng.choice([1e-4, 1.0, 300.0], (nrows, 1))).astype(np.float32)
    x[1, :32] = 0
    got = qmm.quantize_row_q8_1(dev(x)).cpu().numpy()
    assert np.array_equal(got, qref.quantize(qref.Q8_1, x))


def test_quantize_q8_1_strided_unaligned_rows(qmm):
    rng = np.random.default_rng(1)
    big = rng.normal(0, 1, (7, 200)).astype(np.float32)
    xb = dev(big)
    view = xb[:, 3:3 + 160]                       # row stride 200, base offset 12 bytes: scalar-load path
    got = qmm.quantize_row_q8_1(view).cpu().numpy()
    assert np.array_equal(got, qref.quantize(qref.Q8_1, big[:, 3:163]))


def test_quantize_empty(qmm):
    x = torch.empty((0, 64), dtype=torch.float32, device="cuda")
    assert qmm.quantize_row_q8_1(x).shape == (0, 72)


# ------------------------------------------------------------------------------------------------ dequantize
@pytest.mark.parametrize("name,t", TYPES)
def test_dequantize_golden_bit_exact(qmm, golden, name, t):
    w = golden["w_" + name]
    want = golden["deq_" + name]
    got = qmm.dequantize_row(t, dev(w), want.shape[1]).cpu().numpy()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("name,t", TYPES)
def test_dequantize_random_bit_patterns(qmm, name, t):
    rng = np.random.default_rng(17)
    k = 1024
    w = qref.random_blocks(t, 40, k, rng)
    got = qmm.dequantize_row(t, dev(w), k).cpu().numpy()
    want = qref.dequantize(t, w, k)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


# ------------------------------------------------------------------------------------------------ MUL_MAT
def check_mul_mat(qmm, t, w, x, m, k):
    got = qmm.mul_mat(t, dev(w), dev(x), m, k)
    torch.cuda.synchronize()
    got = got.cpu().numpy()
    assert np.isfinite(got).all()
    n = x.reshape(-1, k).shape[0]
    # n <= 8: dp4a GEMV with q8_1 activations; n > 8: int8 tensor-core path with per-256 scales (qref.mul_mat_plane)
    tight = qref.mul_mat(t, w, x, m, k, "q8_1") if n < qmm.lib().b200_mmq_min_batch() else qref.mul_mat_plane(t, w, x, m, k)
    cpu = qref.mul_mat(t, w, x, m, k, "cpu")
    assert qref.nmse(tight, got) < NMSE_TIGHT, ("tight", qref.nmse(tight, got))
    assert qref.nmse(cpu, got) < NMSE_HARNESS, ("harness", qref.nmse(cpu, got))
    return got


@pytest.mark.parametrize("name,t", TYPES)
def test_mul_mat_golden(qmm, golden, name, t):
    w, x = golden["w_" + name], golden["mm_x"]
    got = check_mul_mat(qmm, t, w, x, w.shape[0], x.shape[1])
    assert qref.nmse(golden["mm_cpu_" + name], got) < NMSE_HARNESS


@pytest.mark.parametrize("name,t", TYPES)
@pytest.mark.parametrize("n", [1, 2, 3, 4, 5, 8, 9, 17])
def test_mul_mat_small_harness_shapes(qmm, name, t, n):
    # the reference harness' eval shapes: m=16, k=256, n=1..9 (tests/test-backend-ops.cpp:4143-4147)
    rng = np.random.default_rng(100 + n)
    m, k = 16, 256
    w = qref.random_blocks(t, m, k, rng)
    x = rng.uniform(-1, 1, (n, k)).astype(np.float32)
    check_mul_mat(qmm, t, w, x, m, k)


@pytest.mark.parametrize("name,t", TYPES)
@pytest.mark.parametrize("m,k", [(1, 256), (33, 512), (257, 1024 + 256), (1000, 4096), (300, 14336)])
def test_mul_mat_ragged_rows(qmm, name, t, m, k):
    rng = np.random.default_rng(m * 7 + k)
    w = qref.random_blocks(t, m, k, rng)
    for n in (1, 3):
        x = rng.uniform(-1, 1, (n, k)).astype(np.float32)
        check_mul_mat(qmm, t, w, x, m, k)


@pytest.mark.parametrize("name,t", [("q4_0", qref.Q4_0), ("q8_0", qref.Q8_0)])
def test_mul_mat_k_of_one_block(qmm, name, t):
    rng = np.random.default_rng(5)
    for k in (32, 96):
        w = qref.random_blocks(t, 16, k, rng)
        x = rng.uniform(-1, 1, (1, k)).astype(np.float32)
        check_mul_mat(qmm, t, w, x, 16, k)


def test_mul_mat_empty_and_bad_args(qmm):
    w = torch.zeros(144 * 4, dtype=torch.uint8, device="cuda")
    x = torch.zeros((0, 256), dtype=torch.float32, device="cuda")
    assert qmm.mul_mat(qref.Q4_K, w, x, 4, 256).shape == (0, 4)
    with pytest.raises(qmm.B200Error):
        qmm.mul_mat(7, w, torch.zeros((1, 256), device="cuda"), 4, 256)            # Q5_1: not on the path
    with pytest.raises(qmm.B200Error):
        qmm.mul_mat(qref.Q4_K, w, torch.zeros((1, 128), device="cuda"), 4, 128)    # k % 256 != 0
    with pytest.raises(qmm.B200Error):
        qmm.mul_mat(qref.Q4_K, w, torch.zeros((1, 256)), 4, 256)                   # host tensor: no CPU fallback


@pytest.mark.parametrize("name,t", TYPES)
def test_mul_mat_realistic_quantized_weights_vs_exact(qmm, name, t):
    # weights that went through a real quantizer (golden) replicated to a larger matrix; vs the exact f64 product
    rng = np.random.default_rng(9)
    m, k, n = 512, 4096, 2
    w = qref.random_blocks(t, m, k, rng)
    x = rng.normal(0, 1, (n, k)).astype(np.float32)
    got = qmm.mul_mat(t, dev(w), dev(x), m, k).cpu().numpy()
    exact = qref.mul_mat(t, w, x, m, k, "f64")
    assert qref.nmse(exact, got) < NMSE_HARNESS


# ------------------------------------------------------------------------------------------------ MMQ (tcgen05 int8)
@pytest.mark.parametrize("name,t", TYPES)
@pytest.mark.parametrize("m,k,n", [(128, 256, 16), (16, 256, 9), (300, 1024, 33), (257, 2048, 129), (512, 4096, 512),
                                   (1000, 512, 300)])
def test_mmq_prepared_plane(qmm, name, t, m, k, n):
    """tensor-core path against a prepared plane: tight vs the numpy restatement of its arithmetic (catches descriptor /
    swizzle / pipeline bugs), harness bound vs ggml-cpu semantics and vs the exact f64 product"""
    rng = np.random.default_rng(m + k + n)
    w = qref.random_blocks(t, m, k, rng)
    x = rng.uniform(-1, 1, (n, k)).astype(np.float32)
    wd = dev(w)
    plane = qmm.build_plane(t, wd, m, k)
    got = qmm.mul_mat_mmq(t, plane, dev(x), m, k).cpu().numpy()
    assert np.isfinite(got).all()
    assert qref.nmse(qref.mul_mat_plane(t, w, x, m, k), got) < NMSE_TIGHT
    assert qref.nmse(qref.mul_mat(t, w, x, m, k, "cpu"), got) < NMSE_HARNESS
    assert qref.nmse(qref.mul_mat(t, w, x, m, k, "f64"), got) < NMSE_HARNESS


@pytest.mark.parametrize("name,t", TYPES)
@pytest.mark.parametrize("m,k,n", [(512, 4096, 512), (300, 2048, 200), (128, 1280, 260)])
def test_mmq_split_k_is_exact_and_deterministic(qmm, monkeypatch, name, t, m, k, n):
    """few-tile shapes run as two K halves added with red.global.add onto a zeroed dst (B200_MMQ_KSPLIT=2 forces it;
    m = 300 / an odd chunk count exercise the fall-backs; Q4_0 / Q8_0 split their longer virtual k).  Same tight bound as the unsplit path vs the
    plane restatement, and bit-identical from run to run (two addends onto zero commute)."""
    rng = np.random.default_rng(m + k)
    w = qref.random_blocks(t, m, k, rng)
    x = rng.uniform(-1, 1, (n, k)).astype(np.float32)
    plane = qmm.build_plane(t, dev(w), m, k)
    xd = dev(x)
    monkeypatch.setenv("B200_MMQ_KSPLIT", "2")
    qmm.reload_env()
    out = torch.full((n, m), 7.0, device="cuda")           # stale contents must not leak into the sum
    got = qmm.mul_mat_mmq(t, plane, xd, m, k, out=out).clone()
    for _ in range(3):
        out.fill_(-3.0)
        assert torch.equal(qmm.mul_mat_mmq(t, plane, xd, m, k, out=out), got)
    assert qref.nmse(qref.mul_mat_plane(t, w, x, m, k), got.cpu().numpy()) < NMSE_TIGHT
    monkeypatch.setenv("B200_MMQ_KSPLIT", "1")
    qmm.reload_env()
    one = qmm.mul_mat_mmq(t, plane, xd, m, k)
    assert qref.nmse(one.cpu().numpy(), got.cpu().numpy()) < 1e-11


@pytest.mark.parametrize("name,t", [("q4_0", qref.Q4_0), ("q8_0", qref.Q8_0)])
@pytest.mark.parametrize("m,k,n", [(512, 4096, 512), (300, 1024, 33), (4096, 1024, 200)])
def test_mmq_residual_planes_one_pass(qmm, name, t, m, k, n):
    """the 32-block types run W0.X0 + W0.X1 (+ W1.X0 for Q8_0's residual weight plane) as ONE pass over a longer virtual k:
    tight vs the restatement, bit-identical from run to run, stale dst contents must not leak (the few-tile shapes split
    the virtual k in two and add with red.global.add onto a dst the quantizer zeroed)."""
    rng = np.random.default_rng(m + n)
    w = qref.random_blocks(t, m, k, rng)
    x = rng.uniform(-1, 1, (n, k)).astype(np.float32)
    plane = qmm.build_plane(t, dev(w), m, k)
    xd = dev(x)
    out = torch.full((n, m), 9.0, device="cuda")
    got = qmm.mul_mat_mmq(t, plane, xd, m, k, out=out).clone()
    out.fill_(-1.0)
    assert torch.equal(qmm.mul_mat_mmq(t, plane, xd, m, k, out=out), got)
    assert qref.nmse(qref.mul_mat_plane(t, w, x, m, k), got.cpu().numpy()) < NMSE_TIGHT


@pytest.mark.parametrize("name,t", [("q4_K", qref.Q4_K), ("q8_0", qref.Q8_0), ("q6_K", qref.Q6_K)])
@pytest.mark.parametrize("m,k,n", [(9700, 512, 300), (19000, 256, 512), (2500, 768, 1000)])
def test_mmq_persistent_schedule(qmm, monkeypatch, name, t, m, k, n):
    """more tiles than SMs: the persistent kernel walks several items per CTA, the short last round is cut into
    128 x 128 halves (B200_MMQ_TAIL_HALVES=0 keeps whole tiles), n and m are ragged, Q8_0 walks three segments.
    Both schedules and the one-CTA-per-tile kernel (B200_MMQ_PERSIST=0) must give the same bits."""
    rng = np.random.default_rng(m + n)
    w = qref.random_blocks(t, m, k, rng)
    x = rng.uniform(-1, 1, (n, k)).astype(np.float32)
    plane = qmm.build_plane(t, dev(w), m, k)
    xd = dev(x)
    got = qmm.mul_mat_mmq(t, plane, xd, m, k).clone()
    assert qref.nmse(qref.mul_mat_plane(t, w, x, m, k), got.cpu().numpy()) < NMSE_TIGHT
    monkeypatch.setenv("B200_MMQ_TAIL_HALVES", "0")
    qmm.reload_env()
    assert torch.equal(qmm.mul_mat_mmq(t, plane, xd, m, k), got)
    monkeypatch.setenv("B200_MMQ_PERSIST", "0")
    qmm.reload_env()
    assert torch.equal(qmm.mul_mat_mmq(t, plane, xd, m, k), got)


@pytest.mark.parametrize("name,t", [("q4_0", qref.Q4_0), ("q8_0", qref.Q8_0)])
def test_mmq_k_not_multiple_of_chunk(qmm, name, t):
    rng = np.random.default_rng(3)
    m, k, n = 64, 96 + 256, 20
    w = qref.random_blocks(t, m, k, rng)
    x = rng.uniform(-1, 1, (n, k)).astype(np.float32)
    check_mul_mat(qmm, t, w, x, m, k)


def _quantizer_weights(t, wf):
    """weights that went through the reference's own quantizer: ggml_quantize_chunk of the reference build (oracle/_ref,
    travels to the GPU box) for every type; without it the oracle's pinned restatement covers Q4_0 / Q8_0"""
    import refggml
    if refggml.available():
        return refggml.quantize_weights(t, wf)
    if t in (qref.Q4_0, qref.Q8_0):
        return qref.quantize(t, wf)
    pytest.skip("oracle/_ref (reference build) not present: no K-quant quantizer")


@pytest.mark.parametrize("name,t", TYPES)
@pytest.mark.parametrize("dist", ["normal", "heavy_tail", "act_outliers", "act_outliers_x100"])
def test_mmq_realistic_weights_through_quantizer(qmm, name, t, dist):
    """Weights that went through the reference quantizer, from Gaussian and from heavy-tailed (Student-t, df = 4) data,
    and activations with OUTLIER CHANNELS (six channels scaled 30x / 100x, as LLM activations have).  The harness bound
    (5e-4 vs ggml-cpu semantics) must hold on all of them for all five types at the default build: the K-quants /
    IQ4_XS because the activation scale granularity is ggml-cpu's own q8_K (256), Q4_0 / Q8_0 (ggml-cpu: q8_0, per 32)
    through the residual activation plane.  (A single plane per 256 measured 7e-4 .. 1e-3 here, per 512 1.3e-3 .. 2.2e-3:
    profiles/r02/mmq_nmse_schemes_cpu_study.log.)"""
    rng = np.random.default_rng(11)
    m, k, n = 256, 2048, 24
    wf = (rng.standard_t(4, (m, k)) if dist == "heavy_tail" else rng.normal(0, 1, (m, k))).astype(np.float32)
    w = _quantizer_weights(t, wf)
    x = rng.normal(0, 1, (n, k)).astype(np.float32)
    if dist.startswith("act_outliers"):
        x[:, rng.choice(k, 6, replace=False)] *= 100 if dist.endswith("x100") else 30
    got = qmm.mul_mat(t, dev(w), dev(x), m, k).cpu().numpy()
    assert qref.nmse(qref.mul_mat_plane(t, w, x, m, k), got) < NMSE_TIGHT
    cpu = qref.mul_mat(t, w, x, m, k, "cpu")
    assert qref.nmse(cpu, got) < NMSE_HARNESS, qref.nmse(cpu, got)


@pytest.mark.parametrize("name,t", TYPES)
def test_mmq_vs_exact_dequant_product(qmm, name, t):
    """reference-anchored check that does not go through the restatement of our own arithmetic: the tensor-core result
    against (bit-exact dequantize_row_* of quantizer-produced weights) x f64 activations.  2e-4: the int8 plane's own
    rounding measured 4.4e-5 (two-plane types: < 1e-8) plus one activation scale per 256 measured 5e-5 on Gaussian data
    (ggml-cpu itself is at 2.9e-5 / 4.9e-5 against the same product)."""
    rng = np.random.default_rng(5)
    m, k, n = 384, 4096, 40
    w = _quantizer_weights(t, rng.normal(0, 1, (m, k)).astype(np.float32))
    x = rng.normal(0, 1, (n, k)).astype(np.float32)
    got = qmm.mul_mat(t, dev(w), dev(x), m, k).cpu().numpy()
    exact = qref.mul_mat(t, w, x, m, k, "f64")
    assert qref.nmse(exact, got) < 2e-4, qref.nmse(exact, got)


# ------------------------------------------------------------------------------------------------ full-size properties
@pytest.mark.parametrize("name,t", TYPES)
def test_full_size_properties(qmm, name, t):
    """BASELINE size (m=4096, k=14336; tests/test-backend-ops.cpp:4514-4520): too slow for the scalar oracle on every
    row, so check (a) a sampled set of rows against the oracle, (b) exact power-of-two homogeneity, (c) zero input,
    (d) an independent device path: bit-exact dequantize + fp64 matmul."""
    rng = np.random.default_rng(42)
    m, k = 4096, 14336
    w = qref.random_blocks(t, m, k, rng)
    x = rng.uniform(-1, 1, (1, k)).astype(np.float32)
    wd, xd = dev(w), dev(x)
    got = qmm.mul_mat(t, wd, xd, m, k)
    rows = rng.choice(m, 64, replace=False)
    want = qref.mul_mat(t, w[rows], x, len(rows), k, "q8_1")
    assert qref.nmse(want, got.cpu().numpy()[:, rows]) < NMSE_TIGHT
    twice = qmm.mul_mat(t, wd, xd * 2.0, m, k)
    assert torch.equal(twice, got * 2.0)
    zero = qmm.mul_mat(t, wd, torch.zeros_like(xd), m, k)
    assert torch.count_nonzero(zero) == 0
    deq = qmm.dequantize_row(t, wd, k).double()
    ref = (deq @ xd.double().T).T.float()
    assert qref.nmse(ref.cpu().numpy(), got.cpu().numpy()) < NMSE_HARNESS


# ------------------------------------------------------------------------------------------------ MUL_MAT_ID
@pytest.mark.parametrize("name,t", TYPES)
@pytest.mark.parametrize("n_expert,n_used,n_tokens,bcast", [(4, 1, 1, False), (4, 2, 5, True), (8, 4, 3, False),
                                                            (8, 2, 32, True), (4, 4, 129, False)])
def test_mul_mat_id(qmm, name, t, n_expert, n_used, n_tokens, bcast):
    # shapes of the harness' MUL_MAT_ID cases (tests/test-backend-ops.cpp:4240-4270), smaller m
    rng = np.random.default_rng(n_expert * 100 + n_used * 10 + n_tokens)
    m, k = 64, 256
    ne11 = 1 if bcast else n_used
    w = qref.random_blocks(t, n_expert * m, k, rng)
    b = rng.uniform(-1, 1, (n_tokens, ne11, k)).astype(np.float32)
    # ids is a VIEW of a wider tensor when n_used < n_expert (test-backend-ops.cpp:2105-2110): row stride n_expert
    ids_full = np.stack([rng.permutation(n_expert) for _ in range(n_tokens)]).astype(np.int32)
    ids_dev = dev(ids_full)[:, :n_used]
    got = qmm.mul_mat_id(t, dev(w), dev(b), ids_dev, m, k, n_expert).cpu().numpy()
    ids = ids_full[:, :n_used]
    tight = qref.mul_mat_id(t, w, b, ids, m, k, n_expert, "q8_1")
    cpu = qref.mul_mat_id(t, w, b, ids, m, k, n_expert, "cpu")
    assert qref.nmse(tight, got) < NMSE_TIGHT
    assert qref.nmse(cpu, got) < NMSE_HARNESS


@pytest.mark.parametrize("name,t", TYPES)
@pytest.mark.parametrize("n_expert,n_used,n_tokens,bcast", [(8, 2, 8, True), (8, 2, 33, False), (4, 4, 7, False),
                                                            (16, 1, 5, True), (8, 2, 1, True)])
def test_mul_mat_id_grouped_by_expert(qmm, monkeypatch, name, t, n_expert, n_used, n_tokens, bcast):
    """batched decode: pairs that chose the same expert share one pass over its weights (4 per work item, planned on the
    device).  Forced on (B200_MMID_GROUP=2) and off (=0): both must meet the oracle bounds, agree with each other to fp32
    summation order, and the grouped result must not depend on the atomic order of the plan (bit-identical re-runs).
    ids contain repeated experts across tokens and, with a skewed router, one hot expert."""
    rng = np.random.default_rng(n_expert * 1000 + n_used * 10 + n_tokens)
    m, k = 72, 512
    ne11 = 1 if bcast else n_used
    w = qref.random_blocks(t, n_expert * m, k, rng)
    b = rng.uniform(-1, 1, (n_tokens, ne11, k)).astype(np.float32)
    ids_full = np.stack([rng.permutation(n_expert) for _ in range(n_tokens)]).astype(np.int32)
    ids_full[: n_tokens // 2, 0] = 1                              # a hot expert (rows stay permutations where it matters not)
    for r in range(n_tokens // 2):                                # keep each row's first n_used ids distinct
        row = [1] + [e for e in ids_full[r, 1:] if e != 1]
        ids_full[r, :len(row)] = row
    ids_dev = dev(ids_full)[:, :n_used]
    ids = ids_full[:, :n_used]
    wd, bd = dev(w), dev(b)
    monkeypatch.setenv("B200_MMID_GROUP", "2")
    qmm.reload_env()
    out = torch.full((n_tokens, n_used, m), 5.0, device="cuda")
    got = qmm.mul_mat_id(t, wd, bd, ids_dev, m, k, n_expert, out=out).clone()
    for _ in range(3):
        assert torch.equal(qmm.mul_mat_id(t, wd, bd, ids_dev, m, k, n_expert), got)
    monkeypatch.setenv("B200_MMID_GROUP", "0")
    qmm.reload_env()
    plain = qmm.mul_mat_id(t, wd, bd, ids_dev, m, k, n_expert)
    g = got.cpu().numpy()
    assert qref.nmse(qref.mul_mat_id(t, w, b, ids, m, k, n_expert, "q8_1"), g) < NMSE_TIGHT
    assert qref.nmse(qref.mul_mat_id(t, w, b, ids, m, k, n_expert, "cpu"), g) < NMSE_HARNESS
    assert qref.nmse(plain.cpu().numpy(), g) < 1e-11


def test_launch_counter_counts_kernels(qmm):
    before = qmm.launch_count()
    w = torch.zeros(144 * 16, dtype=torch.uint8, device="cuda")
    qmm.mul_mat(qref.Q4_K, w, torch.ones((1, 256), device="cuda"), 16, 256)
    # one fused launch (quantizer inside the GEMV) or two (separate q8_1 quantizer chained by programmatic launch)
    assert qmm.launch_count() - before in (1, 2)


# ------------------------------------------------------------------------------------------------ several matrices, one x
@pytest.mark.parametrize("n", [1, 2, 4, 5])
def test_mul_mat_multi_equals_separate_calls(qmm, n):
    """b200_mul_mat_multi (q/k/v, gate/up sharing activations): x quantized once, same-type matrices in one launch.
    Mixed types, a row count that is not a multiple of 8 (falls out of the shared launch), IQ4_XS (own activation
    layout) and more than four matrices of one type are all in the list; every result must equal b200_mul_mat's up to
    fp32 summation order (rows per stage, hence the in-warp reduction tree, depend on the launch's total row count)."""
    rng = np.random.default_rng(21)
    k = 4096
    spec = [(12, 4096), (12, 1024), (14, 1024), (12, 4096), (12, 1004), (23, 512), (12, 512), (12, 256), (12, 128),
            (12, 64), (12, 8), (8, 2048), (2, 96)]
    ws = [dev(qref.random_blocks(t, m, k, rng)) for t, m in spec]
    x = dev(rng.uniform(-1, 1, (n, k)).astype(np.float32))
    outs = qmm.mul_mat_multi([t for t, _ in spec], ws, x, [m for _, m in spec], k)
    for (t, m), w, got in zip(spec, ws, outs):
        want = qmm.mul_mat(t, w, x, m, k)
        assert qref.nmse(want.cpu().numpy(), got.cpu().numpy()) < 1e-11, (t, m)
        oracle = qref.mul_mat(t, w.cpu().numpy(), x.cpu().numpy(), m, k, "q8_1")
        assert qref.nmse(oracle, got.cpu().numpy()) < NMSE_TIGHT, (t, m)


def test_mul_mat_multi_large_batch_falls_back_to_tensor_cores(qmm):
    rng = np.random.default_rng(22)
    k, n = 1024, 33
    spec = [(12, 256), (12, 384)]
    ws = [dev(qref.random_blocks(t, m, k, rng)) for t, m in spec]
    x = dev(rng.uniform(-1, 1, (n, k)).astype(np.float32))
    outs = qmm.mul_mat_multi([t for t, _ in spec], ws, x, [m for _, m in spec], k)
    for (t, m), w, got in zip(spec, ws, outs):
        assert torch.equal(got, qmm.mul_mat(t, w, x, m, k))


def test_mul_mat_multi_launches_fewer_kernels(qmm):
    rng = np.random.default_rng(23)
    k = 4096
    spec = [(12, 4096), (12, 1024), (12, 1024)]
    ws = [dev(qref.random_blocks(t, m, k, rng)) for t, m in spec]
    x = dev(rng.uniform(-1, 1, (1, k)).astype(np.float32))
    before = qmm.launch_count()
    got = qmm.mul_mat_multi([t for t, _ in spec], ws, x, [m for _, m in spec], k)
    assert qmm.launch_count() - before == 1          # ONE GEMV for three matrices; its CTAs quantize x themselves (k <= 8192)
    qmm.setenv(B200_FUSED_QUANT_MAX_K=0)
    try:
        before = qmm.launch_count()
        sep = qmm.mul_mat_multi([t for t, _ in spec], ws, x, [m for _, m in spec], k)
        assert qmm.launch_count() - before == 2      # the separate quantizer kernel + one GEMV
    finally:
        qmm.setenv(B200_FUSED_QUANT_MAX_K=None)
    for a, b in zip(got, sep):
        assert torch.equal(a, b)                     # the same quantizer arithmetic either way


def test_mmq_multi_planes_equal_single_calls(qmm):
    """b200_mul_mat_mmq_multi: several planes, one activation matrix, shared grid.  Tiles are independent, so every
    result must be BIT-identical to its own b200_mul_mat_mmq call; a Q8_0 member (two planes) takes the two-pass path
    in the middle of the list, and more than four single-plane members span two launches."""
    rng = np.random.default_rng(31)
    k, n = 2048, 200
    spec = [(12, 512), (12, 136), (14, 128), (8, 256), (23, 384), (2, 64), (12, 1024), (12, 100)]
    ws = [dev(qref.random_blocks(t, m, k, rng)) for t, m in spec]
    planes = [qmm.build_plane(t, w, m, k) for (t, m), w in zip(spec, ws)]
    x = dev(rng.uniform(-1, 1, (n, k)).astype(np.float32))
    outs = [torch.full((n, m), float("nan"), device="cuda") for _, m in spec]
    qmm.MultiMatMulMMQ([t for t, _ in spec], planes, [m for _, m in spec], k, outs, n)(x)
    for (t, m), pl, got in zip(spec, planes, outs):
        want = qmm.mul_mat_mmq(t, pl, x, m, k)
        assert torch.equal(got, want), (t, m)


def test_mul_mat_multi_host_matches_device_path(qmm):
    """b200_mul_mat_multi_host: host activations in, host results out (adjacent results in one copy, a separate one for
    the buffer that is not adjacent)."""
    rng = np.random.default_rng(41)
    k, n = 4096, 2
    spec = [(12, 512), (12, 256), (14, 128), (8, 64)]
    ws = [dev(qref.random_blocks(t, m, k, rng)) for t, m in spec]
    xh = torch.from_numpy(rng.uniform(-1, 1, (n, k)).astype(np.float32)).pin_memory()
    block = torch.full((n * (512 + 256 + 128),), float("nan")).pin_memory()
    outs, off = [], 0
    for _, m in spec[:3]:
        outs.append(block[off:off + n * m].view(n, m))
        off += n * m
    outs.append(torch.full((n, 64), float("nan")).pin_memory())
    qmm.MultiMatMulHost([t for t, _ in spec], ws, [m for _, m in spec], k, outs, n)(xh)
    torch.cuda.synchronize()
    want = qmm.mul_mat_multi([t for t, _ in spec], ws, xh.cuda(), [m for _, m in spec], k)
    for got, w_ in zip(outs, want):
        assert torch.equal(got, w_.cpu())


@pytest.mark.parametrize("pinned", [True, False])
def test_host_activations_zero_copy_equals_staged_copy(qmm, monkeypatch, pinned):
    """pinned host activations are read by the quantizer kernel over PCIe (no staging copy) and results go to pinned
    host memory through a copy kernel; pageable buffers, and B200_HOST_ZEROCOPY=0, take cudaMemcpyAsync.  Same bits either way, for the single and the grouped host entry."""
    rng = np.random.default_rng(43)
    k, n = 4096, 3
    spec = [(12, 512), (14, 128)]
    ws = [dev(qref.random_blocks(t, m, k, rng)) for t, m in spec]
    xh = torch.from_numpy(rng.uniform(-1, 1, (n, k)).astype(np.float32))
    if pinned:
        xh = xh.pin_memory()
    want = qmm.mul_mat_multi([t for t, _ in spec], ws, xh.cuda(), [m for _, m in spec], k)
    for zc in ("1", "0"):
        monkeypatch.setenv("B200_HOST_ZEROCOPY", zc)
        qmm.reload_env()
        outs = [torch.full((n, m), float("nan")) for _, m in spec]
        one = torch.full((n, 512), float("nan"))
        if pinned:                                              # results: copy kernel into pinned memory / cudaMemcpyAsync
            outs, one = [o.pin_memory() for o in outs], one.pin_memory()
        qmm.MultiMatMulHost([t for t, _ in spec], ws, [m for _, m in spec], k, outs, n)(xh)
        qmm.mul_mat_host(12, ws[0], xh, one, 512, k)
        torch.cuda.synchronize()
        for got, w_ in zip(outs, want):
            assert torch.equal(got, w_.cpu())
        assert torch.equal(one, want[0].cpu())


# ------------------------------------------------------------------------------------------------ sequence (one launch)
def _seq_reference(qmm, groups):
    """the same groups through the existing entry point, one b200_mul_mat per matrix, in order"""
    for x, k, mats, _ in groups:
        for t, w, m, out in mats:
            qmm.mul_mat(t, w, x.view(1, k), m, k, out=out.view(1, m))


@pytest.mark.parametrize("seed", [0, 1])
def test_sequence_dependent_chain_bit_identical(qmm, seed):
    """b200_sequence_*: a chain in which every group's activation vector IS the previous group's result (so a missing
    or early group-to-group synchronization shows up as wrong numbers), mixed types inside a group (q/k Q4_K + v Q6_K;
    an IQ4_XS member forces the split into two layout classes), ragged m, k from 256 to 14336, more stages than ring
    slots and fewer stages than CTAs.  Must equal consecutive b200_mul_mat calls BIT FOR BIT, run after run (the plan's
    counters re-arm themselves), and leave no error mark."""
    rng = np.random.default_rng(100 + seed)
    T = qref.WEIGHT_TYPES
    # (k -> [(type, m)]) ; the first matrix of each group feeds the next group: m_first(g) == k(g+1)
    chain = [(4096, [("q4_K", 4096), ("q4_K", 1024), ("q6_K", 1024)]),
             (4096, [("q4_K", 14336), ("q4_K", 14336)]),
             (14336, [("q6_K", 2048)]),
             (2048, [("q8_0", 256), ("iq4_xs", 520), ("q4_0", 264)]),
             (256, [("q4_0", 1024), ("q8_0", 40)]),
             (1024, [("iq4_xs", 4096)]),
             (4096, [("q6_K", 4096), ("q4_K", 8)]),
             (4096, [("q4_K", 512)])]
    scale = 6.0   # keeps activations O(1) down the chain (weights are ~0.02 * N(0,1)-like)

    def build():
        groups, x = [], dev(rng.uniform(-1, 1, (chain[0][0],)).astype(np.float32))
        for k, mats in chain:
            ms = []
            for name, m in mats:
                w = dev(qref.random_blocks(T[name], m, k, rng, scale=0.02 * scale))
                ms.append((T[name], w, m, torch.full((m,), float("nan"), device="cuda")))
            groups.append((x, k, ms, True))
            x = ms[0][3]
        return groups
    state = rng.bit_generator.state
    g_seq = build()
    rng.bit_generator.state = state
    g_ref = build()
    assert torch.equal(g_seq[0][0], g_ref[0][0])
    _seq_reference(qmm, g_ref)
    torch.cuda.synchronize()
    seq = qmm.Sequence(g_seq)
    for rep in range(3):
        for _, _, mats, _ in g_seq:
            for _, _, _, out in mats:
                out.fill_(float("nan"))
        seq()
        torch.cuda.synchronize()
        assert seq.status() == 0
        for (_, _, ma, _), (_, _, mb, _) in zip(g_seq, g_ref):
            for (_, _, _, a), (_, _, _, b) in zip(ma, mb):
                assert torch.isfinite(b).all()
                assert torch.equal(a, b), (rep, a.numel(), float((a - b).abs().max()))


def test_sequence_many_independent_groups_and_limits(qmm):
    """40 groups that only share their input (dependent = 0; more groups than ring halves, alternating k), then the
    argument checks: n_mat, k beyond one CTA's reach, a plan buffer that is too small."""
    rng = np.random.default_rng(7)
    T = qref.WEIGHT_TYPES
    xs = {k: dev(rng.uniform(-1, 1, (k,)).astype(np.float32)) for k in (1024, 4096)}
    names = list(T)
    g_seq, g_ref = [], []
    for i in range(40):
        k = 1024 if i % 3 else 4096
        name, m = names[i % 5], [200, 1536, 72, 4096][i % 4]
        w = dev(qref.random_blocks(T[name], m, k, rng))
        g_seq.append((xs[k], k, [(T[name], w, m, torch.zeros(m, device="cuda"))], False))
        g_ref.append((xs[k], k, [(T[name], w, m, torch.zeros(m, device="cuda"))], False))
    _seq_reference(qmm, g_ref)
    seq = qmm.Sequence(g_seq)
    seq()
    seq()
    torch.cuda.synchronize()
    assert seq.status() == 0
    for (_, _, ma, _), (_, _, mb, _) in zip(g_seq, g_ref):
        assert torch.equal(ma[0][3], mb[0][3])
    w = dev(qref.random_blocks(qref.Q4_K, 8, 65536, rng))
    with pytest.raises(qmm.B200Error):
        qmm.Sequence([(torch.zeros(65536, device="cuda"), 65536, [(qref.Q4_K, w, 8, torch.zeros(8, device="cuda"))], True)])
    L = qmm.lib()
    assert L.b200_sequence_plan_size(0) == 0
    grp = (qmm.qmm.SeqGroup * 1)()
    assert L.b200_sequence_plan_build(grp, 1, None, 0, None) != 0


# ------------------------------------------------------------------------------------------------ MUL_MAT_ID on tensor cores
@pytest.mark.parametrize("name,t", TYPES)
@pytest.mark.parametrize("n_expert,n_used,n_tokens,bcast,m", [(8, 2, 32, True, 256), (4, 2, 129, False, 128), (8, 4, 600, True, 384),
                                                              (16, 1, 40, True, 128), (4, 4, 9, False, 256)])
def test_mul_mat_id_tensor_cores(qmm, name, t, n_expert, n_used, n_tokens, bcast, m):
    """batched MUL_MAT_ID through b200_mul_mat_id_mmq: device-side sort of the pairs by expert (ranges of 1 .. >256 pairs:
    several 256-column items per expert, narrow UMMA N for short ranges, experts nobody chose), gather in the quantizer,
    scatter in the epilogue.  Tight vs the restatement of the tensor-core arithmetic per pair, harness bound vs ggml-cpu
    semantics, bit-identical from run to run, and rows of dst that belong to no valid pair stay untouched."""
    rng = np.random.default_rng(n_expert * 1000 + n_used * 10 + n_tokens)
    k = 512
    ne11 = 1 if bcast else n_used
    w = qref.random_blocks(t, n_expert * m, k, rng)
    b = rng.uniform(-1, 1, (n_tokens, ne11, k)).astype(np.float32)
    # a skewed router (expert 0 is hot), distinct experts per token (ggml's top-k), one expert nobody chooses when >= 8
    E = n_expert - (1 if n_expert >= 8 else 0)
    probs = np.array([0.5] + [0.5 / (E - 1)] * (E - 1))
    ids_full = np.zeros((n_tokens, n_expert), dtype=np.int32)
    for r in range(n_tokens):
        ids_full[r, :n_used] = rng.choice(E, n_used, replace=False, p=probs)
    ids = np.ascontiguousarray(ids_full[:, :n_used])
    ids_dev = dev(ids_full)[:, :n_used]                          # a strided view, like the harness passes
    plane = qmm.build_plane(t, dev(w), n_expert * m, k)
    out = torch.full((n_tokens, n_used, m), 7.5, device="cuda")
    got = qmm.mul_mat_id_mmq(t, plane, dev(b), ids_dev, m, k, n_expert, out=out).clone()
    out.fill_(-2.0)
    assert torch.equal(qmm.mul_mat_id_mmq(t, plane, dev(b), ids_dev, m, k, n_expert, out=out), got)
    g = got.cpu().numpy()
    assert np.isfinite(g).all() and not (g == 7.5).any()
    want = np.zeros_like(g)
    for tok in range(n_tokens):
        for s_ in range(n_used):
            e = ids[tok, s_]
            want[tok, s_] = qref.mul_mat_plane(t, w[e * m:(e + 1) * m], b[tok, s_ % ne11][None], m, k)[0]
    assert qref.nmse(want, g) < NMSE_TIGHT
    assert qref.nmse(qref.mul_mat_id(t, w, b, ids, m, k, n_expert, "cpu"), g) < NMSE_HARNESS
    assert qmm.lib().b200_mul_mat_id_prefers_mmq(m, n_expert, n_used, n_tokens) in (0, 1)


@pytest.mark.gpu
def test_sequence_row_split_two_gpus_bit_identical():
    """b200_sequence_plan_build_rowsplit on 2 GPUs (one process per GPU, torchrun): every rank's result buffer equals the
    whole matmuls computed locally, over repeated runs of the plan (tests/seq_rowsplit_check.py).  Skipped on a 1-GPU box."""
    import os
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(root, "tests", "seq_rowsplit_check.py")]
    for ll in ("1", "0"):          # both exchange protocols: flag-in-data, and mirrored stores + system-scope counters
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300,
                           env=dict(os.environ, SEQ_SPLIT_LL=ll))
        assert r.returncode == 0, r.stdout[-3000:]
        assert r.stdout.count("OK") >= 2, r.stdout[-3000:]


@pytest.mark.gpu
def test_copy_async_kernel_and_fallback(qmm):
    """b200_copy_async / b200_copy_mapped_async: small copies between device memory and PINNED host memory run as a copy
    kernel (any size and alignment, both directions), pageable memory and large copies fall back to cudaMemcpyAsync; the
    bytes must arrive either way and the kernel path must show up in the launch counter."""
    import ctypes as C
    L = qmm.lib()
    rng = np.random.default_rng(5)
    for nbytes, off in [(16, 0), (4096, 0), (4099, 0), (70001, 1), (1 << 20, 4), ((1 << 20) + 64, 0)]:
        src_h = torch.from_numpy(rng.integers(0, 256, nbytes + 16, dtype=np.uint8)).pin_memory()
        dev_buf = torch.zeros(nbytes + 16, dtype=torch.uint8, device="cuda")
        back_h = torch.zeros(nbytes + 16, dtype=torch.uint8).pin_memory()
        b0 = qmm.launch_count()
        qmm.check(L.b200_copy_async(C.c_void_p(dev_buf.data_ptr() + off), C.c_void_p(src_h.data_ptr() + off), nbytes, None))
        qmm.check(L.b200_copy_mapped_async(C.c_void_p(back_h.data_ptr() + off), C.c_void_p(dev_buf.data_ptr() + off), nbytes, None))
        torch.cuda.synchronize()
        assert torch.equal(back_h[off:off + nbytes], src_h[off:off + nbytes])
        assert int(back_h[:off].sum()) == 0 and int(back_h[off + nbytes:].sum()) == 0      # nothing outside the range
        assert (qmm.launch_count() - b0 == 2) == (nbytes <= (1 << 20))                     # kernel path up to 1 MiB
    # pageable host memory: the generic entry falls back to the copy engine
    src_p = torch.from_numpy(rng.integers(0, 256, 5000, dtype=np.uint8))
    dev_buf = torch.zeros(5000, dtype=torch.uint8, device="cuda")
    b0 = qmm.launch_count()
    qmm.check(L.b200_copy_async(C.c_void_p(dev_buf.data_ptr()), C.c_void_p(src_p.data_ptr()), 5000, None))
    torch.cuda.synchronize()
    assert torch.equal(dev_buf.cpu(), src_p) and qmm.launch_count() == b0


@pytest.mark.gpu
def test_small_batches_on_a_cached_plane(qmm):
    """5..8 columns: a caller that holds a plane is told to use it for the single-plane types (b200_mmq_use_cached_plane)
    and the result stays inside the harness bound against ggml-cpu semantics, like the GEMV's."""
    L = qmm.lib()
    assert L.b200_mmq_use_cached_plane(qref.Q4_K, 4) == 0 and L.b200_mmq_use_cached_plane(qref.Q4_K, 5) == 1
    assert L.b200_mmq_use_cached_plane(qref.Q8_0, 8) == 0 and L.b200_mmq_use_cached_plane(qref.Q4_0, 8) == 0
    assert L.b200_mmq_use_cached_plane(qref.Q8_0, 9) == 1 and L.b200_mmq_use_cached_plane(99, 9) == 0
    rng = np.random.default_rng(11)
    m, k = 512, 2048
    for t in (qref.Q4_K, qref.Q6_K, qref.IQ4_XS):
        w = qref.random_blocks(t, m, k, rng)
        wd = dev(w)
        plane = qmm.build_plane(t, wd, m, k)
        for n in (5, 8):
            x = rng.uniform(-1, 1, (n, k)).astype(np.float32)
            want = qref.mul_mat(t, w, x, m, k)                      # ggml-cpu semantics
            got_p = qmm.mul_mat_mmq(t, plane, dev(x), m, k).cpu().numpy()
            got_g = qmm.mul_mat(t, wd, dev(x), m, k).cpu().numpy()
            assert qref.nmse(got_p, want) < 5e-4 and qref.nmse(got_g, want) < 5e-4
            assert qref.nmse(got_p, got_g) < 5e-4


@pytest.mark.gpu
def test_sequence_rows_with_an_odd_number_of_blocks(qmm):
    """k % 64 == 32 (Q4_0 / Q8_0 only): the last 64-element unit of a row has one block; the sequence kernel's in-CTA
    quantizer (rotated chunks, qs padded to whole units) must give the bits of b200_mul_mat."""
    rng = np.random.default_rng(77)
    groups, want = [], []
    for k, t, m in [(96, qref.Q4_0, 40), (1056, qref.Q8_0, 300), (4128, qref.Q4_0, 520), (96, qref.Q8_0, 8)]:
        w = dev(qref.random_blocks(t, m, k, rng))
        x = dev(rng.uniform(-1, 1, (k,)).astype(np.float32))
        out = torch.full((m,), float("nan"), device="cuda")
        groups.append((x, k, [(t, w, m, out)], True))
        want.append(qmm.mul_mat(t, w, x.view(1, k), m, k)[0])
    seq = qmm.Sequence(groups)
    seq()
    torch.cuda.synchronize()
    assert seq.status() == 0
    for (_, _, mats, _), wv in zip(groups, want):
        assert torch.equal(mats[0][3], wv)
