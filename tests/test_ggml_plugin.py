"""The drop-in boundary seen from the reference's side: libggml-b200.so is a loadable ggml backend (ggml_backend_init /
ggml_backend_score, ggml/src/ggml-backend-impl.h:215-251) and the UNMODIFIED reference harness test-backend-ops
(tests/test-backend-ops.cpp:1975-2141 test_mul_mat / test_mul_mat_id, NMSE <= 5e-4 against ggml-cpu) passes against it.

test-backend-ops and libggml-base/-cpu are compiled from the reference sources by oracle/Makefile.ref into oracle/_ref/
(test infrastructure; they travel to the GPU box as prebuilt binaries).  If they are absent the tests are skipped -- the
plug-in can only be built where the reference's ggml headers are."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PLUGIN = os.path.join(ROOT, "llama.cpp.dsp_b200", "lib", "libggml-b200.so")
TBO = os.path.join(ROOT, "oracle", "_ref", "test-backend-ops")

needs_plugin = pytest.mark.skipif(not (os.path.exists(PLUGIN) and os.path.exists(TBO)),
                                  reason="plug-in or oracle/_ref/test-backend-ops not built (needs the reference headers)")


def _run(args, timeout=900):
    env = dict(os.environ, GGML_BACKEND_PATH=PLUGIN)
    r = subprocess.run([TBO] + args, env=env, capture_output=True, text=True, timeout=timeout)
    return r.returncode, re.sub(r"\x1b\[[0-9;]*m", "", r.stdout + r.stderr)


@needs_plugin
def test_plugin_exports_loader_entry_points():
    # symbols only: dlopen needs libggml-base's symbols in the process first, exactly as the registry has them
    C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libggml-base.so"), mode=C.RTLD_GLOBAL)
    L = C.CDLL(PLUGIN)
    for s in ("ggml_backend_init", "ggml_backend_score", "ggml_backend_b200_reg", "ggml_backend_b200_init",
              "ggml_backend_b200_buffer_type", "ggml_backend_b200_get_device_count", "ggml_backend_is_b200"):
        assert hasattr(L, s), s


@needs_plugin
def test_plugin_declines_to_load_without_a_b200():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    rc, out = _run(["test", "-o", "MUL_MAT"], timeout=120)
    assert rc == 0
    assert "libggml-b200.so is not supported on this system" in out      # score 0 -> the registry skips it, no fallback
    assert "B200_0" not in out


@pytest.mark.gpu
@needs_plugin
@pytest.mark.parametrize("op", ["MUL_MAT", "MUL_MAT_ID"])
def test_reference_test_backend_ops_passes(op):
    rc, out = _run(["test", "-o", op, "-b", "B200_0"])
    assert rc == 0, out[-4000:]
    assert "loaded B200 backend" in out
    assert "Backend 2/2: B200_0" in out or "B200_0" in out
    ran = [l for l in out.splitlines() if l.strip().startswith(op + "(") and l.rstrip().endswith("OK")]
    failed = [l for l in out.splitlines() if "FAIL" in l or "NMSE" in l]
    assert not failed, "\n".join(failed[:20])
    # every case of the five on-path types with f32 activations must have RUN (not "not supported")
    for t in ("q4_0", "q8_0", "q4_K", "q6_K", "iq4_xs"):
        mine = [l for l in ran if f"type_a={t},type_b=f32" in l]
        assert len(mine) >= (10 if op == "MUL_MAT" else 2), (t, len(mine))
    m = re.search(r"(\d+)/(\d+) tests passed", out)
    assert m and m.group(1) == m.group(2), out[-2000:]


SPLIT_TEST = os.path.join(ROOT, "oracle", "_ref", "ggml-split-test")
GRAPH_BENCH = os.path.join(ROOT, "oracle", "_ref", "ggml-graph-bench")


@pytest.mark.gpu
@needs_plugin
def test_split_buffer_type_row_split_is_bit_identical():
    """--split-mode row through the public ggml API (oracle/ggml_split_test.cpp): the registry answers
    "ggml_backend_split_buffer_type" (ggml-cuda.cu:3418-3422), a weight allocated in that buffer type is scattered over
    the devices by set_tensor and gathered by get_tensor, and MUL_MAT on it equals MUL_MAT on the unsplit weight bit for
    bit for all five types, n = 1 .. 512, uneven tensor_split, either device as the main one.  Needs two B200s in one
    process: skipped on a one-GPU box (exit code 77)."""
    if not os.path.exists(SPLIT_TEST):
        pytest.skip("oracle/_ref/ggml-split-test not built")
    r = subprocess.run([SPLIT_TEST], env=dict(os.environ, GGML_BACKEND_PATH=PLUGIN), capture_output=True, text=True, timeout=900)
    if r.returncode == 77:
        pytest.skip(r.stdout.strip().splitlines()[-1])
    assert r.returncode == 0, (r.stdout + r.stderr)[-4000:]
    assert "PASS" in r.stdout


@pytest.mark.gpu
@needs_plugin
def test_registry_answers_split_buffer_type_and_host_buffer_type():
    """one-GPU part of the same contract: the proc address exists, the device offers a pinned host buffer type, and a
    small decode token runs through tensor_set_async / graph_compute_async / tensor_get_async (oracle/ggml_graph_bench.cpp)"""
    if not os.path.exists(GRAPH_BENCH):
        pytest.skip("oracle/_ref/ggml-graph-bench not built")
    import json
    for extra in ([], ["--sync"]):
        r = subprocess.run([GRAPH_BENCH, "--backend", "B200_0", "--steps", "2", "--warmup", "1", "--layers", "2", "--vocab", "4096"]
                           + extra, env=dict(os.environ, GGML_BACKEND_PATH=PLUGIN), capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, (r.stdout + r.stderr)[-2000:]
        line = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
        assert line["backend"] == "B200_0" and line["host_buffer"] == "B200_Host" and line["nodes"] == 15
    C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libggml-base.so"), mode=C.RTLD_GLOBAL)
    L = C.CDLL(PLUGIN)
    L.ggml_backend_b200_reg.restype = C.c_void_p
    assert L.ggml_backend_b200_reg()
