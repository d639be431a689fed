"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/b200_qmm.h declares,
and its host-only helpers behave (no compute calls without a GPU)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def qmm():
    import b200pkg
    q = b200pkg.load()
    if not os.path.exists(q.qmm.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return q


def test_header_symbols_all_exported(qmm):
    hdr = open(os.path.join(ROOT, "include", "b200_qmm.h")).read()
    declared = sorted(set(re.findall(r"^B200_API[^;(]*?\b(b200_[a-z0-9_]+)\s*\(", hdr, re.M)))
    assert declared, "no declarations parsed"
    L = qmm.lib()
    missing = [s for s in declared if not hasattr(L, s)]
    assert not missing, missing
    assert sorted(qmm.qmm.EXPORTS) == declared


def test_type_tables_match_ggml(qmm):
    L = qmm.lib()
    # ggml/src/ggml-common.h: 18/34/144/210/136 bytes per 32/32/256/256/256 elements
    for t, blck, size in [(2, 32, 18), (8, 32, 34), (12, 256, 144), (14, 256, 210), (23, 256, 136)]:
        assert L.b200_type_supported(t) == 1
        assert L.b200_block_size(t) == blck and L.b200_type_size(t) == size
        assert L.b200_row_size(t, 14336) == 14336 // blck * size
    for t in (0, 1, 3, 6, 7, 9, 10, 11, 13, 15, 30):
        assert L.b200_type_supported(t) == 0
    assert L.b200_row_size(12, 100) == 0
    assert L.b200_weight_alloc_size(14, 256, 1) >= 210 + 15


def test_row_split_follows_reference_rounding(qmm):
    # get_row_split (ggml/src/ggml-cuda/ggml-cuda.cu:740-753): boundaries rounded down to 128, last takes the rest
    for m in (4096, 1024, 14336, 128256, 8192, 28672, 100):
        for world in (1, 2, 4, 8):
            prev = 0
            for r in range(world):
                lo, hi = qmm.row_split(m, world, r)
                assert lo == prev and hi >= lo
                if r < world - 1:
                    assert hi % 128 == 0
                prev = hi
            assert prev == m
    assert qmm.row_split(128256, 8, 7) == (112128, 128256)
    assert qmm.row_split(1000, 2, 0, [3.0, 1.0]) == (0, 640)


def test_status_strings_and_no_device_is_loud(qmm):
    L = qmm.lib()
    assert L.b200_qmm_abi_version() == 1
    assert b"ok" == L.b200_status_string(0)
    import torch
    if not torch.cuda.is_available():
        assert L.b200_device_ok(0) == 0
        # host tensors are rejected outright: the product has no CPU path
        with pytest.raises(qmm.B200Error):
            qmm.mul_mat(12, torch.zeros(144, dtype=torch.uint8), torch.zeros((1, 256)), 1, 256)


def test_header_compiles_as_c_and_struct_layout_matches_binding(qmm, tmp_path):
    """include/b200_qmm.h is plain C (no CUDA / C++ / torch types) and the one struct that crosses the boundary,
    b200_seq_group, has the layout the ctypes binding assumes (size and field offsets from a C compiler)."""
    import shutil
    import subprocess
    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None:
        pytest.skip("no C compiler")
    src = tmp_path / "layout.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "b200_qmm.h"\n'
                   'int main(void) { printf("%zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(b200_seq_group), '
                   'offsetof(b200_seq_group, x_device), offsetof(b200_seq_group, k), offsetof(b200_seq_group, nmat), '
                   'offsetof(b200_seq_group, type), offsetof(b200_seq_group, W_device), offsetof(b200_seq_group, dst_device), '
                   'offsetof(b200_seq_group, m)); return 0; }\n')
    exe = tmp_path / "layout"
    subprocess.run([cc, "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)],
                   check=True)
    got = [int(v) for v in subprocess.run([str(exe)], check=True, stdout=subprocess.PIPE, text=True).stdout.split()]
    S = qmm.qmm.SeqGroup
    want = [C.sizeof(S), S.x.offset, S.k.offset, S.nmat.offset, S.type.offset, S.W.offset, S.dst.offset, S.m.offset]
    assert got == want, (got, want)
