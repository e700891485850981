"""CPU-side checks of the drop-in boundary: libnmf_b200.so builds, loads without a GPU, exports
every symbol include/nmf.h declares, keeps the reference's file format, and fails loudly (no CPU
fallback) when asked to compute without a device."""
import ctypes
import os
import re
import struct

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def api():
    import nmf_gpu_b200.api as a
    a.lib()
    return a


def declared_functions():
    src = open(os.path.join(ROOT, "include", "nmf.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b([a-z_][a-z0-9_]*)\s*\([^;{}]*\)\s*;", src)))


def test_header_declares_reference_api():
    names = declared_functions()
    for n in ("update_div", "read_matrix", "write_matrix"):      # README.md:31-54 of the reference
        assert n in names
    assert len(names) >= 20


def test_library_exports_every_declared_symbol(api):
    L = ctypes.CDLL(api.lib_path())
    for n in declared_functions():
        assert hasattr(L, n), n
    assert set(declared_functions()) == set(api._SIGNATURES)       # the Python mirror binds all of them


def test_no_oracle_on_the_product_path():
    """the product never links, loads or names the CPU oracle"""
    pkg = os.path.join(ROOT, "nmf_gpu_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cpp", ".hpp", ".cuh", ".inc")):
                text = open(os.path.join(d, f)).read()
                assert "nmf_oracle" not in text and "import oracle" not in text and "from oracle" not in text, f
    import subprocess
    deps = subprocess.run(["ldd", os.path.join(pkg, "libnmf_b200.so")], capture_output=True, text=True).stdout
    assert "oracle" not in deps


def test_matrix_struct_layout(api):
    assert ctypes.sizeof(api.matrix) == 24 and api.matrix.dim.offset == 16
    assert ctypes.sizeof(api.nmf_options) == 64


def test_file_format_roundtrip(api, tmp_path):
    """int32 rows, int32 cols, float32 column-major payload (README.md:33-34, cuda/nmf.cu:188-259)"""
    a = np.asfortranarray(np.random.RandomState(0).rand(7, 5).astype(np.float32))
    a[2, 3] = 0.0
    p = str(tmp_path / "a.bin")
    api.write_matrix(a, p)
    raw = open(p, "rb").read()
    assert struct.unpack("ii", raw[:8]) == (7, 5)
    assert np.array_equal(np.frombuffer(raw[8:], np.float32), a.reshape(-1, order="F"))
    b = api.read_matrix(p)
    assert b.shape == (7, 5) and b[2, 3] == np.float32(2.2204e-16)   # clamp on load, nmf.cu:210-211
    a[2, 3] = b[2, 3]
    assert np.array_equal(a, b)


def test_reads_matrix_export_layout(api, tmp_path, c1):
    """matrix_export.py writes C-order bytes; read_matrix interprets them column-major"""
    np.random.seed(0)
    X = np.random.rand(4096, 350).astype(np.float32)
    p = str(tmp_path / "X.bin")
    with open(p, "wb") as f:
        f.write(struct.pack("ii", *X.shape)); f.write(X.tobytes())
    assert np.array_equal(api.read_matrix(p), np.maximum(c1["X"], np.float32(2.2204e-16)))


def test_read_missing_file_raises(api, tmp_path):
    with pytest.raises(api.NmfError):
        api.read_matrix(str(tmp_path / "nope.bin"))


def test_compute_without_gpu_fails_loudly(api):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    W = np.asfortranarray(np.ones((8, 2), np.float32)); H = np.asfortranarray(np.ones((2, 6), np.float32))
    with pytest.raises(api.NmfError):
        api.update_div(W, H, np.ones((8, 6), np.float32), 0.0, 1)
    with pytest.raises(api.NmfError):
        api.Session(8, 6, 2)


def test_update_div_rejects_non_fortran(api):
    W = np.ones((8, 2), np.float32)            # C-ordered 2-D array is not column-major
    H = np.asfortranarray(np.ones((2, 6), np.float32))
    with pytest.raises(TypeError):
        api.update_div(W, H, np.ones((8, 6), np.float32), 0.0, 1)
