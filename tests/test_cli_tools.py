"""CPU-side checks of the CLI fixture flow: the generator reproduces the reference's
matrix_export.py stream for any block size, and its files read back the way read_matrix loads them."""
import os
import struct
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "examples"))


def test_generator_matches_reference_stream_for_any_block_size(tmp_path):
    import matrix_export as mx
    np.random.seed(5)
    mx.write_rand(str(tmp_path / "a.bin"), 37, 29, block_bytes=4 * 29 * 5)      # 5 rows per block
    np.random.seed(5)
    ref = np.random.rand(37, 29).astype(np.float32)
    raw = open(tmp_path / "a.bin", "rb").read()
    assert raw == struct.pack("ii", 37, 29) + ref.tobytes()


def test_generator_cli_c2_shapes(tmp_path):
    subprocess.check_call([sys.executable, os.path.join(ROOT, "examples", "matrix_export.py"), "--config", "c2",
                           "--cols", "300", "--out", str(tmp_path)])
    for name, shape in (("X", (513, 300)), ("W", (513, 40)), ("H", (40, 300))):
        raw = open(tmp_path / f"{name}.bin", "rb").read()
        assert struct.unpack("ii", raw[:8]) == shape and len(raw) == 8 + 4 * shape[0] * shape[1]
    # same stream as the GPU tests' C2 inputs (numpy legacy seed 1, X then W then H)
    rng = np.random.RandomState(1)
    X = rng.rand(513, 300).astype(np.float32)
    assert open(tmp_path / "X.bin", "rb").read()[8:] == X.tobytes()


def test_generator_planted(tmp_path):
    subprocess.check_call([sys.executable, os.path.join(ROOT, "examples", "matrix_export.py"), "--config", "c3",
                           "--rows", "65", "--cols", "200", "--rank", "8", "--out", str(tmp_path)])
    raw = open(tmp_path / "X.bin", "rb").read()
    X = np.frombuffer(raw[8:], dtype=np.float32).reshape((65, 200), order="F")
    assert struct.unpack("ii", raw[:8]) == (65, 200) and (X > 0).all()
    s = np.linalg.svd(X.astype(np.float64), compute_uv=False)
    assert s[8] / s[0] < 0.05                                     # rank 8 plus 5 % multiplicative noise
