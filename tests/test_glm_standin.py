"""Pins oracle/glm_standin/glm/glm.hpp - the seven glm symbols the reference rasterizer uses, restated because glm is an
un-vendored submodule of the reference - against numpy in float64: column-major storage and constructor, mat3 * mat3,
mat3 * vec3, scalar * mat3, transpose, dot, length, max. The reference CUDA library under oracle/_ref (the golden
implementation of the parity tests) is compiled against this header, so "bit-exact with the reference" means: with the
reference's sources over these glm semantics (FMA contraction inside dot / the products is nvcc's, as it would be with the
real glm, whose operators are the same expressions)."""
import os
import shutil
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="nvcc compiles the stand-in's host path")
def test_glm_standin_matches_numpy(tmp_path):
    exe = tmp_path / "pin"
    subprocess.check_call(["nvcc", "-O1", "-std=c++17", "-I", os.path.join(ROOT, "oracle", "glm_standin"), "-o", str(exe),
                           os.path.join(ROOT, "tests", "glm_pin", "pin.cu")])
    rng = np.random.default_rng(0)
    n = 200
    a, b = rng.normal(size=(n, 9)).astype(np.float32), rng.normal(size=(n, 9)).astype(np.float32)
    v, w = rng.normal(size=(n, 3)).astype(np.float32), rng.normal(size=(n, 3)).astype(np.float32)
    text = f"{n}\n" + "\n".join(" ".join(repr(float(x)) for x in np.concatenate([a[i], b[i], v[i], w[i]])) for i in range(n))
    out = subprocess.run([str(exe)], input=text, capture_output=True, text=True, check=True).stdout
    got = np.array([[float(x) for x in line.split()] for line in out.strip().splitlines()])
    assert got.shape == (n, 9 * 3 + 3 + 2 + 3)
    for i in range(n):
        A = a[i].astype(np.float64).reshape(3, 3).T          # the constructor's scalars are columns: A[:, j] = a[3j : 3j+3]
        B = b[i].astype(np.float64).reshape(3, 3).T
        V, W = v[i].astype(np.float64), w[i].astype(np.float64)
        cols = lambda M: M.T.reshape(-1)                      # noqa: E731  printed column by column
        want = np.concatenate([cols(A @ B), cols(A.T), cols(2.5 * A), A @ V, [V @ W, np.sqrt(V @ V)], np.maximum(V, 0.0)])
        scale = np.maximum(np.abs(want), 1.0)
        assert np.all(np.abs(got[i] - want) <= 4e-6 * scale), (i, np.abs(got[i] - want).max())
    # the product is NOT the transposed convention: a row-major reading of the same scalars must disagree
    A0, B0 = a[0].astype(np.float64).reshape(3, 3), b[0].astype(np.float64).reshape(3, 3)
    assert not np.allclose(got[0][:9], (A0 @ B0).T.reshape(-1), atol=1e-3)
