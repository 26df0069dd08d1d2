"""The C-ABI library loads without a GPU and exports every symbol include/gs_b200.h declares; argument errors are
reported through return codes before any CUDA work."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "gs_b200.h")


def declared_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gs_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_boundary():
    names = declared_functions()
    for required in ("gs_forward_preprocess", "gs_forward_render", "gs_backward", "gs_mark_visible",
                     "gs_compute_relocation", "gs_add_noise"):
        assert required in names


def test_library_exports_every_declared_symbol():
    from diff_gaussian_rasterization import _native
    lib = ctypes.CDLL(_native._LIB_PATH)
    for name in declared_functions():
        assert hasattr(lib, name), f"{name} declared in gs_b200.h but not exported"
    # and the binding lists a prototype for each of them
    assert set(declared_functions()) == set(_native.SIGNATURES)


def test_sizes_and_errors_without_gpu():
    from diff_gaussian_rasterization import _native
    lib = _native._lib
    assert lib.gs_abi_version() >= 100
    assert lib.gs_geom_bytes(0) > 0
    g1, g2 = lib.gs_geom_bytes(1000), lib.gs_geom_bytes(2000)
    assert g2 > g1 >= 1000 * 73
    assert lib.gs_image_bytes(1920, 1080) >= 1920 * 1080 * 8 + 8160 * 12
    assert lib.gs_binning_bytes(1000000, 1920, 1080, 50000) >= 1000000 * 8  # u32 pass-1 keys + u32 sorted ids
    assert lib.gs_backward_scratch_bytes(1000) >= 1000 * 36
    assert lib.gs_error_string(0) == b"ok"
    assert b"invalid argument" in lib.gs_error_string(-1)
    assert lib.gs_stage_count() >= 9 and lib.gs_stage_name(0) == b"preprocess_fwd"
    # bad arguments are rejected with GS_ERR_ARG before touching the device
    rc = lib.gs_forward_preprocess(None, 10, 3, 16, None, None, None, None, None, 1.0, None, None, None, None, None,
                                   0, 0, 1.0, 1.0, 0, None, None, None, None, None)
    assert rc == -1
    rc = lib.gs_compute_relocation(None, 5, None, None, None, None, 51, None, None)
    assert rc == -1
    assert lib.gs_mark_visible(None, -1, None, None, None, None) == -1
    # P == 0 is a no-op success everywhere
    assert lib.gs_add_noise(None, 0, None, None, None, None, None, 1.0, 0, 0) == 0
    assert lib.gs_compute_relocation(None, 0, None, None, None, None, 51, None, None) == 0


def test_missing_library_fails_loudly(tmp_path):
    from diff_gaussian_rasterization import _native
    with pytest.raises(ImportError, match="no CPU fallback|There is no CPU fallback"):
        _native.load_library(str(tmp_path / "lib3dgs_mcmc_b200.so"))
