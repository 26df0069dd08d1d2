"""gs_l1_loss_forward_static: the L1 forward for hosts that cannot allocate device memory (include/gs_b200.h).

Who needs it: a plugin loaded into an engine that keeps the device allocator private, a frame loop where cudaMalloc's
device-wide synchronisation is not acceptable, a process whose GPU memory was budgeted once at start-up. Such a caller
has nowhere to get the workspace gs_l1_loss_forward asks for, so this variant keeps its per-block partial sums in a
16 KB array inside the library image and takes only inputs and the output.

The GPU test drives the entry point through ctypes, exactly as such a caller would, and holds it to the result of
gs_l1_loss_forward on the same inputs (tolerance 2e-6 absolute, the value tolerance of tests/test_loss.py). It makes
two calls back to back, the second on a smaller, different input, so that the second call finds the first call's
partial sums still in the shared array and must not be affected by them.

This file is named to be collected last: with `pytest -x` a failure here cannot hide another test.
"""
import ctypes

import pytest
import torch

import helpers

VAL_TOL = 2e-6          # same as tests/test_loss.py
MAX_N = 8388608         # GS_L1_STATIC_MAX_N in include/gs_b200.h


def _lib():
    from diff_gaussian_rasterization import _native
    return _native._lib


def test_static_l1_rejects_bad_arguments_without_touching_the_device():
    """Argument and size checks come before any CUDA call, so they can be exercised without a GPU."""
    lib = _lib()
    one = ctypes.c_void_p(16)   # never dereferenced: every call below is refused first
    assert lib.gs_l1_loss_forward_static(None, MAX_N + 1, one, one, one) == -3
    assert b"fixed capacity" in lib.gs_error_string(-3)
    assert lib.gs_l1_loss_forward_static(None, 0, one, one, one) == -1
    assert lib.gs_l1_loss_forward_static(None, 16, None, one, one) == -1
    assert lib.gs_l1_loss_forward_static(None, 16, one, one, None) == -1
    assert "GS_L1_STATIC_MAX_N ((size_t)%d)" % MAX_N in open(helpers.ROOT + "/include/gs_b200.h").read()


@pytest.mark.gpu
def test_static_l1_equals_the_workspace_entry_point_on_two_calls_in_a_row():
    """Two consecutive calls on one stream; each must match gs_l1_loss_forward on the same tensors."""
    lib = _lib()
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(20261106)
    # first a full 1080p image (1519 blocks, most of the array), then a small ragged one (2 blocks + scalar tail)
    shapes = [(3, 1080, 1920), (3, 37, 53)]
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream(dev).cuda_stream
        cases = []
        for shape in shapes:
            gt = torch.rand(shape, generator=g)
            img = (gt + 0.1 * torch.randn(shape, generator=g)).clamp(0, 1)
            cases.append((img.to(dev).contiguous(), gt.to(dev).contiguous()))

        # reference values from the existing entry point (caller-allocated workspace)
        want = []
        for img, gt in cases:
            n = img.numel()
            ws = torch.empty(lib.gs_l1_workspace_bytes(n), dtype=torch.uint8, device=dev)
            out = torch.full((2,), -1.0, device=dev)
            assert lib.gs_l1_loss_forward(stream, n, img.data_ptr(), gt.data_ptr(), ws.data_ptr(), out.data_ptr()) == 0
            want.append(out)

        # the two static calls, back to back, no synchronisation in between; outputs are allocated beforehand
        got = [torch.full((2,), -1.0, device=dev) for _ in cases]
        for (img, gt), out in zip(cases, got):
            rc = lib.gs_l1_loss_forward_static(stream, img.numel(), img.data_ptr(), gt.data_ptr(), out.data_ptr())
            assert rc == 0, lib.gs_error_string(rc)
        torch.cuda.synchronize(dev)

    for (img, gt), w, o in zip(cases, want, got):
        ref = torch.abs(img - gt).mean().item()
        assert abs(o[0].item() - w[0].item()) < VAL_TOL
        assert abs(o[0].item() - ref) < VAL_TOL
        assert o[0].item() == w[0].item()      # same launch shape and summation order: identical bits
        assert o[1].item() == 0.0
