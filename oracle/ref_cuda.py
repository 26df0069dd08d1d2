"""ctypes wrapper around oracle/_ref/libref_dgr.so — the UNMODIFIED reference CUDA rasterizer.

TEST INFRASTRUCTURE ONLY. The library is the reference's own cuda_rasterizer/*.cu compiled for sm_100 from
/root/reference (oracle/Makefile, oracle/ref_shim.cu); it is the golden implementation ("O1" in SURVEY.md §8c)
for bit-exact comparisons on the GPU box and the timed reference-CUDA baseline. Only tests/,
__graft_entry__.smoke() and bench.py's baseline legs may import this module; nothing under
3dgs-mcmc_b200/ does.
"""
import ctypes
import os

import torch

_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libref_dgr.so")
_vp, _i, _f = ctypes.c_void_p, ctypes.c_int, ctypes.c_float


def available():
    return os.path.exists(_PATH)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise FileNotFoundError(f"{_PATH} missing: run `make -C oracle ref` where /root/reference is mounted")
        L = ctypes.CDLL(_PATH)
        L.ref_forward.restype = _i
        L.ref_forward.argtypes = [_i, _i, _i, _vp, _i, _i] + [_vp] * 5 + [_f] + [_vp] * 5 + [_f, _f, _i] + [_vp] * 3 + [_i]
        L.ref_backward.restype = _i
        L.ref_backward.argtypes = ([_i, _i, _i, _i, _vp, _i, _i] + [_vp] * 4 + [_f] + [_vp] * 5 + [_f, _f]
                                   + [_vp] * 11 + [_i])
        L.ref_mark_visible.restype = _i
        L.ref_mark_visible.argtypes = [_i] + [_vp] * 4
        L.ref_compute_relocation.restype = _i
        L.ref_compute_relocation.argtypes = [_i] + [_vp] * 4 + [_i] + [_vp] * 2
        L.ref_fetch.restype = ctypes.c_longlong
        L.ref_fetch.argtypes = [_i, _vp, ctypes.c_size_t]
        L.ref_release.restype = None
        _lib = L
    return _lib


def _p(t):
    return None if t is None or t.numel() == 0 else t.data_ptr()


def forward(bg, means3D, colors_precomp, opacities, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
            projmatrix, tanfovx, tanfovy, H, W, sh, degree, campos, prefiltered=False, debug=False):
    """Mirror of the reference glue RasterizeGaussiansCUDA (rasterize_points.cu:36-118) on top of the shim.
    Runs on the legacy default stream like the reference. Returns (R, color, radii, is_used)."""
    dev = means3D.device
    P = means3D.shape[0]
    M = sh.shape[1] if sh is not None and sh.numel() else 0
    color = torch.zeros((3, H, W), dtype=torch.float32, device=dev)
    radii = torch.zeros((P,), dtype=torch.int32, device=dev)
    is_used = torch.zeros((P,), dtype=torch.int32, device=dev)
    R = lib().ref_forward(P, degree, M, _p(bg), W, H, _p(means3D), _p(sh), _p(colors_precomp), _p(opacities),
                          _p(scales), scale_modifier, _p(rotations), _p(cov3D_precomp), _p(viewmatrix), _p(projmatrix),
                          _p(campos), tanfovx, tanfovy, int(prefiltered), color.data_ptr(), is_used.data_ptr(),
                          radii.data_ptr(), int(debug))
    if R < 0:
        raise RuntimeError("reference forward failed")
    return R, color, radii, is_used


def backward(bg, means3D, radii, colors_precomp, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
             projmatrix, tanfovx, tanfovy, dL_dpix, sh, degree, campos, R, debug=False):
    """Mirror of RasterizeGaussiansBackwardCUDA (rasterize_points.cu:120-199): zero-filled gradient tensors,
    then Rasterizer::backward on the buffers kept by the shim from the last forward()."""
    dev = means3D.device
    P = means3D.shape[0]
    H, W = dL_dpix.shape[1], dL_dpix.shape[2]
    M = sh.shape[1] if sh is not None and sh.numel() else 0
    z = lambda *s: torch.zeros(s, dtype=torch.float32, device=dev)
    g = dict(means3D=z(P, 3), means2D=z(P, 3), colors=z(P, 3), conic=z(P, 2, 2), opacity=z(P, 1), cov3D=z(P, 6),
             sh=z(P, M, 3), scales=z(P, 3), rotations=z(P, 4))
    rc = lib().ref_backward(P, degree, M, R, _p(bg), W, H, _p(means3D), _p(sh), _p(colors_precomp), _p(scales),
                            scale_modifier, _p(rotations), _p(cov3D_precomp), _p(viewmatrix), _p(projmatrix),
                            _p(campos), tanfovx, tanfovy, radii.data_ptr(), dL_dpix.contiguous().data_ptr(),
                            g["means2D"].data_ptr(), g["conic"].data_ptr(), g["opacity"].data_ptr(),
                            g["colors"].data_ptr(), g["means3D"].data_ptr(), g["cov3D"].data_ptr(),
                            _p(g["sh"]), g["scales"].data_ptr(), g["rotations"].data_ptr(), int(debug))
    if rc != 0:
        raise RuntimeError("reference backward failed")
    return g


def fetch(which, dtype, device):
    n = lib().ref_fetch(which, None, 0)
    if n < 0:
        raise RuntimeError(f"ref_fetch({which}) failed")
    out = torch.empty(n, dtype=torch.uint8, device=device)
    if n:
        lib().ref_fetch(which, out.data_ptr(), n)
    return out.view(dtype)


def mark_visible(means3D, viewmatrix, projmatrix):
    P = means3D.shape[0]
    present = torch.zeros((P,), dtype=torch.bool, device=means3D.device)
    if P:
        lib().ref_mark_visible(P, means3D.data_ptr(), viewmatrix.data_ptr(), projmatrix.data_ptr(), present.data_ptr())
    return present


def compute_relocation(opacity_old, scale_old, N, binoms, n_max):
    P = opacity_old.shape[0]
    o = torch.zeros((P,), dtype=torch.float32, device=opacity_old.device)
    s = torch.zeros((3 * P,), dtype=torch.float32, device=opacity_old.device)
    if P:
        lib().ref_compute_relocation(P, opacity_old.contiguous().data_ptr(), scale_old.contiguous().data_ptr(),
                                     N.int().contiguous().data_ptr(), binoms.contiguous().data_ptr(), n_max,
                                     o.data_ptr(), s.data_ptr())
    return o, s


# ---- the reference's simple-knn (oracle/_ref/libref_knn.so) ----
_KNN_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libref_knn.so")
_knn_lib = None


def knn_available():
    return os.path.exists(_KNN_PATH)


def knn(points):
    """distCUDA2 of the reference (SimpleKNN::knn) on a CUDA [P,3] tensor -> float32 [P]. Runs on the default stream."""
    global _knn_lib
    if _knn_lib is None:
        if not knn_available():
            raise FileNotFoundError(f"{_KNN_PATH} missing: run `make -C oracle ref` where /root/reference is mounted")
        _knn_lib = ctypes.CDLL(_KNN_PATH)
        _knn_lib.ref_knn.restype = _i
        _knn_lib.ref_knn.argtypes = [_i, _vp, _vp]
    pts = points.contiguous().float()
    out = torch.zeros((pts.shape[0],), dtype=torch.float32, device=pts.device)
    torch.cuda.synchronize()
    rc = _knn_lib.ref_knn(int(pts.shape[0]), pts.data_ptr(), out.data_ptr())
    if rc != 0:
        raise RuntimeError(f"reference simple-knn failed: cudaError {rc}")
    return out
