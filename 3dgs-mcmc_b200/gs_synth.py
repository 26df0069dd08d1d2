"""Seeded synthetic scenes and cameras for tests and benchmarks (no datasets exist on the box).

Camera conventions follow the reference: ``world_view_transform`` is the transposed world-to-view matrix
(scene/cameras.py:63, utils/graphics_utils.py:38-51), ``full_proj_transform`` = view^T-convention product
with the transposed OpenGL-style projection of utils/graphics_utils.py:87-105 (znear 0.01, zfar 100,
scene/cameras.py:57-58), ``camera_center`` = inverse(view)[3, :3]. Scene generators return *activated*
parameters (what ``GaussianRasterizer.forward`` consumes): positive scales, unit quaternions (w,x,y,z),
opacities in (0,1), SH coefficients [P,16,3].

Scenes:
  * ``scene_cube``   — BASELINE configs 1-2: uniform points in [-1.3,1.3]^3 (scene/dataset_readers.py:266),
                       small near-isotropic splats (the 0.1 * 3-NN-distance init of gaussian_model_static.py:168-169).
  * ``scene_garden`` — BASELINE configs 3-4: 70 % of points on a noisy ground disc + objects within radius 3,
                       30 % on a background shell of radius 5-30; log-normal scales, Beta opacities.
"""
import math

import numpy as np

ZNEAR, ZFAR = 0.01, 100.0


def _projection(fovx, fovy, znear=ZNEAR, zfar=ZFAR):
    tx, ty = math.tan(fovx / 2), math.tan(fovy / 2)
    top, right = ty * znear, tx * znear
    P = np.zeros((4, 4), dtype=np.float64)
    P[0, 0] = 2.0 * znear / (2 * right)
    P[1, 1] = 2.0 * znear / (2 * top)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def look_at_camera(eye, target, width, height, fovx, up=(0.0, -1.0, 0.0)):
    """Camera at `eye` looking at `target` (+z forward, +y down, COLMAP style). Returns a dict with the
    float32 arrays / scalars that GaussianRasterizationSettings needs."""
    eye = np.asarray(eye, dtype=np.float64)
    target = np.asarray(target, dtype=np.float64)
    fwd = target - eye
    fwd /= np.linalg.norm(fwd)
    upv = np.asarray(up, dtype=np.float64)
    right = np.cross(upv, fwd)
    if np.linalg.norm(right) < 1e-8:
        right = np.cross(np.array([1.0, 0.0, 0.0]), fwd)
    right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    # world -> view rotation rows are the camera axes
    Rwv = np.stack([right, down, fwd], axis=0)
    view = np.eye(4)
    view[:3, :3] = Rwv
    view[:3, 3] = -Rwv @ eye
    focal = width / (2.0 * math.tan(fovx / 2))
    fovy = 2.0 * math.atan(height / (2.0 * focal))
    proj = _projection(fovx, fovy)
    wvt = view.T  # transposed storage, as the reference passes it
    full = wvt @ proj.T
    return {
        "image_width": int(width),
        "image_height": int(height),
        "tanfovx": math.tan(fovx * 0.5),
        "tanfovy": math.tan(fovy * 0.5),
        "FoVx": fovx,
        "FoVy": fovy,
        "viewmatrix": wvt.astype(np.float32),
        "projmatrix": full.astype(np.float32),
        "campos": np.linalg.inv(wvt)[3, :3].astype(np.float32),
    }


def orbit_cameras(n, radius, width, height, fovx, elevation_deg=20.0, target=(0.0, 0.0, 0.0), phase=0.0):
    cams = []
    el = math.radians(elevation_deg)
    for i in range(n):
        az = phase + 2.0 * math.pi * i / max(n, 1)
        # +y is down in this convention, so "above the ground" is negative y.
        eye = (radius * math.cos(el) * math.cos(az), -radius * math.sin(el), radius * math.cos(el) * math.sin(az))
        cams.append(look_at_camera(eye, target, width, height, fovx))
    return cams


def _random_unit_quats(rng, n):
    q = rng.standard_normal((n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return q.astype(np.float32)


def _sh(rng, n, rest_std=0.05):
    C0 = 0.28209479177387814
    sh = np.zeros((n, 16, 3), dtype=np.float32)
    sh[:, 0, :] = ((rng.random((n, 3)) - 0.5) / C0).astype(np.float32)  # RGB2SH of U(0,1) colours
    sh[:, 1:, :] = (rng.standard_normal((n, 15, 3)) * rest_std).astype(np.float32)
    return sh


def scene_cube(P, seed=0, scale_median=None, opacity_range=(0.05, 0.95)):
    rng = np.random.default_rng(seed)
    xyz = ((rng.random((P, 3)) * 2.0 - 1.0) * 1.3).astype(np.float32)
    if scale_median is None:
        # 0.1 x (typical 3-NN distance of P uniform points in the cube)
        scale_median = 0.1 * 0.8 * (2.6 ** 3 / max(P, 1)) ** (1.0 / 3.0)
    scales = (scale_median * np.exp(rng.standard_normal((P, 3)) * 0.3)).astype(np.float32)
    rots = _random_unit_quats(rng, P)
    lo, hi = opacity_range
    opac = (lo + (hi - lo) * rng.random((P, 1))).astype(np.float32)
    return {"means3D": xyz, "scales": scales, "rotations": rots, "opacities": opac, "shs": _sh(rng, P)}


def scene_garden(P, seed=0, area_scale=1.0):
    rng = np.random.default_rng(seed)
    n_bg = int(0.3 * P)
    n_fg = P - n_bg
    n_ground = int(0.6 * n_fg)
    n_obj = n_fg - n_ground
    # ground disc (y ~ 0, +y down), radius 3
    r = 3.0 * math.sqrt(area_scale) * np.sqrt(rng.random(n_ground))
    th = 2 * math.pi * rng.random(n_ground)
    ground = np.stack([r * np.cos(th), 0.02 * rng.standard_normal(n_ground), r * np.sin(th)], axis=1)
    # a few dozen blob "objects" standing on the ground
    n_blobs = 40
    centers = np.stack([2.5 * math.sqrt(area_scale) * (rng.random(n_blobs) * 2 - 1),
                        -0.3 * rng.random(n_blobs) - 0.1,
                        2.5 * math.sqrt(area_scale) * (rng.random(n_blobs) * 2 - 1)], axis=1)
    which = rng.integers(0, n_blobs, n_obj)
    obj = centers[which] + rng.standard_normal((n_obj, 3)) * np.array([0.15, 0.25, 0.15])
    # background shell
    d = rng.standard_normal((n_bg, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rad = 5.0 + 25.0 * rng.random(n_bg) ** 2
    bg = d * rad[:, None]
    xyz = np.concatenate([ground, obj, bg], axis=0).astype(np.float32)
    perm = rng.permutation(P)
    xyz = xyz[perm]
    is_bg = np.concatenate([np.zeros(n_fg, bool), np.ones(n_bg, bool)])[perm]

    base = 0.01 * np.exp(0.8 * rng.standard_normal(P))
    base = np.where(is_bg, base * (np.linalg.norm(xyz, axis=1) / 3.0), base)  # far splats are larger
    aniso = np.exp(0.5 * rng.standard_normal((P, 3)))
    scales = (base[:, None] * aniso).astype(np.float32)
    rots = _random_unit_quats(rng, P)
    opac = np.clip(rng.beta(0.5, 2.0, (P, 1)), 0.005, 0.99).astype(np.float32)
    return {"means3D": xyz, "scales": scales, "rotations": rots, "opacities": opac, "shs": _sh(rng, P)}


def to_torch(d, device):
    import torch
    out = {}
    for k, v in d.items():
        out[k] = torch.from_numpy(np.ascontiguousarray(v)).to(device) if isinstance(v, np.ndarray) else v
    return out
