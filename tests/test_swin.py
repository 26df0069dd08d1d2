"""Fused sliding-window gather (csrc/swin.cu, diff_gaussian_rasterization/swin.py).

CPU: the float64 oracle (forward + analytic gradient) is pinned against golden vectors produced with the reference's own
utils/tempo_utils.py + torch autograd (tests/golden/make_golden_swin.py). GPU: the kernels, through the autograd API,
against those goldens and against the oracle on a larger seeded case.
Tolerance: FP32 with sinf / cosf / expf and a Rodrigues rotation: 2e-6 relative L2 on outputs, 2e-5 on gradients.
"""
import os

import numpy as np
import pytest
import torch

import helpers  # noqa: F401  (path setup)
from oracle import oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "swin_gather.npz")
OUT_KEYS = ("xyz", "rotation", "feature", "opacity", "scaling")
CASES = ("deform", "skip", "immature_only")
OUT_TOL, GRAD_TOL = 2e-6, 2e-5


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def _load(name):
    d = np.load(GOLD)
    imm = {k: d[f"{name}/imm/{k}"] for k in oracle.SWIN_KEYS}
    mat = {k: d[f"{name}/mat/{k}"] for k in oracle.SWIN_KEYS}
    ups = {k: d[f"{name}/up/{k}"] for k in OUT_KEYS}
    outs = {k: d[f"{name}/out/{k}"] for k in OUT_KEYS}
    grads = {k: d[f"{name}/grad/{k}"] for k in oracle.SWIN_KEYS[:9]}
    return float(d[f"{name}/frame"]), bool(d[f"{name}/deform"]), d[f"{name}/imm_idx"], d[f"{name}/mat_idx"], imm, mat, ups, outs, grads


@pytest.mark.parametrize("name", CASES)
def test_swin_oracle_matches_reference_golden(name):
    frame, deform, ii, mi, imm, mat, ups, outs, grads = _load(name)
    r = oracle.swin_gather(frame, imm, mat, ii, mi, deform, ups)
    for k in OUT_KEYS:
        assert _rel(r[k], outs[k]) < 5e-7, k            # the goldens are FP32
    for k in grads:
        assert _rel(r["grad"][k], grads[k]) < 1e-6, k
        untouched = np.ones(len(imm[k]), bool)
        untouched[ii] = False
        assert not r["grad"][k][untouched].any()


def test_swin_api_has_no_cpu_path():
    from diff_gaussian_rasterization import swin
    frame, deform, ii, mi, imm, mat, *_ = _load("deform")
    t = lambda s: {k: torch.from_numpy(v) for k, v in s.items()}  # noqa: E731
    with pytest.raises(RuntimeError):
        swin.gather_active(frame, t(imm), t(mat), torch.from_numpy(ii), torch.from_numpy(mi))
    assert swin.SET_KEYS == oracle.SWIN_KEYS


def _run_gpu(frame, deform, ii, mi, imm, mat, ups):
    from diff_gaussian_rasterization import swin
    dev = torch.device("cuda:0")
    timm = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in imm.items()}
    tmat = {k: torch.from_numpy(np.ascontiguousarray(v)).to(dev) for k, v in mat.items()}
    for k in oracle.SWIN_KEYS[:9]:
        timm[k].requires_grad_(True)
    res = swin.gather_active(frame, timm, tmat, torch.from_numpy(ii).to(dev), torch.from_numpy(mi).to(dev), deform=deform)
    total = sum((res[k] * torch.from_numpy(ups[k]).to(dev)).sum() for k in OUT_KEYS)
    grads = torch.autograd.grad(total, [timm[k] for k in oracle.SWIN_KEYS[:9]])
    torch.cuda.synchronize()
    return {k: v.detach().cpu().numpy() for k, v in res.items()}, {k: g.cpu().numpy() for k, g in zip(oracle.SWIN_KEYS[:9], grads)}


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_swin_kernels_against_reference_golden(name):
    frame, deform, ii, mi, imm, mat, ups, outs, grads = _load(name)
    res, g = _run_gpu(frame, deform, ii, mi, imm, mat, ups)
    for k in OUT_KEYS:
        assert res[k].shape == outs[k].shape
        assert _rel(res[k], outs[k]) < OUT_TOL, k
    assert np.array_equal(res["feature"], outs["feature"])          # a pure gather: bit-exact
    for k in grads:
        assert _rel(g[k], grads[k]) < GRAD_TOL, k
    assert np.array_equal(g["features_rest"], grads["features_rest"])


@pytest.mark.gpu
def test_swin_kernels_against_oracle_large_case():
    rng = np.random.default_rng(4)
    Ni, Nm, M, frame = 200_003, 90_001, 16, 11.0

    def mk(N):
        return {"xyz": rng.normal(size=(N, 3)), "rotation": rng.normal(size=(N, 4)), "features_dc": rng.normal(size=(N, 1, 3)),
                "features_rest": rng.normal(size=(N, M - 1, 3)) * 0.1, "opacity": rng.normal(size=(N, 1)),
                "scaling": rng.normal(size=(N, 3)) - 3, "rigid_v": rng.normal(size=(N, 3)) * 0.05,
                "rigid_rotvec": rng.normal(size=(N, 3)) * 0.3, "rigid_rotcen": rng.normal(size=(N, 3)),
                "frame_start": rng.integers(0, 30, N).astype(np.float64)}
    imm = {k: v.astype(np.float32) for k, v in mk(Ni).items()}
    mat = {k: v.astype(np.float32) for k, v in mk(Nm).items()}
    ii = np.nonzero(imm["frame_start"] <= frame)[0]
    mi = np.nonzero(mat["frame_start"] <= frame)[0]
    K = len(ii) + len(mi)
    ups = {"xyz": rng.normal(size=(K, 3)), "rotation": rng.normal(size=(K, 4)), "feature": rng.normal(size=(K, M, 3)),
           "opacity": rng.normal(size=(K, 1)), "scaling": rng.normal(size=(K, 3))}
    ups = {k: v.astype(np.float32) for k, v in ups.items()}
    want = oracle.swin_gather(frame, imm, mat, ii, mi, True, ups)
    res, g = _run_gpu(frame, True, ii, mi, imm, mat, ups)
    for k in OUT_KEYS:
        assert _rel(res[k], want[k]) < OUT_TOL, k
    for k in g:
        assert _rel(g[k], want["grad"][k]) < GRAD_TOL, k
