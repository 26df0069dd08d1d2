"""Fused L1 + SSIM loss (csrc/loss.cu, diff_gaussian_rasterization/loss.py).

CPU: the float64 numpy oracle is pinned against golden vectors produced by the reference's own utils/loss_utils.py
(tests/golden/make_golden_loss.py). GPU: the kernels, through the Python API that mirrors loss_utils, against those
goldens, against the oracle on ragged sizes, against a plain torch FP32 restatement of _ssim at 1080p, and through
size-independent properties.
Tolerances (FP32 with 121-term windowed sums): loss values 2e-6 absolute, gradients relative L2 1e-4.
"""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import helpers  # noqa: F401  (path setup)
from oracle import oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "loss_ssim.npz")
VAL_TOL = 2e-6
GRAD_TOL = 1e-4


def _cases():
    d = np.load(GOLD)
    return d, sorted({k.split("/")[0] for k in d.files if "/" in k})


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def test_loss_oracle_matches_reference_golden():
    d, names = _cases()
    lam = float(d["lambda_dssim"])
    assert len(names) == 5
    for n in names:
        r = oracle.photometric_loss(d[f"{n}/img"], d[f"{n}/gt"], lam)
        assert abs(r["l1"] - d[f"{n}/l1"]) < VAL_TOL, n
        assert abs(r["ssim"] - d[f"{n}/ssim"]) < VAL_TOL, n
        assert abs(r["loss"] - d[f"{n}/loss"]) < VAL_TOL, n
        assert _rel(r["grad_loss"], d[f"{n}/grad_loss"]) < 2e-5, n       # the golden itself is FP32 autograd
        assert _rel(r["grad_ssim"], d[f"{n}/grad_ssim"]) < 2e-5, n


def test_window_matches_reference_construction():
    w = oracle.ssim_window()
    g = torch.Tensor([np.exp(-(x - 11 // 2) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)])
    assert np.array_equal(w, (g / g.sum()).numpy())


def test_loss_api_rejects_cpu_tensors():
    from diff_gaussian_rasterization import loss
    with pytest.raises(RuntimeError):
        loss.ssim(torch.rand(3, 8, 8), torch.rand(3, 8, 8))
    with pytest.raises(NotImplementedError):
        loss.ssim(torch.rand(3, 8, 8), torch.rand(3, 8, 8), window_size=7)


def _torch_ref(img, gt, lam):
    """utils/loss_utils.py:21-67 restated with the same torch ops (FP32 reference of the floating-point kernel)."""
    C = img.size(-3)
    g = torch.tensor(oracle.ssim_window(), device=img.device)
    window = (g[:, None] @ g[None, :]).expand(C, 1, 11, 11).contiguous()
    conv = lambda t: F.conv2d(t[None], window, padding=5, groups=C)[0]  # noqa: E731
    mu1, mu2 = conv(img), conv(gt)
    s1 = conv(img * img) - mu1 * mu1
    s2 = conv(gt * gt) - mu2 * mu2
    s12 = conv(img * gt) - mu1 * mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim_map = ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1 * mu1 + mu2 * mu2 + C1) * (s1 + s2 + C2))
    l1 = (img - gt).abs().mean()
    return (1 - lam) * l1 + lam * (1 - ssim_map.mean()), l1, ssim_map.mean()


@pytest.mark.gpu
def test_fused_loss_against_reference_golden():
    from diff_gaussian_rasterization import loss
    dev = torch.device("cuda:0")
    d, names = _cases()
    lam = float(d["lambda_dssim"])
    for n in names:
        img = torch.from_numpy(d[f"{n}/img"]).to(dev).requires_grad_(True)
        gt = torch.from_numpy(d[f"{n}/gt"]).to(dev)
        L = loss.photometric_loss(img, gt, lam)
        (g,) = torch.autograd.grad(L, img)
        assert abs(L.item() - float(d[f"{n}/loss"])) < VAL_TOL, n
        assert _rel(g.cpu().numpy(), d[f"{n}/grad_loss"]) < GRAD_TOL, n
        s = loss.ssim(img, gt)
        (gs,) = torch.autograd.grad(s, img)
        assert abs(s.item() - float(d[f"{n}/ssim"])) < VAL_TOL, n
        assert _rel(gs.cpu().numpy(), d[f"{n}/grad_ssim"]) < GRAD_TOL, n
        assert abs(loss.l1_loss(img, gt).item() - float(d[f"{n}/l1"])) < VAL_TOL


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(3, 1, 1), (3, 5, 200), (1, 17, 15), (3, 100, 130), (4, 33, 16)])
def test_fused_loss_against_oracle_ragged_sizes(shape):
    from diff_gaussian_rasterization import loss
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(sum(shape))
    gt = torch.rand(shape, generator=g)
    img = (gt + 0.1 * torch.randn(shape, generator=g)).clamp(0, 1)
    r = oracle.photometric_loss(img.numpy(), gt.numpy(), 0.2)
    x = img.to(dev).requires_grad_(True)
    L = loss.photometric_loss(x, gt.to(dev), 0.2)
    # upstream gradient other than 1: the loss is usually part of a sum (train_static.py:99-100)
    (gx,) = torch.autograd.grad(3.0 * L, x)
    assert abs(L.item() - r["loss"]) < VAL_TOL
    assert _rel(gx.cpu().numpy() / 3.0, r["grad_loss"]) < GRAD_TOL


@pytest.mark.gpu
def test_fused_loss_full_size_1080p_vs_torch_and_properties():
    from diff_gaussian_rasterization import loss
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(3)
    gt = torch.rand((3, 1080, 1920), generator=g).to(dev)
    img = (gt + 0.05 * torch.randn((3, 1080, 1920), generator=g).to(dev)).clamp(0, 1).requires_grad_(True)
    L = loss.photometric_loss(img, gt, 0.2)
    (gm,) = torch.autograd.grad(L, img)
    ref = img.detach().clone().requires_grad_(True)
    Lr, l1r, sr = _torch_ref(ref, gt, 0.2)
    (gr,) = torch.autograd.grad(Lr, ref)
    assert abs(L.item() - Lr.item()) < VAL_TOL
    assert helpers.rel_l2(gm, gr) < GRAD_TOL
    # properties: identical images -> ssim exactly 1-ish and zero L1; deterministic value; NCHW batch-of-one accepted
    s = loss.ssim(gt.clone().requires_grad_(True), gt)
    assert abs(s.item() - 1.0) < 1e-6
    assert loss.photometric_loss(gt.clone(), gt, 0.2).item() < 1e-6
    a = loss.photometric_loss(img, gt, 0.2).item()
    b = loss.photometric_loss(img, gt, 0.2).item()
    assert a == b
    assert loss.ssim(img[None], gt[None]).item() == loss.ssim(img, gt).item()
    # linearity of the backward in the upstream gradient
    (g2,) = torch.autograd.grad(2.0 * loss.photometric_loss(img, gt, 0.2), img)
    assert helpers.rel_l2(g2, 2.0 * gm) < 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("shape,shift", [((3, 1, 1), 0), ((3, 37, 53), 0), ((3, 64, 64), 1), ((3, 1080, 1920), 0), ((5, 4099), 3)])
def test_fused_l1_loss_vs_oracle_and_torch(shape, shift):
    """loss.l1_loss = torch.abs(network_output - gt).mean() (utils/loss_utils.py:21-22): value against the float64
    oracle and torch, gradient against torch's (upstream gradient other than 1; exact zeros where the images agree).
    shift > 0: tensors on a 4-byte aligned base (scalar path)."""
    from diff_gaussian_rasterization import loss
    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(len(shape) + shape[-1])
    n = int(np.prod(shape))

    def place(t):
        flat = torch.empty(n + shift, device=dev)
        v = flat[shift:].view(shape)
        v.copy_(t)
        return v

    gt_h = torch.rand(shape, generator=g)
    img_h = (gt_h + 0.1 * torch.randn(shape, generator=g)).clamp(0, 1)
    img_h.view(-1)[::7] = gt_h.view(-1)[::7]                      # exact ties: sign(0) = 0
    gt, img = place(gt_h), place(img_h)
    assert shift == 0 or img.data_ptr() % 16 != 0
    x = img.detach().requires_grad_(True)
    L = loss.l1_loss(x, gt)
    (gx,) = torch.autograd.grad(2.5 * L, x)
    xr = img.detach().clone().requires_grad_(True)
    Lr = torch.abs(xr - gt).mean()
    (gr,) = torch.autograd.grad(2.5 * Lr, xr)
    want = oracle.l1_loss(img_h.numpy(), gt_h.numpy())
    assert abs(L.item() - want["l1"]) < 1e-6 * max(want["l1"], 1e-3)
    assert _rel(gx.cpu().numpy() / 2.5, want["grad"]) < 1e-6
    assert abs(L.item() - Lr.item()) < VAL_TOL
    assert helpers.rel_l2(gx, gr) < 1e-6
    assert bool((gx.view(-1)[::7] == 0).all())
    assert loss.l1_loss(x, gt).item() == L.item()                 # fixed summation order
    if len(shape) == 3:
        # the L1 term of the fused photometric loss is the same quantity
        assert abs(loss.photometric_loss(x.detach().contiguous(), gt.contiguous(), 0.0).item() - L.item()) < VAL_TOL


def test_l1_oracle_matches_torch_cpu():
    g = torch.Generator().manual_seed(11)
    gt = torch.rand((3, 19, 23), generator=g)
    img = (gt + 0.1 * torch.randn((3, 19, 23), generator=g)).clamp(0, 1)
    img.view(-1)[::5] = gt.view(-1)[::5]
    x = img.double().requires_grad_(True)
    L = torch.abs(x - gt.double()).mean()
    (gx,) = torch.autograd.grad(L, x)
    r = oracle.l1_loss(img.numpy(), gt.numpy())
    assert abs(r["l1"] - L.item()) < 1e-15
    assert np.array_equal(r["grad"], gx.numpy())


def test_l1_loss_rejects_cpu_tensors():
    from diff_gaussian_rasterization import loss
    with pytest.raises(RuntimeError, match="no CPU path"):
        loss.l1_loss(torch.zeros(3, 4, 4), torch.zeros(3, 4, 4))
