"""Fused activations + regularisers (csrc/activations.cu, diff_gaussian_rasterization/activations.py).

CPU: the float64 oracle is pinned against the reference's own activation functions — they ARE torch ops
(scene/gaussian_model_static.py:43-56: torch.exp, torch.sigmoid, torch.nn.functional.normalize) — and torch autograd.
GPU: the kernels against the oracle and against the torch ops. Tolerance: 1e-6 relative (FP32 expf / division).
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

import helpers
from oracle import oracle

TOL = 2e-6


def _inputs(P, seed=0):
    g = torch.Generator().manual_seed(seed)
    ro = torch.randn((P, 1), generator=g) * 3
    rs = torch.randn((P, 3), generator=g) - 4
    rq = torch.randn((P, 4), generator=g)
    if P > 3:
        rq[3] = 0.0                     # degenerate quaternion: F.normalize clamps the norm at 1e-12
    gs = [torch.randn(t.shape, generator=g) for t in (ro, rs, rq)]
    return ro, rs, rq, gs


def _torch_ref(ro, rs, rq, gs, w_o, w_s, g_reg):
    ro, rs, rq = (t.clone().requires_grad_(True) for t in (ro, rs, rq))
    o, s, q = torch.sigmoid(ro), torch.exp(rs), F.normalize(rq)
    reg = w_o * torch.abs(o).mean() + w_s * torch.abs(s).mean()
    total = (o * gs[0]).sum() + (s * gs[1]).sum() + (q * gs[2]).sum() + g_reg * reg
    d = torch.autograd.grad(total, (ro, rs, rq))
    return o.detach(), s.detach(), q.detach(), reg.detach(), d


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def test_activation_oracle_matches_torch_ops():
    ro, rs, rq, gs = _inputs(501)
    o, s, q, reg, d = _torch_ref(ro.double(), rs.double(), rq.double(), [g.double() for g in gs], 0.01, 0.02, 1.5)
    r = oracle.activate(ro.numpy(), rs.numpy(), rq.numpy(), 0.01, 0.02, gs[0].numpy(), gs[1].numpy(), gs[2].numpy(), 1.5)
    for name, want in (("opacity", o), ("scaling", s), ("rotation", q), ("d_raw_opacity", d[0]), ("d_raw_scaling", d[1]),
                       ("d_raw_rotation", d[2])):
        assert _rel(r[name], want.numpy()) < 1e-12, name
    assert abs(r["reg"] - reg.item()) < 1e-14


def test_activate_rejects_cpu_tensors():
    from diff_gaussian_rasterization.activations import activate
    ro, rs, rq, _ = _inputs(8)
    with pytest.raises(RuntimeError):
        activate(ro, rs, rq)


@pytest.mark.gpu
@pytest.mark.parametrize("P", [1, 255, 4097, 300000])
def test_fused_activations_against_oracle_and_torch(P):
    from diff_gaussian_rasterization.activations import activate
    dev = torch.device("cuda:0")
    ro, rs, rq, gs = _inputs(P, seed=P)
    r = oracle.activate(ro.numpy(), rs.numpy(), rq.numpy(), 0.01, 0.02, gs[0].numpy(), gs[1].numpy(), gs[2].numpy(), 1.5)
    a, b, c = (t.to(dev).requires_grad_(True) for t in (ro, rs, rq))
    o, s, q, reg = activate(a, b, c, 0.01, 0.02)
    total = (o * gs[0].to(dev)).sum() + (s * gs[1].to(dev)).sum() + (q * gs[2].to(dev)).sum() + 1.5 * reg
    d = torch.autograd.grad(total, (a, b, c))
    for name, got in (("opacity", o), ("scaling", s), ("rotation", q), ("d_raw_opacity", d[0]), ("d_raw_scaling", d[1]),
                      ("d_raw_rotation", d[2])):
        assert _rel(got.detach().cpu().numpy(), r[name]) < TOL, name
    assert abs(reg.item() - r["reg"]) < TOL * max(abs(r["reg"]), 1e-3)
    # the reference's torch ops on the GPU
    to, ts, tq, treg, td = _torch_ref(ro.to(dev), rs.to(dev), rq.to(dev), [g.to(dev) for g in gs], 0.01, 0.02, 1.5)
    assert helpers.rel_l2(o, to) < TOL and helpers.rel_l2(s, ts) < TOL and helpers.rel_l2(q, tq) < TOL
    for got, want in zip(d, td):
        assert helpers.rel_l2(got, want) < TOL


@pytest.mark.gpu
def test_fused_activations_without_regularisers_or_some_gradients():
    from diff_gaussian_rasterization.activations import activate
    dev = torch.device("cuda:0")
    ro, rs, rq, gs = _inputs(1000)
    a, b, c = (t.to(dev).requires_grad_(True) for t in (ro, rs, rq))
    o, s, q, reg = activate(a, b, c)
    assert reg.item() == 0.0
    (d_b,) = torch.autograd.grad((s * gs[1].to(dev)).sum(), b)          # only the scaling output is used
    assert helpers.rel_l2(d_b, gs[1].to(dev) * torch.exp(rs.to(dev))) < TOL
