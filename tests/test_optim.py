"""Fused multi-tensor Adam (csrc/optim.cu, diff_gaussian_rasterization/optim.py).

CPU: the numpy oracle is pinned against torch.optim.Adam itself (the reference's optimizer,
scene/gaussian_model_static.py:200), and the host-side FusedAdam keeps torch's state layout. GPU: FusedAdam against
torch.optim.Adam on the reference's six parameter groups, and against the oracle.
Tolerance: the update is FP32 elementwise with one sqrt and one division; per-step relative error <= 2e-6 of the
parameter change (fma contraction differences) plus one ulp of the parameter, stated in ADAM_TOL.
"""
import numpy as np
import pytest
import torch

import helpers  # noqa: F401  (path setup)
from oracle import oracle

ADAM_TOL = 2e-6


def _groups(P, device, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    shapes = {"xyz": (P, 3), "f_dc": (P, 1, 3), "f_rest": (P, 15, 3), "opacity": (P, 1), "scaling": (P, 3), "rotation": (P, 4)}
    lrs = {"xyz": 1.6e-4, "f_dc": 2.5e-3, "f_rest": 2.5e-3 / 20, "opacity": 0.05, "scaling": 5e-3, "rotation": 1e-3}
    params = {k: torch.nn.Parameter(torch.randn(s, generator=g).to(device)) for k, s in shapes.items()}
    return params, [{"params": [params[k]], "lr": lrs[k], "name": k} for k in shapes]


def _set_grads(params, step, scale=1.0):
    g = torch.Generator(device="cpu").manual_seed(100 + step)
    for p in params.values():
        p.grad = (torch.randn(p.shape, generator=g) * scale).to(p.device)


def test_adam_oracle_matches_torch_adam_on_cpu():
    """Pins oracle.adam_step to torch.optim.Adam (CPU, foreach off = the single-tensor reference implementation)."""
    params, groups = _groups(257, "cpu")
    opt = torch.optim.Adam(groups, lr=0.0, eps=1e-15, foreach=False)
    mine = {k: (p.detach().numpy().copy(), np.zeros(p.shape, np.float32), np.zeros(p.shape, np.float32)) for k, p in params.items()}
    for step in range(1, 6):
        _set_grads(params, step)
        for gr in groups:
            k = gr["name"]
            p, m, v = mine[k]
            mine[k] = oracle.adam_step(p, params[k].grad.numpy(), m, v, gr["lr"], step)
        opt.step()
        for k, p in params.items():
            ref = p.detach().numpy()
            err = np.abs(mine[k][0] - ref).max() / max(np.abs(ref).max(), 1e-30)
            assert err <= 2.4e-7, (k, step, err)          # two ulps of the parameter


def test_fused_adam_keeps_torch_state_layout():
    """The reference edits optimizer.state / param_groups by hand (gaussian_model_static.py:302-392)."""
    from diff_gaussian_rasterization.optim import FusedAdam
    params, groups = _groups(5, "cpu")
    opt = FusedAdam(groups, lr=0.0, eps=1e-15)
    assert [g["name"] for g in opt.param_groups] == ["xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation"]
    assert opt.param_groups[0]["eps"] == 1e-15 and opt.param_groups[0]["betas"] == (0.9, 0.999)
    for g in opt.param_groups:
        if g["name"] == "xyz":
            g["lr"] = 1e-5                      # update_learning_rate
    assert opt.state.get(params["xyz"], None) is None
    opt.step()                                  # no gradients: nothing to do, no GPU needed
    opt.zero_grad(set_to_none=True)
    _set_grads(params, 1)
    with pytest.raises(RuntimeError):           # CPU tensors: there is no CPU path
        opt.step()
    with pytest.raises(ValueError):
        FusedAdam(groups, lr=0.0, weight_decay=0.1)
    sd = opt.state_dict()
    assert len(sd["param_groups"]) == 6


@pytest.mark.gpu
@pytest.mark.parametrize("P", [1000, 4097, 300000])
def test_fused_adam_matches_torch_adam_and_oracle(P):
    from diff_gaussian_rasterization.optim import FusedAdam
    dev = torch.device("cuda:0")
    pa, ga = _groups(P, dev)
    pb, gb = _groups(P, dev)
    ref = torch.optim.Adam(ga, lr=0.0, eps=1e-15)
    mine = FusedAdam(gb, lr=0.0, eps=1e-15)
    start = {k: p.detach().clone() for k, p in pa.items()}
    for step in range(1, 5):
        for grp in (ga, gb):                    # the xyz learning rate changes every iteration in the reference
            grp[0]["lr"] = 1.6e-4 * (0.99 ** step)
        _set_grads(pa, step, scale=1e-3 if step == 3 else 1.0)
        _set_grads(pb, step, scale=1e-3 if step == 3 else 1.0)
        ref.step()
        mine.step()
        torch.cuda.synchronize()
        for k in pa:
            moved = (pa[k] - start[k]).abs().max().item()
            err = (pa[k] - pb[k]).abs().max().item()
            ulp = 1.2e-7 * pa[k].abs().max().item()        # one FP32 ulp of the parameter itself
            assert err <= ADAM_TOL * moved + ulp, (k, step, err, moved)
            sa, sb = ref.state[pa[k]], mine.state[pb[k]]
            assert torch.allclose(sa["exp_avg"], sb["exp_avg"], rtol=1e-6, atol=1e-12)
            assert torch.allclose(sa["exp_avg_sq"], sb["exp_avg_sq"], rtol=1e-6, atol=1e-20)
            assert int(sb["step"]) == step
    # against the oracle, one more step on a copy
    k = "f_rest"
    p0 = pb[k].detach().cpu().numpy().copy()
    st = mine.state[pb[k]]
    m0, v0 = st["exp_avg"].cpu().numpy().copy(), st["exp_avg_sq"].cpu().numpy().copy()
    _set_grads(pb, 9)
    g0 = pb[k].grad.cpu().numpy().copy()
    mine.step()
    torch.cuda.synchronize()
    want, wm, wv = oracle.adam_step(p0, g0, m0, v0, gb[2]["lr"], 5)
    got = pb[k].detach().cpu().numpy()
    assert np.abs(got - want).max() <= ADAM_TOL * np.abs(want - p0).max() + 1.2e-7 * np.abs(want).max()


@pytest.mark.gpu
def test_fused_adam_state_surgery_like_the_reference():
    """replace_tensors_to_optimizer (gaussian_model_static.py:354-392) zeroes the moments of some rows and swaps the
    parameter object; cat_tensors_to_optimizer (:302-327) grows them. The next fused step must honour both."""
    from diff_gaussian_rasterization.optim import FusedAdam
    dev = torch.device("cuda:0")
    params, groups = _groups(1000, dev)
    opt = FusedAdam(groups, lr=0.0, eps=1e-15)
    _set_grads(params, 1)
    opt.step()
    group = opt.param_groups[0]
    old = group["params"][0]
    stored = opt.state.get(old)
    idx = torch.arange(0, 1000, 7, device=dev)
    stored["exp_avg"][idx] = 0
    stored["exp_avg_sq"][idx] = 0
    extra = torch.zeros((24, 3), device=dev)
    stored["exp_avg"] = torch.cat((stored["exp_avg"], torch.zeros_like(extra)), 0)
    stored["exp_avg_sq"] = torch.cat((stored["exp_avg_sq"], torch.zeros_like(extra)), 0)
    del opt.state[old]
    new = torch.nn.Parameter(torch.cat((old.detach(), extra), 0).requires_grad_(True))
    group["params"][0] = new
    opt.state[new] = stored
    before = new.detach().clone()
    new.grad = torch.ones_like(new)
    for k, p in params.items():
        if k != "xyz":
            p.grad = None
    opt.step()
    torch.cuda.synchronize()
    assert new.shape == (1024, 3)
    assert bool((new != before).all())
    assert int(opt.state[new]["step"]) == 2
