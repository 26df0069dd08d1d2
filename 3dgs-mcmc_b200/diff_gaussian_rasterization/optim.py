"""Fused multi-tensor Adam for the Gaussian parameters.

Drop-in for the reference's ``torch.optim.Adam(l, lr=0.0, eps=1e-15)`` (scene/gaussian_model_static.py:191-200):
same constructor arguments and ``param_groups`` / ``state`` layout (``state[p] = {"step", "exp_avg", "exp_avg_sq"}``),
so ``update_learning_rate``, ``cat_tensors_to_optimizer``, ``replace_tensors_to_optimizer``, ``state_dict`` /
``load_state_dict`` (gaussian_model_static.py:100,124,206-212,302-392) work unchanged. ``step()`` updates every
parameter tensor with ONE kernel launch (csrc/optim.cu) instead of torch's ~10 foreach kernels; there is no CPU path.
"""
import torch

from . import _native as _C


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False):
        if weight_decay != 0 or amsgrad:
            raise ValueError("FusedAdam implements the reference's configuration: weight_decay=0, amsgrad=False")
        if not 0.0 <= lr or not 0.0 <= eps or not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0:
            raise ValueError("invalid Adam hyper-parameters")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=0, amsgrad=False))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        # tensors that share (betas, eps) go into one launch; the reference has a single such family
        families = {}
        for group in self.param_groups:
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                if p.grad.is_sparse:
                    raise RuntimeError("FusedAdam does not support sparse gradients")
                state = self.state[p]
                if len(state) == 0:
                    state["step"] = torch.tensor(0.0, dtype=torch.float32)
                    state["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    state["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                step = state["step"]
                if torch.is_tensor(step):
                    step += 1
                    step_i = int(step.item()) if step.device.type == "cpu" else int(step.cpu().item())
                else:
                    state["step"] = step = step + 1
                    step_i = int(step)
                fam = families.setdefault((beta1, beta2, group["eps"]), ([], [], [], [], [], []))
                grad = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                for lst, item in zip(fam, (p, grad, state["exp_avg"], state["exp_avg_sq"], group["lr"], step_i)):
                    lst.append(item)
        for (beta1, beta2, eps), (ps, gs, ms, vs, lrs, steps) in families.items():
            _C.adam_step(ps, gs, ms, vs, lrs, steps, beta1, beta2, eps)
        return loss
