"""End-to-end check on the GPU: a miniature of train_static.py:63-140 built ONLY from this repository's pieces
(fused activations, rasterizer, fused L1 + SSIM loss, backward, FusedAdam, MCMC relocate / add_new_gs with
compute_relocation and the optimizer-state surgery of scene/gaussian_model_static.py:302-492, add_noise) fits a target
image: the loss must fall, dead Gaussians must be revived, the population must grow by 5 % steps, everything stays finite.
BASELINE.json config 2's shape in small (cube-initialised Gaussians, one 800x800-like camera set, cap_max growth).
"""
import numpy as np
import pytest
import torch

import helpers  # noqa: F401
import gs_synth

pytestmark = pytest.mark.gpu


def _binoms(dev, n_max=51):
    from scipy.special import comb
    b = torch.zeros((n_max, n_max))
    for n in range(n_max):
        for k in range(n + 1):
            b[n, k] = comb(n, k)
    return b.float().to(dev)


class MiniModel:
    """The slice of GaussianModel the MCMC trainer needs (gaussian_model_static.py:126-146, 302-492)."""
    NAMES = ("xyz", "shs", "opacity", "scaling", "rotation")

    def __init__(self, scene, dev):
        from diff_gaussian_rasterization.optim import FusedAdam
        s = gs_synth.to_torch(scene, dev)
        raw = {"xyz": s["means3D"], "shs": s["shs"] * 0.0 + 0.1, "opacity": torch.logit(torch.full_like(s["opacities"], 0.5)),
               "scaling": torch.log(s["scales"] * 3), "rotation": s["rotations"]}
        self.p = {k: torch.nn.Parameter(v.clone().contiguous()) for k, v in raw.items()}
        lrs = {"xyz": 1.6e-4 * 5, "shs": 2.5e-2, "opacity": 0.05, "scaling": 5e-3, "rotation": 1e-3}
        self.opt = FusedAdam([{"params": [self.p[k]], "lr": lrs[k], "name": k} for k in self.NAMES], lr=0.0, eps=1e-15)
        self.binoms = _binoms(dev)

    def _update_params(self, idxs, ratio):          # :395-410
        from diff_gaussian_rasterization import compute_relocation
        opacity, scaling = torch.sigmoid(self.p["opacity"]), torch.exp(self.p["scaling"])
        N = (ratio[idxs, 0] + 1).clamp_(min=1, max=50)
        new_o, new_s = compute_relocation(opacity[idxs, 0], scaling[idxs], N, self.binoms, 51)
        new_o = torch.clamp(new_o.unsqueeze(-1), max=1.0 - torch.finfo(torch.float32).eps, min=0.005)
        return torch.logit(new_o), torch.log(new_s.reshape(-1, 3))

    def _sample(self, probs, num, alive=None):      # :413-424
        probs = probs / (probs.sum() + torch.finfo(torch.float32).eps)
        idx = torch.multinomial(probs, num, replacement=True)
        if alive is not None:
            idx = alive[idx]
        return idx, torch.bincount(idx, minlength=self.p["xyz"].shape[0]).unsqueeze(-1)

    def _reset_moments(self, inds):                 # replace_tensors_to_optimizer :354-392
        for group in self.opt.param_groups:
            st = self.opt.state.get(group["params"][0], None)
            if st is not None:
                st["exp_avg"][inds] = 0
                st["exp_avg_sq"][inds] = 0

    @torch.no_grad()
    def relocate(self):                             # :427-457
        dead = (torch.sigmoid(self.p["opacity"]) <= 0.005).squeeze(-1)
        if dead.sum() == 0:
            return 0
        dead_idx, alive_idx = dead.nonzero(as_tuple=True)[0], (~dead).nonzero(as_tuple=True)[0]
        src, ratio = self._sample(torch.sigmoid(self.p["opacity"])[alive_idx, 0], dead_idx.shape[0], alive_idx)
        new_o, new_s = self._update_params(src, ratio)
        for k in ("xyz", "shs", "rotation"):
            self.p[k][dead_idx] = self.p[k][src]
        self.p["opacity"][dead_idx] = new_o
        self.p["scaling"][dead_idx] = new_s
        self.p["opacity"][src] = new_o
        self.p["scaling"][src] = new_s
        self._reset_moments(src)
        return int(dead_idx.shape[0])

    @torch.no_grad()
    def add_new(self, cap_max):                     # :460-492 + cat_tensors_to_optimizer :302-327
        cur = self.p["xyz"].shape[0]
        num = max(0, min(cap_max, int(1.05 * cur)) - cur)
        if num <= 0:
            return 0
        src, ratio = self._sample(torch.sigmoid(self.p["opacity"]).squeeze(-1), num)
        new_o, new_s = self._update_params(src, ratio)
        self.p["opacity"][src] = new_o
        self.p["scaling"][src] = new_s
        ext = {"xyz": self.p["xyz"][src], "shs": self.p["shs"][src], "opacity": new_o, "scaling": new_s, "rotation": self.p["rotation"][src]}
        for group in self.opt.param_groups:
            k = group["name"]
            old = group["params"][0]
            st = self.opt.state.get(old, None)
            new = torch.nn.Parameter(torch.cat((old.detach(), ext[k]), 0).contiguous().requires_grad_(True))
            if st is not None:
                st["exp_avg"] = torch.cat((st["exp_avg"], torch.zeros_like(ext[k])), 0)
                st["exp_avg_sq"] = torch.cat((st["exp_avg_sq"], torch.zeros_like(ext[k])), 0)
                del self.opt.state[old]
                self.opt.state[new] = st
            group["params"][0] = new
            self.p[k] = new
        self._reset_moments(src)
        return num


def test_mini_mcmc_training_fits_a_target():
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer, add_noise
    from diff_gaussian_rasterization import loss as fused_loss
    from diff_gaussian_rasterization.activations import activate
    torch.manual_seed(0)
    dev = torch.device("cuda:0")
    W = H = 200
    P0, cap = 6000, 7000
    cams = [gs_synth.to_torch(c, dev) for c in gs_synth.orbit_cameras(4, 4.03, W, H, 0.6911, 20.0)]
    bg = torch.zeros(3, device=dev)

    def render(m, c, raw=None):
        settings = GaussianRasterizationSettings(H, W, c["tanfovx"], c["tanfovy"], bg, 1.0, c["viewmatrix"], c["projmatrix"], 3,
                                                 c["campos"], False, False)
        opacity, scaling, rotation, reg = activate(m.p["opacity"], m.p["scaling"], m.p["rotation"], 0.01, 0.01)
        means2D = torch.zeros_like(m.p["xyz"], requires_grad=True)
        image, radii, used = GaussianRasterizer(settings)(means3D=m.p["xyz"], means2D=means2D, shs=m.p["shs"], opacities=opacity,
                                                         scales=scaling, rotations=rotation)
        return image, reg

    # target: a differently coloured, differently seeded cloud rendered with the same rasterizer
    target_model = MiniModel(gs_synth.scene_cube(P0, seed=11), dev)
    with torch.no_grad():
        target_model.p["shs"][:, 0, :] = torch.rand((P0, 3), device=dev) * 2
        targets = [render(target_model, c)[0].detach() for c in cams]
    m = MiniModel(gs_synth.scene_cube(P0, seed=12), dev)
    with torch.no_grad():
        m.p["opacity"][: P0 // 20] = -8.0            # 5 % dead on arrival: sigmoid(-8) < 0.005
    losses, revived, added = [], 0, 0
    for it in range(1, 61):
        c, gt = cams[it % len(cams)], targets[it % len(cams)]
        image, reg = render(m, c)
        loss = fused_loss.photometric_loss(image, gt, 0.2) + reg
        loss.backward()
        losses.append(float(loss.item()))
        if it % 20 == 0:
            revived += m.relocate()
            added += m.add_new(cap)
        m.opt.step()
        m.opt.zero_grad(set_to_none=True)
        with torch.no_grad():
            add_noise(m.p["xyz"], m.p["scaling"], m.p["rotation"], m.p["opacity"], 5e5 * 1.6e-4 * 5, seed=3, offset=it)
    torch.cuda.synchronize()
    first, last = np.mean(losses[:5]), np.mean(losses[-5:])
    assert np.isfinite(losses).all()
    assert last < 0.8 * first, (first, last)
    assert revived >= P0 // 20                                  # the dead ones were moved onto live ones
    assert added > 0 and m.p["xyz"].shape[0] == min(cap, P0 + added)
    # add_new_gs replaces the Parameters between backward and optimizer.step (train_static.py:123-130), so those three
    # iterations have no gradient to apply - in the reference too
    assert int(m.opt.state[m.p["xyz"]]["step"]) == 60 - 3
    assert m.opt.state[m.p["shs"]]["exp_avg"].shape == m.p["shs"].shape
    for k, p in m.p.items():
        assert torch.isfinite(p).all(), k
    assert float((torch.sigmoid(m.p["opacity"]) <= 0.005).float().mean()) < 0.05
