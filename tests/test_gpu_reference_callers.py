"""The reference's OWN Python callers, executed unchanged on this package (SURVEY.md §8 rows a21, a23 and the north
star's "gaussian_renderer/__init__.py, scene/gaussian_model.py and train_static.py run unchanged").

The files are the reference's, staged byte for byte under baseline/_ref/py by tools/stage_reference.py (git-ignored,
shipped by gpurun); tests/refenv.py makes them importable with `diff_gaussian_rasterization` and `simple_knn` resolving
to THIS repository and stubs for the absent third-party `plyfile` / `open3d`. What runs here:

  * gaussian_renderer.render() and deformable_render()        (gaussian_renderer/__init__.py:19-172)
  * utils.reloc_utils (scipy binomial table, in-place clamp)    (utils/reloc_utils.py:7-15)
  * GaussianModel._update_params / _sample_alives / relocate_gs / add_new_gs and the optimizer surgery
                                                                (scene/gaussian_model_static.py:302-492)
  * train_static.py as __main__: argument parsing, training(), the loop body, training_report  (train_static.py:36-241)

compared with the same callers bound to the reference's own CUDA (oracle/_ref/libref_dgr.so through
oracle/ref_rasterizer.py). The only stand-in is the dataset: DynamicScene reads a SwinGS capture from disk, which does
not exist on the box, so a seeded synthetic scene object with the same methods takes its place.
"""
import math
import os
import random
import runpy
import sys
import types

import numpy as np
import pytest
import torch

import helpers
import gs_synth
import refenv

pytestmark = pytest.mark.gpu

GRAD_TOL = 1e-3


def dev():
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def ref_modules():
    """The staged reference modules (imported once: utils/reloc_utils.py builds its table on the GPU at import)."""
    try:
        refenv.install()
    except RuntimeError as ex:
        pytest.fail(str(ex))
    from oracle import ref_cuda
    if not ref_cuda.available():
        pytest.fail("oracle/_ref/libref_dgr.so did not travel with the snapshot")
    import gaussian_renderer
    import scene.cameras as cameras
    import scene.gaussian_model as gm_swin
    import scene.gaussian_model_static as gm_static
    import utils.graphics_utils as graphics_utils
    import utils.reloc_utils as reloc_utils
    import arguments
    return types.SimpleNamespace(gaussian_renderer=gaussian_renderer, cameras=cameras, gm_swin=gm_swin, gm_static=gm_static,
                                 graphics_utils=graphics_utils, reloc_utils=reloc_utils, arguments=arguments, ref_cuda=ref_cuda)


# ---------------------------------------------------------------------------------------------------------------------
# synthetic stand-ins for the dataset
# ---------------------------------------------------------------------------------------------------------------------
def make_cameras(m, n, W, H, radius=4.03, fovx=0.6911, truth=None):
    """n reference `Camera` objects (scene/cameras.py:21-74) on an orbit, built from R / T / FoV like the reference's
    loaders build them; ground-truth images rendered from `truth` (activated gs_synth scene) with this package."""
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    cams = []
    for i, c in enumerate(gs_synth.orbit_cameras(n, radius, W, H, fovx, 20.0)):
        w2c = np.asarray(c["viewmatrix"], dtype=np.float64).T
        R, T = w2c[:3, :3].T, w2c[:3, 3]          # the loaders' convention: R = camera-to-world rotation, T = w2c translation
        fovy = 2.0 * math.atan(c["tanfovy"])
        image = torch.zeros((3, H, W))
        cam = m.cameras.Camera(colmap_id=i, R=R, T=T, FoVx=fovx, FoVy=fovy, image=image, gt_alpha_mask=None,
                               image_name=f"synthetic_{i:03d}", uid=i, data_device="cuda")
        cam.frame = 0
        if truth is not None:
            st = GaussianRasterizationSettings(H, W, math.tan(fovx * 0.5), math.tan(fovy * 0.5), torch.zeros(3, device=dev()), 1.0,
                                               cam.world_view_transform, cam.full_proj_transform, 3, cam.camera_center, False, False)
            with torch.no_grad():
                img, _, _ = GaussianRasterizer(st)(means3D=truth["means3D"], means2D=torch.zeros_like(truth["means3D"]),
                                                   shs=truth["shs"], opacities=truth["opacities"], scales=truth["scales"],
                                                   rotations=truth["rotations"])
            cam.original_image = img.clamp(0.0, 1.0)
        cams.append(cam)
    return cams


def make_static_model(m, P, seed, randomise=True):
    """GaussianModel initialised by its own create_from_pcd (-> this package's distCUDA2), then perturbed so that
    every parameter family matters."""
    rng = np.random.default_rng(seed)
    pts = rng.uniform(-1.3, 1.3, (P, 3)).astype(np.float32)
    cols = rng.uniform(0, 1, (P, 3)).astype(np.float32)
    pcd = m.graphics_utils.BasicPointCloud(points=pts, colors=cols, normals=np.zeros((P, 3), np.float32))
    g = m.gm_static.GaussianModel(3)
    g.create_from_pcd(pcd, 1.0)
    if randomise:
        gen = torch.Generator(device="cpu").manual_seed(seed)
        with torch.no_grad():
            g._features_rest.copy_((torch.randn(g._features_rest.shape, generator=gen) * 0.15).to(dev()))
            g._opacity.copy_((torch.randn(g._opacity.shape, generator=gen) * 1.5).to(dev()))
            g._rotation.copy_(torch.randn(g._rotation.shape, generator=gen).to(dev()))
            g._scaling.add_((torch.randn(g._scaling.shape, generator=gen) * 0.4 + 0.8).to(dev()))
        g.active_sh_degree = 3
    return g


def pipe_params(**kw):
    p = types.SimpleNamespace(convert_SHs_python=False, compute_cov3D_python=False, debug=False)
    p.__dict__.update(kw)
    return p


def opt_params(m, *argv):
    from argparse import ArgumentParser
    parser = ArgumentParser()
    op = m.arguments.OptimizationParams(parser)
    return op.extract(parser.parse_args(list(argv)))


# ---------------------------------------------------------------------------------------------------------------------
# render() / deformable_render()
# ---------------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("variant", ["rasterizer_sh_and_cov", "convert_SHs_python", "compute_cov3D_python"])
def test_render_unchanged_equals_reference_cuda(ref_modules, variant):
    """gaussian_renderer.render() (reference file, unchanged) through this package == the same function bound to the
    reference's CUDA rasterizer: image / radii / is_used bit for bit, every parameter gradient to 1e-3."""
    from oracle import ref_rasterizer
    m = ref_modules
    W, H = 250, 130
    g = make_static_model(m, 20000, seed=3)
    cam = make_cameras(m, 4, W, H)[1]
    pipe = pipe_params(**({variant: True} if variant != "rasterizer_sh_and_cov" else {}))
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev())
    dL = torch.randn((3, H, W), generator=torch.Generator().manual_seed(11)).to(dev())
    params = (g._xyz, g._features_dc, g._features_rest, g._opacity, g._scaling, g._rotation)

    def run():
        for p in params:
            p.grad = None
        out = m.gaussian_renderer.render(cam, g, pipe, bg)
        (out["render"] * dL).sum().backward()
        torch.cuda.synchronize()
        return out, [p.grad.clone() for p in params], out["viewspace_points"].grad.clone()

    ours, g_ours, vs_ours = run()
    with ref_rasterizer.bound_to(m.gaussian_renderer):
        theirs, g_ref, vs_ref = run()
    assert ours["render"].shape == (3, H, W) and int(ours["visibility_filter"].sum()) > 1000
    assert torch.equal(ours["radii"], theirs["radii"]) and torch.equal(ours["is_used"], theirs["is_used"])
    assert torch.equal(ours["visibility_filter"], theirs["visibility_filter"])
    assert torch.equal(ours["render"], theirs["render"])
    names = ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation")
    for n, a, b in zip(names, g_ours, g_ref):
        assert torch.isfinite(a).all() and float(b.abs().max()) > 0, n
        assert helpers.rel_l2(a, b) < GRAD_TOL, (n, helpers.rel_l2(a, b))
    assert helpers.rel_l2(vs_ours, vs_ref) < GRAD_TOL


def make_swin_model(m, P, seed):
    rng = np.random.default_rng(seed)
    pts = rng.uniform(-1.3, 1.3, (P, 3)).astype(np.float32)
    cols = rng.uniform(0, 1, (P, 3)).astype(np.float32)
    pcd = m.graphics_utils.BasicPointCloud(points=pts, colors=cols, normals=np.zeros((P, 3), np.float32))
    g = m.gm_swin.SwinGaussianModel(3, max_lifespan=30, matured_buffer_size=10, deform=True, dump_path="")
    g.create_from_pcd(pcd, 1.0)
    gen = torch.Generator(device="cpu").manual_seed(seed)
    rnd = lambda t, s: (torch.randn(t.shape, generator=gen) * s).to(dev())  # noqa: E731
    with torch.no_grad():
        g._features_rest.copy_(rnd(g._features_rest, 0.15))
        g._opacity.copy_(rnd(g._opacity, 1.5))
        g._rotation.copy_(rnd(g._rotation, 1.0))
        g._scaling.add_(rnd(g._scaling, 0.4) + 0.8)
        g._rigid_v.copy_(rnd(g._rigid_v, 0.02))
        g._rigid_rotvec.copy_(rnd(g._rigid_rotvec, 0.05))
        g._rigid_rotcen.add_(rnd(g._rigid_rotcen, 0.1))
        g._frame_start.copy_(torch.randint(0, 12, g._frame_start.shape, generator=gen).float().to(dev()))
    g.active_sh_degree = 3
    return g


def test_deformable_render_unchanged_equals_reference_cuda(ref_modules):
    """deformable_render() (gaussian_renderer/__init__.py:105-172) with SwinGaussianModel.get_basic_para_at
    (scene/gaussian_model.py:597-651, utils/tempo_utils.py rigid_deform): same outputs and gradients through this package
    and through the reference's CUDA; the fused gather (diff_gaussian_rasterization.swin) agrees with the reference's
    op sequence on the same model."""
    from oracle import ref_rasterizer
    m = ref_modules
    W, H = 250, 130
    g = make_swin_model(m, 15000, seed=5)
    cam = make_cameras(m, 4, W, H)[2]
    cam.frame = 7
    pipe = pipe_params()
    bg = torch.zeros(3, device=dev())
    dL = torch.randn((3, H, W), generator=torch.Generator().manual_seed(12)).to(dev())
    names = ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation", "_rigid_v", "_rigid_rotvec", "_rigid_rotcen")

    def run():
        for n in names:
            getattr(g, n).grad = None
        out = m.gaussian_renderer.deformable_render(cam, g, pipe, bg)
        (out["render"] * dL).sum().backward()
        torch.cuda.synchronize()
        return out, {n: getattr(g, n).grad.clone() for n in names}

    ours, g_ours = run()
    with ref_rasterizer.bound_to(m.gaussian_renderer):
        theirs, g_ref = run()
    alive = int(((g._frame_start <= 7) & (g._frame_end > 7)).sum())
    assert 0 < alive < 15000 and ours["radii"].shape == (alive,)
    assert torch.equal(ours["render"], theirs["render"]) and torch.equal(ours["radii"], theirs["radii"])
    assert torch.equal(ours["is_used"], theirs["is_used"])
    for n in names:
        assert helpers.rel_l2(g_ours[n], g_ref[n]) < GRAD_TOL, (n, helpers.rel_l2(g_ours[n], g_ref[n]))
    # the fused gather (f2) against the reference's own op sequence on this model
    from diff_gaussian_rasterization import swin
    imm = {"xyz": g._xyz, "rotation": g._rotation, "features_dc": g._features_dc, "features_rest": g._features_rest,
           "opacity": g._opacity, "scaling": g._scaling, "rigid_v": g._rigid_v, "rigid_rotvec": g._rigid_rotvec,
           "rigid_rotcen": g._rigid_rotcen, "frame_start": g._frame_start}
    ii, mi = g.derive_idx_of_active(7)
    fused = swin.gather_active(7.0, imm, None, ii, mi, deform=True)
    want = ours["input_gaussians"]
    for k in ("xyz", "rotation", "feature", "opacity", "scaling"):
        assert torch.allclose(fused[k], want[k].detach(), rtol=2e-5, atol=2e-6), k


# ---------------------------------------------------------------------------------------------------------------------
# MCMC callers
# ---------------------------------------------------------------------------------------------------------------------
def test_reloc_utils_table_and_clamp(ref_modules):
    """utils/reloc_utils.py: the scipy-built binomial table equals the oracle's, and compute_relocation_cuda clamps N in
    place to [1, 50] and hands an int64 tensor to this package."""
    from oracle import oracle
    m = ref_modules
    assert m.reloc_utils.binoms.is_cuda and m.reloc_utils.N_max == 51
    assert np.array_equal(m.reloc_utils.binoms.cpu().numpy(), oracle.binoms_table(51))
    N = torch.tensor([0, 1, 7, 50, 51, 400], device=dev())
    op = torch.tensor([0.5, 0.5, 0.2, 0.9, 0.9, 0.9], device=dev())
    sc = torch.full((6, 3), 0.01, device=dev())
    o, s = m.reloc_utils.compute_relocation_cuda(op, sc, N)
    assert N.tolist() == [1, 1, 7, 50, 50, 50]
    o2, s2 = m.ref_cuda.compute_relocation(op, sc, N, m.reloc_utils.binoms, 51)
    assert torch.equal(o, o2) and torch.equal(s, s2) and s.shape == (18,)


def _model_state(g):
    st = {n: getattr(g, n).detach().clone() for n in ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation")}
    for group in g.optimizer.param_groups:
        s = g.optimizer.state.get(group["params"][0])
        if s:
            st["m_" + group["name"]], st["v_" + group["name"]] = s["exp_avg"].clone(), s["exp_avg_sq"].clone()
    return st


@pytest.mark.parametrize("optimizer", ["torch.optim.Adam", "FusedAdam"])
def test_relocate_and_add_new_gs_unchanged_equal_reference_kernel(ref_modules, optimizer):
    """GaussianModel.relocate_gs / add_new_gs (reference file, unchanged) under a fixed seed: with this package's
    compute_relocation and with the reference's kernel the sampled indices, new opacities / scales, grown tensors and
    the optimizer state after the surgery are identical bit for bit — against torch.optim.Adam and against FusedAdam."""
    from diff_gaussian_rasterization.optim import FusedAdam
    m = ref_modules
    opt = opt_params(m)
    W, H = 160, 96
    cam = make_cameras(m, 4, W, H)[0]
    results = []
    for kernel in ("ours", "reference"):
        g = make_static_model(m, 12000, seed=9)
        with torch.no_grad():
            g._opacity[::7] = -7.0          # opacity sigmoid(-7) < 0.005: dead
        saved = torch.optim.Adam
        try:
            if optimizer == "FusedAdam":
                torch.optim.Adam = FusedAdam
            g.training_setup(opt)
        finally:
            torch.optim.Adam = saved
        assert type(g.optimizer).__name__ == ("FusedAdam" if optimizer == "FusedAdam" else "Adam")
        # two optimizer steps on seeded gradients so that the moments are non-zero before the surgery (no render here:
        # its float atomics would make the two runs differ before the code under test is reached)
        gen = torch.Generator(device="cpu").manual_seed(77)

        def seeded_step():
            for group in g.optimizer.param_groups:
                p_ = group["params"][0]
                p_.grad = (torch.randn(p_.shape, generator=gen) * 1e-3).to(dev())
            g.optimizer.step()
            g.optimizer.zero_grad(set_to_none=True)
        seeded_step()
        seeded_step()
        patched = m.reloc_utils.compute_relocation
        if kernel == "reference":
            m.reloc_utils.compute_relocation = lambda o, s, N, b, n: m.ref_cuda.compute_relocation(o, s, N, b, n)
        try:
            torch.manual_seed(1234)
            with torch.no_grad():
                dead = (g.get_opacity <= 0.005).squeeze(-1)
                n_dead = int(dead.sum())
                g.relocate_gs(dead_mask=dead)
                added = g.add_new_gs(cap_max=13000)
        finally:
            m.reloc_utils.compute_relocation = patched
        torch.cuda.synchronize()
        assert n_dead > 1000 and added == min(13000, int(1.05 * 12000)) - 12000
        assert g._xyz.shape[0] == 12000 + added and g.optimizer.state[g._xyz]["exp_avg"].shape[0] == 12000 + added
        assert int((g.get_opacity <= 0.005).sum()) < n_dead // 4       # the dead ones were moved onto live ones
        # one more step on the grown model: the optimizer accepts the swapped parameters; and a render of it
        seeded_step()
        out = m.gaussian_renderer.render(cam, g, pipe_params(), torch.zeros(3, device=dev()))
        assert torch.isfinite(out["render"]).all() and out["radii"].shape[0] == 12000 + added
        results.append(_model_state(g))
    a, b = results
    assert a.keys() == b.keys() and len(a) == 18
    for k in a:      # same seed, same sampled indices, same relocation bits, same optimizer surgery
        assert torch.equal(a[k], b[k]), k


def test_relocation_outputs_bit_identical_inside_the_caller(ref_modules):
    """_update_params (gaussian_model_static.py:395-411) called as relocate_gs calls it: the tensors it returns with this
    package are bit-identical to those with the reference kernel (no optimizer step in between)."""
    m = ref_modules
    g = make_static_model(m, 30000, seed=2)
    g.training_setup(opt_params(m))
    torch.manual_seed(7)
    with torch.no_grad():
        probs = g.get_opacity.squeeze(-1)
        idx, ratio = g._sample_alives(probs=probs, num=5000)
        ours = g._update_params(idx, ratio=ratio)
        saved = m.reloc_utils.compute_relocation
        m.reloc_utils.compute_relocation = lambda o, s, N, b, n: m.ref_cuda.compute_relocation(o, s, N, b, n)
        try:
            theirs = g._update_params(idx, ratio=ratio)
        finally:
            m.reloc_utils.compute_relocation = saved
    assert int(ratio.max()) >= 2
    for a, b in zip(ours, theirs):
        assert torch.equal(a, b)


# ---------------------------------------------------------------------------------------------------------------------
# train_static.py as __main__
# ---------------------------------------------------------------------------------------------------------------------
class SyntheticDynamicScene:
    """Stand-in for scene.DynamicScene (scene/__init__.py:97-330), whose constructor reads a SwinGS capture from disk:
    same methods as train_static.py uses, cameras and ground truth from a seeded synthetic scene."""
    last = None
    P0, W, H, NCAMS = 20000, 200, 200, 8

    def __init__(self, args, gaussians, load_iteration=None, shuffle=True, resolution_scales=(1.0,)):
        import scene.cameras as cameras
        import utils.graphics_utils as graphics_utils
        m = types.SimpleNamespace(cameras=cameras, graphics_utils=graphics_utils)
        self.model_path, self.gaussians, self.max_frame = args.model_path, gaussians, 1
        truth = gs_synth.to_torch(gs_synth.scene_cube(self.P0, seed=21), dev())
        truth["scales"] = truth["scales"] * 2.0
        self.cams = make_cameras(m, self.NCAMS, self.W, self.H, truth=truth)
        self.test_cams = self.cams[:2]
        rng = np.random.default_rng(4)
        keep = rng.permutation(self.P0)[: self.P0 // 2]
        pts = truth["means3D"].cpu().numpy()[keep] + rng.normal(0, 0.02, (len(keep), 3)).astype(np.float32)
        pcd = graphics_utils.BasicPointCloud(points=pts.astype(np.float32), colors=rng.uniform(0, 1, (len(keep), 3)).astype(np.float32),
                                             normals=np.zeros((len(keep), 3), np.float32))
        self.cameras_extent = 1.3
        gaussians.create_from_pcd(pcd, self.cameras_extent)
        self.saved = []
        SyntheticDynamicScene.last = self

    def clearAll(self):
        pass

    def clearAllTest(self):
        pass

    def getTrainCamerasAt(self, frames, scale=1.0):
        return list(self.cams)

    def getTestCamerasAt(self, frames, scale=1.0):
        return list(self.test_cams)

    def save(self, iteration):
        import gs_io
        g = self.gaussians
        path = os.path.join(self.model_path, f"point_cloud/iteration_{iteration}")
        os.makedirs(path, exist_ok=True)
        gs_io.save_ply(os.path.join(path, "point_cloud.ply"), g._xyz, g._features_dc, g._features_rest, g._opacity, g._scaling, g._rotation)
        self.saved.append(iteration)


def _run_train_static(tmp, iterations, fused_adam):
    import scene
    import utils.loss_utils  # noqa: F401  (train_static.py binds l1_loss from here)
    from diff_gaussian_rasterization.optim import FusedAdam
    saved = (scene.DynamicScene, sys.argv, sys.stdout, torch.optim.Adam, random.getstate())
    losses = []
    l1 = sys.modules["utils.loss_utils"].l1_loss

    def l1_probe(a, b):          # records the L1 term every iteration; same function object contract
        v = l1(a, b)
        losses.append(v.detach())
        return v
    try:
        scene.DynamicScene = SyntheticDynamicScene
        sys.modules["utils.loss_utils"].l1_loss = l1_probe
        if fused_adam:
            torch.optim.Adam = FusedAdam
        sys.argv = ["train_static.py", "-s", str(tmp), "-m", str(tmp / ("fused" if fused_adam else "stock")), "--cap_max", "12000",
                    "--iterations", str(iterations), "--densify_from_iter", "20", "--densification_interval", "25",
                    "--test_iterations", "50", "--quiet", "--opacity_reg", "0.01", "--scale_reg", "0.01"]
        runpy.run_path(os.path.join(refenv.REF_PY, "train_static.py"), run_name="__main__")
    finally:
        scene.DynamicScene, sys.argv, sys.stdout, torch.optim.Adam = saved[:4]
        random.setstate(saved[4])
        sys.modules["utils.loss_utils"].l1_loss = l1
    torch.cuda.synchronize()
    sc = SyntheticDynamicScene.last
    return sc, torch.stack(losses).cpu().numpy()


@pytest.mark.parametrize("fused_adam", [False, True])
def test_train_static_main_runs_unchanged(ref_modules, tmp_path, fused_adam):
    """train_static.py executed as __main__ (its own argument parser, training(), loop body :84-141, training_report) for
    200 iterations on this package: loss falls, the MCMC growth step enlarges the model 5 % at a time up to cap_max, the
    optimizer state follows, the final save goes through. Once with the reference's torch.optim.Adam, once with FusedAdam
    swapped in for it."""
    sc, losses = _run_train_static(tmp_path, 200, fused_adam)
    g = sc.gaussians
    assert len(losses) >= 200 and np.isfinite(losses).all()
    assert losses[-20:].mean() < 0.7 * losses[:10].mean(), (losses[:10].mean(), losses[-20:].mean())
    # densification at iterations 25, 50, ..., 175: 10000 -> 10500 -> 11025 -> 11576 -> 12000 (cap)
    assert g._xyz.shape[0] == 12000
    assert type(g.optimizer).__name__ == ("FusedAdam" if fused_adam else "Adam")
    assert g.optimizer.state[g._xyz]["exp_avg"].shape == g._xyz.shape
    assert sc.saved == [200] and os.path.exists(os.path.join(sc.model_path, "point_cloud/iteration_200/point_cloud.ply"))
    assert all(torch.isfinite(getattr(g, n)).all() for n in ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation"))
