"""Pin the CPU oracle (oracle/oracle_cpu.c) against golden vectors produced by the REFERENCE's own CUDA code
(tests/golden/*.npz, generated on a B200 by tests/golden/make_golden.py from oracle/_ref). No GPU needed.

Integer stages are checked bit-exactly on the golden float inputs of that stage (the CPU's glibc expf / missing FMA
contraction make CPU floats differ from the GPU's in the last ulp, so each stage is fed the reference's own outputs of
the previous stage). Float stages are checked to the tolerances written beside each assert.
"""
import glob
import os

import numpy as np
import pytest

from oracle import oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SCENES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLD, "*.npz")) if os.path.basename(p).startswith(("cube_", "garden_")))


def load(name):
    z = np.load(os.path.join(GOLD, name + ".npz"))
    scene = {k[3:]: z[k] for k in z.files if k.startswith("in_")}
    cam = {}
    for k in z.files:
        if k.startswith("cam_"):
            v = z[k]
            cam[k[4:]] = v if v.ndim else v.item()
    return z, scene, cam


def rel_l2(a, b):
    a = a.astype(np.float64).ravel()
    b = b.astype(np.float64).ravel()
    d = np.linalg.norm(b)
    return np.linalg.norm(a - b) / d if d else np.linalg.norm(a - b)


def test_golden_files_present():
    assert len(SCENES) >= 4 and os.path.exists(os.path.join(GOLD, "relocation.npz"))


@pytest.mark.parametrize("name", SCENES)
def test_preprocess_matches_reference(name):
    z, scene, cam = load(name)
    W, H, deg = cam["image_width"], cam["image_height"], int(z["degree"])
    pre = oracle.preprocess(scene["means3D"], scene["scales"], scene["rotations"], scene["opacities"], scene["shs"],
                            cam["viewmatrix"], cam["projmatrix"], cam["campos"], W, H, cam["tanfovx"], cam["tanfovy"], deg)
    vis = z["radii"] > 0
    # radius = ceil(3 sqrt(lambda)) can flip by one on a last-ulp difference: allow <= 0.5 % of Gaussians
    bad = int((pre["radii"] != z["radii"]).sum())
    assert bad <= max(1, len(vis) // 200), f"{bad} radii differ"
    both = vis & (pre["radii"] > 0)
    assert np.allclose(pre["depths"][both], z["st_depths"][both], rtol=1e-6, atol=1e-6)
    assert np.abs(pre["means2D"][both] - z["st_means2D"][both]).max() < 2e-3          # pixels
    assert rel_l2(pre["cov3D"][both], z["st_cov3D"][both]) < 1e-6
    assert rel_l2(pre["conic_opacity"][both], z["st_conic_opacity"][both]) < 1e-5
    assert np.abs(pre["rgb"][both] - z["st_rgb"][both]).max() < 1e-5
    assert int((pre["clamped"][both] != z["st_clamped"][both]).sum()) <= 1
    same = both & (pre["radii"] == z["radii"])
    assert int((pre["tiles_touched"][same] != z["st_tiles_touched"].astype(np.uint32)[same]).sum()) <= 1
    assert np.array_equal(oracle.mark_visible(scene["means3D"], cam["viewmatrix"]), z["mark_visible"])


@pytest.mark.parametrize("name", SCENES)
def test_binning_bit_exact_on_reference_inputs(name):
    z, scene, cam = load(name)
    W, H = cam["image_width"], cam["image_height"]
    b = oracle.bin(z["st_means2D"], z["st_depths"], z["radii"], z["st_tiles_touched"].astype(np.uint32), W, H)
    assert b["R"] == int(z["R"])
    assert np.array_equal(b["point_offsets"], z["st_point_offsets"].astype(np.uint32))
    assert np.array_equal(b["keys"], z["st_keys"].astype(np.uint64))
    assert np.array_equal(b["point_list"], z["st_point_list"].astype(np.uint32))
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    assert np.array_equal(b["ranges"], z["st_ranges"][:tiles].astype(np.uint32))
    k = b["keys"]
    assert bool((k[1:] >= k[:-1]).all())


@pytest.mark.parametrize("name", SCENES)
def test_render_forward_on_reference_inputs(name):
    z, scene, cam = load(name)
    W, H = cam["image_width"], cam["image_height"]
    P = scene["means3D"].shape[0]
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    r = oracle.render_fwd(W, H, z["st_ranges"][:tiles].astype(np.uint32), z["st_point_list"].astype(np.uint32),
                          z["st_means2D"], z["st_conic_opacity"], z["st_rgb"], z["bg"], P)
    # tolerance from north_star: max abs pixel error 1e-4, allowing a handful of threshold-flip pixels
    err = np.abs(r["color"] - z["color"])
    assert int((err > 1e-4).sum()) <= 3 * max(1, W * H // 2000), f"max err {err.max()}"
    assert int((r["n_contrib"] != z["st_n_contrib"].astype(np.uint32)).sum()) <= max(1, W * H // 1000)
    assert np.abs(r["final_T"] - z["st_final_T"]).max() < 1e-4 or int((np.abs(r["final_T"] - z["st_final_T"]) > 1e-5).sum()) <= 3
    assert int((r["is_used"] != z["is_used"]).sum()) <= 2


@pytest.mark.parametrize("name", SCENES)
def test_backward_on_reference_state(name):
    z, scene, cam = load(name)
    W, H, deg = cam["image_width"], cam["image_height"], int(z["degree"])
    P = scene["means3D"].shape[0]
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    st = dict(ranges=z["st_ranges"][:tiles].astype(np.uint32), point_list=z["st_point_list"].astype(np.uint32), bg=z["bg"],
              means2D=z["st_means2D"], conic_opacity=z["st_conic_opacity"], rgb=z["st_rgb"], final_T=z["st_final_T"],
              n_contrib=z["st_n_contrib"].astype(np.uint32), radii=z["radii"], clamped=z["st_clamped"], cov3D=z["st_cov3D"])
    g = oracle.backward(scene, cam, st, z["dL_dpix"], degree=deg)
    # north_star tolerance: gradient relative L2 <= 1e-3
    for name_, ref in (("means2D", z["grad_means2D"]), ("colors", z["grad_colors"]), ("opacity", z["grad_opacity"]),
                       ("means3D", z["grad_means3D"]), ("cov3D", z["grad_cov3D"]), ("sh", z["grad_sh"]),
                       ("scales", z["grad_scales"]), ("rotations", z["grad_rotations"])):
        assert rel_l2(g[name_], ref) < 1e-3, name_
    assert rel_l2(g["conic"][:, [0, 1, 3]], z["grad_conic"].reshape(P, 4)[:, [0, 1, 3]]) < 1e-3


def test_relocation_matches_reference():
    z = np.load(os.path.join(GOLD, "relocation.npz"))
    o, s = oracle.compute_relocation(z["opacity_old"], z["scale_old"], z["N"], z["binoms"], 51)
    # o' = 1 - powf(1 - o, 1/N): glibc's and CUDA's powf differ by an ulp of a value near 1 (6e-8), which the
    # subtraction turns into an ABSOLUTE error of that size; tolerance 2e-7 absolute. Scales: relative 2e-5.
    assert np.allclose(o, z["opacity_new"], rtol=2e-6, atol=2e-7)
    assert np.allclose(s, z["scale_new"], rtol=2e-5, atol=1e-12)
    assert set(z["N"].tolist()) >= set(range(1, 51))


def test_relocation_known_answers():
    binoms = oracle.binoms_table(51)
    # N = 1: opacity unchanged, denominator = o  ->  scale unchanged
    o, s = oracle.compute_relocation(np.array([0.3, 0.8], np.float32), np.ones((2, 3), np.float32), np.array([1, 1], np.int32), binoms, 51)
    assert np.allclose(o, [0.3, 0.8], rtol=1e-6) and np.allclose(s, 1.0, rtol=1e-6)
    # N = 2: o' = 1 - sqrt(1 - o); denom = 2 o' - o'^2 / sqrt(2)
    o, s = oracle.compute_relocation(np.array([0.5], np.float32), np.full((1, 3), 2.0, np.float32), np.array([2], np.int32), binoms, 51)
    op = 1 - np.sqrt(0.5)
    assert np.allclose(o, op, rtol=1e-6)
    assert np.allclose(s, 2.0 * 0.5 / (2 * op - op * op / np.sqrt(2)), rtol=1e-5)


def test_add_noise_matches_inline_torch():
    """The oracle against the reference's inline torch code (train_static.py:132-140, general_utils.py:78-110)
    restated on CPU tensors."""
    import torch
    rng = np.random.default_rng(2)
    P = 257
    xyz = rng.standard_normal((P, 3)).astype(np.float32)
    sc = rng.normal(-3, 0.7, (P, 3)).astype(np.float32)
    rot = rng.standard_normal((P, 4)).astype(np.float32)
    op = rng.normal(-4, 3, (P, 1)).astype(np.float32)
    noise = rng.standard_normal((P, 3)).astype(np.float32)
    lr = 5e4 * 1.6e-4
    out = oracle.add_noise(xyz, sc, rot, op, noise, lr)

    r = torch.from_numpy(rot)
    q = r / torch.sqrt((r * r).sum(1))[:, None]
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                     2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                     2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], 1).view(P, 3, 3)
    L = R @ torch.diag_embed(torch.exp(torch.from_numpy(sc)))
    cov = L @ L.transpose(1, 2)
    gate = 1 / (1 + torch.exp(-100 * ((1 - torch.sigmoid(torch.from_numpy(op))) - 0.995)))
    n = torch.from_numpy(noise) * gate * lr
    want = torch.from_numpy(xyz) + torch.bmm(cov, n.unsqueeze(-1)).squeeze(-1)
    assert np.allclose(out, want.numpy(), rtol=2e-5, atol=1e-6)
