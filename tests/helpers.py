"""Shared helpers for the test-suite: path setup, scene -> tensor plumbing, and the parity comparison between
the B200 implementation (through its reference-facing Python API / C ABI) and the reference CUDA rasterizer
(oracle/_ref, golden) or the CPU oracle."""
import math
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "3dgs-mcmc_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

import gs_synth  # noqa: E402


def make_case(kind, P, W, H, seed=0, fovx=0.6911, cam_radius=4.03, cam_index=0, ncams=8, elevation=20.0):
    """Numpy scene + one camera. kind: 'cube' | 'garden'."""
    scene = gs_synth.scene_cube(P, seed) if kind == "cube" else gs_synth.scene_garden(P, seed)
    cam = gs_synth.orbit_cameras(ncams, cam_radius, W, H, fovx, elevation_deg=elevation)[cam_index]
    return scene, cam


def torch_inputs(scene, cam, device, bg=(0.0, 0.0, 0.0)):
    import torch
    s = gs_synth.to_torch(scene, device)
    c = gs_synth.to_torch(cam, device)
    c["bg"] = torch.tensor(bg, dtype=torch.float32, device=device)
    return s, c


def native_forward_args(s, c, degree=3, scale_modifier=1.0, colors_precomp=None, cov3D_precomp=None, debug=False):
    """Positional args of _C.rasterize_gaussians / oracle.ref_cuda.forward (same order in both)."""
    import torch
    e = torch.empty(0)
    use_sh = colors_precomp is None
    use_sr = cov3D_precomp is None
    return (c["bg"], s["means3D"], e if use_sh else colors_precomp, s["opacities"],
            s["scales"] if use_sr else e, s["rotations"] if use_sr else e, scale_modifier,
            e if use_sr else cov3D_precomp, c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"],
            c["image_height"], c["image_width"], s["shs"] if use_sh else e, degree, c["campos"], False, debug)


def rel_l2(a, b):
    import torch
    a = a.double().flatten()
    b = b.double().flatten()
    den = torch.linalg.norm(b).item()
    num = torch.linalg.norm(a - b).item()
    if den == 0.0:
        return 0.0 if num == 0.0 else float("inf")
    return num / den


def row_errors(a, b, rows):
    """(max, 99.9th percentile) over the selected rows of |a_i - b_i| / (|b_i| + 1e-3 * rms over rows of |b_i|)."""
    import torch
    P = rows.shape[0]
    a = a.reshape(P, -1)[rows].double()
    b = b.reshape(P, -1)[rows].double()
    if a.shape[0] == 0:
        return 0.0, 0.0
    nb = b.norm(dim=1)
    floor = 1e-3 * float(nb.pow(2).mean().sqrt().item())
    if floor == 0.0:
        return (0.0, 0.0) if float((a - b).abs().max().item()) == 0.0 else (float("inf"), float("inf"))
    err = (a - b).norm(dim=1) / (nb + floor)
    k = max(1, int(round(0.999 * err.numel())))
    return float(err.max().item()), float(err.kthvalue(k).values.item())


def bits_equal_count(a, b):
    """Number of elements whose raw bits differ."""
    import torch
    av = a.contiguous().view(torch.int32) if a.dtype == torch.float32 else a
    bv = b.contiguous().view(torch.int32) if b.dtype == torch.float32 else b
    return int((av != bv).sum().item())


def compare_with_reference(scene, cam, degree=3, bg=(0.0, 0.0, 0.0), seed=1, do_backward=True):
    """Run ours and the reference CUDA rasterizer on identical tensors; return a dict of parity metrics."""
    import torch
    from diff_gaussian_rasterization import _native as C
    from oracle import ref_cuda

    dev = torch.device("cuda:0")
    s, c = torch_inputs(scene, cam, dev, bg)
    args = native_forward_args(s, c, degree)
    P = s["means3D"].shape[0]
    W, H = c["image_width"], c["image_height"]

    R, color, radii, geom, binning, img, is_used = C.rasterize_gaussians(*args)
    torch.cuda.synchronize()
    Rr, color_r, radii_r, is_used_r = ref_cuda.forward(*args)
    torch.cuda.synchronize()

    rep = {"P": P, "W": W, "H": H, "R": R, "R_ref": Rr}
    vis = radii_r > 0
    rep["visible"] = int(vis.sum().item())
    rep["radii_mismatch"] = int((radii != radii_r).sum().item())
    f32, i32, u8, i64 = torch.float32, torch.int32, torch.uint8, torch.int64

    def mine(which, dt):
        return C.debug_fetch(which, P, R, W, H, geom, binning, img, dt)

    def ref(which, dt):
        return ref_cuda.fetch(which, dt, dev)

    both_vis = vis & (radii > 0)
    for name, which, width in (("depths", 0, 1), ("means2D", 3, 2), ("cov3D", 4, 6), ("conic_opacity", 5, 4), ("rgb", 6, 3)):
        a = mine(which, f32).view(P, width)[both_vis]
        b = ref(which, f32).view(P, width)[both_vis]
        rep[f"{name}_bit_mismatch"] = bits_equal_count(a, b)
        rep[f"{name}_max_abs"] = float((a - b).abs().max().item()) if a.numel() else 0.0
    rep["tiles_touched_mismatch"] = int((mine(7, i32) != ref(7, i32)).sum().item())
    # ours scans tiles_touched in depth order (the reference in index order): same total, and the scan must be the
    # running sum over our depth order, which must itself be the stable (depth, index) order of the visible set
    order = mine(9, i32).long()
    tt = mine(7, i32).long()
    rep["point_offsets_mismatch"] = int((torch.cumsum(tt[order], 0) != mine(8, i32).long()).sum().item())
    rep["point_offsets_mismatch"] += int(ref(8, i32)[-1].item() != R) if P else 0
    dbits = mine(0, f32).view(i32).long()
    dkey = torch.where(tt > 0, dbits, torch.full_like(dbits, 0x7FFFFFFF))
    composite = (dkey[order] << 32) | order
    rep["depth_order_mismatch"] = int((composite[1:] <= composite[:-1]).sum().item()) + int(order.unique().numel() != P)
    cl_m = mine(1, u8)[both_vis]
    cl_r = ref(1, u8).view(P, 3)[both_vis]
    cl_r_bits = (cl_r[:, 0] | (cl_r[:, 1] << 1) | (cl_r[:, 2] << 2)).to(u8)
    rep["clamped_mismatch"] = int((cl_m != cl_r_bits).sum().item())

    if R == Rr and R > 0:
        rep["sorted_keys_mismatch"] = int((mine(12, i64) != ref(12, i64)).sum().item())
        rep["sorted_ids_mismatch"] = int((mine(10, i32) != ref(10, i32)).sum().item())
    ntiles = ((W + 15) // 16) * ((H + 15) // 16)
    rep["ranges_mismatch"] = int((mine(22, i32)[: 2 * ntiles] != ref(22, i32)[: 2 * ntiles]).sum().item())
    rep["n_contrib_mismatch"] = int((mine(21, i32) != ref(21, i32)).sum().item())
    rep["final_T_bit_mismatch"] = bits_equal_count(mine(20, f32), ref(20, f32))
    rep["image_bit_mismatch"] = bits_equal_count(color, color_r)
    rep["image_max_abs"] = float((color - color_r).abs().max().item())
    rep["is_used_mismatch"] = int((is_used != is_used_r).sum().item())
    rep["sort_error_flag"] = (int(mine(30, i32)[0].item()) if R > 0 else 0) + int(mine(31, i32)[0].item())

    if do_backward:
        g = torch.Generator(device="cpu").manual_seed(seed)
        dL = torch.randn((3, H, W), generator=g).to(dev)
        e = torch.empty(0)
        bargs = (c["bg"], s["means3D"], radii, e, s["scales"], s["rotations"], 1.0, e, c["viewmatrix"], c["projmatrix"],
                 c["tanfovx"], c["tanfovy"], dL, s["shs"], degree, c["campos"], geom, R, binning, img, False)
        names = ("means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations")
        ours = dict(zip(names, C.rasterize_gaussians_backward(*bargs)))
        torch.cuda.synchronize()
        gr = ref_cuda.backward(c["bg"], s["means3D"], radii_r, e, s["scales"], s["rotations"], 1.0, e, c["viewmatrix"],
                               c["projmatrix"], c["tanfovx"], c["tanfovy"], dL, s["shs"], degree, c["campos"], Rr)
        torch.cuda.synchronize()
        for n in names:
            rep[f"grad_{n}_rel_l2"] = rel_l2(ours[n], gr[n])
            rep[f"grad_{n}_finite"] = bool(torch.isfinite(ours[n]).all().item())
            # Per-Gaussian view of the same comparison: a whole-tensor norm ratio cannot show one small-magnitude row that
            # is wrong by 100 %. Row error = |a_i - b_i| / (|b_i| + 1e-3 * rms_i |b_i|) over the visible Gaussians (the floor
            # keeps rows at the level of the float-atomic summation noise from dominating); reported as max and 99.9th pct.
            rep[f"grad_{n}_row_max"], rep[f"grad_{n}_row_p999"] = row_errors(ours[n], gr[n], both_vis)
            rep[f"grad_{n}_culled_nonzero"] = int((ours[n].reshape(P, -1)[~(radii > 0)] != 0).any(1).sum().item())
    return rep
