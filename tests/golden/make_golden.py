"""Generate the golden vectors under tests/golden/ from the REFERENCE itself.

Runs on a GPU box (gpurun): drives the unmodified reference CUDA rasterizer (oracle/_ref/libref_dgr.so, built from
/root/reference by oracle/Makefile) on small seeded scenes and stores its inputs and every observable output
(per-Gaussian state, sorted keys / ids, tile ranges, image, n_contrib, final_T, is_used, the 9 gradient tensors,
compute_relocation and mark_visible results) as compressed .npz files. The reference has no tests or golden vectors
of its own; these files are the pin for the CPU oracle (tests/test_oracle_golden.py, no GPU needed)
and a second pin for the CUDA path (tests/test_gpu_golden.py).

    gpurun -- 'python tests/golden/make_golden.py gpurun_out/golden'   # then copy *.npz into tests/golden/
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers  # noqa: E402
import gs_synth  # noqa: E402

CASES = {
    # name: (kind, P, W, H, degree, bg, cam_radius, fovx, scale_boost)
    "cube_small_deg3": ("cube", 600, 64, 48, 3, (0.0, 0.0, 0.0), 4.03, 0.6911, 6.0),
    "cube_small_deg1_whitebg": ("cube", 400, 40, 40, 1, (1.0, 1.0, 1.0), 4.03, 0.6911, 8.0),
    "garden_small_deg3": ("garden", 1500, 96, 64, 3, (0.1, 0.2, 0.3), 3.0, 1.2, 1.0),
    "cube_dense_saturating": ("cube", 3000, 32, 32, 0, (0.0, 0.0, 0.0), 4.03, 0.25, 12.0),
}


def build_case(name):
    kind, P, W, H, degree, bg, rad, fovx, boost = CASES[name]
    scene, cam = helpers.make_case(kind, P, W, H, seed=sum(map(ord, name)) % 997,
                                   cam_radius=rad, fovx=fovx)
    scene["scales"] = (scene["scales"] * boost).astype(np.float32)
    if name == "cube_dense_saturating":
        scene["opacities"] = np.clip(scene["opacities"] * 1.6, 0.0, 0.999).astype(np.float32)
    return scene, cam, degree, bg


def main(outdir):
    import torch
    from oracle import ref_cuda
    os.makedirs(outdir, exist_ok=True)
    dev = torch.device("cuda:0")
    f32, i32, u8, i64 = torch.float32, torch.int32, torch.uint8, torch.int64
    for name in CASES:
        scene, cam, degree, bg = build_case(name)
        s, c = helpers.torch_inputs(scene, cam, dev, bg)
        args = helpers.native_forward_args(s, c, degree)
        P, W, H = scene["means3D"].shape[0], cam["image_width"], cam["image_height"]
        R, color, radii, is_used = ref_cuda.forward(*args)
        torch.cuda.synchronize()
        st = {
            "depths": ref_cuda.fetch(0, f32, dev), "clamped": ref_cuda.fetch(1, u8, dev).view(P, 3),
            "means2D": ref_cuda.fetch(3, f32, dev).view(P, 2), "cov3D": ref_cuda.fetch(4, f32, dev).view(P, 6),
            "conic_opacity": ref_cuda.fetch(5, f32, dev).view(P, 4), "rgb": ref_cuda.fetch(6, f32, dev).view(P, 3),
            "tiles_touched": ref_cuda.fetch(7, i32, dev), "point_offsets": ref_cuda.fetch(8, i32, dev),
            "point_list": ref_cuda.fetch(10, i32, dev), "keys": ref_cuda.fetch(12, i64, dev),
            "final_T": ref_cuda.fetch(20, f32, dev), "n_contrib": ref_cuda.fetch(21, i32, dev),
            "ranges": ref_cuda.fetch(22, i32, dev).view(-1, 2),
        }
        g = torch.Generator(device="cpu").manual_seed(7)
        dL = torch.randn((3, H, W), generator=g).to(dev)
        e = torch.empty(0)
        grads = ref_cuda.backward(c["bg"], s["means3D"], radii, e, s["scales"], s["rotations"], 1.0, e, c["viewmatrix"],
                                  c["projmatrix"], c["tanfovx"], c["tanfovy"], dL, s["shs"], degree, c["campos"], R)
        vis_mark = ref_cuda.mark_visible(s["means3D"], c["viewmatrix"], c["projmatrix"])
        torch.cuda.synchronize()
        # culled Gaussians leave their geometry slots uninitialised in the reference: zero them for a stable file
        vis = (radii > 0).cpu().numpy()
        out = {f"in_{k}": v for k, v in scene.items()}
        out.update({f"cam_{k}": np.asarray(v) for k, v in cam.items()})
        out["bg"] = np.asarray(bg, np.float32)
        out["degree"] = np.int32(degree)
        out["R"] = np.int64(R)
        out["color"] = color.cpu().numpy()
        out["radii"] = radii.cpu().numpy()
        out["is_used"] = is_used.cpu().numpy()
        out["mark_visible"] = vis_mark.cpu().numpy()
        out["dL_dpix"] = dL.cpu().numpy()
        for k, v in st.items():
            a = v.cpu().numpy()
            if k in ("depths", "clamped", "means2D", "cov3D", "conic_opacity", "rgb"):
                a = a.copy()
                a[~vis] = 0
            out[f"st_{k}"] = a
        for k, v in grads.items():
            out[f"grad_{k}"] = v.cpu().numpy()
        np.savez_compressed(os.path.join(outdir, f"{name}.npz"), **out)
        print(name, "P", P, "R", R, "visible", int(vis.sum()), "used", int(out["is_used"].sum()))

    # compute_relocation golden: every N in 1..50 and a spread of opacities
    from oracle import oracle
    rng = np.random.default_rng(11)
    K = 400
    N = np.concatenate([np.arange(1, 51), rng.integers(1, 51, K - 50)]).astype(np.int32)
    op = np.concatenate([np.array([0.005, 0.0051, 0.5, 0.99, 0.999], np.float32),
                         rng.uniform(0.005, 0.99, K - 5).astype(np.float32)])
    sc = np.exp(rng.normal(-4.0, 1.0, (K, 3))).astype(np.float32)
    binoms = oracle.binoms_table(51)
    o, s2 = ref_cuda.compute_relocation(torch.from_numpy(op).to(dev), torch.from_numpy(sc).to(dev),
                                        torch.from_numpy(N).to(dev), torch.from_numpy(binoms).to(dev), 51)
    torch.cuda.synchronize()
    np.savez_compressed(os.path.join(outdir, "relocation.npz"), opacity_old=op, scale_old=sc, N=N, binoms=binoms,
                        opacity_new=o.cpu().numpy(), scale_new=s2.cpu().numpy())
    print("relocation", K)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/golden")
