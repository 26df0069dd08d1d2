"""GPU-box report: parity of the B200 implementation against the reference CUDA rasterizer (oracle/_ref)
on seeded synthetic scenes, plus quick fwd/bwd timings of both. Writes gpurun_out/parity_report.json.
Usage: python tools/parity_report.py [--big]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import helpers  # noqa: E402
import torch  # noqa: E402


def timed(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    big = "--big" in sys.argv
    from diff_gaussian_rasterization import _native as C
    from oracle import ref_cuda
    cases = [("cube", 2000, 64, 48, 4.03), ("cube", 20000, 256, 256, 4.03), ("cube", 100000, 800, 800, 4.03),
             ("garden", 300000, 1297, 840, 6.0)]
    if big:
        cases.append(("garden", 3000000, 1920, 1080, 6.0))
    out = []
    for kind, P, W, H, rad in cases:
        t0 = time.time()
        scene, cam = helpers.make_case(kind, P, W, H, seed=0, cam_radius=rad, fovx=0.6911 if kind == "cube" else 1.2)
        rep = helpers.compare_with_reference(scene, cam)
        rep["case"] = f"{kind}-{P}-{W}x{H}"
        # timings
        dev = torch.device("cuda:0")
        s, c = helpers.torch_inputs(scene, cam, dev)
        args = helpers.native_forward_args(s, c, 3)
        dL = torch.randn((3, H, W), device=dev)
        e = torch.empty(0)

        def ours():
            R, color, radii, geom, binning, img, is_used = C.rasterize_gaussians(*args)
            C.rasterize_gaussians_backward(c["bg"], s["means3D"], radii, e, s["scales"], s["rotations"], 1.0, e,
                                           c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], dL, s["shs"], 3,
                                           c["campos"], geom, R, binning, img, False)

        def ours_fwd():
            C.rasterize_gaussians(*args)

        def theirs():
            R, color, radii, is_used = ref_cuda.forward(*args)
            ref_cuda.backward(c["bg"], s["means3D"], radii, e, s["scales"], s["rotations"], 1.0, e, c["viewmatrix"],
                              c["projmatrix"], c["tanfovx"], c["tanfovy"], dL, s["shs"], 3, c["campos"], R)

        def theirs_fwd():
            ref_cuda.forward(*args)

        rep["ms_ours_fwd_bwd"] = timed(ours)
        rep["ms_ours_fwd"] = timed(ours_fwd)
        rep["ms_ref_fwd_bwd"] = timed(theirs)
        rep["ms_ref_fwd"] = timed(theirs_fwd)
        rep["wall_s"] = time.time() - t0
        print(json.dumps(rep), flush=True)
        out.append(rep)
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/parity_report.json", "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
