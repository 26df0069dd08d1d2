#!/usr/bin/env python
"""Where the end-to-end step's time beyond the device-timed loop goes (measurement aid): the public-API step of bench.py's e2e
leg with pieces removed one at a time. usage: python tools/e2e_gap.py [P W H]"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3dgs-mcmc_b200")); sys.path.insert(0, ROOT)
import numpy as np, torch, gs_synth
from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
from diff_gaussian_rasterization.loss import l1_loss

P, W, H = (int(a) for a in sys.argv[1:4]) if len(sys.argv) > 3 else (3_000_000, 1920, 1080)
dev = torch.device("cuda:0")
scene = gs_synth.scene_garden(P, 0)
cams = gs_synth.orbit_cameras(16, 6.0, W, H, 1.2, 20.0)
s = gs_synth.to_torch(scene, dev)
cam_host = [{k: (torch.from_numpy(np.ascontiguousarray(v)).pin_memory() if isinstance(v, np.ndarray) else v) for k, v in c.items()} for c in cams]
params = {k: s[k].clone().requires_grad_(True) for k in ("means3D", "shs", "opacities", "scales", "rotations")}
bg = torch.zeros(3, device=dev)
gen = torch.Generator().manual_seed(5)
gt_host = [torch.rand((3, H, W), generator=gen).pin_memory() for _ in range(2)]
gt_dev = [g.to(dev) for g in gt_host]
dL = (torch.randn((3, H, W), generator=gen) / (3 * H * W)).to(dev)
view_dev, proj_dev, campos_dev = torch.empty((4, 4), device=dev), torch.empty((4, 4), device=dev), torch.empty(3, device=dev)
copy_stream = torch.cuda.Stream()
loss_host = torch.zeros(1).pin_memory()


loss_ready = [torch.cuda.Event() for _ in range(2)]
gt_ready = [torch.cuda.Event() for _ in range(2)]
gt_free = [torch.cuda.Event() for _ in range(2)]
for e_ in gt_free + gt_ready:
    e_.record()


def step(i, loss=True, h2d=True, cam_copy=True, d2h=True, fresh_means2d=True, drain=False, gt_events=False, ready_only=False, state={}):
    ch = cam_host[i % 16]
    if h2d:
        with torch.cuda.stream(copy_stream):
            if gt_events and not ready_only:
                copy_stream.wait_event(gt_free[(i + 1) % 2])
            gt_dev[(i + 1) % 2].copy_(gt_host[(i + 1) % 2], non_blocking=True)
            if gt_events:
                gt_ready[(i + 1) % 2].record(copy_stream)
    if cam_copy:
        view_dev.copy_(ch["viewmatrix"], non_blocking=True); proj_dev.copy_(ch["projmatrix"], non_blocking=True); campos_dev.copy_(ch["campos"], non_blocking=True)
    st = GaussianRasterizationSettings(H, W, ch["tanfovx"], ch["tanfovy"], bg, 1.0, view_dev, proj_dev, 3, campos_dev, False, False)
    if fresh_means2d or "m2d" not in state:
        state["m2d"] = torch.zeros_like(params["means3D"], requires_grad=True)
    image = GaussianRasterizer(st)(means3D=params["means3D"], means2D=state["m2d"], shs=params["shs"], opacities=params["opacities"],
                                   scales=params["scales"], rotations=params["rotations"])[0]
    if loss:
        if gt_events:
            torch.cuda.current_stream().wait_event(gt_ready[i % 2])
        l = l1_loss(image, gt_dev[i % 2]); l.backward()
        if gt_events and not ready_only:
            gt_free[i % 2].record()
        if d2h:
            loss_host.copy_(l.detach().reshape(1), non_blocking=True)
            if drain:
                loss_ready[i % 2].record()
                if i > 0:
                    loss_ready[(i - 1) % 2].synchronize()   # read the previous step's loss, as bench.py's e2e leg does
                    float(loss_host[0])
    else:
        image.backward(dL)
    for p_ in params.values():
        p_.grad = None
    state["m2d"].grad = None


def timeit(**kw):
    for i in range(4):
        step(i, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(20):
        step(4 + i, **kw)
    e1.record(); torch.cuda.synchronize()
    return round(e0.elapsed_time(e1) / 20, 4)


view_dev.copy_(cam_host[0]["viewmatrix"]); proj_dev.copy_(cam_host[0]["projmatrix"]); campos_dev.copy_(cam_host[0]["campos"])
print(json.dumps({"full": timeit(), "no_loss": timeit(loss=False), "no_h2d_image": timeit(h2d=False), "no_camera_copy": timeit(cam_copy=False),
                  "no_loss_d2h": timeit(d2h=False), "reused_means2D": timeit(fresh_means2d=False),
                  "with_loss_drain": timeit(drain=True), "with_gt_events": timeit(gt_events=True), "both": timeit(drain=True, gt_events=True),
                  "ready_event_only_plus_drain": timeit(drain=True, gt_events=True, ready_only=True),
                  "bare (no loss, no copies)": timeit(loss=False, h2d=False, cam_copy=False, fresh_means2d=False)}))
