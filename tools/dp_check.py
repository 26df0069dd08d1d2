#!/usr/bin/env python
"""Multi-GPU check of the view-sharded gradient exchange (run under torchrun on the GPU box):

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/dp_check.py [--P 300000]

Every rank renders its own view, then the summed gradients are obtained three ways and compared:
  (a) fused backward + NCCL all-reduce of the full 59-float arena (dp.GradientArena)  -- the baseline
  (b) dp.FactoredExchange(transport="nccl")
  (c) dp.FactoredExchange(transport="symm")   (NVLink symmetric memory, own kernels)
Prints one JSON line from rank 0 with the relative L2 differences, whether (b)/(c) are identical across ranks, and the
per-step device time of the backward+exchange for each variant.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "3dgs-mcmc_b200"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def train_replica_check(args, C, dp, s, cam_t, dev, rank, world, P, W, H):
    """Every rank holds a replica of the raw parameters and of the Adam state, renders a different view per step, and
    applies the summed gradients: after K steps the replicas must agree bit for bit."""
    from diff_gaussian_rasterization import GaussianRasterizationSettings, add_noise
    from diff_gaussian_rasterization import loss as fused_loss
    from diff_gaussian_rasterization.activations import activate
    from diff_gaussian_rasterization.optim import FusedAdam
    raw = {"xyz": s["means3D"].clone(), "shs": s["shs"].clone(), "opacity": torch.logit(s["opacities"].clamp(1e-4, 1 - 1e-4)),
           "scaling": torch.log(s["scales"]), "rotation": s["rotations"].clone()}
    raw = {k: torch.nn.Parameter(v) for k, v in raw.items()}
    lrs = {"xyz": 1.6e-4, "shs": 2.5e-3, "opacity": 0.05, "scaling": 5e-3, "rotation": 1e-3}
    opt = FusedAdam([{"params": [raw[k]], "lr": lrs[k], "name": k} for k in raw], lr=0.0, eps=1e-15)
    vp = dp.ViewParallelRasterizer(P - 1000, 16, dev, capacity=P + 64)   # allocated for the cap, then grown: no new rendezvous
    vp.resize(P)
    bg = torch.zeros(3, device=dev)
    gts = [torch.rand((3, H, W), generator=torch.Generator().manual_seed(500 + v)).to(dev) for v in range(len(cam_t))]
    losses = []
    for step in range(args.train_steps):
        view = (step * world + rank) % len(cam_t)
        c = cam_t[view]
        cv = torch.stack([cam_t[(step * world + r) % len(cam_t)]["campos"] for r in range(world)])
        settings = GaussianRasterizationSettings(H, W, c["tanfovx"], c["tanfovy"], bg, 1.0, c["viewmatrix"], c["projmatrix"], 3,
                                                 c["campos"], False, False)
        opacity, scaling, rotation, reg = activate(raw["opacity"], raw["scaling"], raw["rotation"], 0.01, 0.01)
        means2D = torch.zeros_like(raw["xyz"], requires_grad=True)
        image = vp(settings, cv, raw["xyz"], means2D, opacity, raw["shs"], scaling, rotation)[0]
        loss = fused_loss.photometric_loss(image, gts[view], 0.2) / world + reg
        loss.backward()
        opt.step()
        opt.zero_grad(set_to_none=True)
        with torch.no_grad():
            add_noise(raw["xyz"], raw["scaling"], raw["rotation"], raw["opacity"], 5e5 * lrs["xyz"], seed=7, offset=step)
        losses.append(float(loss.item()))
    torch.cuda.synchronize()
    ident, moved = {}, {}
    start = {"xyz": s["means3D"], "shs": s["shs"], "rotation": s["rotations"]}
    for k, p_ in raw.items():
        bits = p_.detach().contiguous().view(torch.int32).long().flatten()
        h = torch.stack([bits.sum(), (bits * (torch.arange(bits.numel(), device=dev) % 1000003 + 1)).sum()])
        hs = [torch.zeros_like(h) for _ in range(world)]
        dist.all_gather(hs, h)
        ident[k] = all(bool((x == hs[0]).all()) for x in hs)
        if k in start:
            moved[k] = float((p_.detach() - start[k]).abs().max())
    return {"steps": args.train_steps, "identical_on_all_ranks": ident, "max_abs_change": moved, "transport": vp.exchange.transport,
            "rank0_losses": losses}


def mcmc_lockstep_check(args, dp, dev, rank, world):
    """Multi-rank MCMC lock step (SURVEY.md 5 last row; north_star "allreducing the per-Gaussian gradients and MCMC
    statistics"): every rank runs the REFERENCE's own GaussianModel (scene/gaussian_model_static.py, staged unmodified
    under baseline/_ref/py) as a replica - its activations, `relocate_gs`, `add_new_gs`, optimizer surgery - renders its own
    view through dp.ViewParallelRasterizer, and applies the summed gradients with FusedAdam and the reference's inline noise
    (train_static.py:132-140, torch.randn_like on the identically seeded device RNG). Every `interval` steps: relocation +
    5 % growth (torch.multinomial / bincount, :412-422, :458-492) inside dp.lockstep_sampling() - the sampled indices are the
    one thing that is broadcast, because torch's CUDA multinomial is not reproducible from launch to launch - then
    ViewParallelRasterizer.resize(P) and dp.assert_replicas_in_sync. The replicas must stay bit-identical throughout."""
    import types
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import refenv
    refenv.install()
    import gs_synth
    import scene.gaussian_model_static as gm_static
    import utils.graphics_utils as graphics_utils
    from scene.gaussian_model import build_scaling_rotation
    from arguments import OptimizationParams
    from diff_gaussian_rasterization import GaussianRasterizationSettings
    from diff_gaussian_rasterization import loss as fused_loss
    from diff_gaussian_rasterization.optim import FusedAdam

    P0, cap, W, H = args.mcmc_P, int(args.mcmc_P * 1.18), args.width, args.height
    ap = argparse.ArgumentParser()
    opt = OptimizationParams(ap).extract(ap.parse_args([]))
    truth = gs_synth.to_torch(gs_synth.scene_garden(P0, 1), dev)
    cams = [gs_synth.to_torch(c, dev) for c in gs_synth.orbit_cameras(16, 6.0, W, H, 1.2, 20.0)]
    bg = torch.zeros(3, device=dev)
    from diff_gaussian_rasterization import GaussianRasterizer
    gts = []
    for c in cams:
        st = GaussianRasterizationSettings(H, W, c["tanfovx"], c["tanfovy"], bg, 1.0, c["viewmatrix"], c["projmatrix"], 3, c["campos"], False, False)
        with torch.no_grad():
            gts.append(GaussianRasterizer(st)(means3D=truth["means3D"], means2D=torch.zeros_like(truth["means3D"]), shs=truth["shs"],
                                              opacities=truth["opacities"], scales=truth["scales"], rotations=truth["rotations"])[0].clamp(0, 1))
    rng = np.random.default_rng(4)
    pts = (truth["means3D"].cpu().numpy() + rng.normal(0, 0.02, (P0, 3))).astype(np.float32)
    pcd = graphics_utils.BasicPointCloud(points=pts, colors=rng.uniform(0, 1, (P0, 3)).astype(np.float32), normals=np.zeros((P0, 3), np.float32))
    torch.manual_seed(0)                                  # safe_state() seeds every process alike (utils/general_utils.py:133-136)
    g = gm_static.GaussianModel(3)
    g.create_from_pcd(pcd, 1.0)
    with torch.no_grad():
        g._opacity[::9] = -7.0                             # some dead from the start, so that relocation has work
    saved = torch.optim.Adam
    torch.optim.Adam = FusedAdam
    try:
        g.training_setup(opt)
    finally:
        torch.optim.Adam = saved
    g.active_sh_degree = 3
    vp = dp.ViewParallelRasterizer(P0, 16, dev, capacity=cap)
    params = lambda: [g._xyz, g._features_dc, g._features_rest, g._opacity, g._scaling, g._rotation]  # noqa: E731
    state = lambda: [t for grp in g.optimizer.param_groups for t in (g.optimizer.state[grp["params"][0]]["exp_avg"],  # noqa: E731
                                                                      g.optimizer.state[grp["params"][0]]["exp_avg_sq"])]
    relocations = grown = 0
    sizes, losses = [P0], []
    names = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
    trace = []

    def differing(tag, tensors, labels):
        """--mcmc-debug: which of the given replicated tensors differ between the ranks right now."""
        if not args.mcmc_debug:
            return
        bad = []
        for lab, t in zip(labels, tensors):
            h = dp.replica_fingerprint([t]).to(dev)
            hs = [torch.zeros_like(h) for _ in range(world)]
            dist.all_gather(hs, h)
            if not all(bool((x == hs[0]).all()) for x in hs):
                bad.append(lab)
        if bad:
            trace.append({"step": it, "at": tag, "differ": bad})
    for it in range(1, args.mcmc_steps + 1):
        xyz_lr = g.update_learning_rate(it)
        view = (it * world + rank) % len(cams)
        c = cams[view]
        cv = torch.stack([cams[(it * world + r) % len(cams)]["campos"] for r in range(world)])
        settings = GaussianRasterizationSettings(H, W, c["tanfovx"], c["tanfovy"], bg, 1.0, c["viewmatrix"], c["projmatrix"], 3, c["campos"], False, False)
        means2D = torch.zeros_like(g.get_xyz, requires_grad=True)
        image = vp(settings, cv, g.get_xyz, means2D, g.get_opacity, g.get_features, g.get_scaling, g.get_rotation)[0]
        loss = fused_loss.photometric_loss(image, gts[view], opt.lambda_dssim) / world
        loss = loss + opt.opacity_reg * torch.abs(g.get_opacity).mean() + opt.scale_reg * torch.abs(g.get_scaling).mean()
        loss.backward()
        losses.append(float(loss.item()))
        with torch.no_grad():
            differing("grads after backward", [t.grad for t in params()], names)
            if it % args.mcmc_interval == 0:
                dead = (g.get_opacity <= 0.005).squeeze(-1)
                n_dead = int(dead.sum())
                differing("before relocation", params() + state(), names + tuple(f"state{i}" for i in range(12)))
                with dp.lockstep_sampling():              # rank 0's multinomial draws, broadcast (dp.py: why seeding alone fails)
                    g.relocate_gs(dead_mask=dead)
                    differing("after relocate_gs", params() + state(), names + tuple(f"state{i}" for i in range(12)))
                    added = g.add_new_gs(cap_max=cap)
                differing("after add_new_gs", params() + state(), names + tuple(f"state{i}" for i in range(12)))
                relocations += n_dead
                grown += int(added)
                vp.resize(g.get_xyz.shape[0])
                sizes.append(int(g.get_xyz.shape[0]))
                if not args.mcmc_debug:
                    dp.assert_replicas_in_sync(params() + state(), what=f"replicas after relocation/growth at step {it}")
            g.optimizer.step()
            g.optimizer.zero_grad(set_to_none=True)
            L = build_scaling_rotation(g.get_scaling, g.get_rotation)
            cov = L @ L.transpose(1, 2)
            noise = torch.randn_like(g._xyz) * (1 / (1 + torch.exp(-100 * ((1 - g.get_opacity) - 0.995)))) * opt.noise_lr * xyz_lr
            g._xyz.add_(torch.bmm(cov, noise.unsqueeze(-1)).squeeze(-1))
            differing("after Adam + noise", params(), names)
            if args.mcmc_debug and len(trace) >= 6:
                break
    torch.cuda.synchronize()
    if args.mcmc_debug:
        return {"debug_trace": trace}
    ident = {}
    for name, t in zip(names, params()):
        h = dp.replica_fingerprint([t]).to(dev)
        hs = [torch.zeros_like(h) for _ in range(world)]
        dist.all_gather(hs, h)
        ident[name] = all(bool((x == hs[0]).all()) for x in hs)
    dp.assert_replicas_in_sync(params() + state(), what="replicas at the end of the run")
    return {"steps": args.mcmc_steps, "interval": args.mcmc_interval, "relocations": relocations, "grown_by": grown, "sizes": sizes,
            "ragged_sizes": [s_ for s_ in sizes if s_ % 4], "identical_on_all_ranks": ident, "transport": vp.exchange.transport,
            "multicast": bool(getattr(vp.exchange, "_multicast", False)), "model": "reference GaussianModel (baseline/_ref/py), unmodified",
            "rank0_first_last_loss": [losses[0], losses[-1]]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--P", type=int, default=300000)
    ap.add_argument("--width", type=int, default=1297)
    ap.add_argument("--height", type=int, default=840)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--transports", default="nccl,symm:dense,symm:sparse,symm:records,symm-p2p:records",
                    help="comma list of transport[-p2p][:layout]; -p2p = peer pointers instead of the multicast mapping")
    ap.add_argument("--breakdown", action="store_true")
    ap.add_argument("--train-steps", type=int, default=0,
                    help="also run this many full training iterations (activations, render, loss, exchange, Adam, noise) and "
                         "check that the replicated parameters stay bit-identical on all ranks")
    ap.add_argument("--mcmc-steps", type=int, default=0,
                    help="run this many training iterations of the reference's own GaussianModel as replicas, with relocation + "
                         "growth + resize every --mcmc-interval steps, and check that they stay bit-identical")
    ap.add_argument("--mcmc-interval", type=int, default=5)
    ap.add_argument("--mcmc-debug", action="store_true", help="report where the replicas first differ instead of raising")
    ap.add_argument("--mcmc-P", type=int, default=100003)
    ap.add_argument("--skip-exchange", action="store_true", help="only the --train-steps / --mcmc-steps checks")
    args = ap.parse_args()
    import gs_synth
    from diff_gaussian_rasterization import _native as C, dp

    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=dev)
    P, W, H = args.P, args.width, args.height
    scene = gs_synth.scene_garden(P, 0)
    cams = gs_synth.orbit_cameras(16, 6.0, W, H, 1.2, 20.0)
    s = gs_synth.to_torch(scene, dev)
    cam_t = [gs_synth.to_torch(c, dev) for c in cams]
    c = cam_t[rank % len(cam_t)]
    campos_views = torch.stack([cam_t[r % len(cam_t)]["campos"] for r in range(world)])
    bg = torch.zeros(3, device=dev)
    g = torch.Generator(device="cpu").manual_seed(100 + rank)
    dL = torch.randn((3, H, W), generator=g).to(dev)
    e = torch.empty(0)
    R, color, radii, geom, binning, img, used = C.rasterize_gaussians(
        bg, s["means3D"], e, s["opacities"], s["scales"], s["rotations"], 1.0, e, c["viewmatrix"], c["projmatrix"],
        c["tanfovx"], c["tanfovy"], H, W, s["shs"], 3, c["campos"], False, False)

    def timed(fn):
        for _ in range(3):
            fn()
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / args.iters], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    arena = dp.GradientArena(P, 16, dev)

    def baseline():
        C.rasterize_gaussians_backward(bg, s["means3D"], radii, e, s["scales"], s["rotations"], 1.0, e, c["viewmatrix"],
                                       c["projmatrix"], c["tanfovx"], c["tanfovy"], dL, s["shs"], 3, c["campos"], geom, R,
                                       binning, img, False, out_buffers=arena.backward_buffers())
        arena.all_reduce()

    out = {"world": world, "P": P, "W": W, "H": H}
    out["ms_full_all_reduce"] = timed(baseline)
    want = {k: arena.views[k].clone() for k in ("means3D", "sh", "opacity", "scales", "rotations")}

    def rel(a, b):
        return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))

    for transport in ([] if args.skip_exchange else args.transports.split(",")):
        try:
            tname, _, layout = transport.partition(":")
            ex = dp.FactoredExchange(P, 16, dev, transport=tname.split("-")[0],
                                     multicast=False if tname.endswith("-p2p") else None, layout=layout or None)

            def step():
                return ex.backward(C, bg, s["means3D"], radii, s["scales"], s["rotations"], 1.0, c["viewmatrix"],
                                   c["projmatrix"], c["tanfovx"], c["tanfovy"], dL, s["shs"], 3, c["campos"],
                                   campos_views, geom, R, binning, img)
            out[f"ms_factored_{transport}"] = timed(step)
            grads = step()
            torch.cuda.synchronize()
            diffs = {k: rel(grads[k], want[k]) for k in want}
            # identical on every rank? compare a checksum of the raw bits
            ident = True
            for k in want:
                bits = grads[k].contiguous().view(torch.int32).long()
                h = torch.stack([bits.sum(), (bits * torch.arange(1, bits.numel() + 1, device=dev).view_as(bits) % 1000003).sum()])
                hs = [torch.zeros_like(h) for _ in range(world)]
                dist.all_gather(hs, h)
                ident = ident and all(bool((x == hs[0]).all()) for x in hs)
            out[f"factored_{transport}"] = {"rel_l2_vs_full_all_reduce": diffs, "identical_on_all_ranks": ident,
                                            "multicast": bool(getattr(ex, "_multicast", False)),
                                            "transport": ex.transport, "layout": ex.layout}
            if args.breakdown and ex.layout == "records":
                scratch = C.backward_render(bg, P, R, dL, geom, binning, img)
                bd = {"layout": "records"}
                bd["backward_render"] = timed(lambda: C.backward_render(bg, P, R, dL, geom, binning, img))
                own = ex._gathered_all[0]
                dst, mc = (ex._gather_ptrs, ex._multicast) if ex.transport == "symm" else ([own.data_ptr()], False)

                def push_only():
                    C.backward_geometry_push(s["means3D"], radii, s["scales"], s["rotations"], 1.0, c["viewmatrix"], c["projmatrix"],
                                             c["tanfovx"], c["tanfovy"], H, W, s["shs"], 3, c["campos"], geom, scratch, dst,
                                             rank * ex.slot_stride if ex.transport == "symm" else 0, mc)
                    if ex.transport == "symm":
                        ex._hdl.barrier(channel=0)
                bd["geometry_push+barrier"] = timed(push_only)
                bd["geometry_with_sh_dense_outputs"] = timed(lambda: C.backward_geometry(
                    s["means3D"], radii, s["scales"], s["rotations"], 1.0, e, c["viewmatrix"], c["projmatrix"], c["tanfovx"],
                    c["tanfovy"], H, W, s["shs"], 3, c["campos"], geom, scratch, out_buffers=arena.backward_buffers()))
                flat_out, outs = ex._fresh_outputs()
                flat_out.zero_()
                bd["combine_prezeroed"] = timed(lambda: C.dp_grad_record_combine(s["means3D"], campos_views, own, ex.slot_stride, 3, 16,
                                                                                 outs=outs, outputs_are_zero=True))
                bd["zero_fill_708MB_at_3M"] = timed(lambda: outs["sh"].zero_())
                out[f"breakdown_{transport}"] = bd
            elif args.breakdown:
                scratch = C.backward_render(bg, P, R, dL, geom, binning, img)
                bd = {}
                bd["backward_render"] = timed(lambda: C.backward_render(bg, P, R, dL, geom, binning, img))
                bd["push"] = timed(lambda: ex.push_color_grads(C, radii, geom, scratch) and None)
                bd["geometry_no_sh"] = timed(lambda: C.backward_geometry(
                    s["means3D"], radii, s["scales"], s["rotations"], 1.0, e, c["viewmatrix"], c["projmatrix"], c["tanfovx"],
                    c["tanfovy"], H, W, s["shs"], 3, c["campos"], geom, scratch, out_buffers=ex.out_buffers(), write_sh=False))
                bd["geometry_with_sh"] = timed(lambda: C.backward_geometry(
                    s["means3D"], radii, s["scales"], s["rotations"], 1.0, e, c["viewmatrix"], c["projmatrix"], c["tanfovx"],
                    c["tanfovy"], H, W, s["shs"], 3, c["campos"], geom, scratch, out_buffers=arena.backward_buffers()))
                bd["combine"] = timed(lambda: C.dp_sh_grad_combine(s["means3D"], campos_views, ex.gathered, ex.slot_stride, 3, 16,
                                                                   out=ex.grads["sh"], layout=ex.layout))
                bd["layout"] = ex.layout
                if ex.transport == "symm":
                    bd["slab_all_reduce_alone"] = timed(lambda: C.dp_slab_all_reduce(dev, ex._slab_ptrs, ex._n_slab, rank, world, ex._multicast))
                    bd["barrier"] = timed(lambda: ex._hdl.barrier(channel=0))

                    def ar():
                        ex._hdl.barrier(channel=0)
                        C.dp_slab_all_reduce(dev, ex._slab_ptrs, ex._n_slab, rank, world, ex._multicast)
                        ex._hdl.barrier(channel=1)
                    bd["barrier+slab_all_reduce+barrier"] = timed(ar)
                else:
                    bd["nccl_all_reduce_slab"] = timed(lambda: dist.all_reduce(ex.slab))
                    bd["nccl_all_gather"] = timed(lambda: dist.all_gather_into_tensor(ex.gathered.view(-1), ex._send))
                out[f"breakdown_{transport}"] = bd
            del ex
        except Exception as ex_:  # report, keep going with the next transport
            out[f"factored_{transport}"] = {"error": repr(ex_)[:400]}
    if args.train_steps > 0:
        out["train_replicas"] = train_replica_check(args, C, dp, s, cam_t, dev, rank, world, P, W, H)
    if args.mcmc_steps > 0:
        out["mcmc_lockstep"] = mcmc_lockstep_check(args, dp, dev, rank, world)
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
