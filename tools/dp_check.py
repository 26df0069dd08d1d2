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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--P", type=int, default=300000)
    ap.add_argument("--width", type=int, default=1297)
    ap.add_argument("--height", type=int, default=840)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--transports", default="nccl,symm,symm-p2p")
    ap.add_argument("--breakdown", action="store_true")
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

    for transport in args.transports.split(","):
        try:
            ex = dp.FactoredExchange(P, 16, dev, transport=transport.split("-")[0],
                                     multicast=False if transport.endswith("-p2p") else None)

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
                                            "transport": ex.transport}
            if args.breakdown:
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
                                                                   out=ex.grads["sh"]))
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
    if rank == 0:
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
