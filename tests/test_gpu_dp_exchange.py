"""GPU tests of the view-sharded gradient exchange (csrc/dp_exchange.cu, diff_gaussian_rasterization/dp.py) on ONE
device: the views of a step are rendered one after the other and the factored exchange (masked dL/dRGB per view +
rebuilt SH gradient + summed 11-float slab) is compared with the plain sum of the per-view gradients of the fused
backward — which the other GPU tests compare with the reference CUDA rasterizer. The 2-GPU transports (NVLink
symmetric memory and NCCL) are checked by tools/dp_check.py under torchrun; the last test here runs it on two ranks
when the box has two GPUs.
"""
import pytest
import torch

import helpers
import gs_synth

pytestmark = pytest.mark.gpu

NAMES = ("means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations")


def native():
    from diff_gaussian_rasterization import _native
    return _native


def _views(kind, P, W, H, V, degree, rad, fovx, seed=0):
    dev = torch.device("cuda:0")
    scene = gs_synth.scene_cube(P, seed) if kind == "cube" else gs_synth.scene_garden(P, seed)
    if degree < 3:
        scene["shs"] = scene["shs"][:, : (degree + 1) ** 2].copy()
    cams = gs_synth.orbit_cameras(max(V, 4), rad, W, H, fovx, 20.0)[:V]
    s = gs_synth.to_torch(scene, dev)
    cs = []
    for cam in cams:
        c = gs_synth.to_torch(cam, dev)
        c["bg"] = torch.tensor((0.1, 0.0, 0.3), dtype=torch.float32, device=dev)
        cs.append(c)
    g = torch.Generator(device="cpu").manual_seed(7)
    dLs = [torch.randn((3, H, W), generator=g).to(dev) for _ in range(V)]
    return s, cs, dLs


def _forward(C, s, c, degree):
    return C.rasterize_gaussians(*helpers.native_forward_args(s, c, degree))


def _fused_backward(C, s, c, degree, dL, fwd):
    R, color, radii, geom, binning, img, is_used = fwd
    e = torch.empty(0)
    return dict(zip(NAMES, C.rasterize_gaussians_backward(
        c["bg"], s["means3D"], radii, e, s["scales"], s["rotations"], 1.0, e, c["viewmatrix"], c["projmatrix"],
        c["tanfovx"], c["tanfovy"], dL, s["shs"], degree, c["campos"], geom, R, binning, img, False)))


@pytest.mark.parametrize("kind,P,W,H,rad,fovx", [("cube", 20000, 250, 130, 4.03, 0.6911),
                                                  ("garden", 100003, 640, 360, 6.0, 1.2)])
def test_split_backward_equals_fused_backward(kind, P, W, H, rad, fovx):
    """gs_backward_render + gs_backward_geometry write the same bits as gs_backward."""
    C = native()
    s, cs, dLs = _views(kind, P, W, H, 1, 3, rad, fovx)
    c, dL = cs[0], dLs[0]
    fwd = _forward(C, s, c, 3)
    R, _, radii, geom, binning, img, _ = fwd
    fused = _fused_backward(C, s, c, 3, dL, fwd)
    scratch = C.backward_render(c["bg"], P, R, dL, geom, binning, img)
    e = torch.empty(0)
    split = dict(zip(NAMES, C.backward_geometry(s["means3D"], radii, s["scales"], s["rotations"], 1.0, e, c["viewmatrix"],
                                                c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W, s["shs"], 3,
                                                c["campos"], geom, scratch)))
    torch.cuda.synchronize()
    for n in NAMES:
        # float atomics in the tile backward reorder sums from run to run: compare to the run-to-run noise floor
        assert helpers.rel_l2(split[n], fused[n]) < 2e-5, n
    # without the SH rows: every other output unchanged, SH slot None
    nosh = dict(zip(NAMES, C.backward_geometry(s["means3D"], radii, s["scales"], s["rotations"], 1.0, e, c["viewmatrix"],
                                               c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W, s["shs"], 3,
                                               c["campos"], geom, scratch, write_sh=False)))
    torch.cuda.synchronize()
    assert nosh["sh"] is None
    for n in NAMES:
        if n != "sh":
            assert helpers.bits_equal_count(nosh[n], split[n]) == 0, n


def _decode_sparse_slot(slot, P):
    """[P, 3] colour gradients from one sparse slot (include/gs_b200.h GS_DP_SPARSE): per 256 Gaussians 8 x (mask, before)
    words, then the non-zero triples in index order. Also checks the header's running counts."""
    import numpy as np
    raw = slot.cpu().numpy()
    words = raw.view(np.uint32)
    out = np.zeros((P, 3), np.float32)
    for c in range((P + 255) // 256):
        base = c * 784
        seen = 0
        for w in range(8):
            mask, before = int(words[base + 2 * w]), int(words[base + 2 * w + 1])
            assert before == seen
            for lane in range(32):
                if (mask >> lane) & 1:
                    i = c * 256 + 32 * w + lane
                    assert i < P
                    out[i] = raw[base + 16 + 3 * seen: base + 19 + 3 * seen]
                    assert out[i].any()
                    seen += 1
    return torch.from_numpy(out).to(slot.device)


@pytest.mark.parametrize("kind,P,W,H,V,degree,rad,fovx", [
    ("cube", 20001, 250, 130, 3, 3, 4.03, 0.6911),      # P not a multiple of 4: scalar tail of the push kernel
    ("cube", 5000, 96, 64, 2, 1, 4.03, 0.6911),         # lower degree, M = 4 rows
    ("cube", 5000, 96, 64, 2, 0, 4.03, 0.6911),
    ("garden", 200000, 800, 450, 4, 3, 6.0, 1.2),
])
@pytest.mark.parametrize("layout", ["dense", "sparse"])
def test_factored_sh_gradient_equals_sum_of_per_view_gradients(kind, P, W, H, V, degree, rad, fovx, layout):
    """Masked dL/dRGB per view + gs_dp_sh_grad_combine == sum over views (in order) of the per-view dL_dsh that the
    same scratch gives through the ordinary backward, bit for bit - in the dense [P][3] layout and in the sparse one
    (bit masks + non-zero triples only), whose slot is decoded here to check what travels."""
    C = native()
    dev = torch.device("cuda:0")
    s, cs, dLs = _views(kind, P, W, H, V, degree, rad, fovx)
    M = s["shs"].shape[1]
    slot_stride = (C.dp_slot_floats(P, layout) + 63) // 64 * 64
    assert slot_stride >= 3 * P
    gathered = torch.full((V, slot_stride), float("nan"), device=dev)
    e = torch.empty(0)
    total = torch.zeros((P, M, 3), device=dev)
    campos = torch.stack([c["campos"] for c in cs])
    for v, (c, dL) in enumerate(zip(cs, dLs)):
        R, _, radii, geom, binning, img, _ = _forward(C, s, c, degree)
        scratch = C.backward_render(c["bg"], P, R, dL, geom, binning, img)
        C.dp_color_grad_push(P, radii, geom, scratch, [gathered.data_ptr()], v * slot_stride, False, layout=layout)
        outs = dict(zip(NAMES, C.backward_geometry(s["means3D"], radii, s["scales"], s["rotations"], 1.0, e,
                                                   c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W,
                                                   s["shs"], degree, c["campos"], geom, scratch)))
        total = total + outs["sh"]
        # the pushed colour gradient is the backward's dL_dcolors with the clamp mask applied
        g = gathered[v, : 3 * P].view(P, 3) if layout == "dense" else _decode_sparse_slot(gathered[v], P)
        assert torch.isfinite(g).all()
        changed = g != outs["colors"]
        assert bool(((g == 0) | ~changed).all())
        assert bool(((g != 0).any(1) | (outs["sh"].abs().amax((1, 2)) == 0)).all())   # nothing non-zero was dropped
    sh = C.dp_sh_grad_combine(s["means3D"], campos, gathered, slot_stride, degree, M, layout=layout)
    torch.cuda.synchronize()
    assert sh.shape == total.shape
    assert float(total.abs().max()) > 0
    assert helpers.bits_equal_count(sh + 0.0, total + 0.0) == 0, helpers.rel_l2(sh, total)


def _decode_records(slot, P):
    """{index: 14 floats} from one slot in the records layout (include/gs_b200.h GS_DP_RECORDS); checks the running counts."""
    import numpy as np
    raw = slot.cpu().numpy()
    words = raw.view(np.uint32)
    out = {}
    for c in range((P + 127) // 128):
        base = c * 1800
        seen = 0
        for w in range(4):
            mask, before = int(words[base + 2 * w]), int(words[base + 2 * w + 1])
            assert before == seen
            for lane in range(32):
                if (mask >> lane) & 1:
                    i = c * 128 + 32 * w + lane
                    assert i < P
                    out[i] = raw[base + 8 + 14 * seen: base + 22 + 14 * seen].copy()
                    seen += 1
    return out


@pytest.mark.parametrize("kind,P,W,H,V,degree,rad,fovx", [
    ("cube", 20001, 250, 130, 3, 3, 4.03, 0.6911),      # ragged P: a partial last chunk
    ("cube", 5000, 96, 64, 2, 1, 4.03, 0.6911),
    ("cube", 5000, 96, 64, 2, 0, 4.03, 0.6911),
    ("cube", 5000, 96, 64, 5, 2, 4.03, 0.6911),         # V not a multiple of the view batch, 27-float rows
    ("garden", 200000, 800, 450, 4, 3, 6.0, 1.2),
])
def test_gradient_records_equal_sum_of_per_view_gradients(kind, P, W, H, V, degree, rad, fovx):
    """The records exchange on one GPU: every view's per-Gaussian backward pushes its chunk records into its slot
    (gs_backward_geometry_push), gs_dp_grad_record_combine adds the views in order. All five parameter gradients must equal
    the sum, in view order, of what the ordinary backward gives per view - bit for bit - and a record must exist exactly for
    the Gaussians whose gradient is not all zero."""
    C = native()
    dev = torch.device("cuda:0")
    s, cs, dLs = _views(kind, P, W, H, V, degree, rad, fovx)
    M = s["shs"].shape[1]
    stride = (C.dp_slot_floats(P, "records") + 63) // 64 * 64
    gathered = torch.full((V, stride), float("nan"), device=dev)
    e = torch.empty(0)
    keys = ("means3D", "opacity", "scales", "rotations", "sh")
    total = None
    campos = torch.stack([c["campos"] for c in cs])
    for v, (c, dL) in enumerate(zip(cs, dLs)):
        R, _, radii, geom, binning, img, _ = _forward(C, s, c, degree)
        scratch = C.backward_render(c["bg"], P, R, dL, geom, binning, img)
        m2d = C.backward_geometry_push(s["means3D"], radii, s["scales"], s["rotations"], 1.0, c["viewmatrix"], c["projmatrix"],
                                       c["tanfovx"], c["tanfovy"], H, W, s["shs"], degree, c["campos"], geom, scratch,
                                       [gathered.data_ptr()], v * stride, False)
        outs = dict(zip(NAMES, C.backward_geometry(s["means3D"], radii, s["scales"], s["rotations"], 1.0, e,
                                                   c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W,
                                                   s["shs"], degree, c["campos"], geom, scratch)))
        assert helpers.bits_equal_count(m2d + 0.0, outs["means2D"] + 0.0) == 0
        total = {k: outs[k].clone() for k in keys} if total is None else {k: total[k] + outs[k] for k in keys}
        if P <= 20001:
            recs = _decode_records(gathered[v], P)
            nonzero = torch.zeros(P, dtype=torch.bool, device=dev)
            for k in keys:
                nonzero |= (outs[k].reshape(P, -1) != 0).any(1)
            assert sorted(recs) == nonzero.nonzero().flatten().tolist()
            i = next(iter(recs))
            import numpy as np
            want = np.concatenate([outs[k][i].cpu().numpy().reshape(-1) for k in ("means3D", "opacity", "scales", "rotations")])
            assert np.array_equal(recs[i][3:], want)
    got = C.dp_grad_record_combine(s["means3D"], campos, gathered, stride, degree, M)
    # ... and into outputs the caller zeroed itself
    pre = {k: torch.zeros_like(v) for k, v in got.items()}
    C.dp_grad_record_combine(s["means3D"], campos, gathered, stride, degree, M, outs=pre, outputs_are_zero=True)
    torch.cuda.synchronize()
    assert float(total["sh"].abs().max()) > 0 and float(total["rotations"].abs().max()) > 0
    for k in keys:
        assert got[k].shape == total[k].shape
        assert helpers.bits_equal_count(got[k] + 0.0, total[k] + 0.0) == 0, (k, helpers.rel_l2(got[k], total[k]))
        assert helpers.bits_equal_count(pre[k] + 0.0, total[k] + 0.0) == 0, k


@pytest.mark.parametrize("layout", ["records", "sparse", "dense"])
def test_factored_exchange_world_1_equals_fused_backward(layout):
    """dp.FactoredExchange without a process group: same gradients as the fused backward, in every layout."""
    from diff_gaussian_rasterization import dp
    C = native()
    P, W, H = 30000, 320, 200
    s, cs, dLs = _views("cube", P, W, H, 1, 3, 4.03, 0.6911)
    c, dL = cs[0], dLs[0]
    fwd = _forward(C, s, c, 3)
    R, _, radii, geom, binning, img, _ = fwd
    fused = _fused_backward(C, s, c, 3, dL, fwd)
    ex = dp.FactoredExchange(P, 16, torch.device("cuda:0"), layout=layout)
    assert ex.world == 1 and ex.layout == layout
    assert dp.FactoredExchange(P, 16, torch.device("cuda:0")).layout == "records"     # the default
    for _ in range(3):                                   # the buffers are reused step after step
        grads = ex.backward(C, c["bg"], s["means3D"], radii, s["scales"], s["rotations"], 1.0, c["viewmatrix"],
                            c["projmatrix"], c["tanfovx"], c["tanfovy"], dL, s["shs"], 3, c["campos"],
                            c["campos"].view(1, 3), geom, R, binning, img)
    torch.cuda.synchronize()
    for n in ("means3D", "sh", "opacity", "scales", "rotations"):
        assert helpers.rel_l2(grads[n].reshape(fused[n].shape), fused[n]) < 2e-5, n    # two runs of the tile backward: atomics
    assert helpers.rel_l2(ex.local["means2D"], fused["means2D"]) < 2e-5


@pytest.mark.parametrize("layout", ["records", "sparse"])
@pytest.mark.parametrize("P", [30001, 30002, 30003])
def test_ragged_gaussian_counts_through_arena_and_exchange(P, layout):
    """MCMC growth makes P arbitrary (int(1.05 * P)). Every field block of GradientArena and of the exchange slab must stay
    16-byte aligned for any P (the kernels' float4 paths), through resize() as well: same gradients as plain outputs."""
    from diff_gaussian_rasterization import dp
    C = native()
    W, H = 320, 200
    s, cs, dLs = _views("cube", P, W, H, 1, 3, 4.03, 0.6911)
    c, dL = cs[0], dLs[0]
    fwd = _forward(C, s, c, 3)
    R, _, radii, geom, binning, img, _ = fwd
    e = torch.empty(0)
    scratch = C.backward_render(c["bg"], P, R, dL, geom, binning, img)
    plain = dict(zip(NAMES, C.backward_geometry(s["means3D"], radii, s["scales"], s["rotations"], 1.0, e, c["viewmatrix"],
                                                c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W, s["shs"], 3, c["campos"],
                                                geom, scratch)))
    arena = dp.GradientArena(P, 16, torch.device("cuda:0"))
    for v in arena.views.values():
        assert v.data_ptr() % 16 == 0
    got = dict(zip(NAMES, C.backward_geometry(s["means3D"], radii, s["scales"], s["rotations"], 1.0, e, c["viewmatrix"],
                                              c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W, s["shs"], 3, c["campos"],
                                              geom, scratch, out_buffers=arena.backward_buffers())))
    torch.cuda.synchronize()
    for n in NAMES:
        assert helpers.bits_equal_count(got[n] + 0.0, plain[n] + 0.0) == 0, n
    # the exchange, allocated for a larger cap and resized down to the ragged P (and up again)
    ex = dp.FactoredExchange(P - 7, 16, torch.device("cuda:0"), capacity=P + 1000, layout=layout)
    ex.resize(P + 999).resize(P)
    for v in ex.grads.values():
        assert v.data_ptr() % 16 == 0 and v.shape[0] == P
    grads = ex.backward(C, c["bg"], s["means3D"], radii, s["scales"], s["rotations"], 1.0, c["viewmatrix"],
                        c["projmatrix"], c["tanfovx"], c["tanfovy"], dL, s["shs"], 3, c["campos"],
                        c["campos"].view(1, 3), geom, R, binning, img)
    torch.cuda.synchronize()
    for n in ("means3D", "sh", "opacity", "scales", "rotations"):
        assert helpers.rel_l2(grads[n], plain[n]) < 2e-5, n


def test_dp_entry_points_reject_bad_arguments():
    C = native()
    dev = torch.device("cuda:0")
    x = torch.zeros(64, device=dev)
    with pytest.raises(RuntimeError):
        C.dp_slab_all_reduce(dev, [x.data_ptr()], 6, 0, 2)           # n_ptrs != world
    with pytest.raises(RuntimeError):
        C.dp_slab_all_reduce(dev, [x.data_ptr(), x.data_ptr()], 6, 0, 2)   # not a multiple of 4 floats
    with pytest.raises(RuntimeError):
        C.dp_sh_grad_combine(torch.zeros((8, 3), device=dev), torch.zeros((1, 3), device=dev), x, 8, 3, 16)  # stride < 3P
    with pytest.raises(RuntimeError):
        C.dp_sh_grad_combine(torch.zeros((8, 3), device=dev), torch.zeros((1, 3), device=dev), x, 32, 3, 16, layout="sparse")  # < one chunk
    with pytest.raises(RuntimeError):                                  # a records slot is 1800 floats per 128 Gaussians
        C.dp_grad_record_combine(torch.zeros((8, 3), device=dev), torch.zeros((1, 3), device=dev), torch.zeros(1024, device=dev), 1024, 3, 16)
    with pytest.raises(RuntimeError):                                  # more views than the header staging holds
        C.dp_grad_record_combine(torch.zeros((8, 3), device=dev), torch.zeros((33, 3), device=dev), torch.zeros(33 * 1856, device=dev), 1856, 3, 16)
    assert C.dp_slot_floats(129, "records") == 3600 and C.dp_slot_floats(257, "sparse") == 1568 and C.dp_slot_floats(5, "dense") == 16


def test_slab_all_reduce_single_device_two_buffers():
    """The peer-pointer all-reduce kernel with both 'ranks' on one device: each call reduces one slice of two buffers."""
    C = native()
    dev = torch.device("cuda:0")
    n = 4 * 1031
    g = torch.Generator(device="cpu").manual_seed(1)
    a = torch.randn(n, generator=g).to(dev)
    b = torch.randn(n, generator=g).to(dev)
    want = a + b
    ptrs = [a.data_ptr(), b.data_ptr()]
    C.dp_slab_all_reduce(dev, ptrs, n, 0, 2)
    C.dp_slab_all_reduce(dev, ptrs, n, 1, 2)
    torch.cuda.synchronize()
    assert helpers.bits_equal_count(a, want) == 0
    assert helpers.bits_equal_count(b, want) == 0


@pytest.mark.parametrize("kind,P,W,H,degree,rad,fovx", [
    ("cube", 20001, 250, 130, 3, 4.03, 0.6911),          # ragged last warp: the float-by-float tail of the split staging
    ("cube", 5003, 96, 64, 1, 4.03, 0.6911),
    ("cube", 5000, 96, 64, 0, 4.03, 0.6911),
    ("cube", 4999, 96, 64, 2, 4.03, 0.6911),
    ("garden", 200000, 800, 450, 3, 6.0, 1.2),
])
def test_split_sh_tensors_equal_the_concatenated_layout(kind, P, W, H, degree, rad, fovx):
    """features_dc [P,1,3] / features_rest [P,15,3] read as they lie (gs_forward_preprocess_split_sh,
    gs_backward_geometry_split_sh, and the records push / combine) give the same bits as the reference's
    cat((features_dc, features_rest), dim=1) layout (scene/gaussian_model_static.py:139-144): image, radii, every gradient."""
    C = native()
    dev = torch.device("cuda:0")
    s, cs, dLs = _views(kind, P, W, H, 2, degree, rad, fovx)
    if s["shs"].shape[1] < 16:       # degree-3 storage with a lower active degree, as in the first 3000 training iterations
        g = torch.Generator(device="cpu").manual_seed(11)
        pad = torch.randn((P, 16 - s["shs"].shape[1], 3), generator=g).to(dev) * 0.1
        s["shs"] = torch.cat((s["shs"], pad), 1).contiguous()
    dc, rest = s["shs"][:, :1].contiguous(), s["shs"][:, 1:].contiguous()
    e = torch.empty(0)
    stride = (C.dp_slot_floats(P, "records") + 63) // 64 * 64
    slots = {name: torch.zeros((2, stride), device=dev) for name in ("cat", "split")}
    for v, (c, dL) in enumerate(zip(cs, dLs)):
        fwd_cat = _forward(C, s, c, degree)
        fwd_split = C.rasterize_gaussians(c["bg"], s["means3D"], e, s["opacities"], s["scales"], s["rotations"], 1.0, e,
                                          c["viewmatrix"], c["projmatrix"], c["tanfovx"], c["tanfovy"], H, W, (dc, rest), degree,
                                          c["campos"], False, False)
        assert fwd_cat[0] == fwd_split[0]
        assert helpers.bits_equal_count(fwd_cat[1], fwd_split[1]) == 0          # image
        assert torch.equal(fwd_cat[2], fwd_split[2]) and torch.equal(fwd_cat[6], fwd_split[6])   # radii, is_used
        R, _, radii, geom, binning, img, _ = fwd_split
        scratch = C.backward_render(c["bg"], P, R, dL, geom, binning, img)
        args = (s["means3D"], radii, s["scales"], s["rotations"], 1.0, e, c["viewmatrix"], c["projmatrix"], c["tanfovx"],
                c["tanfovy"], H, W)
        cat = dict(zip(NAMES, C.backward_geometry(*args, s["shs"], degree, c["campos"], geom, scratch)))
        split = dict(zip(NAMES, C.backward_geometry(*args, (dc, rest), degree, c["campos"], geom, scratch)))
        for n in NAMES:
            if n != "sh":
                assert helpers.bits_equal_count(split[n] + 0.0, cat[n] + 0.0) == 0, n
        assert float(cat["sh"].abs().max()) > 0
        assert split["sh"][0].shape == (P, 1, 3) and split["sh"][1].shape == (P, 15, 3)
        assert helpers.bits_equal_count(torch.cat(split["sh"], 1) + 0.0, cat["sh"] + 0.0) == 0
        # records: the same chunk records whichever layout the SH coefficients were read from
        pargs = (s["means3D"], radii, s["scales"], s["rotations"], 1.0, c["viewmatrix"], c["projmatrix"], c["tanfovx"],
                 c["tanfovy"], H, W)
        C.backward_geometry_push(*pargs, s["shs"], degree, c["campos"], geom, scratch, [slots["cat"].data_ptr()], v * stride)
        C.backward_geometry_push(*pargs, (dc, rest), degree, c["campos"], geom, scratch, [slots["split"].data_ptr()], v * stride)
    torch.cuda.synchronize()
    if P <= 20001:                                   # (the pad floats closing a chunk's run are not part of the record)
        for v in range(2):
            ra, rb = _decode_records(slots["cat"][v], P), _decode_records(slots["split"][v], P)
            assert sorted(ra) == sorted(rb) and all((ra[i].view("i4") == rb[i].view("i4")).all() for i in ra)
    campos = torch.stack([c["campos"] for c in cs])
    whole = C.dp_grad_record_combine(s["means3D"], campos, slots["cat"], stride, degree, 16)
    parts = C.dp_grad_record_combine(s["means3D"], campos, slots["split"], stride, degree, 16, split_sh=True)
    assert parts["sh"].shape == (P, 1, 3) and parts["sh_rest"].shape == (P, 15, 3)
    assert helpers.bits_equal_count(torch.cat((parts["sh"], parts["sh_rest"]), 1) + 0.0, whole["sh"] + 0.0) == 0
    for n in ("means3D", "opacity", "scales", "rotations"):
        assert helpers.bits_equal_count(parts[n] + 0.0, whole[n] + 0.0) == 0, n


def test_split_sh_through_the_public_rasterizer():
    """GaussianRasterizer(shs=(features_dc, features_rest)): gradients arrive on the two leaves, equal (up to the tile
    backward's atomics) to those of the concatenated call; dp.ViewParallelRasterizer takes the pair as well."""
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer, dp
    P, W, H = 30000, 320, 200
    s, cs, dLs = _views("cube", P, W, H, 1, 3, 4.03, 0.6911)
    c, dL = cs[0], dLs[0]
    st = GaussianRasterizationSettings(H, W, c["tanfovx"], c["tanfovy"], c["bg"], 1.0, c["viewmatrix"], c["projmatrix"], 3,
                                       c["campos"], False, False)

    def leaves():
        return {k: s[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "scales", "rotations")}

    a = leaves()
    shs = s["shs"].clone().requires_grad_(True)
    img_a = GaussianRasterizer(st)(means3D=a["means3D"], means2D=torch.zeros_like(a["means3D"]), opacities=a["opacities"], shs=shs,
                                   scales=a["scales"], rotations=a["rotations"])[0]
    (img_a * dL).sum().backward()
    results = {}
    for name in ("single", "view_parallel"):
        b = leaves()
        dc, rest = s["shs"][:, :1].clone().requires_grad_(True), s["shs"][:, 1:].clone().requires_grad_(True)
        if name == "single":
            img_b = GaussianRasterizer(st)(means3D=b["means3D"], means2D=torch.zeros_like(b["means3D"]), opacities=b["opacities"],
                                           shs=(dc, rest), scales=b["scales"], rotations=b["rotations"])[0]
        else:
            vp = dp.ViewParallelRasterizer(P, 16, torch.device("cuda:0"))
            img_b = vp(st, c["campos"].view(1, 3), b["means3D"], torch.zeros_like(b["means3D"]), b["opacities"], (dc, rest),
                       b["scales"], b["rotations"])[0]
        assert helpers.bits_equal_count(img_a, img_b) == 0
        (img_b * dL).sum().backward()
        assert dc.grad.shape == (P, 1, 3) and rest.grad.shape == (P, 15, 3)
        assert helpers.rel_l2(torch.cat((dc.grad, rest.grad), 1), shs.grad) < 2e-5, name
        for k in b:
            assert helpers.rel_l2(b[k].grad, a[k].grad) < 2e-5, (name, k)
        results[name] = True
    assert len(results) == 2


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on one node (gpurun --gpus 2)")
def test_two_rank_exchange_matches_full_all_reduce_every_transport():
    """The cross-DEVICE part the single-GPU tests cannot reach: two ranks under torchrun (tools/dp_check.py), every
    transport x layout (NCCL collectives; symmetric memory with dense / sparse / records layouts, multicast and peer
    pointers) against dp.GradientArena + one NCCL all-reduce of all 59 floats per Gaussian. Tolerance 1e-4 relative L2
    per tensor (both arms sum the tile backward with float atomics, run-to-run noise ~1e-6), and the summed gradients
    must be bit-identical on both ranks. Also a short replicated training run: the replicas stay in lock step."""
    import json
    import os
    import socket
    import subprocess
    import sys
    sock = socket.socket()
    sock.bind(("127.0.0.1", 0))
    port = sock.getsockname()[1]
    sock.close()
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(helpers.ROOT, "tools", "dp_check.py"), "--P", "100003", "--width", "640",
           "--height", "400", "--iters", "2", "--train-steps", "6"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-4000:]
    line = [ln for ln in out.stdout.splitlines() if ln.startswith("{")][-1]
    res = json.loads(line)
    assert res["world"] == 2
    checked = 0
    for key, val in res.items():
        if not key.startswith("factored_"):
            continue
        assert "error" not in val, (key, val)
        assert val["identical_on_all_ranks"], key
        for name, r in val["rel_l2_vs_full_all_reduce"].items():
            assert r < 1e-4, (key, name, r)
        checked += 1
    assert checked >= 5
    tr = res["train_replicas"]
    assert tr["steps"] == 6 and all(tr["identical_on_all_ranks"].values()), tr
