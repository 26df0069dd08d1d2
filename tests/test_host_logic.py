"""Host-side behaviour of the drop-in Python API (no GPU): argument validation mirrors the reference wrapper
(diff_gaussian_rasterization/__init__.py:171-220), there is no CPU fallback, the DP gradient arena and the
synthetic camera conventions are what the rest of the suite assumes."""
import math
import os
import socket
import subprocess
import sys
import textwrap

import numpy as np
import pytest
import torch

import gs_synth
from diff_gaussian_rasterization import (GaussianRasterizationSettings, GaussianRasterizer, compute_relocation,
                                         rasterize_gaussians)
from diff_gaussian_rasterization import _native, dp


def settings(W=32, H=16):
    cam = gs_synth.orbit_cameras(1, 4.0, W, H, 0.7)[0]
    return GaussianRasterizationSettings(H, W, cam["tanfovx"], cam["tanfovy"], torch.zeros(3), 1.0,
                                         torch.from_numpy(cam["viewmatrix"]), torch.from_numpy(cam["projmatrix"]), 3,
                                         torch.from_numpy(cam["campos"]), False, False)


def test_settings_fields_match_reference_order():
    assert GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix", "projmatrix",
        "sh_degree", "campos", "prefiltered", "debug")


def test_exactly_one_colour_source_and_one_covariance_source():
    r = GaussianRasterizer(settings())
    P = 4
    m, o = torch.zeros(P, 3), torch.ones(P, 1)
    sh, col = torch.zeros(P, 16, 3), torch.zeros(P, 3)
    sc, rot, cov = torch.ones(P, 3), torch.ones(P, 4), torch.zeros(P, 6)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(m, m, o, shs=None, colors_precomp=None, scales=sc, rotations=rot)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(m, m, o, shs=sh, colors_precomp=col, scales=sc, rotations=rot)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(m, m, o, shs=sh, scales=sc, rotations=None)
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(m, m, o, shs=sh, scales=sc, rotations=rot, cov3D_precomp=cov)


def test_no_cpu_fallback():
    r = GaussianRasterizer(settings())
    P = 4
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        r(torch.zeros(P, 3), torch.zeros(P, 3), torch.ones(P, 1), shs=torch.zeros(P, 16, 3), scales=torch.ones(P, 3),
          rotations=torch.ones(P, 4))
    with pytest.raises(RuntimeError, match="num_points, 3"):
        rasterize_gaussians(torch.zeros(P, 4), torch.zeros(P, 3), torch.zeros(P, 16, 3), torch.Tensor([]), torch.ones(P, 1),
                            torch.ones(P, 3), torch.ones(P, 4), torch.Tensor([]), settings())


def test_compute_relocation_empty_is_a_noop_without_gpu():
    o, s = compute_relocation(torch.zeros(0), torch.zeros(0, 3), torch.zeros(0, dtype=torch.int64), torch.zeros(51, 51), 51)
    assert o.shape == (0,) and s.shape == (0,)


def test_gradient_arena_layout():
    a = dp.GradientArena(7, 16, "cpu")
    assert a.floats_per_gaussian() == 59
    bufs = a.backward_buffers()
    assert set(bufs) == {"means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations"}
    for name in ("means3D", "sh", "opacity", "scales", "rotations"):
        assert bufs[name].is_contiguous()
        bufs[name].fill_(1.0)
    assert float(a.flat.sum()) == 7 * 59          # the five parameter gradients tile the arena (pad floats stay zero)
    assert dp.shard_views(8, 1, 4) == [1, 5]
    assert a.all_reduce() is a.flat                # no process group: no-op


def test_factored_exchange_capacity_and_resize():
    ex = dp.FactoredExchange(10, 4, "cpu", capacity=24)
    assert ex.world == 1 and ex.P == 10 and ex.slot_stride >= 3 * 24
    ptr = ex.flat.data_ptr()
    ex.resize(17)
    assert ex.flat.data_ptr() == ptr                      # same buffer, new views
    assert ex.grads["means3D"].shape == (17, 3) and ex.grads["rotations"].shape == (17, 4) and ex.grads["sh"].shape == (17, 4, 3)
    # every field block starts on a 16-byte boundary for any P (17 * 3 = 51 -> 52, + 17 -> 69 -> 72, + 51 -> 123 -> 124)
    assert ex.grads["rotations"].data_ptr() == ptr + 4 * 124
    for P in range(1, 40):
        ex2 = dp.FactoredExchange(P, 4, "cpu", capacity=40)
        offs = [(v.data_ptr() - ex2.flat.data_ptr()) // 4 for k, v in ex2.grads.items() if k != "sh"]
        assert all(o % 4 == 0 for o in offs) and offs == sorted(offs)
        assert ex2._n_slab <= ex2._slab_capacity and ex2._n_slab >= offs[-1] + 4 * P
        a2 = dp.GradientArena(P, 16, "cpu")
        assert all((v.data_ptr() - a2.flat.data_ptr()) % 16 == 0 for v in a2.views.values())
    assert ex._n_slab % 64 == 0 and ex._n_slab >= 11 * 17
    with pytest.raises(ValueError):
        ex.resize(25)
    with pytest.raises(ValueError):
        dp.FactoredExchange(10, 4, "cpu", capacity=5)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_arena_all_reduce_world_size_2_gloo(tmp_path):
    """The N>1 exchange step on CPU: two gloo ranks fill their arenas differently; after all_reduce both hold the sum."""
    script = tmp_path / "worker.py"
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script.write_text(textwrap.dedent(f"""
        import os, sys
        sys.path.insert(0, {os.path.join(root, '3dgs-mcmc_b200')!r})
        import torch, torch.distributed as dist
        from diff_gaussian_rasterization import dp
        dist.init_process_group("gloo")
        r = dist.get_rank()
        a = dp.GradientArena(11, 16, "cpu")
        for i, (name, v) in enumerate(sorted(a.views.items())):
            v.fill_(float((r + 1) * (i + 1)))
        a.all_reduce()
        for i, (name, v) in enumerate(sorted(a.views.items())):
            assert torch.all(v == 3.0 * (i + 1)), name
        a.all_reduce(average=True)
        for i, (name, v) in enumerate(sorted(a.views.items())):
            assert torch.all(v == 3.0 * (i + 1)), name
        assert dp.shard_views(6, r, 2) == [r, r + 2, r + 4]
        dist.destroy_process_group()
        open(os.path.join({str(tmp_path)!r}, f"ok{{r}}"), "w").write("ok")
    """))
    port = _free_port()
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                         capture_output=True, text=True, timeout=240)
    assert out.returncode == 0, out.stdout + out.stderr
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def test_factored_exchange_world_size_2_gloo(tmp_path):
    """dp.FactoredExchange over gloo (transport "nccl" = plain torch.distributed collectives) with the CUDA kernels
    replaced by numpy stand-ins: after the exchange both ranks hold the summed 11-float slab and the SH gradient
    sum_v basis(dir_v) (x) g_v rebuilt from the gathered colour gradients, bit-identical on both ranks."""
    script = tmp_path / "worker.py"
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script.write_text(textwrap.dedent(f"""
        import ctypes, os, sys
        import numpy as np
        sys.path.insert(0, {os.path.join(root, '3dgs-mcmc_b200')!r})
        import torch, torch.distributed as dist
        from diff_gaussian_rasterization import dp
        dist.init_process_group("gloo")
        r, world = dist.get_rank(), dist.get_world_size()
        P, M = 37, 4                                   # degree 1
        rng = np.random.default_rng(5)
        means = rng.normal(size=(P, 3)).astype(np.float32)
        campos = np.array([[4.0, 0.5, 0.0], [-3.0, 1.0, 2.0]], np.float32)
        g_all = rng.normal(size=(world, P, 3)).astype(np.float32)
        g_all[1, ::3] = 0.0                            # Gaussians culled in view 1

        def basis(d):                                  # degree-1 real SH basis of unit vectors d[P,3]
            c0, c1 = 0.28209479177387814, 0.4886025119029199
            return np.stack([np.full(len(d), c0, np.float32), -c1 * d[:, 1], c1 * d[:, 2], -c1 * d[:, 0]], 1).astype(np.float32)

        class FakeNative:                              # stands in for _native on a box without a GPU (test only)
            def dp_color_grad_push(self, P_, radii, geom, scratch, dst_ptrs, slot_offset, multicast, layout=None):
                src = np.ascontiguousarray(g_all[r].reshape(-1))
                for p in dst_ptrs:
                    ctypes.memmove(p + 4 * slot_offset, src.ctypes.data, src.nbytes)
            def dp_sh_grad_combine(self, means3D, campos_views, gathered, slot_stride, degree, M_, out=None, layout=None):
                acc = np.zeros((P, M_, 3), np.float32)
                for v in range(campos_views.shape[0]):
                    g = gathered.view(-1).numpy()[v * slot_stride: v * slot_stride + 3 * P].reshape(P, 3)
                    d = means3D.numpy() - campos_views[v].numpy()
                    d = d / np.sqrt((d * d).sum(1, keepdims=True))
                    acc += basis(d)[:, :, None] * g[:, None, :]
                out.copy_(torch.from_numpy(acc))
                return out

        ex = dp.FactoredExchange(P, M, "cpu", transport="auto")
        assert ex.transport == "nccl" and ex.world == 2 and ex.slot_stride % 64 == 0 and ex.slot_stride >= 3 * P
        assert set(ex.out_buffers()) == {{"means2D", "colors", "cov3D", "means3D", "opacity", "scales", "rotations"}}
        for i, name in enumerate(("means3D", "opacity", "scales", "rotations")):
            ex.grads[name].fill_(float((r + 1) * (i + 1)))       # what the per-Gaussian backward would have written
        native = FakeNative()
        pending = ex.push_color_grads(native, None, None, None)
        grads = ex.finish(native, torch.from_numpy(means), torch.from_numpy(campos), 1, pending)
        for i, name in enumerate(("means3D", "opacity", "scales", "rotations")):
            assert torch.all(grads[name] == 3.0 * (i + 1)), name
        want = np.zeros((P, M, 3), np.float32)
        for v in range(world):
            d = means - campos[v]
            d = d / np.sqrt((d * d).sum(1, keepdims=True))
            want += basis(d)[:, :, None] * g_all[v][:, None, :]
        assert np.array_equal(grads["sh"].numpy(), want)
        both = [torch.zeros_like(grads["sh"]) for _ in range(world)]
        dist.all_gather(both, grads["sh"])
        assert torch.equal(both[0], both[1])
        rx, slab = ex.bytes_per_step()
        assert rx == (world - 1) * 12 * P and slab == 44 * P
        dist.destroy_process_group()
        open(os.path.join({str(tmp_path)!r}, f"ok{{r}}"), "w").write("ok")
    """))
    port = _free_port()
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                         capture_output=True, text=True, timeout=240)
    assert out.returncode == 0, out.stdout + out.stderr
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def test_factored_exchange_auto_layout_follows_the_final_transport():
    from diff_gaussian_rasterization import dp
    f = dp.FactoredExchange._auto_layout
    assert f(1, "nccl", "cuda") == "records" and f(8, "symm", "cuda") == "records"
    assert f(8, "nccl", "cuda") == "dense" and f(2, "symm", "cpu") == "dense"
    ex = dp.FactoredExchange(1000, 16, "cpu", layout="records")         # world 1: upper bound of what a view would send
    assert ex.bytes_per_step() == (0, 0)


def test_factored_exchange_rendezvous_failure_on_one_rank_world_size_2_gloo(tmp_path):
    """The symmetric-memory set-up fails on rank 1 only (forced). Both ranks must end on the same transport AND on the
    layout / slot stride / buffer size that belong to it - a rank left on the records layout would enter the all-gather
    with a different size and hang or corrupt the run."""
    script = tmp_path / "worker.py"
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script.write_text(textwrap.dedent(f"""
        import os, sys, warnings
        sys.path.insert(0, {os.path.join(root, '3dgs-mcmc_b200')!r})
        import torch, torch.distributed as dist
        from diff_gaussian_rasterization import dp
        dist.init_process_group("gloo")
        r, world = dist.get_rank(), dist.get_world_size()
        X = dp.FactoredExchange
        X._pick_transport = lambda self, t: "symm"                  # every rank believes it can do symmetric memory
        X._auto_layout = staticmethod(lambda world, transport, dev: "records" if transport == "symm" else "dense")
        tried = []
        def setup(self, total, slab_floats, group, multicast):
            tried.append((self.layout, self.slot_stride, total))
            if r == 1:
                raise RuntimeError("forced: no peer access on this rank")
            self.flat, self._hdl, self._multicast = torch.zeros(total), object(), False
        X._setup_symmetric = setup
        P, cap = 1001, 1100
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            ex = X(P, 16, "cpu", capacity=cap)
        # the rendezvous was attempted once, with the records layout's double-buffered gather and no slab
        assert len(tried) == 1 and tried[0][0] == "records" and tried[0][2] == 2 * world * tried[0][1]
        assert ex.transport == "nccl" and ex._hdl is None and ex.layout == "dense" and ex._nbuf == 1
        assert ex.slot_stride == (3 * cap + 63) // 64 * 64
        slab = (11 * cap + 12 + 63) // 64 * 64
        assert ex.flat.numel() == slab + world * ex.slot_stride and ex._send.numel() == ex.slot_stride
        mine = torch.tensor([ex.slot_stride, ex.flat.numel(), ex._nbuf], dtype=torch.int64)
        both = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(both, mine)
        assert torch.equal(both[0], both[1])
        ex._send.fill_(float(r + 1))                                # the fallback's gather runs with matching sizes
        dist.all_gather_into_tensor(ex.gathered.view(-1), ex._send)
        assert torch.all(ex.gathered[0] == 1.0) and torch.all(ex.gathered[1] == 2.0)
        assert ex.bytes_per_step() == ((world - 1) * 12 * P, 44 * P)
        dist.destroy_process_group()
        open(os.path.join({str(tmp_path)!r}, f"ok{{r}}"), "w").write("ok")
    """))
    port = _free_port()
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                         capture_output=True, text=True, timeout=240)
    assert out.returncode == 0, out.stdout + out.stderr
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def test_replica_guard_world_size_2_gloo(tmp_path):
    """dp.assert_replicas_in_sync: passes while the replicas are bit-identical (also after a resize to a ragged P), raises
    on EVERY rank as soon as one rank's copy differs in a single bit - the guard view-sharded MCMC training calls after
    each relocation / growth step."""
    script = tmp_path / "worker.py"
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script.write_text(textwrap.dedent(f"""
        import os, sys
        sys.path.insert(0, {os.path.join(root, '3dgs-mcmc_b200')!r})
        import torch, torch.distributed as dist
        from diff_gaussian_rasterization import dp
        dist.init_process_group("gloo")
        r = dist.get_rank()
        g = torch.Generator().manual_seed(3)
        xyz, op = torch.randn((1001, 3), generator=g), torch.randn((1001, 1), generator=g)
        idx = torch.randint(0, 1001, (64,), generator=g)
        assert dp.assert_replicas_in_sync([xyz, op, idx])
        ex = dp.FactoredExchange(1001, 16, "cpu", capacity=1100)
        ex.resize(1051)                               # int(1.05 * 1001): ragged
        assert all(v.data_ptr() % 16 == 0 for k, v in ex.grads.items() if k != "sh")
        if r == 1:
            xyz.view(torch.int32)[500, 1] ^= 1        # one mantissa bit on one rank
        try:
            dp.assert_replicas_in_sync([xyz, op, idx])
            raised = False
        except RuntimeError as ex_:
            raised = "out of lock step" in str(ex_)
        assert raised
        # a different Gaussian count (one rank grew, the other did not) is caught as well
        try:
            dp.assert_replicas_in_sync([op[: 1001 - r]])
            raised = False
        except RuntimeError:
            raised = True
        assert raised
        dist.destroy_process_group()
        open(os.path.join({str(tmp_path)!r}, f"ok{{r}}"), "w").write("ok")
    """))
    port = _free_port()
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                         capture_output=True, text=True, timeout=240)
    assert out.returncode == 0, out.stdout + out.stderr
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def test_replica_fingerprint_sees_every_bit_of_any_dtype():
    """Values are never converted: a one-bit change in fp16 / bf16 / fp64 / int8 state (all of them collapse under
    .to(int64) for magnitudes below 1) changes the fingerprint."""
    from diff_gaussian_rasterization import dp
    g = torch.Generator().manual_seed(4)
    for dtype, word in ((torch.float16, torch.int16), (torch.bfloat16, torch.int16), (torch.float64, torch.int64),
                        (torch.float32, torch.int32), (torch.int8, torch.int8)):
        base = (torch.rand(257, generator=g) * 0.9).to(dtype) if dtype != torch.int8 else torch.arange(257).to(dtype)
        other = base.clone()
        other.view(word)[100] ^= 1
        assert not torch.equal(dp.replica_fingerprint([base]), dp.replica_fingerprint([other])), dtype
        assert torch.equal(dp.replica_fingerprint([base]), dp.replica_fingerprint([base.clone()]))
    flags = torch.zeros(64, dtype=torch.bool)
    flipped = flags.clone(); flipped[7] = True
    assert not torch.equal(dp.replica_fingerprint([flags]), dp.replica_fingerprint([flipped]))


def test_lockstep_sampling_world_size_2_gloo(tmp_path):
    """dp.lockstep_sampling: ranks whose multinomial draws would differ (here: differently seeded generators, on CUDA:
    the unordered cumsum inside torch.multinomial) end up with rank 0's indices; outside the context torch.multinomial is
    the stock function again. The caller is the reference's _sample_alives (scene/gaussian_model_static.py:412-422)."""
    script = tmp_path / "worker.py"
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script.write_text(textwrap.dedent(f"""
        import os, sys
        sys.path.insert(0, {os.path.join(root, '3dgs-mcmc_b200')!r})
        import torch, torch.distributed as dist
        from diff_gaussian_rasterization import dp
        dist.init_process_group("gloo")
        r = dist.get_rank()
        stock, stock_method = torch.multinomial, torch.Tensor.multinomial
        torch.manual_seed(10 + r)                      # the ranks would draw differently
        probs = torch.linspace(0.1, 1.0, 5000)
        with dp.lockstep_sampling():
            idx = torch.multinomial(probs, 700, replacement=True)
            ratio = torch.bincount(idx, minlength=5000)
            idx_m = probs.multinomial(700, replacement=True)            # the method form is in lock step too
            import threading
            seen = []
            t = threading.Thread(target=lambda: seen.append(torch.multinomial(probs, 50, replacement=True)))
            t.start(); t.join()                                          # another thread: stock draw, no collective
            try:
                with dp.lockstep_sampling():
                    nested = True
            except RuntimeError:
                nested = False
            assert not nested
        assert torch.multinomial is stock and torch.Tensor.multinomial is stock_method and len(seen) == 1
        both = [torch.zeros_like(idx_m) for _ in range(2)]
        dist.all_gather(both, idx_m)
        assert torch.equal(both[0], both[1])
        both = [torch.zeros_like(idx) for _ in range(2)]
        dist.all_gather(both, idx)
        assert torch.equal(both[0], both[1])
        assert dp.assert_replicas_in_sync([idx, ratio])
        free = torch.multinomial(probs, 700, replacement=True)
        both = [torch.zeros_like(free) for _ in range(2)]
        dist.all_gather(both, free)
        assert not torch.equal(both[0], both[1])        # without the context the differently seeded ranks disagree
        dist.destroy_process_group()
        open(os.path.join({str(tmp_path)!r}, f"ok{{r}}"), "w").write("ok")
    """))
    port = _free_port()
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                         capture_output=True, text=True, timeout=240)
    assert out.returncode == 0, out.stdout + out.stderr
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def test_camera_conventions():
    cam = gs_synth.look_at_camera((0.0, 0.0, -4.0), (0.0, 0.0, 0.0), 64, 48, 0.8)
    V = cam["viewmatrix"]            # transposed storage: element [4*i+j] of the transposed matrix
    m = V.reshape(-1)
    p = np.array([0.1, -0.2, 0.3], np.float32)
    view = np.array([m[0] * p[0] + m[4] * p[1] + m[8] * p[2] + m[12],
                     m[1] * p[0] + m[5] * p[1] + m[9] * p[2] + m[13],
                     m[2] * p[0] + m[6] * p[1] + m[10] * p[2] + m[14]])
    assert abs(view[2] - 4.3) < 1e-5            # +z is forward
    assert np.allclose(cam["campos"], [0, 0, -4], atol=1e-6)
    assert np.allclose(cam["projmatrix"], V @ (np.linalg.inv(V) @ cam["projmatrix"]), atol=1e-5)
    assert math.isclose(cam["tanfovx"], math.tan(0.4), rel_tol=1e-6)
    # aspect: tanfovy / tanfovx = H / W
    assert math.isclose(cam["tanfovy"] / cam["tanfovx"], 48 / 64, rel_tol=1e-6)


def test_scene_generators_are_deterministic_and_well_formed():
    a, b = gs_synth.scene_garden(2000, seed=4), gs_synth.scene_garden(2000, seed=4)
    for k in a:
        assert np.array_equal(a[k], b[k]) and a[k].dtype == np.float32
    assert a["shs"].shape == (2000, 16, 3) and a["opacities"].shape == (2000, 1)
    assert np.allclose(np.linalg.norm(a["rotations"], axis=1), 1.0, atol=1e-5)
    assert (a["scales"] > 0).all() and (a["opacities"] >= 0.005).all() and (a["opacities"] <= 0.99).all()
    c = gs_synth.scene_cube(1000, seed=1)
    assert np.abs(c["means3D"]).max() <= 1.3


def test_binning_workspace_size_classes():
    """_stable_size rounds the binning workspace up to 1/8-octave steps so that views with slightly different instance
    counts reuse one cached block (no cudaMalloc inside the training loop): never smaller than asked, at most 12.5 %
    larger, idempotent, monotone, and small sizes pass through."""
    from diff_gaussian_rasterization import _native as C
    prev = 0
    for n in [0, 1, 4096, (1 << 20), (1 << 20) + 1, 3 << 20, 123456789, 780_000_000, (1 << 33) + 12345]:
        s = C._stable_size(n)
        assert s >= n and s <= n + (n >> 3) + 1
        assert C._stable_size(s) == s
        assert s >= prev
        prev = s
    assert C._stable_size(1000) == 1000
    # neighbouring instance counts (camera to camera: a few per cent) land in at most two classes
    base = 32_000_000 * 24
    classes = {C._stable_size(int(base * (1 + 0.01 * k))) for k in range(6)}
    assert len(classes) <= 2


def test_binning_workspace_high_water_capacity():
    """One allocation size per device: a slightly larger view is served from the headroom (no new size -> no cudaMalloc in
    the training loop), a much larger one raises the mark once, a much smaller scene lowers it again."""
    from diff_gaussian_rasterization import _native as C
    key = "test-device"
    a = C._binning_capacity(key, 400_000_000)
    assert a >= 500_000_000 * 0.99 and C._stable_size(a) == a
    for k in range(1, 24):                      # views up to 24 % larger: same size
        assert C._binning_capacity(key, int(400_000_000 * (1 + 0.01 * k))) == a
    b = C._binning_capacity(key, 600_000_000)
    assert b > a and C._binning_capacity(key, 400_000_000) == b
    c = C._binning_capacity(key, 1_000_000)     # a tiny scene afterwards does not keep allocating half a gigabyte
    assert c < 2_000_000


def test_l1_loss_symbols_and_python_api_exist():
    from diff_gaussian_rasterization import _native as C, loss
    for sym in ("gs_l1_workspace_bytes", "gs_l1_loss_forward", "gs_l1_loss_backward"):
        assert sym in C.SIGNATURES
    assert callable(loss.l1_loss) and "l1_loss" in loss.__all__
    assert C._lib.gs_l1_workspace_bytes(1) > 0 and C._lib.gs_l1_workspace_bytes(3 * 1080 * 1920) >= 8 * 1519


def test_bench_presets_cover_baseline_configs(monkeypatch):
    """bench.py --config names one preset per BASELINE.json config that fits the bench contract (2-5; config 1 is the CPU
    oracle's parity case), and the default is the headline metric."""
    import importlib
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    monkeypatch.syspath_prepend(root)
    bench = importlib.import_module("bench")
    assert {"headline", "cpu100k", "lego300k", "garden1297", "large6m", "swin"} <= set(bench.CONFIGS)
    monkeypatch.setattr(sys, "argv", ["bench.py"])
    args = bench.parse()
    assert (args.P, args.width, args.height, args.gpus) == (3_000_000, 1920, 1080, 1) and args.warmup >= 3
    assert bench.metric_name(args) == bench.METRIC and "3M Gaussians" in bench.METRIC
    monkeypatch.setattr(sys, "argv", ["bench.py", "--config", "swin", "--gpus", "8"])
    args = bench.parse()
    assert (args.width, args.height) == (1352, 1014) and args.P == 800_000
