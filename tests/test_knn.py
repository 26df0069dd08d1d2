"""distCUDA2 (csrc/knn.cu, simple_knn/_C.py).

CPU: the brute-force C oracle against scipy's k-d tree, and the drop-in import path. GPU: our kernel against the
reference's own simple-knn compiled for sm_100 (oracle/_ref/libref_knn.so) — bit-exact, the result is an exact
nearest-neighbour query — on uniform, strongly non-uniform, degenerate and tiny inputs, and at 3M points.
"""
import numpy as np
import pytest
import torch

import helpers  # noqa: F401  (path setup)
import gs_synth
from oracle import oracle


def _kdtree_mean_dist2(pts):
    from scipy.spatial import cKDTree
    d, _ = cKDTree(pts.astype(np.float64)).query(pts.astype(np.float64), k=4)
    return (d[:, 1:] ** 2).mean(1)


def test_knn_oracle_matches_kdtree():
    rng = np.random.default_rng(0)
    pts = rng.normal(size=(700, 3)).astype(np.float32)
    got = oracle.knn_mean_dist2(pts)
    want = _kdtree_mean_dist2(pts)
    assert np.allclose(got, want, rtol=2e-6, atol=0)


def test_knn_oracle_counts_coincident_points():
    pts = np.array([[0, 0, 0], [0, 0, 0], [1, 0, 0], [0, 2, 0], [5, 5, 5]], np.float32)
    got = oracle.knn_mean_dist2(pts)
    assert got[0] == np.float32((0 + 1 + 4) / 3.0)        # its twin at distance 0 counts
    assert got[4] == np.float32((59 + 66 + 75) / 3.0)   # (1,0,0): 16+25+25=66; (0,2,0): 25+9+25=59; origin: 75 (x2, one used)


def test_simple_knn_drop_in_import_and_no_cpu_path():
    from simple_knn._C import distCUDA2        # scene/gaussian_model_static.py:24
    with pytest.raises(RuntimeError):
        distCUDA2(torch.zeros((4, 3)))


def _ours(pts):
    from simple_knn._C import distCUDA2
    return distCUDA2(pts)


def _ref(pts):
    from oracle import ref_cuda
    if not ref_cuda.knn_available():
        pytest.fail("oracle/_ref/libref_knn.so did not travel with the snapshot (build it with `make -C oracle ref`)")
    return ref_cuda.knn(pts)


def _cases():
    rng = np.random.default_rng(1)
    cube = gs_synth.scene_cube(100000, 0)["means3D"]
    garden = gs_synth.scene_garden(300000, 0)["means3D"]           # dense disc + sparse shell out to radius 30
    dup = np.repeat(rng.normal(size=(2000, 3)).astype(np.float32), 3, axis=0)          # every point three times
    planar = rng.normal(size=(5000, 3)).astype(np.float32); planar[:, 2] = 0.25       # degenerate axis
    line = np.zeros((3000, 3), np.float32); line[:, 0] = rng.uniform(-1, 1, 3000)     # two degenerate axes
    clustered = np.concatenate([rng.normal(size=(20000, 3)) * 1e-3, rng.normal(size=(50, 3)) * 100]).astype(np.float32)
    tiny4 = rng.normal(size=(4, 3)).astype(np.float32)
    return {"cube": cube, "garden": garden, "duplicates": dup, "planar": planar, "line": line, "clustered": clustered,
            "four_points": tiny4}


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["cube", "garden", "duplicates", "planar", "line", "clustered", "four_points"])
def test_knn_bit_exact_vs_reference(name):
    dev = torch.device("cuda:0")
    pts = torch.from_numpy(np.ascontiguousarray(_cases()[name])).to(dev)
    got = _ours(pts)
    want = _ref(pts)
    torch.cuda.synchronize()
    assert got.shape == want.shape
    bad = int((got.view(torch.int32) != want.view(torch.int32)).sum().item())
    assert bad == 0, (name, bad, float((got - want).abs().max()))


@pytest.mark.gpu
def test_knn_vs_cpu_oracle_and_edge_sizes():
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(2)
    pts = (rng.normal(size=(6000, 3)) * np.array([1.0, 0.1, 10.0])).astype(np.float32)
    got = _ours(torch.from_numpy(pts).to(dev)).cpu().numpy()
    assert np.array_equal(got, oracle.knn_mean_dist2(pts))
    assert _ours(torch.zeros((0, 3), device=dev)).shape == (0,)
    for P in (1, 2, 3):                       # fewer than three neighbours: FLT_MAX stays in the sum, as in the reference
        p = torch.from_numpy(rng.normal(size=(P, 3)).astype(np.float32)).to(dev)
        a, b = _ours(p), _ref(p)
        assert bool(((a == b) | (torch.isinf(a) & torch.isinf(b))).all())


@pytest.mark.gpu
def test_knn_full_size_3M_bit_exact_and_faster_than_reference():
    dev = torch.device("cuda:0")
    pts = torch.from_numpy(gs_synth.scene_garden(3_000_000, 0)["means3D"]).to(dev)
    _ours(pts[:1000])
    torch.cuda.synchronize()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record()
    got = _ours(pts)
    e1.record()
    torch.cuda.synchronize()
    want = _ref(pts)
    e2.record()
    torch.cuda.synchronize()
    assert int((got.view(torch.int32) != want.view(torch.int32)).sum().item()) == 0
    ours_ms, ref_ms = e0.elapsed_time(e1), e1.elapsed_time(e2)
    print(f"distCUDA2 at 3M points: ours {ours_ms:.2f} ms, reference simple-knn {ref_ms:.2f} ms")
    assert ours_ms < ref_ms
