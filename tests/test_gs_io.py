"""On-disk / wire formats (3dgs-mcmc_b200/gs_io.py). Host-side, CPU only.

stream_dump: byte-identical to golden bytes produced by the reference's own utils/stream_utils.py (and to the reference
function itself where /root/reference is mounted). PLY: plyfile is absent from the image, so the layout is checked
through the header text, a round trip and the reference's attribute order (parity with plyfile itself: unpinned).
"""
import contextlib
import io
import json
import os
import sys

import numpy as np
import pytest
import torch

import helpers  # noqa: F401  (path setup)
import gs_io

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "stream_dump.npz")
KEYS = ("start_frame", "end_frame", "xyz", "f_dc", "f_rest", "scaling", "rotation", "opacity")


def _case(name):
    d = np.load(GOLD)
    return int(d[f"{name}/degree"]), {k: torch.from_numpy(d[f"{name}/in/{k}"]) for k in KEYS}, d[f"{name}/bytes"].tobytes(), \
        d[f"{name}/format_json"].tobytes()


@pytest.mark.parametrize("name", ["deg0", "deg1", "deg3"])
def test_stream_dump_bytes_match_reference_golden(name, tmp_path):
    degree, params, want_bytes, want_json = _case(name)
    path = str(tmp_path / "stream.bin")
    assert gs_io.stream_dump(params, path, shs_degree=degree) == params["xyz"].shape[0]
    gs_io.stream_dump(params, path, shs_degree=degree)          # the golden appended twice as well
    assert open(path, "rb").read() == want_bytes
    assert open(tmp_path / "format.json", "rb").read() == want_json
    # load: arrays and the reference's list-of-tuples form
    arrays = gs_io.stream_load_arrays(str(tmp_path / "format.json"), path)
    N = params["xyz"].shape[0]
    assert arrays["xyz"].shape == (2 * N, 3)
    assert np.array_equal(arrays["xyz"][:N], params["xyz"].numpy())
    assert np.array_equal(arrays["f_rest"][:N], params["f_rest"].reshape(N, -1).numpy())
    assert np.array_equal(arrays["start_frame"][:N, 0], params["start_frame"].numpy().astype(np.uint32))
    rows = gs_io.stream_load(str(tmp_path / "format.json"), path)
    assert len(rows) == 2 * N and len(rows[0]) == 2 + 3 + 3 + 3 * ((degree + 1) ** 2 - 1) + 3 + 4 + 1
    assert rows[0][0] == int(params["start_frame"][0]) and rows[0][2] == float(params["xyz"][0, 0])


@pytest.mark.skipif(not os.path.exists("/root/reference/utils/stream_utils.py"), reason="reference tree not mounted")
def test_stream_dump_matches_reference_function_directly(tmp_path):
    sys.path.insert(0, "/root/reference")
    from utils.stream_utils import stream_dump as ref_dump, stream_load as ref_load
    g = torch.Generator().manual_seed(1)
    N, degree = 301, 2
    p = {"start_frame": torch.randint(0, 9, (N,), generator=g).float(), "end_frame": torch.randint(9, 20, (N,), generator=g).float(),
         "xyz": torch.randn((N, 3), generator=g), "f_dc": torch.randn((N, 1, 3), generator=g), "f_rest": torch.randn((N, 8, 3), generator=g),
         "scaling": torch.randn((N, 3), generator=g), "rotation": torch.randn((N, 4), generator=g), "opacity": torch.randn((N, 1), generator=g)}
    (tmp_path / "a").mkdir(); (tmp_path / "b").mkdir()
    with contextlib.redirect_stdout(io.StringIO()):
        ref_dump({k: v.clone() for k, v in p.items()}, str(tmp_path / "a" / "s.bin"), shs_degree=degree)
        gs_io.stream_dump(p, str(tmp_path / "b" / "s.bin"), shs_degree=degree)
        assert open(tmp_path / "a" / "s.bin", "rb").read() == open(tmp_path / "b" / "s.bin", "rb").read()
        assert json.load(open(tmp_path / "a" / "format.json")) == json.load(open(tmp_path / "b" / "format.json"))
        assert ref_load(str(tmp_path / "a" / "format.json"), str(tmp_path / "a" / "s.bin")) == \
            gs_io.stream_load(str(tmp_path / "b" / "format.json"), str(tmp_path / "b" / "s.bin"))


def test_stream_dump_rejects_what_the_reference_rejects(tmp_path):
    degree, params, *_ = _case("deg1")
    bad = dict(params)
    bad["xyz"] = params["xyz"][:-1]
    with pytest.raises(AssertionError):
        gs_io.stream_dump(bad, str(tmp_path / "s.bin"), shs_degree=degree)
    neg = dict(params)
    neg["start_frame"] = params["start_frame"] - 100.0
    with pytest.raises(ValueError):                      # struct.error('I' format requires 0 <= number) in the reference
        gs_io.stream_dump(neg, str(tmp_path / "s.bin"), shs_degree=degree)


def test_ply_round_trip_and_header(tmp_path):
    g = torch.Generator().manual_seed(2)
    N, M = 41, 16
    p = {"xyz": torch.randn((N, 3), generator=g), "features_dc": torch.randn((N, 1, 3), generator=g),
         "features_rest": torch.randn((N, M - 1, 3), generator=g), "opacity": torch.randn((N, 1), generator=g),
         "scaling": torch.randn((N, 3), generator=g), "rotation": torch.randn((N, 4), generator=g)}
    path = str(tmp_path / "point_cloud" / "iteration_7" / "point_cloud.ply")
    gs_io.save_ply(path, **p)
    raw = open(path, "rb").read()
    header, _, body = raw.partition(b"end_header\n")
    lines = header.decode().splitlines()
    assert lines[:3] == ["ply", "format binary_little_endian 1.0", f"element vertex {N}"]
    props = [ln.split()[-1] for ln in lines[3:]]
    assert props == gs_io.ply_attributes(3, 45) and len(props) == 62          # gaussian_model_static.py:214-227
    assert all(ln.startswith("property float ") for ln in lines[3:])
    assert len(body) == N * 62 * 4
    rec = np.frombuffer(body, dtype="<f4").reshape(N, 62)
    assert np.array_equal(rec[:, :3], p["xyz"].numpy()) and not rec[:, 3:6].any()
    # f_rest is channel-major on disk: column 9 + c * 15 + k  <->  features_rest[:, k, c]
    assert np.array_equal(rec[:, 9 + 1 * 15 + 4], p["features_rest"][:, 4, 1].numpy())
    back = gs_io.load_ply(path, max_sh_degree=3)
    for k, v in p.items():
        assert back[k].dtype == np.float32 and np.array_equal(back[k], v.numpy()), k


def test_ply_loader_reads_ascii_and_big_endian(tmp_path):
    names = gs_io.ply_attributes(3, 0)
    vals = np.arange(2 * len(names), dtype=np.float32).reshape(2, len(names))
    head = "ply\nformat {}\ncomment made by a viewer\nelement vertex 2\n" + "".join(f"property float {n}\n" for n in names) + "end_header\n"
    a = tmp_path / "a.ply"
    a.write_text(head.format("ascii 1.0") + "\n".join(" ".join(str(float(x)) for x in row) for row in vals) + "\n")
    b = tmp_path / "b.ply"
    b.write_bytes(head.format("binary_big_endian 1.0").encode() + vals.astype(">f4").tobytes())
    for path in (a, b):
        got = gs_io.load_ply(str(path), max_sh_degree=0)
        assert np.array_equal(got["xyz"], vals[:, :3]) and got["features_rest"].shape == (2, 0, 3)
        assert np.array_equal(got["rotation"], vals[:, -4:])


def test_ply_bytes_against_a_literal_golden(tmp_path):
    """Byte parity of save_ply without plyfile (absent from the image) and without gs_io's own helpers: the header is
    written out literally as plyfile emits it for the structured array of gaussian_model_static.py:229-249 (native byte
    order on x86 = binary_little_endian, one `property float <name>` per 'f4' field, names in the order of
    construct_list_of_attributes :214-226), and the rows are packed one float at a time in that order, with the
    transpose(1, 2).flatten(1) of the two SH tensors spelled out as index arithmetic. SH degree 1, two Gaussians."""
    import struct
    N, K = 2, 3                                   # K = (degree + 1)^2 - 1 rest coefficients
    xyz = [[0.5, -1.25, 2.0], [3.0, 4.5, -6.0]]
    dc = [[[0.125, 0.25, 0.375]], [[-0.5, -0.625, -0.75]]]                                   # [N][1][3]
    rest = [[[(100 * n + 10 * k + c) / 8.0 for c in range(3)] for k in range(K)] for n in range(N)]   # [N][K][3]
    opacity = [[-2.0], [1.5]]
    scaling = [[-3.0, -3.5, -4.0], [-1.0, -1.5, -2.5]]
    rotation = [[1.0, 0.0, 0.0, 0.0], [0.5, -0.5, 0.5, -0.5]]
    golden_header = (
        "ply\n"
        "format binary_little_endian 1.0\n"
        "element vertex 2\n"
        "property float x\nproperty float y\nproperty float z\n"
        "property float nx\nproperty float ny\nproperty float nz\n"
        "property float f_dc_0\nproperty float f_dc_1\nproperty float f_dc_2\n"
        "property float f_rest_0\nproperty float f_rest_1\nproperty float f_rest_2\n"
        "property float f_rest_3\nproperty float f_rest_4\nproperty float f_rest_5\n"
        "property float f_rest_6\nproperty float f_rest_7\nproperty float f_rest_8\n"
        "property float opacity\n"
        "property float scale_0\nproperty float scale_1\nproperty float scale_2\n"
        "property float rot_0\nproperty float rot_1\nproperty float rot_2\nproperty float rot_3\n"
        "end_header\n").encode("ascii")
    body = b""
    for n in range(N):
        row = list(xyz[n]) + [0.0, 0.0, 0.0]
        row += [dc[n][0][c] for c in range(3)]                       # transpose(1, 2).flatten(1) of [1][3]: channel major
        row += [rest[n][k][c] for c in range(3) for k in range(K)]   # f_rest_{c * K + k} = features_rest[n, k, c]
        row += opacity[n] + scaling[n] + rotation[n]
        assert len(row) == 26
        body += b"".join(struct.pack("<f", v) for v in row)
    path = str(tmp_path / "point_cloud.ply")
    t = lambda v: torch.tensor(v, dtype=torch.float32)
    gs_io.save_ply(path, t(xyz), t(dc), t(rest), t(opacity), t(scaling), t(rotation))
    assert open(path, "rb").read() == golden_header + body
