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
