"""On-disk / wire formats of the reference, vectorised. Host-side numpy; no GPU involved.

* ``stream_dump`` / ``stream_load`` — the big-endian record stream of utils/stream_utils.py:11-83
  (``!II fff fff fff*((d+1)^2-1) fff ffff f`` per Gaussian, appended to ``filename``, with ``format.json`` beside it).
  The reference packs one ``struct.pack`` per row in a Python loop (seconds per 100K Gaussians); here the whole block
  is one structured-array view written with a single ``write``. Output bytes and ``format.json`` are identical
  (tests/test_gs_io.py compares with golden bytes produced by the reference function itself).
* ``save_ply`` / ``load_ply`` — the ``point_cloud.ply`` layout of scene/gaussian_model_static.py:214-296
  (x y z nx ny nz f_dc_* f_rest_* opacity scale_* rot_*, float32, binary little endian) without the ``plyfile``
  package (absent from this image; the layout is restated from the PLY specification and the reference's attribute
  list, so byte parity with ``plyfile`` itself is unpinned — covered by round trips and a header check).
"""
import json
import os

import numpy as np

ENDIAN = "!"


def _np(t):
    if hasattr(t, "detach"):
        t = t.detach().cpu().numpy()
    return np.asarray(t)


def stream_format(shs_degree=1):
    """The FORMAT dict of utils/stream_utils.py:16-25 (insertion order matters: it is the record layout)."""
    return {"start_frame": "I", "end_frame": "I", "xyz": "fff", "f_dc": "fff",
            "f_rest": "fff" * ((shs_degree + 1) ** 2 - 1), "scaling": "fff", "rotation": "ffff", "opacity": "f"}


def _record_dtype(fmt):
    fields = []
    for name, code in fmt.items():
        kind = {"I": ">u4", "f": ">f4"}[code[0]] if code else ">f4"
        if code and any(c != code[0] for c in code):
            raise ValueError(f"mixed field codes in {name!r}: {code!r}")
        fields.append((name, kind, (len(code),)))
    return np.dtype(fields)


def stream_dump(params_dict, filename, shs_degree=1):
    """Append the Gaussians of params_dict (keys of stream_format(); tensors or arrays with N rows) to `filename` and
    (re)write format.json in its directory — byte for byte what utils/stream_utils.py:11-65 produces."""
    fmt = stream_format(shs_degree)
    arrays = {k: _np(params_dict[k]) for k in fmt}
    N = arrays["start_frame"].shape[0]
    if not all(a.shape[0] == N for a in arrays.values()):
        raise AssertionError("Batch size mismatch")
    rec = np.zeros(N, dtype=_record_dtype(fmt))
    for k, code in fmt.items():
        a = arrays[k].reshape(N, -1)
        if a.shape[1] != len(code):
            raise ValueError(f"{k}: expected {len(code)} values per Gaussian, got {a.shape[1]}")
        if code[:1] == "I":
            if (a < 0).any() or (a > 0xFFFFFFFF).any():
                raise ValueError(f"{k}: 'I' format requires 0 <= number <= 4294967295")     # struct.error in the reference
            rec[k] = np.trunc(a).astype(np.uint32)      # int(x.item()) truncates
        else:
            rec[k] = a.astype(np.float32)
    out_fmt = dict(fmt)
    out_fmt["ENDIAN"] = ENDIAN
    with open(os.path.join(os.path.dirname(filename), "format.json"), "w") as f:
        json.dump(out_fmt, f, indent=4)
    with open(filename, "ab") as f:
        f.write(rec.tobytes())
    return N


def stream_load_arrays(fmtjson, filename):
    """dict of arrays (one per field, native byte order) read from a stream file."""
    with open(fmtjson, "r") as f:
        fmt = json.load(f)
    endian = fmt.pop("ENDIAN")
    if endian not in ("!", ">"):
        raise ValueError(f"unsupported byte order {endian!r}")
    dt = _record_dtype(fmt)
    data = np.fromfile(filename, dtype=np.uint8)
    n = data.size // dt.itemsize
    rec = data[: n * dt.itemsize].view(dt)
    return {k: rec[k].astype(rec[k].dtype.newbyteorder("=")) for k in fmt}


def stream_load(fmtjson, filename):
    """List of per-Gaussian tuples, like utils/stream_utils.py:67-83."""
    arrays = stream_load_arrays(fmtjson, filename)
    cols = [a.reshape(a.shape[0], -1) for a in arrays.values()]
    n = cols[0].shape[0] if cols else 0
    rows = [[] for _ in range(n)]
    for c in cols:
        for row, vals in zip(rows, c.tolist()):
            row.extend(vals)
    return [tuple(r) for r in rows]


# ----------------------------------------------------------------------------------------------------------------------
# PLY
# ----------------------------------------------------------------------------------------------------------------------
def ply_attributes(n_dc, n_rest, n_scale=3, n_rot=4):
    """construct_list_of_attributes (scene/gaussian_model_static.py:214-227)."""
    names = ["x", "y", "z", "nx", "ny", "nz"]
    names += [f"f_dc_{i}" for i in range(n_dc)]
    names += [f"f_rest_{i}" for i in range(n_rest)]
    names.append("opacity")
    names += [f"scale_{i}" for i in range(n_scale)]
    names += [f"rot_{i}" for i in range(n_rot)]
    return names


def save_ply(path, xyz, features_dc, features_rest, opacity, scaling, rotation):
    """save_ply (gaussian_model_static.py:229-250): features_dc [N,1,3] and features_rest [N,M-1,3] are written channel
    major (transpose(1, 2).flatten(1)), normals as zeros, everything float32, binary little endian."""
    xyz = _np(xyz).astype(np.float32)
    N = xyz.shape[0]
    f_dc = np.transpose(_np(features_dc), (0, 2, 1)).reshape(N, -1)
    f_rest = np.transpose(_np(features_rest), (0, 2, 1)).reshape(N, -1)
    attrs = np.concatenate((xyz, np.zeros_like(xyz), f_dc, f_rest, _np(opacity).reshape(N, -1), _np(scaling).reshape(N, -1),
                            _np(rotation).reshape(N, -1)), axis=1).astype("<f4")
    names = ply_attributes(f_dc.shape[1], f_rest.shape[1], _np(scaling).reshape(N, -1).shape[1], _np(rotation).reshape(N, -1).shape[1])
    assert attrs.shape[1] == len(names)
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    header = "ply\nformat binary_little_endian 1.0\nelement vertex %d\n" % N
    header += "".join(f"property float {n}\n" for n in names) + "end_header\n"
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        f.write(np.ascontiguousarray(attrs).tobytes())


def _read_ply_vertices(path):
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError("not a PLY file")
        fmt, count, props, in_vertex = None, 0, [], False
        while True:
            line = f.readline()
            if not line:
                raise ValueError("PLY header is not terminated")
            tok = line.decode("ascii").split()
            if not tok or tok[0] == "comment":
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                in_vertex = tok[1] == "vertex"
                if in_vertex:
                    count = int(tok[2])
            elif tok[0] == "property" and in_vertex:
                if tok[1] == "list":
                    raise ValueError("list properties are not supported in the vertex element")
                props.append((tok[2], tok[1]))
            elif tok[0] == "end_header":
                break
        kinds = {"float": "f4", "float32": "f4", "double": "f8", "float64": "f8", "uchar": "u1", "uint8": "u1", "char": "i1",
                 "int8": "i1", "short": "i2", "int16": "i2", "ushort": "u2", "uint16": "u2", "int": "i4", "int32": "i4",
                 "uint": "u4", "uint32": "u4"}
        if fmt == "ascii":
            table = np.loadtxt(f, max_rows=count, ndmin=2)
            return {name: table[:, i] for i, (name, _) in enumerate(props)}
        order = {"binary_little_endian": "<", "binary_big_endian": ">"}[fmt]
        dt = np.dtype([(name, order + kinds[kind]) for name, kind in props])
        rec = np.frombuffer(f.read(count * dt.itemsize), dtype=dt, count=count)
        return {name: rec[name] for name, _ in props}


def load_ply(path, max_sh_degree=3):
    """load_ply (gaussian_model_static.py:252-296) as float32 arrays in the model's layouts:
    xyz [N,3], features_dc [N,1,3], features_rest [N,M-1,3], opacity [N,1], scaling [N,3], rotation [N,4]."""
    v = _read_ply_vertices(path)
    xyz = np.stack((v["x"], v["y"], v["z"]), axis=1)
    N = xyz.shape[0]
    features_dc = np.stack((v["f_dc_0"], v["f_dc_1"], v["f_dc_2"]), axis=1).reshape(N, 3, 1)

    def ordered(prefix):
        names = sorted((n for n in v if n.startswith(prefix)), key=lambda x: int(x.split("_")[-1]))
        return np.stack([v[n] for n in names], axis=1) if names else np.zeros((N, 0))
    extra = ordered("f_rest_")
    assert extra.shape[1] == 3 * (max_sh_degree + 1) ** 2 - 3
    features_rest = extra.reshape(N, 3, (max_sh_degree + 1) ** 2 - 1)
    f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)  # noqa: E731
    return {"xyz": f32(xyz), "features_dc": f32(np.transpose(features_dc, (0, 2, 1))),
            "features_rest": f32(np.transpose(features_rest, (0, 2, 1))), "opacity": f32(np.asarray(v["opacity"])[..., None]),
            "scaling": f32(ordered("scale_")), "rotation": f32(ordered("rot"))}
