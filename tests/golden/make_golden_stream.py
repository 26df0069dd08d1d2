#!/usr/bin/env python
"""Golden bytes for the record stream, produced by the REFERENCE's own utils/stream_utils.py:stream_dump imported from
/root/reference (CPU). Run once: python tests/golden/make_golden_stream.py  -> tests/golden/stream_dump.npz"""
import io
import os
import sys
import tempfile
import contextlib

import numpy as np
import torch

sys.path.insert(0, "/root/reference")
from utils.stream_utils import stream_dump  # noqa: E402


def params(N, degree, seed):
    g = torch.Generator().manual_seed(seed)
    M = (degree + 1) ** 2
    return {"start_frame": torch.randint(0, 30, (N,), generator=g).float(), "end_frame": torch.randint(30, 60, (N,), generator=g).float(),
            "xyz": torch.randn((N, 3), generator=g), "f_dc": torch.randn((N, 1, 3), generator=g),
            "f_rest": torch.randn((N, M - 1, 3), generator=g), "scaling": torch.randn((N, 3), generator=g),
            "rotation": torch.randn((N, 4), generator=g), "opacity": torch.randn((N, 1), generator=g)}


def main():
    out = {}
    for name, (N, degree) in {"deg1": (57, 1), "deg3": (23, 3), "deg0": (5, 0)}.items():
        p = params(N, degree, 3 + N)
        with tempfile.TemporaryDirectory() as d, contextlib.redirect_stdout(io.StringIO()):
            path = os.path.join(d, "stream.bin")
            stream_dump({k: v.clone() for k, v in p.items()}, path, shs_degree=degree)
            stream_dump({k: v.clone() for k, v in p.items()}, path, shs_degree=degree)      # appends
            out[f"{name}/bytes"] = np.frombuffer(open(path, "rb").read(), dtype=np.uint8)
            out[f"{name}/format_json"] = np.frombuffer(open(os.path.join(d, "format.json"), "rb").read(), dtype=np.uint8)
        out[f"{name}/degree"] = np.int32(degree)
        for k, v in p.items():
            out[f"{name}/in/{k}"] = v.numpy()
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "stream_dump.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
