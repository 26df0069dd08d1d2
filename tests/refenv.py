"""Import environment for running the reference's OWN Python callers (staged unmodified under baseline/_ref/py by
tools/stage_reference.py) against this repository's packages. Used only by tests/ and tools/dp_check.py.

* sys.path: baseline/_ref/py (the reference's `gaussian_renderer`, `scene`, `utils`, `arguments`, `train_static`),
  after 3dgs-mcmc_b200/ so that `diff_gaussian_rasterization` and `simple_knn` are THIS repository's.
* stub modules for the two third-party imports the image lacks: `plyfile` (PLY I/O is not on the path under test) and
  `open3d` (only used by train_swin's ARAP neighbour search). Any USE of a stub raises.
"""
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "3dgs-mcmc_b200")
REF_PY = os.path.join(ROOT, "baseline", "_ref", "py")


def staged():
    return os.path.isfile(os.path.join(REF_PY, "gaussian_renderer", "__init__.py"))


class _Absent:
    def __init__(self, name):
        self._name = name

    def __call__(self, *a, **k):
        raise RuntimeError(f"{self._name} is a stub: the real module is not installed in this image")

    def __getattr__(self, item):
        raise AttributeError(f"{self._name}.{item}: stub module, not installed in this image")


def install():
    """Make the staged reference importable. Returns REF_PY. Raises if the files did not travel."""
    if not staged():
        raise RuntimeError(f"{REF_PY} is missing: run `python tools/stage_reference.py` where /root/reference is mounted "
                           "(__graft_entry__.build() does); gpurun ships the directory to the GPU box")
    for p in (ROOT, PKG):
        if p not in sys.path:
            sys.path.insert(0, p)
    if REF_PY not in sys.path:
        sys.path.append(REF_PY)          # after PKG: the rasterizer / knn packages are ours
    if "plyfile" not in sys.modules:
        try:
            import plyfile  # noqa: F401
        except ImportError:
            m = types.ModuleType("plyfile")
            m.PlyData, m.PlyElement = _Absent("plyfile.PlyData"), _Absent("plyfile.PlyElement")
            m.__stub__ = True
            sys.modules["plyfile"] = m
    if "open3d" not in sys.modules:
        try:
            import open3d  # noqa: F401
        except ImportError:
            m = types.ModuleType("open3d")
            m.__stub__ = True

            def _absent_attr(item):     # AttributeError (not RuntimeError): inspect / hasattr probe modules for dunders
                raise AttributeError(f"open3d.{item}: stub module, not installed in this image")
            m.__getattr__ = _absent_attr
            sys.modules["open3d"] = m
    # this repository's rasterizer must be the one the reference code binds
    import diff_gaussian_rasterization as dgr
    assert os.path.abspath(dgr.__file__).startswith(PKG), dgr.__file__
    return REF_PY
