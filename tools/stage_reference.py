#!/usr/bin/env python
"""Stage the reference's own Python callers, UNMODIFIED, under the git-ignored baseline/_ref/py/ so that the GPU box
(which has no /root/reference) can execute them against this package:

    gaussian_renderer/, scene/, utils/, arguments/, train_static.py, train_swin.py   (*.py only)

Nothing is edited. The files never enter the repository history (.gitignore: baseline/_ref/); `gpurun` ships the
directory like oracle/_ref. tests/refenv.py puts it on sys.path together with stub modules for the two third-party
imports that are absent from this image (plyfile, open3d); `diff_gaussian_rasterization` and `simple_knn` resolve to
this repository's packages. Run by __graft_entry__.build() whenever the reference tree is mounted.
"""
import filecmp
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEFAULT_REF = "/root/reference"
DIRS = ("gaussian_renderer", "scene", "utils", "arguments")
FILES = ("train_static.py", "train_swin.py")


def stage(ref=DEFAULT_REF, dst=None, verbose=True):
    dst = dst or os.path.join(ROOT, "baseline", "_ref", "py")
    if not os.path.isdir(os.path.join(ref, "gaussian_renderer")):
        if verbose:
            print(f"stage_reference: {ref} absent - keeping whatever is staged under {dst}")
        return False
    n = 0
    for d in DIRS:
        for base, _dirs, files in os.walk(os.path.join(ref, d)):
            rel = os.path.relpath(base, ref)
            for f in files:
                if f.endswith(".py"):
                    os.makedirs(os.path.join(dst, rel), exist_ok=True)
                    src, out = os.path.join(base, f), os.path.join(dst, rel, f)
                    if not (os.path.exists(out) and filecmp.cmp(src, out, shallow=False)):
                        shutil.copyfile(src, out)
                    n += 1
    for f in FILES:
        os.makedirs(dst, exist_ok=True)
        shutil.copyfile(os.path.join(ref, f), os.path.join(dst, f))
        n += 1
    if verbose:
        print(f"stage_reference: {n} reference Python files staged unmodified under {dst}")
    return True


if __name__ == "__main__":
    stage(*(sys.argv[1:2] or [DEFAULT_REF]))
