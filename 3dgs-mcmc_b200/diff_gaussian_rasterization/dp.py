"""View-sharded data parallelism for the rasterizer path: one process per GPU, Gaussians replicated, each rank
renders its own camera view(s), and the per-Gaussian gradients are summed across ranks with ONE collective.

The reference is single-process / single-GPU (no torch.distributed anywhere); this is the multi-GPU
exchange step named in BASELINE.json's north_star. The backward kernel writes its five parameter gradients
(means3D 3 + SH 3M + opacity 1 + scales 3 + rotations 4 floats per Gaussian; 59 at M = 16) straight into one
contiguous FP32 arena, so a single NCCL all-reduce (708 MB at 3M Gaussians) covers them; forward and render
backward need no communication.
"""
import torch
import torch.distributed as dist


class GradientArena:
    """Contiguous gradient storage laid out as [means3D | sh | opacity | scales | rotations] blocks."""

    FIELDS = (("means3D", 3), ("sh", None), ("opacity", 1), ("scales", 3), ("rotations", 4))

    def __init__(self, P, M, device):
        self.P, self.M = int(P), int(M)
        widths = [w if w is not None else 3 * self.M for _, w in self.FIELDS]
        self.flat = torch.zeros(self.P * sum(widths), dtype=torch.float32, device=device)
        self.views = {}
        off = 0
        for (name, _), w in zip(self.FIELDS, widths):
            block = self.flat[off: off + self.P * w]
            self.views[name] = block.view(self.P, self.M, 3) if name == "sh" else block.view(self.P, w)
            off += self.P * w
        # gradients that are not parameters of the model (not reduced): kept to avoid per-step allocation
        self.local = {"means2D": torch.empty((self.P, 3), dtype=torch.float32, device=device),
                      "colors": torch.empty((self.P, 3), dtype=torch.float32, device=device),
                      "cov3D": torch.empty((self.P, 6), dtype=torch.float32, device=device)}

    def backward_buffers(self):
        """Output buffers for _native.rasterize_gaussians_backward(out_buffers=...)."""
        out = dict(self.local)
        out.update(self.views)
        return out

    def all_reduce(self, average=False, group=None):
        """Sum (or average) the arena over all ranks. No-op without an initialised process group."""
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
            return self.flat
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
        if average:
            self.flat.div_(dist.get_world_size(group))
        return self.flat

    def floats_per_gaussian(self):
        return self.flat.numel() // max(self.P, 1)


def shard_views(num_views, rank, world_size):
    """Indices of the camera views rank `rank` renders in a step of `num_views` views (round-robin)."""
    return list(range(rank, num_views, world_size))
