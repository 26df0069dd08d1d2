"""View-sharded data parallelism for the rasterizer path: one process per GPU, Gaussians replicated, each rank
renders its own camera view(s), and the per-Gaussian gradients are summed across ranks with ONE collective.

The reference is single-process / single-GPU (no torch.distributed anywhere); this is the multi-GPU
exchange step named in BASELINE.json's north_star. The backward kernel writes its five parameter gradients
(means3D 3 + SH 3M + opacity 1 + scales 3 + rotations 4 floats per Gaussian; 59 at M = 16) straight into one
contiguous FP32 arena, so a single NCCL all-reduce (708 MB at 3M Gaussians) covers them; forward and render
backward need no communication.
"""
import contextlib
import os
import threading

import torch
import torch.distributed as dist


def _round_up(n, m):
    return (n + m - 1) // m * m


class GradientArena:
    """Contiguous gradient storage laid out as [means3D | sh | opacity | scales | rotations] blocks."""

    FIELDS = (("means3D", 3), ("sh", None), ("opacity", 1), ("scales", 3), ("rotations", 4))

    def __init__(self, P, M, device):
        self.P, self.M = int(P), int(M)
        widths = [w if w is not None else 3 * self.M for _, w in self.FIELDS]
        # every field block starts on a 16-byte boundary whatever P is (the kernels' float4 paths; MCMC growth makes P
        # arbitrary): up to 3 pad floats per block, zero and all-reduced along with the rest
        starts, off = [], 0
        for w in widths:
            starts.append(off)
            off = _round_up(off + self.P * w, 4)
        self.flat = torch.zeros(off, dtype=torch.float32, device=device)
        self.views = {}
        for (name, _), w, off in zip(self.FIELDS, widths, starts):
            block = self.flat[off: off + self.P * w]
            self.views[name] = block.view(self.P, self.M, 3) if name == "sh" else block.view(self.P, w)
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
        return sum(w if w is not None else 3 * self.M for _, w in self.FIELDS)


def replica_fingerprint(tensors):
    """Two 64-bit integers that depend on every bit of the given tensors (sum of the raw words and a position-weighted
    sum): equal on two ranks iff - up to collisions - their replicas are bit-identical."""
    acc = None
    for t in tensors:
        t = t.detach().contiguous()
        if t.numel() == 0:
            continue
        # the raw storage words, never converted values: t.to(int64) would map every fp16 / bf16 / fp64 value in (-1, 1)
        # to 0 and hide a divergence there
        raw = {1: torch.uint8, 2: torch.int16, 4: torch.int32, 8: torch.int64}.get(t.element_size())
        if t.dtype == torch.bool:
            bits = t.to(torch.uint8)
        elif raw is None or t.is_complex():
            bits = t.flatten().view(torch.uint8)
        else:
            bits = t.view(raw)
        bits = bits.flatten().long()
        w = torch.arange(bits.numel(), device=bits.device) % 1000003 + 1
        h = torch.stack([bits.sum(), (bits * w).sum(), torch.tensor(bits.numel(), device=bits.device)])
        acc = h if acc is None else acc * 31 + h
    return acc if acc is not None else torch.zeros(3, dtype=torch.int64)


def assert_replicas_in_sync(tensors, group=None, what="replicated Gaussian parameters"):
    """View-sharded training keeps a full replica of the model (and of its MCMC state) on every rank and relies on the
    replicas evolving in lock step: the summed gradients are bit-identical on all ranks, the noise comes from the device
    RNG seeded alike, and relocation / growth draw inside lockstep_sampling(). A rank that diverged (different seed, a
    skipped step, a draw outside lockstep_sampling) would silently train a different
    model and - after add_new_gs - hand resize() a different P, corrupting the symmetric buffers. Call this after every
    relocation / growth step (it costs two tiny all-reduces): raises RuntimeError on every rank if any replica differs."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return True
    h = replica_fingerprint(tensors).to(tensors[0].device)
    lo, hi = h.clone(), h.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
    if not bool((lo == hi).all()):
        raise RuntimeError(f"{what} differ between ranks (rank {dist.get_rank(group)}: fingerprint {h.tolist()}, "
                           f"min {lo.tolist()}, max {hi.tolist()}): replicas are out of lock step")
    return True


_lockstep_guard = threading.Lock()


@contextlib.contextmanager
def lockstep_sampling(group=None, src=0):
    """Make the MCMC draws of relocate_gs / add_new_gs (scene/gaussian_model_static.py:412-422: torch.multinomial over the
    alive opacities) the same on every rank.

    Seeding every process alike is not enough: on CUDA torch.multinomial normalises with a floating-point cumsum whose
    summation order is not fixed from launch to launch (it is on torch's own list of nondeterministic ops), so two
    replicas holding bit-identical probabilities can draw different indices - measured on 2 x B200: the replicas of
    tools/dp_check.py --mcmc-steps agreed through every training step and parted at the first relocation. Inside this
    context every rank still makes each draw (the device RNG streams stay aligned for the noise that follows) and rank
    `src`'s indices - K int64 per draw - are broadcast over them: this is the "MCMC statistics" half of the north star's
    exchange. The reference's callers run unchanged inside the context. Both torch.multinomial and Tensor.multinomial are
    covered; other RNG calls are not (the reference's relocation draws through torch.multinomial only). One context at a time
    per process; draws made by other threads while it is open are left alone."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        yield
        return
    real, real_method = torch.multinomial, torch.Tensor.multinomial
    src_global = dist.get_global_rank(group, src) if group is not None else src
    owner = threading.get_ident()

    def in_step(out):
        # Only the thread that opened the context is in lock step with the other ranks. A draw from any other thread (an
        # evaluator, a data loader) gets the stock function: broadcasting there would enter a collective the peers never call.
        if threading.get_ident() == owner:
            dist.broadcast(out, src=src_global, group=group)
        return out

    def multinomial(*args, **kwargs):
        return in_step(real(*args, **kwargs))

    def multinomial_method(self, *args, **kwargs):
        return in_step(real_method(self, *args, **kwargs))

    if not _lockstep_guard.acquire(blocking=False):
        raise RuntimeError("lockstep_sampling is already active in this process (it is neither re-entrant nor shareable "
                           "between threads)")
    torch.multinomial, torch.Tensor.multinomial = multinomial, multinomial_method
    try:
        yield
    finally:
        torch.multinomial, torch.Tensor.multinomial = real, real_method
        _lockstep_guard.release()


def shard_views(num_views, rank, world_size):
    """Indices of the camera views rank `rank` renders in a step of `num_views` views (round-robin)."""
    return list(range(rank, num_views, world_size))


# ----------------------------------------------------------------------------------------------------------------------
# Factored exchange: all-gather 3 floats + all-reduce 11 floats per Gaussian instead of all-reducing 59
# ----------------------------------------------------------------------------------------------------------------------
class FactoredExchange:
    """Backward + cross-rank gradient sum for view-sharded training, one camera view per rank.

    The SH gradient of one view is an outer product, dL_dsh[i, k, :] = basis_k(dir_i) * dL_dRGB_i
    (cuda_rasterizer/backward.cu:20-139), and dir_i only needs the Gaussian's position (replicated) and the view's
    camera position. So the ranks exchange the clamp-masked dL_dRGB (3 floats per Gaussian per view) and each rank
    rebuilds sum_v dL_dsh_v itself (csrc/dp_exchange.cu: sh_grad_combine_kernel); only means3D / opacity / scales /
    rotations (11 floats per Gaussian) are all-reduced. Per rank and step at P = 3M, world 8: 252 MB received by the
    gather + a 132 MB all-reduce, against a 708 MB all-reduce; the summed gradients are the same on every rank bit
    for bit (views are added in rank order).

    layout (what travels; all three give the same sums, views added in rank order):
      "records" (default) where a view's colour gradient is zero (culled, or never blended: 92 % of the Gaussians per view
              at the bench scene) so are its other gradients, so the per-Gaussian backward itself stores - per chunk of 128
              Gaussians - a bit mask and the 14-float records of the non-zero ones into every rank's gather buffer, and each
              rank adds the views in rank order (gs_backward_geometry_push + gs_dp_grad_record_combine). No all-reduce, no
              slab, ONE barrier per step (the gather buffer is double buffered), outputs are fresh tensors (zeroed on a side
              stream while the tile backward runs) that autograd takes without a copy.
      "sparse" colour gradients as bit masks + non-zero triples, the other 11 floats all-reduced in a symmetric slab.
      "dense"  colour gradients as [P][3] floats, the other 11 floats all-reduced.

    transport:
      "symm"  torch symmetric memory over NVLink: the kernel that masks dL_dRGB stores it straight into every rank's
              gather buffer (NVSwitch multicast store when available, else peer stores) - the all-gather is fused
              into its producer - and the slab is summed by our own two-shot kernel (multimem.ld_reduce/st or peer
              loads in rank order). Cross-rank ordering = two symmetric-memory barriers per step.
      "nccl"  (any torch.distributed backend) all_gather_into_tensor + all_reduce; the gather overlaps the
              per-Gaussian backward kernel.
      "auto"  "symm" when the process group is NCCL and symmetric memory can be set up, else "nccl".
    """

    SLAB_FIELDS = (("means3D", 3), ("opacity", 1), ("scales", 3), ("rotations", 4))

    def __init__(self, P, M, device, group=None, transport="auto", multicast=None, capacity=None, layout=None):
        """capacity: number of Gaussians the (symmetric) buffers are sized for, >= P; the MCMC trainer grows the model
        by 5 % every densification step up to cap_max (gaussian_model_static.py:460-492) - allocate once for cap_max and
        call resize(P) after each growth instead of a new rendezvous.
        multicast: None = use the NVSwitch multicast mapping when the symmetric allocation has one and world > 2
        (measured on 8xB200 at P = 3M: 2.73 ms backward+exchange with multicast against 2.94 ms with peer pointers at
        world 8, but 2.32 against 2.12 ms at world 2); True / False force it."""
        self.P, self.M = int(P), int(M)
        self.capacity = int(capacity) if capacity is not None else self.P
        if self.capacity < self.P:
            raise ValueError(f"capacity {self.capacity} < P {self.P}")
        self.device = torch.device(device)
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if self.world > 1 else 0
        # "classic": slab all-reduce beside the SH rebuild (default). "early_reduce": slab all-reduce while the gather is
        # still in flight, SH rebuild alone afterwards - measured equal within noise on 8xB200 at P = 3M (2.67 vs 2.66 ms
        # backward + exchange): the reduce then competes with the gather for NVLink ingress.
        self.schedule = os.environ.get("GS_DP_SCHEDULE", "classic")
        if self.schedule not in ("early_reduce", "classic"):
            raise ValueError(f"GS_DP_SCHEDULE must be early_reduce or classic, got {self.schedule!r}")
        requested_layout = os.environ.get("GS_DP_LAYOUT") if layout is None else layout
        if requested_layout not in (None, "records", "sparse", "dense"):
            raise ValueError(f"layout must be records, sparse or dense, got {requested_layout!r}")
        self._step = 0
        self._hdl = None
        # Order matters: the transport is settled first (local pick -> agreement of all ranks -> rendezvous -> second
        # agreement), and only the FINAL transport decides the layout, the slot stride, the buffering depth and the buffer
        # size. Every rank therefore derives the same sizes, also when the symmetric set-up fails on some or all ranks.
        self.transport = self._pick_transport(transport)
        auto = transport == "auto" and self.world > 1
        if auto:
            # Every rank takes part in the agreement, whatever it picked locally: a rank whose symmetric-memory import
            # failed must not skip a collective the others enter. Only if ALL ranks can do "symm" is the rendezvous tried.
            self.transport = "symm" if self._all_ranks_agree(self.transport == "symm") else "nccl"
        slab_floats, total = self._plan(requested_layout)
        if self.transport == "symm":
            try:
                self._setup_symmetric(total, slab_floats, group, multicast)
            except Exception as ex:  # no peer access / no symmetric-memory support on this box
                if transport == "symm":
                    raise
                import warnings
                warnings.warn(f"FactoredExchange: symmetric memory unavailable ({ex!r}); using torch.distributed collectives")
                self.transport = "nccl"
            if auto and not self._all_ranks_agree(self.transport == "symm"):
                self.transport = "nccl"          # the rendezvous failed somewhere: everyone falls back together
        if self.transport != "symm":
            self._hdl = None
            slab_floats, total = self._plan(requested_layout)   # the fallback's own layout and sizes, on a plain buffer
            self.flat = torch.zeros(total, dtype=torch.float32, device=self.device)
            self._multicast = False
            self._send = torch.zeros(self.slot_stride, dtype=torch.float32, device=self.device)
        self._slab_capacity = slab_floats
        self._gathered_all = self.flat[slab_floats:].view(self._nbuf, self.world, self.slot_stride)
        self.gathered = self._gathered_all[0]
        if self.device.type == "cuda":
            self._zero_stream = torch.cuda.Stream(device=self.device, priority=-1)
            self._ev_zero, self._ev_cleared = torch.cuda.Event(), None
        self.resize(self.P)

    def resize(self, P):
        """Re-slice the buffers for P <= capacity Gaussians (no allocation of symmetric memory, no rendezvous). Every rank
        must call it with the same P before the next backward."""
        P = int(P)
        if P > self.capacity:
            raise ValueError(f"P {P} exceeds the capacity {self.capacity} this exchange was allocated for")
        self.P = P
        self.grads = {}
        if self.layout == "records":                              # no slab, no persistent outputs: nothing to re-slice
            self.local = {}
            return self
        off = 0
        for name, w in self.SLAB_FIELDS:                          # every block 16-byte aligned for any P
            self.grads[name] = self.flat[off: off + P * w].view(P, w)
            off = _round_up(off + P * w, 4)
        self._n_slab = _round_up(off, 64)                         # the all-reduced prefix of the slab
        self.slab = self.flat[: self._n_slab]
        self.grads["sh"] = torch.zeros((P, self.M, 3), dtype=torch.float32, device=self.device)
        self.local = {"means2D": torch.empty((P, 3), dtype=torch.float32, device=self.device),
                      "colors": torch.empty((P, 3), dtype=torch.float32, device=self.device),
                      "cov3D": torch.empty((P, 6), dtype=torch.float32, device=self.device)}
        return self

    @staticmethod
    def _auto_layout(world, transport, device_type):
        """Layout when the caller names none. Records need the gather to be cheap for absent Gaussians, i.e. peer stores:
        through an NCCL all-gather whole 56 B / Gaussian slots would travel (169 MB per rank and step at 3M, against 36 MB
        gathered + 132 MB all-reduced), so that transport keeps the dense gather + slab all-reduce."""
        return "records" if (world == 1 or transport == "symm") and device_type == "cuda" else "dense"

    def _plan(self, requested_layout):
        """Set layout / slot_stride / _nbuf for the CURRENT self.transport; return (slab_floats, total floats)."""
        self.layout = requested_layout or self._auto_layout(self.world, self.transport, self.device.type)
        # fixed: slot r of a gather buffer starts at r * slot_stride floats (include/gs_b200.h: gs_dp_slot_floats)
        per_layout = {"records": (self.capacity + 127) // 128 * 1800, "sparse": (self.capacity + 255) // 256 * 784,
                      "dense": 3 * self.capacity}
        self.slot_stride = _round_up(per_layout[self.layout], 64)
        if self.layout == "records":
            slab_floats = 0
            self._nbuf = 2 if self.world > 1 else 1               # double buffered: one barrier per step is enough
        else:
            slab_floats = _round_up(11 * self.capacity + 12, 64)  # + the pad that aligns every field block (resize)
            self._nbuf = 1
        return slab_floats, slab_floats + self._nbuf * self.world * self.slot_stride

    def _setup_symmetric(self, total, slab_floats, group, multicast):
        import torch.distributed._symmetric_memory as symm_mem
        self.flat = symm_mem.empty(total, dtype=torch.float32, device=self.device)
        self._hdl = symm_mem.rendezvous(self.flat, (group or dist.group.WORLD).group_name)
        self.flat.zero_()
        self._multicast = bool(self._hdl.multicast_ptr) and (multicast if multicast is not None else self.world > 2)
        self._ev_push = torch.cuda.Event()
        self._side = torch.cuda.Stream(device=self.device)
        self._ev_fork, self._ev_join = torch.cuda.Event(), torch.cuda.Event()
        self._slab_ptrs = [int(self._hdl.multicast_ptr)] if self._multicast else [int(p) for p in self._hdl.buffer_ptrs]
        gather_off = 4 * slab_floats
        self._gather_ptrs = [p + gather_off for p in self._slab_ptrs]

    def _all_ranks_agree(self, ok):
        """MIN over the ranks of a local yes / no; a collective every rank of the group must call."""
        t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=self.device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self.group)
        return int(t.item()) == 1

    def _pick_transport(self, transport):
        if transport not in ("auto", "symm", "nccl"):
            raise ValueError(f"unknown transport {transport!r}")
        if self.world == 1:
            return "nccl"                       # nothing to exchange; the collectives are skipped
        if transport != "auto":
            return transport
        try:
            if dist.get_backend(self.group) != "nccl" or self.device.type != "cuda":
                return "nccl"
            import torch.distributed._symmetric_memory  # noqa: F401
            return "symm"
        except Exception:
            return "nccl"

    # -- pieces (also used by the tests) ------------------------------------------------------------------------------
    def out_buffers(self):
        out = dict(self.local)
        out.update({k: v for k, v in self.grads.items() if k != "sh"})
        return out

    def bytes_per_step(self):
        """(bytes this rank receives in the gather, bytes of the all-reduced slab). "dense": exact. "sparse" / "records":
        the upper bound (every Gaussian present: headers + one triple / one 14-float record each); what actually travels
        over symmetric memory is the headers plus the non-zero entries only, while a torch.distributed all-gather moves
        whole slots (slot_stride floats per view)."""
        peers = self.world - 1
        if self.layout == "records":
            per_view = self.slot_stride * 4 if self.transport != "symm" else (self.P + 127) // 128 * 32 + 56 * self.P
            return peers * per_view, 0
        if self.layout == "sparse":
            per_view = self.slot_stride * 4 if self.transport != "symm" else (self.P + 255) // 256 * 64 + 12 * self.P
            return peers * per_view, 11 * self.P * 4
        return peers * 3 * self.P * 4, 11 * self.P * 4

    def push_color_grads(self, native, radii, geom, scratch):
        """Stage 1 (right after the tile backward): this rank's masked dL_dRGB goes to slot `rank` of every rank's
        gather buffer. "symm": the kernel's own NVLink stores; "nccl": an asynchronous all-gather that overlaps the
        per-Gaussian backward kernel. Returns a handle for finish()."""
        if self.world == 1:
            native.dp_color_grad_push(self.P, radii, geom, scratch, [self.gathered.data_ptr()], 0, False, layout=self.layout)
            return None
        if self.transport == "symm":
            # on the side stream: the NVLink stores (36 MB out, (world-1) x 36 MB in at P = 3M) drain while the
            # per-Gaussian backward kernel, which is HBM bound, runs on the caller's stream
            main = torch.cuda.current_stream(self.device)
            self._ev_fork.record(main)
            self._side.wait_event(self._ev_fork)
            with torch.cuda.stream(self._side):
                native.dp_color_grad_push(self.P, radii, geom, scratch, self._gather_ptrs, self.rank * self.slot_stride,
                                          self._multicast, layout=self.layout)
                self._ev_push.record(self._side)
            return None
        native.dp_color_grad_push(self.P, radii, geom, scratch, [self._send.data_ptr()], 0, False, layout=self.layout)
        return dist.all_gather_into_tensor(self.gathered.view(-1), self._send, group=self.group, async_op=True)

    def finish(self, native, means3D, campos_views, degree, pending=None, sh_out=None):
        """Stage 2 (after the per-Gaussian backward wrote the slab): sum the slab over the ranks and rebuild the summed
        SH gradient from the gathered colour gradients."""
        reduce_work = None
        symm = self.world > 1 and self.transport == "symm"
        if symm and self.schedule == "early_reduce":
            # The slab is complete as soon as the per-Gaussian kernel ends, long before the gather (its NVLink ingress takes
            # ~2x as long as that kernel at world 8): reduce the slab while the gather is still in flight, then rebuild
            # the SH gradient with all SMs to itself.
            main = torch.cuda.current_stream(self.device)
            self._hdl.barrier(channel=0)                 # every rank's slab is written
            native.dp_slab_all_reduce(self.device, self._slab_ptrs, self._n_slab, self.rank, self.world, self._multicast)
            main.wait_event(self._ev_push)
            self._hdl.barrier(channel=1)                 # every rank's pushes have landed; all slab slices too
        elif symm:
            torch.cuda.current_stream(self.device).wait_event(self._ev_push)
            self._hdl.barrier(channel=0)                 # every rank's pushes and slab are written
            # the slab all-reduce (NVLink bound) runs beside the SH rebuild (HBM bound)
            main = torch.cuda.current_stream(self.device)
            self._ev_fork.record(main)
            self._side.wait_event(self._ev_fork)
            with torch.cuda.stream(self._side):
                native.dp_slab_all_reduce(self.device, self._slab_ptrs, self._n_slab, self.rank, self.world,
                                          self._multicast)
                self._ev_join.record(self._side)
        elif self.world > 1:
            reduce_work = dist.all_reduce(self.slab, group=self.group, async_op=True)
            pending.wait()
        sh_grad = native.dp_sh_grad_combine(means3D, campos_views, self.gathered, self.slot_stride, degree, self.M,
                                            out=self.grads["sh"] if sh_out is None else sh_out, layout=self.layout)
        if symm and self.schedule == "early_reduce":
            self._hdl.barrier(channel=2)                 # peers are done reading this rank's gather slots
        elif symm:
            torch.cuda.current_stream(self.device).wait_event(self._ev_join)
            self._hdl.barrier(channel=1)                 # all slices landed; peers are done reading this rank's buffers
        elif self.world > 1:
            reduce_work.wait()
        if sh_out is not None:
            return dict(self.grads, sh=sh_grad)
        return self.grads

    def backward(self, native, bg, means3D, radii, scales, rotations, scale_modifier, viewmatrix, projmatrix, tan_fovx,
                 tan_fovy, dL_dout_color, sh, degree, campos, campos_views, geom, R, binning, img, sh_out=None):
        """Rasterizer backward of this rank's view followed by the exchange. sh_out: tensor to receive the summed SH
        gradient instead of the exchange's own (reused) buffer. campos_views[world, 3]: camera positions
        of the step's views in rank order (every rank knows the step's view list; row `rank` equals `campos`).
        Returns {"means3D", "sh", "opacity", "scales", "rotations"}: gradients summed over all ranks."""
        H, W = int(dL_dout_color.size(1)), int(dL_dout_color.size(2))
        if self.layout == "records":
            return self._backward_records(native, bg, means3D, radii, scales, rotations, scale_modifier, viewmatrix, projmatrix,
                                          tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos, campos_views, geom, R, binning,
                                          img, H, W)
        e = torch.empty(0)
        scratch = native.backward_render(bg, self.P, R, dL_dout_color, geom, binning, img)
        pending = self.push_color_grads(native, radii, geom, scratch)
        native.backward_geometry(means3D, radii, scales, rotations, scale_modifier, e, viewmatrix, projmatrix, tan_fovx,
                                 tan_fovy, H, W, sh, degree, campos, geom, scratch, out_buffers=self.out_buffers(),
                                 write_sh=False)
        return self.finish(native, means3D, campos_views, degree, pending, sh_out=sh_out)


    OUT_FIELDS = (("means3D", 3), ("opacity", 1), ("scales", 3), ("rotations", 4))

    def _fresh_outputs(self, split_sh=False):
        """The five summed-gradient tensors as views of ONE fresh allocation (returned with them; zero-filled by the caller).
        Fresh, because autograd keeps what backward returns as .grad without a copy only if nobody else holds it."""
        P, M = self.P, self.M
        widths = [w for _, w in self.OUT_FIELDS] + ([3, 3 * (M - 1)] if split_sh else [3 * M])
        starts, off = [], 0
        for w in widths:
            starts.append(off)
            off = _round_up(off + P * w, 4)
        flat = torch.empty(off, dtype=torch.float32, device=self.device)
        outs = {name: flat[o: o + P * w].view(P, w) for (name, w), o in zip(self.OUT_FIELDS, starts)}
        if split_sh:                                        # gradients of features_dc / features_rest, as the model stores them
            outs["sh"] = flat[starts[-2]: starts[-2] + P * 3].view(P, 1, 3)
            outs["sh_rest"] = flat[starts[-1]: starts[-1] + P * 3 * (M - 1)].view(P, M - 1, 3)
        else:
            outs["sh"] = flat[starts[-1]: starts[-1] + P * 3 * M].view(P, M, 3)
        return flat, outs

    def _zero_beside_tile_backward(self, native, flat):
        """Zero-fill `flat` (708 MB at P = 3M) on the side stream, ordered after the accumulator clear at the head of the
        backward (the event gs_backward_render_ev records): the fill then runs under the compute-bound tile kernel, which
        leaves HBM idle, instead of sharing bandwidth with the clear (measured: the clear took 0.116 ms instead of 0.023)."""
        native.stream_wait_event(self._zero_stream, self._ev_cleared)
        with torch.cuda.stream(self._zero_stream):
            flat.zero_()
            self._ev_zero.record(self._zero_stream)

    def _backward_records(self, native, bg, means3D, radii, scales, rotations, scale_modifier, viewmatrix, projmatrix, tan_fovx,
                          tan_fovy, dL_dout_color, sh, degree, campos, campos_views, geom, R, binning, img, H, W):
        flat, outs = self._fresh_outputs(split_sh=isinstance(sh, (tuple, list)))
        if self._ev_cleared is None:
            self._ev_cleared = native.new_event()
        scratch = native.backward_render(bg, self.P, R, dL_dout_color, geom, binning, img, cleared_event=self._ev_cleared)
        self._zero_beside_tile_backward(native, flat)
        buf = self._step % self._nbuf
        self._step += 1
        gathered = self._gathered_all[buf]
        base = buf * self.world * self.slot_stride
        symm = self.world > 1 and self.transport == "symm"
        if symm:
            dst, off, mc = self._gather_ptrs, base + self.rank * self.slot_stride, self._multicast
        elif self.world > 1:
            dst, off, mc = [self._send.data_ptr()], 0, False
        else:
            dst, off, mc = [gathered.data_ptr()], 0, False
        # the per-Gaussian backward stores its chunk records straight into every rank's gather buffer
        self.local = {"means2D": native.backward_geometry_push(means3D, radii, scales, rotations, scale_modifier, viewmatrix,
                                                               projmatrix, tan_fovx, tan_fovy, H, W, sh, degree, campos, geom,
                                                               scratch, dst, off, mc)}
        if symm:
            # every rank's records have landed. The one barrier of the step: rank r overwrites this buffer again two steps from
            # now, after the barrier of the next step, which no rank enters before its combine below is done.
            self._hdl.barrier(channel=0)
        elif self.world > 1:
            dist.all_gather_into_tensor(gathered.view(-1), self._send, group=self.group)
        torch.cuda.current_stream(self.device).wait_event(self._ev_zero)
        return native.dp_grad_record_combine(means3D, campos_views, gathered, self.slot_stride, degree, self.M, outs=outs,
                                             outputs_are_zero=True)


# ----------------------------------------------------------------------------------------------------------------------
# Public API: a GaussianRasterizer whose backward returns gradients already summed over the ranks
# ----------------------------------------------------------------------------------------------------------------------
class _RasterizeGaussiansViewParallel(torch.autograd.Function):
    """Same forward as diff_gaussian_rasterization._RasterizeGaussians for the shs + scales/rotations path; the
    backward runs this rank's view through FactoredExchange, so the gradients autograd receives for means3D, sh,
    opacities, scales and rotations are the SUM over all ranks' views (do not all-reduce them again). means2D gets
    this rank's own screen-space gradient (the densification statistic is per view, as in the reference)."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, sh_rest, opacities, scales, rotations, raster_settings, campos_views, exchange):
        """sh_rest None: sh is the concatenated [P,M,3] tensor; else sh = features_dc [P,1,3], sh_rest = features_rest
        [P,15,3] read as they lie (records layout only)."""
        from . import _native as C
        s = raster_settings
        e = torch.empty(0)
        if sh_rest is not None and exchange.layout != "records":
            raise RuntimeError("split SH tensors need the records exchange layout")
        R, color, radii, geom, binning, img, is_used = C.rasterize_gaussians(
            s.bg, means3D, e, opacities, scales, rotations, s.scale_modifier, e, s.viewmatrix, s.projmatrix, s.tanfovx,
            s.tanfovy, s.image_height, s.image_width, sh if sh_rest is None else (sh, sh_rest), s.sh_degree, s.campos,
            s.prefiltered, s.debug)
        ctx.raster_settings, ctx.num_rendered, ctx.exchange, ctx.split = s, R, exchange, sh_rest is not None
        ctx.save_for_backward(means3D, scales, rotations, radii, sh, campos_views, geom, binning, img,
                              *([sh_rest] if sh_rest is not None else []))
        ctx.mark_non_differentiable(radii, is_used)
        return color, radii, is_used

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii, _grad_is_used):
        from . import _native as C
        s, ex = ctx.raster_settings, ctx.exchange
        means3D, scales, rotations, radii, sh, campos_views, geom, binning, img = ctx.saved_tensors[:9]
        if ex.layout == "records":
            # fresh tensors already: autograd keeps them as .grad without a copy
            g = ex.backward(C, s.bg, means3D, radii, scales, rotations, s.scale_modifier, s.viewmatrix, s.projmatrix,
                            s.tanfovx, s.tanfovy, grad_out_color, (sh, ctx.saved_tensors[9]) if ctx.split else sh, s.sh_degree,
                            s.campos, campos_views, geom, ctx.num_rendered, binning, img)
            return (g["means3D"], ex.local["means2D"], g["sh"], g.get("sh_rest"), g["opacity"], g["scales"], g["rotations"],
                    None, None, None)
        g = ex.backward(C, s.bg, means3D, radii, scales, rotations, s.scale_modifier, s.viewmatrix, s.projmatrix,
                        s.tanfovx, s.tanfovy, grad_out_color, sh, s.sh_degree, s.campos, campos_views, geom,
                        ctx.num_rendered, binning, img, sh_out=torch.empty_like(sh))
        # the exchange's buffers are reused next step: autograd gets its own tensors (copies of the small ones, a fresh
        # tensor for the SH gradient)
        return (g["means3D"].clone(), ex.local["means2D"].clone(), g["sh"], None, g["opacity"].clone(), g["scales"].clone(),
                g["rotations"].clone(), None, None, None)


class ViewParallelRasterizer(torch.nn.Module):
    """GaussianRasterizer for view-sharded training: construct once per model size, call once per step with this
    rank's settings and the camera positions of all ranks' views (rank order)."""

    def __init__(self, P, M, device, group=None, transport="auto", capacity=None):
        super().__init__()
        self.exchange = FactoredExchange(P, M, device, group=group, transport=transport, capacity=capacity)

    def resize(self, P):
        """After the model grew (add_new_gs): same buffers, P <= capacity Gaussians."""
        self.exchange.resize(P)
        return self

    def forward(self, raster_settings, campos_views, means3D, means2D, opacities, shs, scales, rotations):
        """shs: the concatenated [P,M,3] tensor, or the pair (features_dc [P,1,3], features_rest [P,15,3]) as the model
        stores them (no 576 MB cat; the gradients come back per tensor)."""
        sh, sh_rest = shs if isinstance(shs, (tuple, list)) else (shs, None)
        return _RasterizeGaussiansViewParallel.apply(means3D, means2D, sh, sh_rest, opacities, scales, rotations,
                                                     raster_settings, campos_views, self.exchange)
