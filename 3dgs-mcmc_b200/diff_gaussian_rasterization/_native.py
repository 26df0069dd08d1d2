"""ctypes binding of lib3dgs_mcmc_b200.so (C ABI: include/gs_b200.h).

This module plays the role of the reference's pybind extension ``diff_gaussian_rasterization._C``
(submodules/diff-gaussian-rasterization/ext.cpp:15-20): it exposes the same four callables with the
same positional arguments and return tuples as ``RasterizeGaussiansCUDA``,
``RasterizeGaussiansBackwardCUDA``, ``markVisible`` and ``ComputeRelocationCUDA``
(rasterize_points.cu:36-248), plus the additive ``add_noise``. Torch is used for device memory and
the current stream only; all compute is in the CUDA library. There is no CPU or eager fallback: a
missing library or a non-CUDA tensor raises.
"""
import ctypes
import os

import torch

_LIB_NAME = "lib3dgs_mcmc_b200.so"
_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), _LIB_NAME)

_c_void_p = ctypes.c_void_p
_c_int = ctypes.c_int
_c_uint = ctypes.c_uint
_c_float = ctypes.c_float
_c_size_t = ctypes.c_size_t
_c_ull = ctypes.c_ulonglong

# name -> (restype, argtypes); must list every symbol include/gs_b200.h declares.
SIGNATURES = {
    "gs_abi_version": (_c_int, []),
    "gs_error_string": (ctypes.c_char_p, [_c_int]),
    "gs_geom_bytes": (_c_size_t, [_c_int]),
    "gs_image_bytes": (_c_size_t, [_c_int, _c_int]),
    "gs_binning_bytes": (_c_size_t, [_c_uint, _c_int, _c_int, _c_int]),
    "gs_backward_scratch_bytes": (_c_size_t, [_c_int]),
    "gs_forward_preprocess": (_c_int, [_c_void_p, _c_int, _c_int, _c_int] + [_c_void_p] * 5 + [_c_float]
                              + [_c_void_p] * 5 + [_c_int, _c_int, _c_float, _c_float, _c_int]
                              + [_c_void_p] * 5),
    "gs_host_event_create": (_c_void_p, []),
    "gs_host_event_wait": (_c_int, [_c_void_p]),
    "gs_host_event_destroy": (_c_int, [_c_void_p]),
    "gs_forward_render": (_c_int, [_c_void_p, _c_int, _c_uint, _c_int, _c_int] + [_c_void_p] * 8),
    "gs_backward": (_c_int, [_c_void_p, _c_int, _c_int, _c_int, _c_uint, _c_int, _c_int] + [_c_void_p] * 5
                    + [_c_float] + [_c_void_p] * 5 + [_c_float, _c_float] + [_c_void_p] * 14),
    "gs_backward_render": (_c_int, [_c_void_p, _c_int, _c_uint, _c_int, _c_int] + [_c_void_p] * 6),
    "gs_backward_render_ev": (_c_int, [_c_void_p, _c_int, _c_uint, _c_int, _c_int] + [_c_void_p] * 7),
    "gs_stream_wait_event": (_c_int, [_c_void_p, _c_void_p]),
    "gs_backward_geometry": (_c_int, [_c_void_p, _c_int, _c_int, _c_int, _c_int, _c_int] + [_c_void_p] * 3 + [_c_float]
                             + [_c_void_p] * 5 + [_c_float, _c_float] + [_c_void_p] * 11),
    "gs_dp_slot_floats": (_c_size_t, [_c_int, _c_int]),
    "gs_backward_geometry_push": (_c_int, [_c_void_p, _c_int, _c_int, _c_int, _c_int, _c_int] + [_c_void_p] * 4 + [_c_float]
                                  + [_c_void_p] * 4 + [_c_float, _c_float] + [_c_void_p] * 5 + [_c_int, _c_int, _c_size_t]),
    "gs_dp_grad_record_combine": (_c_int, [_c_void_p, _c_int, _c_int, _c_int, _c_int, _c_void_p, _c_void_p, _c_void_p,
                                           _c_size_t, _c_int] + [_c_void_p] * 6),
    "gs_forward_preprocess_split_sh": (_c_int, [_c_void_p, _c_int, _c_int] + [_c_void_p] * 5 + [_c_float] + [_c_void_p] * 4
                                       + [_c_int, _c_int, _c_float, _c_float, _c_int] + [_c_void_p] * 5),
    "gs_backward_geometry_split_sh": (_c_int, [_c_void_p, _c_int, _c_int, _c_int, _c_int] + [_c_void_p] * 4 + [_c_float]
                                      + [_c_void_p] * 4 + [_c_float, _c_float] + [_c_void_p] * 12),
    "gs_dp_color_grad_push": (_c_int, [_c_void_p, _c_int, _c_void_p, _c_void_p, _c_void_p, _c_void_p, _c_int, _c_int,
                                       _c_size_t, _c_int]),
    "gs_dp_sh_grad_combine": (_c_int, [_c_void_p, _c_int, _c_int, _c_int, _c_int, _c_void_p, _c_void_p, _c_void_p,
                                       _c_size_t, _c_int, _c_void_p]),
    "gs_dp_slab_all_reduce": (_c_int, [_c_void_p, _c_void_p, _c_int, _c_int, _c_size_t, _c_int, _c_int]),
    "gs_loss_workspace_bytes": (_c_size_t, [_c_int, _c_int, _c_int]),
    "gs_photometric_loss_forward": (_c_int, [_c_void_p, _c_int, _c_int, _c_int] + [_c_void_p] * 4),
    "gs_photometric_loss_backward": (_c_int, [_c_void_p, _c_int, _c_int, _c_int] + [_c_void_p] * 3 + [_c_float, _c_float]
                                     + [_c_void_p] * 2),
    "gs_l1_workspace_bytes": (_c_size_t, [_c_size_t]),
    "gs_l1_loss_forward": (_c_int, [_c_void_p, _c_size_t] + [_c_void_p] * 4),
    "gs_l1_loss_backward": (_c_int, [_c_void_p, _c_size_t] + [_c_void_p] * 4),
    "gs_l1_loss_forward_static": (_c_int, [_c_void_p, _c_size_t] + [_c_void_p] * 3),
    "gs_activation_workspace_bytes": (_c_size_t, [_c_int]),
    "gs_activate_forward": (_c_int, [_c_void_p, _c_int] + [_c_void_p] * 8),
    "gs_activate_backward": (_c_int, [_c_void_p, _c_int] + [_c_void_p] * 6 + [_c_float, _c_float] + [_c_void_p] * 4),
    "gs_knn_workspace_bytes": (_c_size_t, [_c_int]),
    "gs_knn_mean_dist2": (_c_int, [_c_void_p, _c_int, _c_void_p, _c_void_p, _c_void_p]),
    "gs_swin_gather_forward": (_c_int, [_c_void_p, _c_int, _c_int, _c_int, _c_float, _c_int] + [_c_void_p] * 9),
    "gs_swin_gather_backward": (_c_int, [_c_void_p, _c_int, _c_int, _c_float, _c_int] + [_c_void_p] * 4),
    "gs_adam_step": (_c_int, [_c_void_p, _c_int] + [_c_void_p] * 7 + [ctypes.c_double] * 3),
    "gs_mark_visible": (_c_int, [_c_void_p, _c_int] + [_c_void_p] * 4),
    "gs_compute_relocation": (_c_int, [_c_void_p, _c_int] + [_c_void_p] * 4 + [_c_int] + [_c_void_p] * 2),
    "gs_add_noise": (_c_int, [_c_void_p, _c_int] + [_c_void_p] * 5 + [_c_float, _c_ull, _c_ull]),
    "gs_profile_enable": (_c_int, [_c_int]),
    "gs_profile_collect": (_c_int, [_c_void_p, _c_void_p]),
    "gs_stage_count": (_c_int, []),
    "gs_stage_name": (ctypes.c_char_p, [_c_int]),
    "gs_launch_count": (_c_ull, []),
    "gs_debug_forward_work": (_c_int, [_c_void_p]),
    "gs_debug_render_work": (_c_int, [_c_void_p, _c_int, _c_uint, _c_int, _c_int] + [_c_void_p] * 5),
    "gs_debug_fetch": (ctypes.c_longlong, [_c_int, _c_int, _c_uint, _c_int, _c_int] + [_c_void_p] * 4 + [_c_size_t]),
}


def load_library(path=_LIB_PATH):
    """Load the CUDA library and attach prototypes. Fails loudly when it has not been built."""
    if not os.path.exists(path):
        raise ImportError(
            f"{_LIB_NAME} not found at {path}. Build it with `bash 3dgs-mcmc_b200/build.sh` "
            "(or `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback.")
    lib = ctypes.CDLL(path)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype = restype
        fn.argtypes = argtypes
    return lib


_lib = load_library()
_count_pool = {}   # device index -> free _CountSlot objects


def _check(rc, what):
    if rc != 0:
        msg = _lib.gs_error_string(rc)
        raise RuntimeError(f"{what} failed: {msg.decode() if msg else rc} (code {rc})")


def _ptr(t):
    """Device pointer of a tensor, or NULL for an absent optional (empty tensor), as in the reference where
    empty CPU tensors yield null data pointers (rasterize_points.cu:86,149; forward.cu:205,241)."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _f32_cuda(t, name, device):
    if t is None or t.numel() == 0:
        return t
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (no CPU path exists)")
    if t.device != device:
        raise RuntimeError(f"{name} is on {t.device}, expected {device}")
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32, got {t.dtype}")
    return t.contiguous()


def _stream(device):
    return torch.cuda.current_stream(device).cuda_stream


class _CountSlot:
    """One pinned host word + one host-visible event: where a forward's instance count lands. Checked out per CALL
    (not per device), so two forwards in flight on two streams of one device - a prefetching evaluator beside the
    trainer - can never read each other's count."""
    __slots__ = ("buf", "event")

    def __init__(self):
        self.buf = torch.zeros(1, dtype=torch.int32, pin_memory=True)
        self.event = _lib.gs_host_event_create()
        if not self.event:
            raise RuntimeError("gs_host_event_create failed")


def _acquire_count_slot(device):
    key = device.index if device.index is not None else torch.cuda.current_device()
    pool = _count_pool.setdefault(key, [])
    try:
        return key, pool.pop()           # list.pop / append are atomic under the GIL
    except IndexError:
        return key, _CountSlot()


def _release_count_slot(key, slot):
    _count_pool[key].append(slot)


def _stable_size(nbytes):
    """Round a workspace size up to 1/8-octave steps. The instance count changes with every camera; rounding makes the
    binning workspace fall into a handful of size classes, so torch's caching allocator can hand back the block the
    previous step released instead of calling cudaMalloc (a device-wide sync) whenever a slightly larger view shows up."""
    if nbytes <= (1 << 20):
        return nbytes
    step = 1 << (nbytes.bit_length() - 4)
    return (nbytes + step - 1) // step * step


_binning_high_water = {}   # device index -> bytes


def _binning_capacity(key, need):
    """Bytes to allocate for a binning workspace that needs `need`: the device's high-water mark, raised with 25 % headroom
    whenever a view needs more (and dropped when a view needs less than a quarter of it). The instance count changes with
    every camera; a per-call exact size would make torch's caching allocator call cudaMalloc - a device-wide stall of
    10-30 ms at these sizes - each time a larger view shows up; with one size per device the block released by the previous
    step is simply handed back."""
    cap = _binning_high_water.get(key, 0)
    if need > cap or need < cap // 4:
        cap = _stable_size(need + need // 4)
        _binning_high_water[key] = cap
    return cap


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                        prefiltered, debug):
    """Same contract as RasterizeGaussiansCUDA (rasterize_points.cu:36-118): returns
    (num_rendered, out_color[3,H,W], radii[P] i32, geomBuffer, binningBuffer, imgBuffer, is_used[P] i32)."""
    if means3D.ndimension() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise RuntimeError("means3D must be a CUDA tensor (no CPU path exists)")
    device = means3D.device
    P, H, W = int(means3D.size(0)), int(image_height), int(image_width)
    byte_opts = dict(dtype=torch.uint8, device=device)
    if P == 0:
        z = torch.zeros
        return (0, z((3, H, W), dtype=torch.float32, device=device), z((0,), dtype=torch.int32, device=device),
                torch.empty(0, **byte_opts), torch.empty(0, **byte_opts), torch.empty(0, **byte_opts),
                z((0,), dtype=torch.int32, device=device))

    background = _f32_cuda(background, "bg", device)
    means3D = _f32_cuda(means3D, "means3D", device)
    colors = _f32_cuda(colors, "colors_precomp", device)
    opacity = _f32_cuda(opacity, "opacities", device)
    scales = _f32_cuda(scales, "scales", device)
    rotations = _f32_cuda(rotations, "rotations", device)
    cov3D_precomp = _f32_cuda(cov3D_precomp, "cov3D_precomp", device)
    viewmatrix = _f32_cuda(viewmatrix, "viewmatrix", device)
    projmatrix = _f32_cuda(projmatrix, "projmatrix", device)
    split = isinstance(sh, (tuple, list))     # (features_dc [P,1,3], features_rest [P,15,3]) read as they lie (additive)
    if split:
        sh_dc, sh_rest = _split_sh(sh, P, device)
    else:
        sh = _f32_cuda(sh, "sh", device)
    campos = _f32_cuda(campos, "campos", device)
    M = 16 if split else (int(sh.size(1)) if sh is not None and sh.numel() != 0 else 0)

    with torch.cuda.device(device):
        stream = _stream(device)
        out_color = torch.empty((3, H, W), dtype=torch.float32, device=device)
        radii = torch.empty((P,), dtype=torch.int32, device=device)
        is_used = torch.empty((P,), dtype=torch.int32, device=device)
        geom = torch.empty(_lib.gs_geom_bytes(P), **byte_opts)
        img = torch.empty(_lib.gs_image_bytes(W, H), **byte_opts)
        slot_key, slot = _acquire_count_slot(device)
        count, count_ev = slot.buf, slot.event

        if split:
            _check(_lib.gs_forward_preprocess_split_sh(
                stream, P, int(degree), _ptr(means3D), sh_dc.data_ptr(), sh_rest.data_ptr(), _ptr(opacity), _ptr(scales),
                float(scale_modifier), _ptr(rotations), _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos), W, H,
                float(tan_fovx), float(tan_fovy), int(bool(prefiltered)), radii.data_ptr(), is_used.data_ptr(),
                geom.data_ptr(), count.data_ptr(), count_ev), "gs_forward_preprocess_split_sh")
        else:
            _check(_lib.gs_forward_preprocess(
                stream, P, int(degree), M, _ptr(means3D), _ptr(sh), _ptr(colors), _ptr(opacity), _ptr(scales),
                float(scale_modifier), _ptr(rotations), _ptr(cov3D_precomp), _ptr(viewmatrix), _ptr(projmatrix),
                _ptr(campos), W, H, float(tan_fovx), float(tan_fovy), int(bool(prefiltered)),
                radii.data_ptr(), is_used.data_ptr(), geom.data_ptr(), count.data_ptr(), count_ev),
                "gs_forward_preprocess")
        # The one host wait of the forward (the reference blocks the stream here, rasterizer_impl.cu:282): the
        # binning workspace is sized from the instance count. Only the preprocess kernel is waited for; the depth
        # sort and scan enqueued behind it keep the GPU busy meanwhile.
        _check(_lib.gs_host_event_wait(count_ev), "gs_host_event_wait")
        rendered = int(count[0])
        _release_count_slot(slot_key, slot)
        if debug:
            _debug_sync("preprocess")

        binning = torch.empty(_binning_capacity(slot_key, _lib.gs_binning_bytes(rendered, W, H, P)), **byte_opts)
        _check(_lib.gs_forward_render(
            stream, P, rendered, W, H, _ptr(background), _ptr(colors), radii.data_ptr(), geom.data_ptr(),
            binning.data_ptr(), img.data_ptr(), out_color.data_ptr(), is_used.data_ptr()), "gs_forward_render")
        if debug:
            _debug_sync("render")
            flag = torch.zeros(1, dtype=torch.int32, device=device)
            _lib.gs_debug_fetch(30, P, rendered, W, H, geom.data_ptr(), binning.data_ptr(), img.data_ptr(),
                                flag.data_ptr(), 4)
            if int(flag.item()) != 0:
                raise RuntimeError("radix sort look-back timed out")
    return rendered, out_color, radii, geom, binning, img, is_used


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                                 viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos,
                                 geomBuffer, R, binningBuffer, imageBuffer, debug, out_buffers=None):
    """Same contract as RasterizeGaussiansBackwardCUDA (rasterize_points.cu:120-199): returns
    (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations).
    Additive: out_buffers (dict name -> preallocated contiguous tensor, names means2D/colors/opacity/means3D/cov3D/
    sh/scales/rotations) lets the caller supply the outputs, e.g. views of one all-reduce arena (dp.GradientArena)."""
    device = means3D.device
    P = int(means3D.size(0))
    H, W = int(dL_dout_color.size(1)), int(dL_dout_color.size(2))
    if isinstance(sh, (tuple, list)):
        # split SH tensors: the two halves of the backward; dL_dsh comes back as (dL_dfeatures_dc, dL_dfeatures_rest)
        if P == 0:
            z = lambda *shape: torch.zeros(shape, dtype=torch.float32, device=device)  # noqa: E731
            return z(0, 3), z(0, 3), z(0, 1), z(0, 3), z(0, 6), (z(0, 1, 3), z(0, 15, 3)), z(0, 3), z(0, 4)
        scratch = backward_render(background, P, R, dL_dout_color, geomBuffer, binningBuffer, imageBuffer)
        outs = backward_geometry(means3D, radii, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix,
                                 tan_fovx, tan_fovy, H, W, sh, degree, campos, geomBuffer, scratch, out_buffers=out_buffers)
        if debug:
            _debug_sync("backward")
        return outs
    M = int(sh.size(1)) if sh is not None and sh.numel() != 0 else 0
    opts = dict(dtype=torch.float32, device=device)
    shapes = [(P, 3), (P, 3), (P, 1), (P, 3), (P, 6), (P, M, 3), (P, 3), (P, 4)]
    if P == 0:
        return tuple(torch.zeros(s, **opts) for s in shapes)

    background = _f32_cuda(background, "bg", device)
    means3D = _f32_cuda(means3D, "means3D", device)
    colors = _f32_cuda(colors, "colors_precomp", device)
    scales = _f32_cuda(scales, "scales", device)
    rotations = _f32_cuda(rotations, "rotations", device)
    cov3D_precomp = _f32_cuda(cov3D_precomp, "cov3D_precomp", device)
    viewmatrix = _f32_cuda(viewmatrix, "viewmatrix", device)
    projmatrix = _f32_cuda(projmatrix, "projmatrix", device)
    dL_dout_color = _f32_cuda(dL_dout_color, "dL_dout_color", device)
    sh = _f32_cuda(sh, "sh", device)
    campos = _f32_cuda(campos, "campos", device)
    radii = radii.contiguous()

    with torch.cuda.device(device):
        outs = _backward_outputs(P, M, device, out_buffers, want_sh=True)
        dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations = outs
        scratch = torch.empty(_lib.gs_backward_scratch_bytes(P), dtype=torch.uint8, device=device)
        _check(_lib.gs_backward(
            _stream(device), P, int(degree), M, int(R), W, H, _ptr(background), _ptr(means3D), _ptr(sh), _ptr(colors),
            _ptr(scales), float(scale_modifier), _ptr(rotations), _ptr(cov3D_precomp), _ptr(viewmatrix),
            _ptr(projmatrix), _ptr(campos), float(tan_fovx), float(tan_fovy), radii.data_ptr(),
            geomBuffer.data_ptr(), binningBuffer.data_ptr(), imageBuffer.data_ptr(), dL_dout_color.data_ptr(),
            scratch.data_ptr(), dL_dmeans2D.data_ptr(), dL_dcolors.data_ptr(), dL_dopacity.data_ptr(),
            dL_dmeans3D.data_ptr(), dL_dcov3D.data_ptr(), _ptr(dL_dsh), dL_dscales.data_ptr(),
            dL_drotations.data_ptr()), "gs_backward")
        if debug:
            _debug_sync("backward")
    return dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations


_BWD_NAMES = ("means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations")


def _backward_outputs(P, M, device, out_buffers, want_sh):
    """The eight gradient tensors of the backward, taken from out_buffers where given. Every one is fully written by
    the kernels (no zero fill). want_sh=False leaves the SH slot None (factored exchange)."""
    opts = dict(dtype=torch.float32, device=device)
    shapes = [(P, 3), (P, 3), (P, 1), (P, 3), (P, 6), (P, M, 3), (P, 3), (P, 4)]
    outs = []
    for name, shape in zip(_BWD_NAMES, shapes):
        if name == "sh" and not want_sh:
            outs.append(None)
            continue
        t = out_buffers.get(name) if out_buffers else None
        if t is None:
            t = torch.empty(shape, **opts)
        elif tuple(t.shape) != shape or t.dtype != torch.float32 or not t.is_contiguous() or t.device != device:
            raise RuntimeError(f"out_buffers['{name}'] must be a contiguous float32 {shape} tensor on {device}")
        outs.append(t)
    return outs


def new_event():
    """A CUDA event without timing owned by the library (gs_host_event_create), as an opaque handle."""
    ev = _lib.gs_host_event_create()
    if not ev:
        raise RuntimeError("gs_host_event_create failed")
    return ev


def stream_wait_event(stream, event):
    """Make the torch stream wait for a library event (gs_stream_wait_event)."""
    _check(_lib.gs_stream_wait_event(ctypes.c_void_p(stream.cuda_stream), event), "gs_stream_wait_event")


def backward_render(background, P, R, dL_dout_color, geomBuffer, binningBuffer, imageBuffer, cleared_event=None):
    """First half of the backward (gs_backward_render): zero the per-Gaussian accumulators and run the tile kernel.
    Returns the scratch tensor that backward_geometry / dp_color_grad_push read. cleared_event (new_event()): recorded
    between the two, so that side-stream work can be ordered after the clear and beside the tile kernel."""
    device = dL_dout_color.device
    H, W = int(dL_dout_color.size(1)), int(dL_dout_color.size(2))
    background = _f32_cuda(background, "bg", device)
    dL_dout_color = _f32_cuda(dL_dout_color, "dL_dout_color", device)
    with torch.cuda.device(device):
        scratch = torch.empty(max(_lib.gs_backward_scratch_bytes(P), 16), dtype=torch.uint8, device=device)
        _check(_lib.gs_backward_render_ev(_stream(device), int(P), int(R), W, H, _ptr(background), geomBuffer.data_ptr(),
                                          binningBuffer.data_ptr(), imageBuffer.data_ptr(), dL_dout_color.data_ptr(),
                                          scratch.data_ptr(), cleared_event), "gs_backward_render")
    return scratch


def backward_geometry(means3D, radii, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix, tan_fovx,
                      tan_fovy, H, W, sh, degree, campos, geomBuffer, scratch, out_buffers=None, write_sh=True):
    """Second half of the backward (gs_backward_geometry). write_sh=False: the SH gradient is not written (the caller
    rebuilds the cross-view sum with dp_sh_grad_combine); the returned tuple has None in its place."""
    device = means3D.device
    P = int(means3D.size(0))
    split = isinstance(sh, (tuple, list))
    M = 16 if split else (int(sh.size(1)) if sh is not None and sh.numel() != 0 else 0)
    means3D = _f32_cuda(means3D, "means3D", device)
    scales = _f32_cuda(scales, "scales", device)
    rotations = _f32_cuda(rotations, "rotations", device)
    cov3D_precomp = _f32_cuda(cov3D_precomp, "cov3D_precomp", device)
    viewmatrix = _f32_cuda(viewmatrix, "viewmatrix", device)
    projmatrix = _f32_cuda(projmatrix, "projmatrix", device)
    campos = _f32_cuda(campos, "campos", device)
    radii = radii.contiguous()
    if split:
        if not write_sh:
            raise RuntimeError("split SH tensors: write_sh=False is the records exchange (backward_geometry_push)")
        sh_dc, sh_rest = _split_sh(sh, P, device)
        with torch.cuda.device(device):
            outs = _backward_outputs(P, M, device, out_buffers, want_sh=False)
            dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, _, dL_dscales, dL_drotations = outs
            buf = (out_buffers or {}).get("sh")
            d_dc, d_rest = _split_sh(buf, P, device) if buf is not None else (
                torch.empty((P, 1, 3), dtype=torch.float32, device=device), torch.empty((P, 15, 3), dtype=torch.float32, device=device))
            if P:
                _check(_lib.gs_backward_geometry_split_sh(
                    _stream(device), P, int(degree), int(W), int(H), _ptr(means3D), sh_dc.data_ptr(), sh_rest.data_ptr(),
                    _ptr(scales), float(scale_modifier), _ptr(rotations), _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos),
                    float(tan_fovx), float(tan_fovy), radii.data_ptr(), geomBuffer.data_ptr(), scratch.data_ptr(),
                    dL_dmeans2D.data_ptr(), dL_dcolors.data_ptr(), dL_dopacity.data_ptr(), dL_dmeans3D.data_ptr(),
                    dL_dcov3D.data_ptr(), d_dc.data_ptr(), d_rest.data_ptr(), dL_dscales.data_ptr(), dL_drotations.data_ptr()),
                    "gs_backward_geometry_split_sh")
        return (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, (d_dc, d_rest), dL_dscales, dL_drotations)
    sh = _f32_cuda(sh, "sh", device)
    with torch.cuda.device(device):
        outs = _backward_outputs(P, M, device, out_buffers, want_sh=write_sh)
        dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations = outs
        if P:
            _check(_lib.gs_backward_geometry(
                _stream(device), P, int(degree), M, int(W), int(H), _ptr(means3D), _ptr(sh), _ptr(scales),
                float(scale_modifier), _ptr(rotations), _ptr(cov3D_precomp), _ptr(viewmatrix), _ptr(projmatrix),
                _ptr(campos), float(tan_fovx), float(tan_fovy), radii.data_ptr(), geomBuffer.data_ptr(),
                scratch.data_ptr(), dL_dmeans2D.data_ptr(), dL_dcolors.data_ptr(), dL_dopacity.data_ptr(),
                dL_dmeans3D.data_ptr(), dL_dcov3D.data_ptr(), _ptr(dL_dsh), dL_dscales.data_ptr(),
                dL_drotations.data_ptr()), "gs_backward_geometry")
    return tuple(outs)


def _split_sh(pair, P, device):
    """(features_dc [P,1,3], features_rest [P,15,3]) as contiguous float32 CUDA tensors on 16-byte aligned bases."""
    dc, rest = pair
    for name, t, shape in (("features_dc", dc, (P, 1, 3)), ("features_rest", rest, (P, 15, 3))):
        if not isinstance(t, torch.Tensor) or t.dtype != torch.float32 or t.device != device or tuple(t.shape) != shape:
            raise RuntimeError(f"split SH: {name} must be a float32 {shape} tensor on {device} (degree-3 storage, M = 16)")
    dc, rest = dc.contiguous(), rest.contiguous()
    if dc.data_ptr() % 16 or rest.data_ptr() % 16:
        raise RuntimeError("split SH: features_dc / features_rest must start on 16-byte boundaries")
    return dc, rest


def _ptr_array(ptrs):
    arr = (ctypes.c_void_p * len(ptrs))(*[int(p) for p in ptrs])
    return arr


DP_LAYOUTS = {"dense": 0, "sparse": 1, "records": 2}


def backward_geometry_push(means3D, radii, scales, rotations, scale_modifier, viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W,
                           sh, degree, campos, geomBuffer, scratch, dst_ptrs, slot_offset, multicast=False, means2D_out=None):
    """Second half of the backward for the "records" exchange (gs_backward_geometry_push): the parameter gradients leave as
    sparse chunk records at float offset slot_offset of every buffer in dst_ptrs; returns dL_dmeans2D [P, 3]."""
    device = means3D.device
    P = int(means3D.size(0))
    if isinstance(sh, (tuple, list)):
        sh, sh_rest = _split_sh(sh, P, device)
        M = 16
    else:
        sh, sh_rest, M = _f32_cuda(sh, "sh", device), None, int(sh.size(1))
    means3D = _f32_cuda(means3D, "means3D", device)
    scales = _f32_cuda(scales, "scales", device)
    rotations = _f32_cuda(rotations, "rotations", device)
    viewmatrix = _f32_cuda(viewmatrix, "viewmatrix", device)
    projmatrix = _f32_cuda(projmatrix, "projmatrix", device)
    campos = _f32_cuda(campos, "campos", device)
    radii = radii.contiguous()
    arr = _ptr_array(dst_ptrs)
    with torch.cuda.device(device):
        if means2D_out is None:
            means2D_out = torch.empty((P, 3), dtype=torch.float32, device=device)
        elif tuple(means2D_out.shape) != (P, 3) or means2D_out.dtype != torch.float32 or not means2D_out.is_contiguous():
            raise RuntimeError(f"means2D_out must be a contiguous float32 {(P, 3)} tensor")
        if P:
            _check(_lib.gs_backward_geometry_push(
                _stream(device), P, int(degree), M, int(W), int(H), _ptr(means3D), _ptr(sh), _ptr(sh_rest), _ptr(scales),
                float(scale_modifier), _ptr(rotations), _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos), float(tan_fovx),
                float(tan_fovy), radii.data_ptr(), geomBuffer.data_ptr(), scratch.data_ptr(), means2D_out.data_ptr(),
                ctypes.cast(arr, _c_void_p), len(dst_ptrs), 1 if multicast else 0, int(slot_offset)), "gs_backward_geometry_push")
    return means2D_out


def dp_grad_record_combine(means3D, campos_views, records, slot_stride, degree, M, outs=None, outputs_are_zero=False,
                           split_sh=False):
    """Gradients summed over the V views whose chunk records sit in records[V, slot_stride] (gs_dp_grad_record_combine).
    outs: dict of contiguous float32 tensors "means3D" [P,3], "opacity" [P,1], "scales" [P,3], "rotations" [P,4], "sh" [P,M,3]
    (allocated when None); outputs_are_zero=True promises they hold zeros already. split_sh (or an "sh_rest" entry in outs):
    "sh" is the [P,1,3] gradient of features_dc and "sh_rest" the [P,M-1,3] one of features_rest."""
    device = means3D.device
    P, V = int(means3D.size(0)), int(campos_views.size(0))
    means3D = _f32_cuda(means3D, "means3D", device)
    campos_views = _f32_cuda(campos_views, "campos_views", device)
    records = _f32_cuda(records, "records", device)
    if records.numel() < V * slot_stride:
        raise RuntimeError("records is smaller than V * slot_stride floats")
    shapes = {"means3D": (P, 3), "opacity": (P, 1), "scales": (P, 3), "rotations": (P, 4)}
    split = split_sh or (outs is not None and "sh_rest" in outs)
    if split:
        shapes.update(sh=(P, 1, 3), sh_rest=(P, int(M) - 1, 3))     # gradients of features_dc / features_rest
    else:
        shapes["sh"] = (P, int(M), 3)
    if outs is None:
        outs = {k: torch.empty(shp, dtype=torch.float32, device=device) for k, shp in shapes.items()}
        outputs_are_zero = False
    for k, shp in shapes.items():
        t = outs[k]
        if tuple(t.shape) != shp or t.dtype != torch.float32 or not t.is_contiguous() or t.device != device:
            raise RuntimeError(f"outs[{k!r}] must be a contiguous float32 {shp} tensor on {device}")
    if P:
        with torch.cuda.device(device):
            _check(_lib.gs_dp_grad_record_combine(
                _stream(device), P, int(degree), int(M), V, means3D.data_ptr(), campos_views.data_ptr(), records.data_ptr(),
                int(slot_stride), 1 if outputs_are_zero else 0, outs["means3D"].data_ptr(), outs["opacity"].data_ptr(),
                outs["scales"].data_ptr(), outs["rotations"].data_ptr(), outs["sh"].data_ptr(),
                outs["sh_rest"].data_ptr() if split else None), "gs_dp_grad_record_combine")
    return outs


def dp_slot_floats(P, layout="sparse"):
    """Floats one view's slot of the gather buffer needs for P Gaussians (gs_dp_slot_floats)."""
    return int(_lib.gs_dp_slot_floats(int(P), DP_LAYOUTS[layout]))


def dp_color_grad_push(P, radii, geomBuffer, scratch, dst_ptrs, slot_offset, multicast=False, layout="dense"):
    """Masked dL/dRGB of this rank's view -> float offset slot_offset of every buffer in dst_ptrs (raw device
    pointers: this rank's own buffer and its peers', or one multicast pointer). layout "dense": [P][3] floats; "sparse":
    bit masks + the non-zero triples only (include/gs_b200.h). gs_dp_color_grad_push."""
    device = radii.device
    arr = _ptr_array(dst_ptrs)
    with torch.cuda.device(device):
        _check(_lib.gs_dp_color_grad_push(_stream(device), int(P), radii.data_ptr(), geomBuffer.data_ptr(),
                                          scratch.data_ptr(), ctypes.cast(arr, _c_void_p), len(dst_ptrs),
                                          1 if multicast else 0, int(slot_offset), DP_LAYOUTS[layout]),
               "gs_dp_color_grad_push")


def dp_sh_grad_combine(means3D, campos_views, color_grads, slot_stride, degree, M, out=None, layout="dense"):
    """dL_dsh[P, M, 3] summed over the V views whose masked colour gradients sit in color_grads[V, slot_stride]
    in the given layout (gs_dp_sh_grad_combine)."""
    device = means3D.device
    P = int(means3D.size(0))
    V = int(campos_views.size(0))
    means3D = _f32_cuda(means3D, "means3D", device)
    campos_views = _f32_cuda(campos_views, "campos_views", device)
    color_grads = _f32_cuda(color_grads, "color_grads", device)
    if color_grads.numel() < V * slot_stride:
        raise RuntimeError("color_grads is smaller than V * slot_stride floats")
    if out is None:
        out = torch.empty((P, M, 3), dtype=torch.float32, device=device)
    elif tuple(out.shape) != (P, M, 3) or out.dtype != torch.float32 or not out.is_contiguous():
        raise RuntimeError(f"out must be a contiguous float32 {(P, M, 3)} tensor")
    if P:
        with torch.cuda.device(device):
            _check(_lib.gs_dp_sh_grad_combine(_stream(device), P, int(degree), int(M), V, means3D.data_ptr(),
                                              campos_views.data_ptr(), color_grads.data_ptr(), int(slot_stride),
                                              DP_LAYOUTS[layout], out.data_ptr()), "gs_dp_sh_grad_combine")
    return out


def dp_slab_all_reduce(device, slab_ptrs, n_floats, rank, world, multicast=False):
    """In-place two-shot all-reduce of a symmetric slab over NVLink (gs_dp_slab_all_reduce); the caller brackets it
    with cross-rank barriers."""
    arr = _ptr_array(slab_ptrs)
    with torch.cuda.device(device):
        _check(_lib.gs_dp_slab_all_reduce(_stream(device), ctypes.cast(arr, _c_void_p), len(slab_ptrs),
                                          1 if multicast else 0, int(n_floats), int(rank), int(world)),
               "gs_dp_slab_all_reduce")


def mark_visible(means3D, viewmatrix, projmatrix):
    """present[i] = view-space depth of means3D[i] > 0.2 (rasterize_points.cu:201-220)."""
    device = means3D.device
    P = int(means3D.size(0))
    present = torch.zeros((P,), dtype=torch.bool, device=device)
    if P != 0:
        means3D = _f32_cuda(means3D, "means3D", device)
        viewmatrix = _f32_cuda(viewmatrix, "viewmatrix", device)
        projmatrix = _f32_cuda(projmatrix, "projmatrix", device)
        with torch.cuda.device(device):
            _check(_lib.gs_mark_visible(_stream(device), P, _ptr(means3D), _ptr(viewmatrix), _ptr(projmatrix),
                                        present.data_ptr()), "gs_mark_visible")
    return present


def compute_relocation(opacity_old, scale_old, N, binoms, n_max):
    """(final_opacity[P], final_scale[3P]) as ComputeRelocationCUDA (rasterize_points.cu:222-248)."""
    device = opacity_old.device
    P = int(opacity_old.size(0))
    final_opacity = torch.zeros((P,), dtype=torch.float32, device=device)
    final_scale = torch.zeros((3 * P,), dtype=torch.float32, device=device)
    if P != 0:
        opacity_old = _f32_cuda(opacity_old, "opacity_old", device)
        scale_old = _f32_cuda(scale_old, "scale_old", device)
        binoms = _f32_cuda(binoms, "binoms", device)
        if N.dtype != torch.int32 or not N.is_cuda:
            raise RuntimeError("N must be an int32 CUDA tensor")
        N = N.contiguous()
        with torch.cuda.device(device):
            _check(_lib.gs_compute_relocation(_stream(device), P, _ptr(opacity_old), _ptr(scale_old), N.data_ptr(),
                                              _ptr(binoms), int(n_max), final_opacity.data_ptr(),
                                              final_scale.data_ptr()), "gs_compute_relocation")
    return final_opacity, final_scale


def add_noise(xyz, scaling_raw, rotation_raw, opacity_raw, lr_scale, noise=None, seed=0, offset=0):
    """In place: xyz += Sigma * (n * gate(opacity) * lr_scale). See gs_add_noise in include/gs_b200.h."""
    device = xyz.device
    P = int(xyz.size(0))
    if P == 0:
        return xyz
    if not (xyz.is_cuda and xyz.dtype == torch.float32 and xyz.is_contiguous()):
        raise RuntimeError("xyz must be a contiguous float32 CUDA tensor (updated in place)")
    scaling_raw = _f32_cuda(scaling_raw, "scaling", device)
    rotation_raw = _f32_cuda(rotation_raw, "rotation", device)
    opacity_raw = _f32_cuda(opacity_raw, "opacity", device)
    noise = _f32_cuda(noise, "noise", device)
    with torch.cuda.device(device):
        _check(_lib.gs_add_noise(_stream(device), P, xyz.data_ptr(), _ptr(scaling_raw), _ptr(rotation_raw),
                                 _ptr(opacity_raw), _ptr(noise), float(lr_scale), int(seed), int(offset)),
               "gs_add_noise")
    return xyz


def _chw(t, name, device):
    t = _f32_cuda(t, name, device)
    if t.dim() == 4 and t.size(0) == 1:
        t = t[0]
    if t.dim() != 3:
        raise RuntimeError(f"{name} must be [C, H, W] (or [1, C, H, W]), got {tuple(t.shape)}")
    return t


def photometric_loss_forward(image, gt):
    """(means, ws): means = device float[2] {mean |image - gt|, mean ssim_map}; ws keeps the derivative maps for
    photometric_loss_backward (gs_photometric_loss_forward)."""
    device = image.device
    if not image.is_cuda:
        raise RuntimeError("image must be a CUDA tensor (no CPU path exists)")
    image, gt = _chw(image, "image", device), _chw(gt, "gt", device)
    if image.shape != gt.shape:
        raise RuntimeError(f"image {tuple(image.shape)} and gt {tuple(gt.shape)} differ in shape")
    C, H, W = (int(x) for x in image.shape)
    with torch.cuda.device(device):
        ws = torch.empty(_lib.gs_loss_workspace_bytes(C, H, W), dtype=torch.uint8, device=device)
        means = torch.empty(2, dtype=torch.float32, device=device)
        _check(_lib.gs_photometric_loss_forward(_stream(device), C, H, W, image.data_ptr(), gt.data_ptr(), ws.data_ptr(),
                                                means.data_ptr()), "gs_photometric_loss_forward")
    return means, ws


def photometric_loss_backward(image, gt, ws, c_l1, c_ssim, grad_scale=None):
    """dL/dimage = grad_scale * (c_l1 * sign(image - gt) + c_ssim * d(sum ssim_map)/dimage) (gs_photometric_loss_backward)."""
    device = image.device
    shape = image.shape
    image, gt = _chw(image, "image", device), _chw(gt, "gt", device)
    C, H, W = (int(x) for x in image.shape)
    if grad_scale is not None:
        grad_scale = _f32_cuda(grad_scale.reshape(1), "grad_scale", device)
    with torch.cuda.device(device):
        out = torch.empty((C, H, W), dtype=torch.float32, device=device)
        _check(_lib.gs_photometric_loss_backward(_stream(device), C, H, W, image.data_ptr(), gt.data_ptr(), ws.data_ptr(),
                                                 float(c_l1), float(c_ssim), _ptr(grad_scale), out.data_ptr()),
               "gs_photometric_loss_backward")
    return out.view(shape)


def l1_loss_forward(x, y):
    """Device float[2] whose element 0 is mean |x - y| (gs_l1_loss_forward); x, y: CUDA tensors of one shape."""
    device = x.device
    if not x.is_cuda:
        raise RuntimeError("network_output must be a CUDA tensor (no CPU path exists)")
    x, y = _f32_cuda(x, "network_output", device), _f32_cuda(y, "gt", device)
    if x.shape != y.shape:
        raise RuntimeError(f"network_output {tuple(x.shape)} and gt {tuple(y.shape)} differ in shape")
    n = int(x.numel())
    if n == 0:
        raise RuntimeError("l1_loss of empty tensors")
    with torch.cuda.device(device):
        ws = torch.empty(_lib.gs_l1_workspace_bytes(n), dtype=torch.uint8, device=device)
        means = torch.empty(2, dtype=torch.float32, device=device)
        _check(_lib.gs_l1_loss_forward(_stream(device), n, x.data_ptr(), y.data_ptr(), ws.data_ptr(), means.data_ptr()),
               "gs_l1_loss_forward")
    return means


def l1_loss_backward(x, y, grad_scale=None):
    """grad_scale * sign(x - y) / numel (gs_l1_loss_backward), shaped like x."""
    device = x.device
    x, y = _f32_cuda(x, "network_output", device), _f32_cuda(y, "gt", device)
    if grad_scale is not None:
        grad_scale = _f32_cuda(grad_scale.reshape(1), "grad_scale", device)
    with torch.cuda.device(device):
        out = torch.empty_like(x)
        _check(_lib.gs_l1_loss_backward(_stream(device), int(x.numel()), x.data_ptr(), y.data_ptr(), _ptr(grad_scale),
                                        out.data_ptr()), "gs_l1_loss_backward")
    return out


def activate_forward(raw_opacity, raw_scaling, raw_rotation):
    """(opacity, scaling, rotation, means): sigmoid / exp / normalize in one pass; means = device float[2]
    {mean |opacity|, mean |scaling|} (gs_activate_forward)."""
    device = raw_opacity.device
    if not raw_opacity.is_cuda:
        raise RuntimeError("raw_opacity must be a CUDA tensor (no CPU path exists)")
    raw_opacity = _f32_cuda(raw_opacity, "raw_opacity", device)
    raw_scaling = _f32_cuda(raw_scaling, "raw_scaling", device)
    raw_rotation = _f32_cuda(raw_rotation, "raw_rotation", device)
    P = int(raw_opacity.size(0))
    if tuple(raw_scaling.shape) != (P, 3) or tuple(raw_rotation.shape) != (P, 4) or raw_opacity.numel() != P:
        raise RuntimeError("expected raw_opacity [P,1], raw_scaling [P,3], raw_rotation [P,4]")
    with torch.cuda.device(device):
        opacity, scaling, rotation = torch.empty_like(raw_opacity), torch.empty_like(raw_scaling), torch.empty_like(raw_rotation)
        means = torch.zeros(2, dtype=torch.float32, device=device)
        if P:
            ws = torch.empty(_lib.gs_activation_workspace_bytes(P), dtype=torch.uint8, device=device)
            _check(_lib.gs_activate_forward(_stream(device), P, raw_opacity.data_ptr(), raw_scaling.data_ptr(),
                                            raw_rotation.data_ptr(), opacity.data_ptr(), scaling.data_ptr(),
                                            rotation.data_ptr(), ws.data_ptr(), means.data_ptr()), "gs_activate_forward")
    return opacity, scaling, rotation, means


def activate_backward(raw_opacity, raw_scaling, raw_rotation, g_opacity, g_scaling, g_rotation, c_opacity, c_scaling,
                      reg_grad):
    """Gradients w.r.t. the raw parameters (gs_activate_backward); g_* / reg_grad may be None."""
    device = raw_opacity.device
    P = int(raw_opacity.size(0))
    g_opacity = _f32_cuda(g_opacity, "g_opacity", device)
    g_scaling = _f32_cuda(g_scaling, "g_scaling", device)
    g_rotation = _f32_cuda(g_rotation, "g_rotation", device)
    if reg_grad is not None:
        reg_grad = _f32_cuda(reg_grad.reshape(1), "reg_grad", device)
    with torch.cuda.device(device):
        d_o, d_s, d_q = torch.empty_like(raw_opacity), torch.empty_like(raw_scaling), torch.empty_like(raw_rotation)
        if P:
            _check(_lib.gs_activate_backward(_stream(device), P, raw_opacity.data_ptr(), raw_scaling.data_ptr(),
                                             raw_rotation.data_ptr(), _ptr(g_opacity), _ptr(g_scaling), _ptr(g_rotation),
                                             float(c_opacity), float(c_scaling), _ptr(reg_grad), d_o.data_ptr(),
                                             d_s.data_ptr(), d_q.data_ptr()), "gs_activate_backward")
    return d_o, d_s, d_q


def knn_mean_dist2(points):
    """Mean squared distance of every point to its 3 nearest other points (gs_knn_mean_dist2) -> float32 [P]."""
    device = points.device
    if not points.is_cuda:
        raise RuntimeError("points must be a CUDA tensor (no CPU path exists)")
    points = _f32_cuda(points, "points", device)
    if points.dim() != 2 or points.size(1) != 3:
        raise RuntimeError(f"points must be [P, 3], got {tuple(points.shape)}")
    P = int(points.size(0))
    with torch.cuda.device(device):
        means = torch.zeros((P,), dtype=torch.float32, device=device)
        if P:
            ws = torch.empty(_lib.gs_knn_workspace_bytes(P), dtype=torch.uint8, device=device)
            _check(_lib.gs_knn_mean_dist2(_stream(device), P, points.data_ptr(), means.data_ptr(), ws.data_ptr()),
                   "gs_knn_mean_dist2")
    return means


SWIN_SET_KEYS = ("xyz", "rotation", "features_dc", "features_rest", "opacity", "scaling", "rigid_v", "rigid_rotvec",
                 "rigid_rotcen", "frame_start")


def _swin_set(tensors, device, what):
    out = []
    for key, t in zip(SWIN_SET_KEYS, tensors):
        t = _f32_cuda(t, f"{what}.{key}", device)
        out.append(t)
    return out


def _idx64(t, device, name):
    if t.device != device or t.dtype != torch.int64:
        raise RuntimeError(f"{name} must be an int64 tensor on {device}")
    return t.contiguous()


def swin_gather_forward(frame, deform, imm_idx, mat_idx, imm_set, mat_set):
    """Rasterizer inputs of the Gaussians alive at `frame` (gs_swin_gather_forward). imm_set / mat_set: the ten tensors
    of SWIN_SET_KEYS. Returns (xyz, rotation, feature, opacity, scaling) for the Ki + Km active Gaussians."""
    device = imm_set[0].device
    if not imm_set[0].is_cuda:
        raise RuntimeError("swin_gather needs CUDA tensors (no CPU path exists)")
    imm_idx, mat_idx = _idx64(imm_idx, device, "imm_idx"), _idx64(mat_idx, device, "mat_idx")
    Ki, Km = int(imm_idx.numel()), int(mat_idx.numel())
    M = int(imm_set[3].size(1)) + 1
    imm = _swin_set(imm_set, device, "immature")
    mat = _swin_set(mat_set, device, "matured") if Km else None
    K = Ki + Km
    opts = dict(dtype=torch.float32, device=device)
    with torch.cuda.device(device):
        outs = (torch.empty((K, 3), **opts), torch.empty((K, 4), **opts), torch.empty((K, M, 3), **opts),
                torch.empty((K, 1), **opts), torch.empty((K, 3), **opts))
        if K:
            ia = _ptr_array([t.data_ptr() for t in imm])
            ma = _ptr_array([t.data_ptr() for t in mat]) if mat else None
            _check(_lib.gs_swin_gather_forward(
                _stream(device), Ki, Km, M, float(frame), 1 if deform else 0, _ptr(imm_idx), _ptr(mat_idx),
                ctypes.cast(ia, _c_void_p), ctypes.cast(ma, _c_void_p) if ma is not None else None,
                *[o.data_ptr() for o in outs]), "gs_swin_gather_forward")
    return outs


def swin_gather_backward(frame, deform, imm_idx, imm_set, out_grads):
    """Gradients w.r.t. the first nine tensors of the immature set (dense, zero outside the gathered rows)."""
    device = imm_set[0].device
    imm_idx = _idx64(imm_idx, device, "imm_idx")
    Ki = int(imm_idx.numel())
    M = int(imm_set[3].size(1)) + 1
    imm = _swin_set(imm_set, device, "immature")
    grads = [_f32_cuda(g, "out_grad", device) if g is not None else None for g in out_grads]
    with torch.cuda.device(device):
        d = [torch.zeros_like(t) for t in imm[:9]]
        if Ki:
            ia = _ptr_array([t.data_ptr() for t in imm])
            ga = _ptr_array([g.data_ptr() if g is not None and g.numel() else 0 for g in grads])
            da = _ptr_array([t.data_ptr() for t in d])
            _check(_lib.gs_swin_gather_backward(_stream(device), Ki, M, float(frame), 1 if deform else 0, imm_idx.data_ptr(),
                                                ctypes.cast(ia, _c_void_p), ctypes.cast(ga, _c_void_p),
                                                ctypes.cast(da, _c_void_p)), "gs_swin_gather_backward")
    return d


def adam_step(params, grads, exp_avgs, exp_avg_sqs, lrs, steps, beta1, beta2, eps):
    """One multi-tensor Adam launch (gs_adam_step) over up to 8 contiguous float32 CUDA tensors; updates params,
    exp_avgs and exp_avg_sqs in place. lrs / steps: per-tensor learning rate and (already incremented) step count."""
    n = len(params)
    if n == 0:
        return
    if n > 8:
        for i in range(0, n, 8):
            adam_step(params[i:i + 8], grads[i:i + 8], exp_avgs[i:i + 8], exp_avg_sqs[i:i + 8], lrs[i:i + 8],
                      steps[i:i + 8], beta1, beta2, eps)
        return
    device = params[0].device
    for group in (params, grads, exp_avgs, exp_avg_sqs):
        for t in group:
            if not t.is_cuda or t.device != device or t.dtype != torch.float32 or not t.is_contiguous():
                raise RuntimeError("adam_step needs contiguous float32 CUDA tensors on one device (no CPU path exists)")
    for p, g, m, v in zip(params, grads, exp_avgs, exp_avg_sqs):
        if not (p.numel() == g.numel() == m.numel() == v.numel()):
            raise RuntimeError("adam_step: param / grad / state sizes differ")
    arr = lambda ts: ctypes.cast(_ptr_array([t.data_ptr() for t in ts]), _c_void_p)  # noqa: E731
    numel = (ctypes.c_ulonglong * n)(*[t.numel() for t in params])
    lr = (ctypes.c_double * n)(*[float(x) for x in lrs])
    st = (ctypes.c_int * n)(*[int(x) for x in steps])
    with torch.cuda.device(device):
        _check(_lib.gs_adam_step(_stream(device), n, arr(params), arr(grads), arr(exp_avgs), arr(exp_avg_sqs),
                                 ctypes.cast(numel, _c_void_p), ctypes.cast(lr, _c_void_p), ctypes.cast(st, _c_void_p),
                                 float(beta1), float(beta2), float(eps)), "gs_adam_step")


def debug_fetch(which, P, R, W, H, geom, binning, img, dtype):
    """Copy one internal workspace array into a fresh tensor (tests only; numbering in gs_b200.h)."""
    nbytes = _lib.gs_debug_fetch(which, P, R, W, H, _ptr(geom), _ptr(binning), _ptr(img), None, 0)
    if nbytes < 0:
        raise RuntimeError(f"gs_debug_fetch({which}) failed")
    out = torch.empty(nbytes, dtype=torch.uint8, device=geom.device if geom.numel() else img.device)
    if nbytes:
        _lib.gs_debug_fetch(which, P, R, W, H, _ptr(geom), _ptr(binning), _ptr(img), out.data_ptr(), nbytes)
    return out.view(dtype)


def debug_forward_work(counters):
    """Arm (tensor of 5 int64 on the device, zeroed) or disarm (None) the tallying forward kernel (gs_debug_forward_work)."""
    _check(_lib.gs_debug_forward_work(counters.data_ptr() if counters is not None else None), "gs_debug_forward_work")


def debug_render_work(P, R, W, H, geom, binning, img):
    """(walked, evaluated): int32 [tiles] each - (patch, splat) positions the backward walks per tile and those that
    survive the patch cull (gs_debug_render_work). Measurement only."""
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    dev = img.device
    walked = torch.zeros(tiles, dtype=torch.int32, device=dev)
    evaluated = torch.zeros(tiles, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _check(_lib.gs_debug_render_work(_stream(dev), int(P), int(R), int(W), int(H), geom.data_ptr(), binning.data_ptr(),
                                         img.data_ptr(), walked.data_ptr(), evaluated.data_ptr()), "gs_debug_render_work")
    return walked, evaluated


def _debug_sync(stage):
    # debug=True in the reference syncs and throws after every stage (auxiliary.h:166-173).
    try:
        torch.cuda.synchronize()
    except RuntimeError as ex:
        raise RuntimeError(f"[CUDA ERROR] in rasterizer stage '{stage}': {ex}") from ex


def profile_enable(max_marks):
    """Arm (or, with 0, disarm) the library's per-stage CUDA-event timing. Measurement only."""
    _check(_lib.gs_profile_enable(int(max_marks)), "gs_profile_enable")


def profile_collect():
    """-> {stage_name: (ms_sum, calls)} since the last collect; synchronises on the last recorded mark."""
    n = _lib.gs_stage_count()
    ms = (ctypes.c_float * n)()
    calls = (ctypes.c_int * n)()
    _check(_lib.gs_profile_collect(ctypes.cast(ms, _c_void_p), ctypes.cast(calls, _c_void_p)), "gs_profile_collect")
    return {_lib.gs_stage_name(i).decode(): (float(ms[i]), int(calls[i])) for i in range(n)}


def launch_count():
    return int(_lib.gs_launch_count())
