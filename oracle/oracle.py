"""numpy front-end of the plain-C CPU oracle (oracle/oracle_cpu.c -> liboracle_cpu.so).

TEST INFRASTRUCTURE ONLY — the checker, never the product. Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this module. See oracle_cpu.c for the reference file:line each
function follows and for the parity-pin status.

The stages are separate on purpose so that integer stages can be checked bit-exactly on identical float inputs:
  preprocess() -> bin() -> render_fwd() -> render_bwd() -> preprocess_bwd()
``forward()`` / ``backward()`` chain them the way Rasterizer::forward / ::backward do
(cuda_rasterizer/rasterizer_impl.cu:198-338, :342-436).
"""
import ctypes
import os
import subprocess

import numpy as np

_DIR = os.path.dirname(os.path.abspath(__file__))
_PATH = os.path.join(_DIR, "liboracle_cpu.so")
_lib = None

_fp = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_ip = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_up = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_bp = np.ctypeslib.ndpointer(dtype=np.uint8, flags="C_CONTIGUOUS")
_kp = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")
_vp, _i, _f = ctypes.c_void_p, ctypes.c_int, ctypes.c_float


def build():
    subprocess.check_call(["make", "-C", _DIR, "liboracle_cpu.so"], stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_PATH):
            build()
        L = ctypes.CDLL(_PATH)
        L.orc_preprocess.argtypes = [_i, _i, _i, _fp, _vp, _f, _vp, _fp, _vp, _vp, _vp, _fp, _fp, _vp, _i, _i, _f, _f,
                                     _ip, _fp, _fp, _fp, _fp, _fp, _up, _bp]
        L.orc_mark_visible.argtypes = [_i, _fp, _fp, _bp]
        L.orc_bin.argtypes = [_i, _i, _i, _fp, _fp, _ip, _up, ctypes.c_uint32, _kp, _up, _up]
        L.orc_render_fwd.argtypes = [_i, _i, _up, _up, _fp, _fp, _fp, _fp, _fp, _fp, _up, _ip, _i]
        L.orc_render_bwd.argtypes = [_i, _i, _up, _up, _fp, _fp, _fp, _fp, _fp, _up, _fp, _fp, _fp, _fp, _fp, _i]
        L.orc_preprocess_bwd.argtypes = [_i, _i, _i, _fp, _ip, _vp, _bp, _vp, _vp, _f, _fp, _fp, _fp, _vp, _i, _i, _f, _f,
                                         _fp, _fp, _fp, _fp, _fp, _fp, _fp, _fp]
        L.orc_compute_relocation.argtypes = [_i, _fp, _fp, _ip, _fp, _i, _fp, _fp]
        L.orc_add_noise.argtypes = [_i, _fp, _fp, _fp, _fp, _fp, _f]
        L.orc_knn_mean_dist2.argtypes = [_i, _fp, _fp]
        L.orc_knn_mean_dist2.restype = None
        L.orc_num_threads.restype = _i
        L.orc_set_num_threads.argtypes = [_i]
        L.orc_set_num_threads.restype = None
        for n in ("orc_preprocess", "orc_mark_visible", "orc_bin", "orc_render_fwd", "orc_render_bwd",
                  "orc_preprocess_bwd", "orc_compute_relocation", "orc_add_noise"):
            getattr(L, n).restype = None
        _lib = L
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _opt(a):
    """Optional float array -> void* (NULL when absent)."""
    if a is None or a.size == 0:
        return None, None
    a = _f32(a)
    return a, a.ctypes.data_as(ctypes.c_void_p)


def num_threads():
    return int(lib().orc_num_threads())


def set_num_threads(n=None):
    """Use n OpenMP threads (default: every core this process may run on), whatever OMP_NUM_THREADS says - torchrun
    exports OMP_NUM_THREADS=1 to its ranks. Returns the thread count now in effect."""
    import os
    if n is None:
        n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    lib().orc_set_num_threads(int(n))
    return num_threads()


def preprocess(means3D, scales, rotations, opacities, shs, view, proj, campos, W, H, tanfovx, tanfovy, degree=3,
               scale_modifier=1.0, cov3D_precomp=None, colors_precomp=None):
    means3D = _f32(means3D)
    P = means3D.shape[0]
    k_sc, p_sc = _opt(scales)
    k_ro, p_ro = _opt(rotations)
    k_sh, p_sh = _opt(shs)
    k_c3, p_c3 = _opt(cov3D_precomp)
    k_cp, p_cp = _opt(colors_precomp)
    k_cam, p_cam = _opt(campos)
    M = k_sh.shape[1] if k_sh is not None else 0
    out = dict(radii=np.zeros(P, np.int32), means2D=np.zeros((P, 2), np.float32), depths=np.zeros(P, np.float32),
               cov3D=np.zeros((P, 6), np.float32), rgb=np.zeros((P, 3), np.float32),
               conic_opacity=np.zeros((P, 4), np.float32), tiles_touched=np.zeros(P, np.uint32),
               clamped=np.zeros((P, 3), np.uint8))
    lib().orc_preprocess(P, degree, M, means3D, p_sc, scale_modifier, p_ro, _f32(opacities).reshape(-1), p_sh, p_c3, p_cp,
                         _f32(view).reshape(-1), _f32(proj).reshape(-1), p_cam, W, H, tanfovx, tanfovy, out["radii"],
                         out["means2D"], out["depths"], out["cov3D"], out["rgb"], out["conic_opacity"],
                         out["tiles_touched"], out["clamped"])
    return out


def mark_visible(means3D, view):
    means3D = _f32(means3D)
    present = np.zeros(means3D.shape[0], np.uint8)
    lib().orc_mark_visible(means3D.shape[0], means3D, _f32(view).reshape(-1), present)
    return present.astype(bool)


def bin(means2D, depths, radii, tiles_touched, W, H):
    """-> dict(point_offsets, keys (sorted u64), point_list (sorted ids u32), ranges [tiles,2], R)."""
    P = radii.shape[0]
    offsets = np.cumsum(tiles_touched.astype(np.uint64)).astype(np.uint32)
    R = int(offsets[-1]) if P else 0
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    keys = np.zeros(max(R, 1), np.uint64)
    vals = np.zeros(max(R, 1), np.uint32)
    ranges = np.zeros((tiles, 2), np.uint32)
    lib().orc_bin(P, W, H, _f32(means2D), _f32(depths), np.ascontiguousarray(radii, np.int32), offsets, R, keys, vals,
                  ranges.reshape(-1))
    return dict(point_offsets=offsets, keys=keys[:R], point_list=vals[:R], ranges=ranges, R=R)


def render_fwd(W, H, ranges, point_list, means2D, conic_opacity, colors, bg, P, tile_step=1):
    out_color = np.zeros((3, H, W), np.float32)
    final_T = np.zeros(H * W, np.float32)
    n_contrib = np.zeros(H * W, np.uint32)
    is_used = np.zeros(P, np.int32)
    pl = np.ascontiguousarray(point_list, np.uint32) if len(point_list) else np.zeros(1, np.uint32)
    lib().orc_render_fwd(W, H, np.ascontiguousarray(ranges, np.uint32).reshape(-1), pl, _f32(means2D), _f32(conic_opacity),
                         _f32(colors), _f32(bg), out_color.reshape(-1), final_T, n_contrib, is_used, tile_step)
    return dict(color=out_color, final_T=final_T, n_contrib=n_contrib, is_used=is_used)


def render_bwd(W, H, ranges, point_list, bg, means2D, conic_opacity, colors, final_T, n_contrib, dL_dpix, P, tile_step=1):
    g = dict(dL_dmean2D=np.zeros((P, 3), np.float32), dL_dconic=np.zeros((P, 4), np.float32),
             dL_dopacity=np.zeros(P, np.float32), dL_dcolors=np.zeros((P, 3), np.float32))
    pl = np.ascontiguousarray(point_list, np.uint32) if len(point_list) else np.zeros(1, np.uint32)
    lib().orc_render_bwd(W, H, np.ascontiguousarray(ranges, np.uint32).reshape(-1), pl, _f32(bg), _f32(means2D),
                         _f32(conic_opacity), _f32(colors), _f32(final_T), np.ascontiguousarray(n_contrib, np.uint32),
                         _f32(dL_dpix).reshape(-1), g["dL_dmean2D"], g["dL_dconic"], g["dL_dopacity"], g["dL_dcolors"],
                         tile_step)
    return g


def preprocess_bwd(means3D, radii, shs, clamped, scales, rotations, scale_modifier, cov3D, view, proj, campos, W, H,
                   tanfovx, tanfovy, degree, dL_dmean2D, dL_dconic, dL_dcolors):
    means3D = _f32(means3D)
    P = means3D.shape[0]
    k_sh, p_sh = _opt(shs)
    k_sc, p_sc = _opt(scales)
    k_ro, p_ro = _opt(rotations)
    k_cam, p_cam = _opt(campos)
    M = k_sh.shape[1] if k_sh is not None else 0
    g = dict(dL_dmeans3D=np.zeros((P, 3), np.float32), dL_dcov3D=np.zeros((P, 6), np.float32),
             dL_dsh=np.zeros((P, max(M, 1), 3), np.float32), dL_dscales=np.zeros((P, 3), np.float32),
             dL_drots=np.zeros((P, 4), np.float32))
    lib().orc_preprocess_bwd(P, degree, M, means3D, np.ascontiguousarray(radii, np.int32), p_sh,
                             np.ascontiguousarray(clamped, np.uint8).reshape(-1), p_sc, p_ro, scale_modifier, _f32(cov3D),
                             _f32(view).reshape(-1), _f32(proj).reshape(-1), p_cam, W, H, tanfovx, tanfovy,
                             _f32(dL_dmean2D), _f32(dL_dconic), _f32(dL_dcolors), g["dL_dmeans3D"], g["dL_dcov3D"],
                             g["dL_dsh"].reshape(-1), g["dL_dscales"], g["dL_drots"])
    if M == 0:
        g["dL_dsh"] = np.zeros((P, 0, 3), np.float32)
    return g


def forward(scene, cam, bg=(0.0, 0.0, 0.0), degree=3, scale_modifier=1.0, tile_step=1):
    """Whole forward on numpy inputs (scene/cam dicts as produced by gs_synth). Returns a dict with every stage's
    arrays (the CPU analogue of geomBuffer / binningBuffer / imgBuffer) plus color, radii, is_used."""
    W, H = cam["image_width"], cam["image_height"]
    P = scene["means3D"].shape[0]
    pre = preprocess(scene["means3D"], scene["scales"], scene["rotations"], scene["opacities"], scene["shs"],
                     cam["viewmatrix"], cam["projmatrix"], cam["campos"], W, H, cam["tanfovx"], cam["tanfovy"], degree,
                     scale_modifier)
    b = bin(pre["means2D"], pre["depths"], pre["radii"], pre["tiles_touched"], W, H)
    bg = np.asarray(bg, np.float32)
    r = render_fwd(W, H, b["ranges"], b["point_list"], pre["means2D"], pre["conic_opacity"], pre["rgb"], bg, P, tile_step)
    st = dict(pre)
    st.update(b)
    st.update(r)
    st["bg"] = bg
    return st


def backward(scene, cam, st, dL_dpix, degree=3, scale_modifier=1.0, tile_step=1):
    """Whole backward from the state of forward(). Returns the 8 API gradients (reference order of
    rasterize_points.cu:198) in a dict."""
    W, H = cam["image_width"], cam["image_height"]
    P = scene["means3D"].shape[0]
    rb = render_bwd(W, H, st["ranges"], st["point_list"], st["bg"], st["means2D"], st["conic_opacity"], st["rgb"],
                    st["final_T"], st["n_contrib"], dL_dpix, P, tile_step)
    pb = preprocess_bwd(scene["means3D"], st["radii"], scene["shs"], st["clamped"], scene["scales"], scene["rotations"],
                        scale_modifier, st["cov3D"], cam["viewmatrix"], cam["projmatrix"], cam["campos"], W, H,
                        cam["tanfovx"], cam["tanfovy"], degree, rb["dL_dmean2D"], rb["dL_dconic"], rb["dL_dcolors"])
    return dict(means2D=rb["dL_dmean2D"], colors=rb["dL_dcolors"], opacity=rb["dL_dopacity"].reshape(P, 1),
                means3D=pb["dL_dmeans3D"], cov3D=pb["dL_dcov3D"], sh=pb["dL_dsh"], scales=pb["dL_dscales"],
                rotations=pb["dL_drots"], conic=rb["dL_dconic"])


def compute_relocation(opacity_old, scale_old, N, binoms, n_max):
    o = _f32(opacity_old).reshape(-1)
    P = o.shape[0]
    new_o = np.zeros(P, np.float32)
    new_s = np.zeros(3 * P, np.float32)
    lib().orc_compute_relocation(P, o, _f32(scale_old).reshape(-1), np.ascontiguousarray(N, np.int32), _f32(binoms).reshape(-1),
                                 n_max, new_o, new_s)
    return new_o, new_s


def add_noise(xyz, scaling_raw, rotation_raw, opacity_raw, noise, lr_scale):
    out = _f32(xyz).copy()
    lib().orc_add_noise(out.shape[0], out, _f32(scaling_raw), _f32(rotation_raw), _f32(opacity_raw).reshape(-1), _f32(noise),
                        lr_scale)
    return out


def binoms_table(n_max=51):
    """C(n,k) as float32, the table utils/reloc_utils.py:7-11 builds with scipy.special.comb."""
    from scipy.special import comb
    t = np.zeros((n_max, n_max), np.float32)
    for n in range(n_max):
        for k in range(n + 1):
            t[n, k] = comb(n, k)
    return t


# ----------------------------------------------------------------------------------------------------------------------
# SURVEY.md §8 f1 ("next" rows): the optimizer step either side of the rasterizer
# ----------------------------------------------------------------------------------------------------------------------
def adam_step(param, grad, exp_avg, exp_avg_sq, lr, step, beta1=0.9, beta2=0.999, eps=1e-15):
    """One torch.optim.Adam update in numpy FP32 (the reference's optimizer: scene/gaussian_model_static.py:191-200,
    eps=1e-15, betas (0.9, 0.999), no weight decay / amsgrad; stepped at train_static.py:129-130). Follows the op
    sequence of torch's non-capturable path: lerp, mul+addcmul, sqrt / bias_correction2_sqrt + eps, addcdiv with
    step_size = lr / bias_correction1, the scalars computed in double and rounded to FP32 where they meet tensors.
    Returns (param, exp_avg, exp_avg_sq). Pinned against torch.optim.Adam by tests/test_oracle_golden.py."""
    f = np.float32
    p, g, m, v = (np.asarray(a, dtype=np.float32) for a in (param, grad, exp_avg, exp_avg_sq))
    w1, w2 = f(1.0 - beta1), f(1.0 - beta2)
    m = (m + w1 * (g - m)).astype(np.float32)
    v = (v * f(beta2) + (w2 * g) * g).astype(np.float32)
    bc1 = 1.0 - beta1 ** step
    bc2_sqrt = (1.0 - beta2 ** step) ** 0.5
    denom = (np.sqrt(v) / f(bc2_sqrt) + f(eps)).astype(np.float32)
    p = (p + f(-(lr / bc1)) * (m / denom)).astype(np.float32)
    return p, m, v


def ssim_window():
    """utils/loss_utils.py:27-29: exp(-(x - 5)^2 / (2 * 1.5^2)) for x in 0..10 as FP32, divided by its sum. torch's
    CPU sum of these 11 floats equals the correctly rounded sum (3.7592328), which a float64 accumulation reproduces; a
    sequential FP32 accumulation gives 3.7592325 (checked by tests/test_loss.py against torch itself)."""
    g = np.array([np.exp(-(x - 5) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)], dtype=np.float32)
    return (g / np.float32(g.astype(np.float64).sum())).astype(np.float32)


def _win_conv(a, w):
    """Per-plane 11x11 windowed sum with zero padding 5 (F.conv2d(..., padding=5, groups=C), loss_utils.py:48-56),
    separable, in float64. a: [C, H, W]."""
    C, H, W = a.shape
    pad = np.zeros((C, H + 10, W + 10), np.float64)
    pad[:, 5:5 + H, 5:5 + W] = a
    tmp = np.zeros((C, H + 10, W), np.float64)
    for k in range(11):
        tmp += w[k] * pad[:, :, k:k + W]
    out = np.zeros((C, H, W), np.float64)
    for k in range(11):
        out += w[k] * tmp[:, k:k + H, :]
    return out


def l1_loss(network_output, gt):
    """utils/loss_utils.py:21-22, torch.abs(network_output - gt).mean(), in float64 numpy. Returns dict(l1, grad): grad is
    sign(network_output - gt) / n with sign(0) = 0, torch's derivative of abs().mean()."""
    x = np.asarray(network_output, np.float64)
    y = np.asarray(gt, np.float64)
    d = x - y
    return {"l1": float(np.abs(d).mean()), "grad": np.sign(d) / d.size}


def photometric_loss(img, gt, lambda_dssim=0.2):
    """The training loss of train_static.py:94-97 in float64 numpy:
    (1 - lambda) * l1_loss(img, gt) + lambda * (1 - ssim(img, gt)), l1_loss / ssim as utils/loss_utils.py:21-22, :40-67
    (C1 = 0.01^2, C2 = 0.03^2, mean over all elements). Returns dict(l1, ssim, loss, grad_loss, grad_ssim); the
    gradients (w.r.t. img) are the analytic derivative of the same expressions. Pinned against the reference's own
    functions by tests/test_oracle_golden.py::test_loss_oracle_matches_reference_golden."""
    x = np.asarray(img, np.float64)
    y = np.asarray(gt, np.float64)
    w = ssim_window().astype(np.float64)
    n = x.size
    mu, nu = _win_conv(x, w), _win_conv(y, w)
    e11, e22, e12 = _win_conv(x * x, w), _win_conv(y * y, w), _win_conv(x * y, w)
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    A1 = 2 * mu * nu + C1
    A2 = 2 * (e12 - mu * nu) + C2
    B1 = mu * mu + nu * nu + C1
    B2 = (e11 - mu * mu) + (e22 - nu * nu) + C2
    S = A1 * A2 / (B1 * B2)
    # partial derivatives of S w.r.t. the windowed sums that depend on x
    dS_de11 = -S / B2
    dS_de12 = 2 * A1 / (B1 * B2)
    dS_dmu = 2 * nu * (A2 - A1) / (B1 * B2) - 2 * mu * S * (1 / B1 - 1 / B2)
    grad_sum_S = _win_conv(dS_dmu, w) + 2 * x * _win_conv(dS_de11, w) + y * _win_conv(dS_de12, w)
    grad_ssim = grad_sum_S / n
    l1 = np.abs(x - y).mean()
    grad_l1 = np.sign(x - y) / n
    ssim_v = S.mean()
    lam = float(lambda_dssim)
    return {"l1": l1, "ssim": ssim_v, "loss": (1 - lam) * l1 + lam * (1 - ssim_v),
            "grad_loss": (1 - lam) * grad_l1 - lam * grad_ssim, "grad_ssim": grad_ssim}


def activate(raw_opacity, raw_scaling, raw_rotation, opacity_reg=0.0, scale_reg=0.0, g_opacity=None, g_scaling=None,
             g_rotation=None, g_reg=1.0):
    """The reference's parameter activations (scene/gaussian_model_static.py:43-56: sigmoid, exp, F.normalize with eps
    1e-12) and the MCMC regularisers (train_static.py:99-100) in float64 numpy, with the analytic gradient w.r.t. the
    raw parameters for upstream gradients g_* (None = 0) and g_reg. Pinned against the same torch ops by
    tests/test_activations.py::test_activation_oracle_matches_torch_ops."""
    ro, rs, rq = (np.asarray(a, np.float64) for a in (raw_opacity, raw_scaling, raw_rotation))
    P = ro.shape[0]
    o = 1.0 / (1.0 + np.exp(-ro))
    s = np.exp(rs)
    n = np.maximum(np.sqrt((rq * rq).sum(1, keepdims=True)), 1e-12)
    q = rq / n
    reg = opacity_reg * np.abs(o).mean() + scale_reg * np.abs(s).mean()
    go = (0.0 if g_opacity is None else np.asarray(g_opacity, np.float64)) + g_reg * opacity_reg / P * np.sign(o)
    gs = (0.0 if g_scaling is None else np.asarray(g_scaling, np.float64)) + g_reg * scale_reg / (3 * P) * np.sign(s)
    d_o = go * o * (1 - o)
    d_s = gs * s
    if g_rotation is None:
        d_q = np.zeros_like(rq)
    else:
        gq = np.asarray(g_rotation, np.float64)
        d_q = (gq - q * (q * gq).sum(1, keepdims=True)) / n
    return {"opacity": o, "scaling": s, "rotation": q, "reg": reg, "d_raw_opacity": d_o, "d_raw_scaling": d_s,
            "d_raw_rotation": d_q}


def knn_mean_dist2(points):
    """distCUDA2 by brute force (oracle_cpu.c: orc_knn_mean_dist2) -> float32 [P]. O(P^2): small cases only."""
    pts = _f32(points).reshape(-1, 3)
    out = np.zeros(pts.shape[0], np.float32)
    lib().orc_knn_mean_dist2(pts.shape[0], pts.reshape(-1), out)
    return out


SWIN_KEYS = ("xyz", "rotation", "features_dc", "features_rest", "opacity", "scaling", "rigid_v", "rigid_rotvec", "rigid_rotcen",
             "frame_start")


def _quat_mul(q1, q2):
    """utils/tempo_utils.py:42-47 (w, x, y, z)."""
    w1, x1, y1, z1 = (q1[:, i] for i in range(4))
    w2, x2, y2, z2 = (q2[:, i] for i in range(4))
    return np.stack((w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                     w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2), 1)


def swin_gather(frame, imm, mat, imm_idx, mat_idx, deform=True, upstream=None):
    """SwinGaussianModel.get_basic_para_at (scene/gaussian_model.py:598-649) in float64 numpy: gather both parameter
    sets, age-based rigid_deform (utils/tempo_utils.py:49-84; rotvec2quat :4-17, rotvec2mat :19-40), F.normalize /
    sigmoid / exp. With `upstream` (dict of gradients of the five outputs) also returns the analytic gradients w.r.t.
    the nine immature parameter tensors under "grad". Pinned against goldens made with the reference's own
    tempo_utils + autograd by tests/test_swin.py::test_swin_oracle_matches_reference_golden."""
    f64 = lambda a: np.asarray(a, np.float64)  # noqa: E731
    ii, mi = np.asarray(imm_idx, np.int64), np.asarray(mat_idx, np.int64)
    Ki = len(ii)

    def cat(k):
        a = f64(imm[k])[ii]
        return np.concatenate((a, f64(mat[k])[mi]), 0) if len(mi) else a
    age = (frame - cat("frame_start"))[:, None]
    xyz, rot = cat("xyz"), cat("rotation")
    if deform:
        v, rvec, cen = cat("rigid_v"), cat("rigid_rotvec"), cat("rigid_rotcen")
        rv = rvec * age
        a = np.sqrt((rv * rv).sum(1, keepdims=True))
        inv = 1.0 / np.maximum(a, 1e-12)
        u = rv * inv
        s, c, sh, ch = np.sin(a), np.cos(a), np.sin(a / 2), np.cos(a / 2)
        pt = xyz - cen
        up_, uup = np.cross(u, pt), np.cross(u, np.cross(u, pt))
        pos = pt + s * up_ + (1 - c) * uup + cen + v * age
        q = np.concatenate((ch, u * sh), 1)
        rr = _quat_mul(q, rot)
    else:
        pos, rr = xyz, rot
    norm = np.sqrt((rr * rr).sum(1, keepdims=True))
    ninv = 1.0 / np.maximum(norm, 1e-12)
    out = {"xyz": pos, "rotation": rr * ninv,
           "feature": np.concatenate((cat("features_dc"), cat("features_rest")), 1),
           "opacity": 1.0 / (1.0 + np.exp(-cat("opacity"))), "scaling": np.exp(cat("scaling"))}
    if upstream is None:
        return out
    g = {k: f64(upstream[k])[:Ki] for k in upstream}
    sl = slice(0, Ki)
    gx, gn = g["xyz"], g["rotation"]
    nh = (rr * ninv)[sl]
    h = np.where(norm[sl] > 1e-12, (gn - nh * (nh * gn).sum(1, keepdims=True)) * ninv[sl], gn * ninv[sl])
    grads = {}
    r = rot[sl]
    if deform:
        u_, s_, c_, sh_, ch_, pt_, a_, inv_, age_ = u[sl], s[sl], c[sl], sh[sl], ch[sl], pt[sl], a[sl], inv[sl], age[sl]
        ug = np.cross(u_, gx)
        d_pt = gx - s_ * ug + (1 - c_) * np.cross(u_, ug)
        d_s = (gx * np.cross(u_, pt_)).sum(1, keepdims=True)
        d_c = -(gx * np.cross(u_, np.cross(u_, pt_))).sum(1, keepdims=True)
        gu, gp, upd = (gx * u_).sum(1, keepdims=True), (gx * pt_).sum(1, keepdims=True), (u_ * pt_).sum(1, keepdims=True)
        d_u = s_ * np.cross(pt_, gx) + (1 - c_) * (gx * upd + pt_ * gu - 2 * u_ * gp)
        q_ = q[sl]
        w1, x1, y1, z1 = (q_[:, i:i + 1] for i in range(4))
        w2, x2, y2, z2 = (r[:, i:i + 1] for i in range(4))
        hw, hx, hy, hz = (h[:, i:i + 1] for i in range(4))
        dq = np.concatenate((hw * w2 + hx * x2 + hy * y2 + hz * z2, -hw * x2 + hx * w2 - hy * z2 + hz * y2,
                             -hw * y2 + hx * z2 + hy * w2 - hz * x2, -hw * z2 - hx * y2 + hy * x2 + hz * w2), 1)
        d_rot = np.concatenate((hw * w1 + hx * x1 + hy * y1 + hz * z1, -hw * x1 + hx * w1 + hy * z1 - hz * y1,
                                -hw * y1 - hx * z1 + hy * w1 + hz * x1, -hw * z1 + hx * y1 - hy * x1 + hz * w1), 1)
        d_ch, d_sh = dq[:, :1], (u_ * dq[:, 1:]).sum(1, keepdims=True)
        d_u = d_u + sh_ * dq[:, 1:]
        d_a = d_s * c_ - d_c * s_ - 0.5 * d_ch * sh_ + 0.5 * d_sh * ch_
        d_rv = np.where(a_ > 1e-12, (d_u - u_ * (u_ * d_u).sum(1, keepdims=True)) * inv_ + d_a * u_, d_u * inv_)
        per = {"xyz": d_pt, "rotation": d_rot, "rigid_v": gx * age_, "rigid_rotvec": d_rv * age_, "rigid_rotcen": gx - d_pt}
    else:
        z3 = np.zeros_like(gx)
        per = {"xyz": gx, "rotation": h, "rigid_v": z3, "rigid_rotvec": z3, "rigid_rotcen": z3}
    o = out["opacity"][sl]
    per["opacity"] = g["opacity"] * o * (1 - o)
    per["scaling"] = g["scaling"] * out["scaling"][sl]
    per["features_dc"] = g["feature"][:, :1]
    per["features_rest"] = g["feature"][:, 1:]
    for k in SWIN_KEYS[:9]:
        full = np.zeros(np.asarray(imm[k]).shape, np.float64)
        full[ii] = per[k]
        grads[k] = full
    out["grad"] = grads
    return out
