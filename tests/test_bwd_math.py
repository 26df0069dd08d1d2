"""The matrix formulation of the per-Gaussian backward that csrc/preprocess_bwd.cu implements, restated in float64 numpy
and checked against the CPU oracle (oracle/oracle_cpu.c, which follows the reference's scalar expansions,
cuda_rasterizer/backward.cu:20-139, :144-274, :278-341) on random inputs. This pins the DERIVATION; the kernel itself is
checked against the reference CUDA on the GPU box (tests/test_gpu_parity.py).

Formulation (column vectors, world -> view rotation Wv, view-space mean t, J the 2x3 projection Jacobian, M = J Wv):

  cov2D      S = M Sigma M^T + 0.3 I,   conic K = S^-1
  conic      G_K = [[gA, gB], [gB, gC]]          (the tile backward accumulates HALF of d/dB for the off-diagonal parameter,
                                                  backward.cu:517, so gB already is the entry of the symmetric gradient matrix)
  cov2D      G_S = -adj(S) G_K adj(S) / (det^2 + 1e-7)          (dK = -K dS K, with the reference's regularised 1/det^2)
  Sigma      G_Sigma = M^T G_S M                                  (dL_dcov3D: diagonal as is, off-diagonals doubled)
  M          G_M = 2 G_S M Sigma,  G_J = G_M Wv^T                 (2x3 products instead of per-entry expansions)
  t          through J's four non-constant entries; clamped axes pass no gradient
  SH         s_k = <sh_k, g>;  dL/ddir = sum_k s_k grad B_k(dir)  (gradients of the basis polynomials)
  normalise  dL/dv = (g - d <d, g>) / |v|
  Sigma = A A^T, A = R diag(s):  G_A = 2 G_Sigma A;  dL/ds_j = <G_A[:, j], R[:, j]>;  G_R = G_A diag(s)
  quaternion (r, u):  a = axial(G_R - G_R^T);  dL/dr = 2 <u, a>;  dL/du = 2 [(G_R + G_R^T) u - 2 tr(G_R) u + r a]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "3dgs-mcmc_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

import gs_synth  # noqa: E402
from oracle import oracle  # noqa: E402

C0 = 0.28209479177387814
C1 = 0.4886025119029199
C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
      1.445305721320277, -0.5900435899266435]


def basis_gradients(d, deg):
    """grad_d B_k(d) for k = 0..(deg+1)^2-1 as rows [k, 3] (B_k = the real SH basis polynomials in x, y, z)."""
    x, y, z = d
    xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
    g = [(0.0, 0.0, 0.0)]
    if deg > 0:
        g += [(0.0, -C1, 0.0), (0.0, 0.0, C1), (-C1, 0.0, 0.0)]
    if deg > 1:
        g += [(C2[0] * y, C2[0] * x, 0.0), (0.0, C2[1] * z, C2[1] * y), (-2 * C2[2] * x, -2 * C2[2] * y, 4 * C2[2] * z),
              (C2[3] * z, 0.0, C2[3] * x), (2 * C2[4] * x, -2 * C2[4] * y, 0.0)]
    if deg > 2:
        g += [(C3[0] * 6 * xy, C3[0] * 3 * (xx - yy), 0.0),
              (C3[1] * yz, C3[1] * xz, C3[1] * xy),
              (C3[2] * -2 * xy, C3[2] * (4 * zz - xx - 3 * yy), C3[2] * 8 * yz),
              (C3[3] * -6 * xz, C3[3] * -6 * yz, C3[3] * (6 * zz - 3 * xx - 3 * yy)),
              (C3[4] * (4 * zz - 3 * xx - yy), C3[4] * -2 * xy, C3[4] * 8 * xz),
              (C3[5] * 2 * xz, C3[5] * -2 * yz, C3[5] * (xx - yy)),
              (C3[6] * 3 * (xx - yy), C3[6] * -6 * xy, 0.0)]
    return np.array(g, dtype=np.float64)


def rotation(q):
    r, x, y, z = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y)],
                     [2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x)],
                     [2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)]])


def backward_one(mean, scale, q, sh, clamp_bits, deg, view, proj, campos, W, H, tanx, tany, g_mean2d, g_conic, g_rgb, mod=1.0):
    """Matrix-form backward of one visible Gaussian (float64). view / proj: the transposed 4x4s the rasterizer takes."""
    Vm = view.reshape(4, 4).T.astype(np.float64)          # math convention: t = Vm @ [p, 1]
    Pm = proj.reshape(4, 4).T.astype(np.float64)
    Wv = Vm[:3, :3]
    fx, fy = W / (2 * tanx), H / (2 * tany)
    R = rotation(q)
    s = mod * scale
    A = R * s[None, :]
    Sigma = A @ A.T
    t = Vm[:3, :3] @ mean + Vm[:3, 3]
    limx, limy = 1.3 * tanx, 1.3 * tany
    tx_over, ty_over = t[0] / t[2], t[1] / t[2]
    on_x, on_y = float(-limx <= tx_over <= limx), float(-limy <= ty_over <= limy)
    tx, ty, tz = np.clip(tx_over, -limx, limx) * t[2], np.clip(ty_over, -limy, limy) * t[2], t[2]
    J = np.array([[fx / tz, 0, -fx * tx / tz ** 2], [0, fy / tz, -fy * ty / tz ** 2]])
    M = J @ Wv
    S = M @ Sigma @ M.T + 0.3 * np.eye(2)
    det = S[0, 0] * S[1, 1] - S[0, 1] ** 2
    adj = np.array([[S[1, 1], -S[0, 1]], [-S[0, 1], S[0, 0]]])
    G_K = np.array([[g_conic[0], g_conic[1]], [g_conic[1], g_conic[2]]])
    G_S = -(adj @ G_K @ adj) / (det * det + 1e-7)
    G_Sigma = M.T @ G_S @ M
    dcov = np.array([G_Sigma[0, 0], 2 * G_Sigma[0, 1], 2 * G_Sigma[0, 2], G_Sigma[1, 1], 2 * G_Sigma[1, 2], G_Sigma[2, 2]])
    G_M = 2 * G_S @ M @ Sigma
    G_J = G_M @ Wv.T
    iz2, iz3 = 1 / tz ** 2, 1 / tz ** 3
    g_t = np.array([on_x * -fx * iz2 * G_J[0, 2], on_y * -fy * iz2 * G_J[1, 2],
                    -fx * iz2 * G_J[0, 0] - fy * iz2 * G_J[1, 1] + 2 * fx * tx * iz3 * G_J[0, 2] + 2 * fy * ty * iz3 * G_J[1, 2]])
    dmean = Wv.T @ g_t
    # screen-space mean -> 3D mean through the perspective division
    hom = Pm @ np.append(mean, 1.0)
    iw = 1.0 / (hom[3] + 1e-7)
    for axis, gm in ((0, g_mean2d[0]), (1, g_mean2d[1])):
        dmean += gm * (Pm[axis, :3] * iw - Pm[3, :3] * hom[axis] * iw * iw)
    # colour -> SH coefficients and view direction
    g = np.array([0.0 if (clamp_bits >> c) & 1 else g_rgb[c] for c in range(3)])
    v = mean - campos
    n = np.linalg.norm(v)
    d = v / n
    ncoef = (deg + 1) ** 2
    x, y, z = d
    basis = np.zeros(16)
    basis[0] = C0
    if deg > 0:
        basis[1:4] = [-C1 * y, C1 * z, -C1 * x]
    if deg > 1:
        basis[4:9] = [C2[0] * x * y, C2[1] * y * z, C2[2] * (2 * z * z - x * x - y * y), C2[3] * x * z, C2[4] * (x * x - y * y)]
    if deg > 2:
        basis[9:16] = [C3[0] * y * (3 * x * x - y * y), C3[1] * x * y * z, C3[2] * y * (4 * z * z - x * x - y * y),
                       C3[3] * z * (2 * z * z - 3 * x * x - 3 * y * y), C3[4] * x * (4 * z * z - x * x - y * y),
                       C3[5] * z * (x * x - y * y), C3[6] * x * (x * x - 3 * y * y)]
    dsh = np.zeros((sh.shape[0], 3))
    dsh[:ncoef] = basis[:ncoef, None] * g[None, :]
    s_k = sh[:ncoef].astype(np.float64) @ g                    # <sh_k, g>
    g_dir = s_k @ basis_gradients(d, deg)
    dmean += (g_dir - d * (d @ g_dir)) / n
    # Sigma = A A^T -> scale, quaternion
    G_full = np.array([[dcov[0], 0.5 * dcov[1], 0.5 * dcov[2]], [0.5 * dcov[1], dcov[3], 0.5 * dcov[4]],
                       [0.5 * dcov[2], 0.5 * dcov[4], dcov[5]]])
    G_A = 2 * G_full @ A
    dscale = np.einsum("ij,ij->j", G_A, R)      # w.r.t. (mod * scale), as the reference returns it (backward.cu:318-321)
    G_R = G_A * s[None, :]
    r_, u = q[0], q[1:]
    a = np.array([G_R[2, 1] - G_R[1, 2], G_R[0, 2] - G_R[2, 0], G_R[1, 0] - G_R[0, 1]])
    dq = np.concatenate([[2 * (u @ a)], 2 * ((G_R + G_R.T) @ u - 2 * np.trace(G_R) * u + r_ * a)])
    return dmean, dcov, dsh, dscale, dq


def test_matrix_formulation_matches_the_oracle():
    P, W, H = 400, 320, 200
    scene = gs_synth.scene_cube(P, seed=5)
    scene["scales"] = (scene["scales"] * 3).astype(np.float32)
    scene["rotations"] = (scene["rotations"] * np.linspace(0.6, 1.4, P, dtype=np.float32)[:, None]).astype(np.float32)
    cam = gs_synth.orbit_cameras(4, 4.03, W, H, 1.4)[1]          # wide field of view: some means sit beyond the 1.3 tan clamp
    for deg in (3, 1, 0):
        pre = oracle.preprocess(scene["means3D"], scene["scales"], scene["rotations"], scene["opacities"], scene["shs"],
                                cam["viewmatrix"], cam["projmatrix"], cam["campos"], W, H, cam["tanfovx"], cam["tanfovy"], deg, 1.3)
        rng = np.random.default_rng(deg)
        g2d = rng.standard_normal((P, 3)).astype(np.float32)       # the reference's layouts: float3 (z unused) ...
        gcon = rng.standard_normal((P, 4)).astype(np.float32)      # ... and float4 (x = A, y = B, w = C; z unused)
        gcol = rng.standard_normal((P, 3)).astype(np.float32)
        want = oracle.preprocess_bwd(scene["means3D"], pre["radii"], scene["shs"], pre["clamped"], scene["scales"], scene["rotations"],
                                     1.3, pre["cov3D"], cam["viewmatrix"], cam["projmatrix"], cam["campos"], W, H, cam["tanfovx"],
                                     cam["tanfovy"], deg, g2d, gcon, gcol)
        vis = np.nonzero(pre["radii"] > 0)[0]
        assert len(vis) > 100
        got = {k: np.zeros_like(v, dtype=np.float64) for k, v in want.items()}
        clamp = pre["clamped"].reshape(P, -1)
        for i in vis:
            bits = int(clamp[i, 0]) | (int(clamp[i, 1]) << 1) | (int(clamp[i, 2]) << 2) if clamp.shape[1] == 3 else int(clamp[i, 0])
            dm, dc, dsh, dsc, dq = backward_one(scene["means3D"][i].astype(np.float64), scene["scales"][i].astype(np.float64),
                                                scene["rotations"][i].astype(np.float64), scene["shs"][i], bits, deg,
                                                np.asarray(cam["viewmatrix"]), np.asarray(cam["projmatrix"]),
                                                np.asarray(cam["campos"], np.float64), W, H, cam["tanfovx"], cam["tanfovy"],
                                                g2d[i, :2].astype(np.float64), gcon[i, [0, 1, 3]].astype(np.float64), gcol[i].astype(np.float64), mod=1.3)
            got["dL_dmeans3D"][i], got["dL_dcov3D"][i], got["dL_dsh"][i], got["dL_dscales"][i], got["dL_drots"][i] = dm, dc, dsh, dsc, dq
        for k in want:
            num = np.linalg.norm(got[k] - want[k].astype(np.float64))
            den = np.linalg.norm(want[k].astype(np.float64))
            assert den > 0 and num / den < 2e-4, (deg, k, num / den)
