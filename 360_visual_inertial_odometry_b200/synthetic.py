"""Seeded synthetic inputs for the two hot paths.

The same bytes are fed to the CPU oracle and to the CUDA path.  Nothing here is on a
timed path; it only manufactures inputs (BA windows, ERP frames, track points).
"""
import numpy as np

# config/default_config.yaml:57-61 (extrinsics.T_BC) of the reference
T_BC_DEFAULT = np.array(
    [[-0.0013741, -0.99974421, -0.02257504, 0.01065397],
     [-0.02183404, -0.02253969, 0.9995075, 0.00614827],
     [-0.99976066, 0.00186632, -0.02179749, 0.01690583],
     [0.0, 0.0, 0.0, 1.0]], dtype=np.float64)


def _rot_y(a):
    c, s = np.cos(a), np.sin(a)
    R = np.zeros(a.shape + (3, 3))
    R[..., 0, 0] = c; R[..., 0, 2] = s; R[..., 1, 1] = 1.0; R[..., 2, 0] = -s; R[..., 2, 2] = c
    return R


def _so3_exp(w):
    th = np.linalg.norm(w, axis=-1)[..., None, None]
    ths = np.where(th < 1e-12, 1.0, th)
    K = np.zeros(w.shape[:-1] + (3, 3))
    K[..., 0, 1] = -w[..., 2]; K[..., 0, 2] = w[..., 1]
    K[..., 1, 0] = w[..., 2]; K[..., 1, 2] = -w[..., 0]
    K[..., 2, 0] = -w[..., 1]; K[..., 2, 1] = w[..., 0]
    K = K / ths
    return np.eye(3) + np.sin(th) * K + (1.0 - np.cos(th)) * (K @ K)


def erp_project(Pc, cols, rows):
    """ERP projection used by the factors (Factors.cpp:402-406)."""
    x, y, z = Pc[..., 0], Pc[..., 1], Pc[..., 2]
    L = np.sqrt(x * x + y * y + z * z)
    theta = np.arctan2(x, z)
    phi = -np.arcsin(y / L)
    return cols * (0.5 + theta / (2.0 * np.pi)), rows * (0.5 - phi / np.pi)


def make_ba_window(K=10, M=2000, degree=None, seed=20000, cols=960, rows=480,
                   loop_radius=1.0, yaw_spread_deg=60.0, shell=(2.0, 10.0),
                   px_sigma=1.0, gross_frac=0.01, fixed_point_frac=0.10,
                   flagged_outlier_frac=0.002, boundary_margin=20.0, mid_iteration=False,
                   sort_by_landmark=True):
    """Local-window BA problem (BASELINE configs 3 and 5).

    degree=None: every landmark is seen by all K poses (C3).  Otherwise landmark j is
    seen by `degree` consecutive poses starting at a start pose that increases with j
    (co-visibility band, C5).  Poses/points are rounded through fp32 exactly as the
    reference's Frame/MapPoint storage does (Optimizer.cpp:777,784-785)."""
    rng = np.random.default_rng(seed)
    # body poses on an arc of radius loop_radius, looking roughly tangentially
    if degree is None:
        ang = np.deg2rad(yaw_spread_deg) * (np.arange(K) / max(K - 1, 1) - 0.5)
    else:
        ang = 2.0 * np.pi * np.arange(K) / K
    centers = loop_radius * np.stack([np.cos(ang), np.zeros(K), np.sin(ang)], axis=1)
    R_wb_true = _rot_y(-ang) @ _so3_exp(rng.normal(0, 0.02, (K, 3)))
    T_bc = T_BC_DEFAULT.astype(np.float32)
    T_cb = np.linalg.inv(T_bc.astype(np.float32)).astype(np.float32).astype(np.float64)
    # landmarks in a shell around the trajectory
    if degree is None:
        starts = np.zeros(M, dtype=np.int64); deg = K
        base = np.zeros((M, 3))
    else:
        deg = int(degree)
        starts = np.minimum((np.arange(M) * (K - deg + 1)) // M, K - deg)
        mid = (starts + deg // 2) % K
        base = centers[mid]
    d = rng.uniform(shell[0], shell[1], M)
    dirs = rng.normal(size=(M, 3)); dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    # keep landmarks off the poles of most views (ERP pole = +-Y axis)
    dirs[:, 1] *= 0.5; dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
    P_true = base + d[:, None] * dirs
    # observation list: landmark-major
    point_idx = np.repeat(np.arange(M), deg)
    pose_idx = (starts[:, None] + np.arange(deg)[None, :]).reshape(-1)
    R_bw = np.transpose(R_wb_true, (0, 2, 1))
    Pb = np.einsum("oij,oj->oi", R_bw[pose_idx], P_true[point_idx] - centers[pose_idx])
    Pc = Pb @ T_cb[:3, :3].T + T_cb[:3, 3]
    u, v = erp_project(Pc, float(cols), float(rows))
    u = u + rng.normal(0, px_sigma, u.shape); v = v + rng.normal(0, px_sigma, v.shape)
    gross = rng.random(u.shape) < gross_frac
    u = u + gross * rng.normal(0, 30.0, u.shape); v = v + gross * rng.normal(0, 30.0, v.shape)
    # pixel coordinates are fp32 in Feature (Optimizer.cpp:826)
    u = u.astype(np.float32).astype(np.float64); v = v.astype(np.float32).astype(np.float64)
    keep = ~((u < boundary_margin) | (u > cols - boundary_margin) |
             (v < boundary_margin) | (v > rows - boundary_margin))   # Optimizer.cpp:816
    pose_idx, point_idx, u, v = pose_idx[keep], point_idx[keep], u[keep], v[keep]
    O = int(u.shape[0])
    # initial estimates: truth o exp(noise), rounded through fp32
    dR = _so3_exp(rng.normal(0, np.deg2rad(0.5), (K, 3)))
    dt = rng.normal(0, 0.02, (K, 3))
    Twb = np.tile(np.eye(4), (K, 1, 1))
    Twb[:, :3, :3] = R_wb_true @ dR
    Twb[:, :3, 3] = centers + np.einsum("kij,kj->ki", R_wb_true, dt)
    Twb = Twb.astype(np.float32).astype(np.float64)
    points = (P_true + rng.normal(0, 0.05, (M, 3))).astype(np.float32).astype(np.float64)
    pose_fixed = np.zeros(K, np.uint8); pose_fixed[0] = 1            # Optimizer.cpp:854-858
    point_fixed = (rng.random(M) < fixed_point_frac).astype(np.uint8)
    outlier = (rng.random(O) < flagged_outlier_frac).astype(np.uint8)
    pose_delta = np.zeros((K, 6))
    if mid_iteration:
        pose_delta = rng.normal(0, 1e-2, (K, 6))
        pose_delta[0] = 0.0
    if not sort_by_landmark:
        perm = rng.permutation(O)
        pose_idx, point_idx, u, v, outlier = (pose_idx[perm], point_idx[perm], u[perm],
                                              v[perm], outlier[perm])
    return dict(K=K, M=M, O=O, pose_idx=pose_idx.astype(np.int32),
                point_idx=point_idx.astype(np.int32), obs_u=np.ascontiguousarray(u),
                obs_v=np.ascontiguousarray(v), outlier=outlier,
                Twb_init=np.ascontiguousarray(Twb), Tcb=np.ascontiguousarray(T_cb),
                cols=float(cols), rows=float(rows), info=np.eye(2).reshape(-1) / 1.0,
                huber_a=1.0, pose_fixed=pose_fixed, point_fixed=point_fixed,
                pose_delta=np.ascontiguousarray(pose_delta), points=np.ascontiguousarray(points))


# ----------------------------------------------------------------------------- ERP frames
def pixel_to_bearing(u, v, W, H):
    """Camera::PixelToBearing (Camera.cpp:22-47) in float64."""
    lon = (u / W - 0.5) * 2.0 * np.pi
    lat = -(v / H - 0.5) * np.pi
    cl = np.cos(lat)
    return np.stack([cl * np.sin(lon), -np.sin(lat), cl * np.cos(lon)], axis=-1)


def sphere_texture_params(seed, n_waves=256, sigma_omega=60.0):
    rng = np.random.default_rng(seed)
    omega = rng.normal(0, sigma_omega, (n_waves, 3))
    amp = rng.uniform(0.5, 1.0, n_waves)
    psi = rng.uniform(0, 2 * np.pi, n_waves)
    return omega, amp, psi


def camera_pose(t):
    """Known trajectory of SURVEY.md section 8-D: small circle, slow yaw, pitch wobble."""
    c = 0.25 * np.array([np.cos(2 * np.pi * t / 200.0), 0.0, np.sin(2 * np.pi * t / 200.0)])
    yaw = np.deg2rad(0.3 * t)
    pitch = np.deg2rad(0.1) * np.sin(2 * np.pi * t / 50.0)
    cy, sy, cp, sp = np.cos(yaw), np.sin(yaw), np.cos(pitch), np.sin(pitch)
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rx = np.array([[1, 0, 0], [0, cp, -sp], [0, sp, cp]])
    return Ry @ Rx, c


def render_erp_frame(W, H, t, seed=360, n_waves=256, sigma_omega=60.0, rho=5.0, rows_per_chunk=64,
                     xp=np, device=None):
    """Textured-sphere ERP frame t (uint8 HxW).  xp=np renders on the host; passing the
    torch module renders on `device` (used by bench.py to manufacture inputs quickly)."""
    omega, amp, psi = sphere_texture_params(seed, n_waves, sigma_omega)
    R_wc, c = camera_pose(float(t))
    norm = float(np.sum(amp))
    if xp is np:
        out = np.empty((H, W), np.uint8)
        uu = np.arange(W, dtype=np.float64)
        for r0 in range(0, H, rows_per_chunk):
            vv = np.arange(r0, min(H, r0 + rows_per_chunk), dtype=np.float64)
            U, V = np.meshgrid(uu, vv)
            b = pixel_to_bearing(U, V, float(W), float(H)) @ R_wc.T
            bc = b @ c
            s = -bc + np.sqrt(bc * bc - (c @ c - rho * rho))
            X = (c + s[..., None] * b) / rho
            f = np.cos(X @ omega.T + psi) @ amp
            out[r0:r0 + vv.shape[0]] = np.clip(127.5 + 127.5 * 10.0 * f / norm, 0, 255).astype(np.uint8)
        return out
    torch = xp
    dt = torch.float32
    om = torch.tensor(omega, dtype=dt, device=device); am = torch.tensor(amp, dtype=dt, device=device)
    ps = torch.tensor(psi, dtype=dt, device=device)
    Rt = torch.tensor(R_wc.T, dtype=dt, device=device); ct = torch.tensor(c, dtype=dt, device=device)
    out = torch.empty((H, W), dtype=torch.uint8, device=device)
    uu = torch.arange(W, dtype=dt, device=device)
    for r0 in range(0, H, rows_per_chunk):
        vv = torch.arange(r0, min(H, r0 + rows_per_chunk), dtype=dt, device=device)
        V, U = torch.meshgrid(vv, uu, indexing="ij")
        lon = (U / W - 0.5) * (2.0 * np.pi); lat = -(V / H - 0.5) * np.pi
        cl = torch.cos(lat)
        b = torch.stack([cl * torch.sin(lon), -torch.sin(lat), cl * torch.cos(lon)], dim=-1) @ Rt
        bc = b @ ct
        s = -bc + torch.sqrt(bc * bc - (float(c @ c) - rho * rho))
        X = (ct + s[..., None] * b) / rho
        f = torch.cos(X @ om.T + ps) @ am
        out[r0:r0 + vv.shape[0]] = torch.clamp(127.5 + 127.5 * 10.0 * f / norm, 0, 255).to(torch.uint8)
    return out


def make_track_points(W, H, n, seed, seam_frac=0.05, margin=20):
    """Track points U over the valid band, 1/8-px grid, with a fixed
    fraction within 104 px of the left/right seam so the border branches are hit."""
    rng = np.random.default_rng(seed)
    x = rng.uniform(margin, W - margin, n)
    y = rng.uniform(0.15 * H, 0.85 * H, n)
    ns = int(n * seam_frac)
    if ns > 0:
        side = rng.random(ns) < 0.5
        x[:ns] = np.where(side, rng.uniform(margin, 104, ns), rng.uniform(W - 104, W - margin, ns))
    pts = np.stack([x, y], axis=1)
    return (np.round(pts * 8.0) / 8.0).astype(np.float32)


def make_rotation_tracks(W, H, n, seed, angle_deg=1.2, noise_px=0.3, outlier_frac=0.2):
    """Tracked pairs under a pure camera rotation (what RejectOutliersRotationRANSAC models,
    FeatureTracker.cpp:253-328) + pixel noise + gross outliers.  returns prev, cur (float32), is_outlier."""
    rng = np.random.default_rng(seed)
    prev = np.stack([rng.uniform(20, W - 20, n), rng.uniform(0.15 * H, 0.85 * H, n)], 1).astype(np.float32)
    p = prev.astype(np.float64)
    lon = (p[:, 0] / W - 0.5) * 2 * np.pi
    lat = -(p[:, 1] / H - 0.5) * np.pi
    b = np.stack([np.cos(lat) * np.sin(lon), -np.sin(lat), np.cos(lat) * np.cos(lon)], 1)
    ax = np.array([0.2, 0.9, 0.1]); ax /= np.linalg.norm(ax)
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    a = np.deg2rad(angle_deg)
    R = np.eye(3) + np.sin(a) * K + (1 - np.cos(a)) * K @ K
    rb = b @ R.T
    th = np.arctan2(rb[:, 0], rb[:, 2]); ph = -np.arcsin(rb[:, 1] / np.linalg.norm(rb, axis=1))
    cur = np.stack([W * (0.5 + th / (2 * np.pi)), H * (0.5 - ph / np.pi)], 1) + rng.normal(0, noise_px, (n, 2))
    out = rng.random(n) < outlier_frac
    cur[out] += rng.normal(0, 40, (int(out.sum()), 2))
    return prev, cur.astype(np.float32), out


def draw_ransac_samples(n, iters, seed):
    """iters x 3 distinct indices, drawn like FeatureTracker.cpp:283-289 (uniform, redraw on repeats)."""
    rng = np.random.default_rng(seed)
    out = np.zeros((iters, 3), np.int32)
    for it in range(iters):
        idx = []
        while len(idx) < 3:
            k = int(rng.integers(0, n))
            if k not in idx:
                idx.append(k)
        out[it] = idx
    return out


def make_triangulation_pairs(n, seed, baseline=0.5):
    """Two camera poses `baseline` apart looking at n points 2-10 m away: unit bearings in both camera
    frames and the world-to-camera transforms (row-major 4x4, float32), plus the true points."""
    rng = np.random.default_rng(seed)
    def rot(axis, ang):
        axis = np.asarray(axis, np.float64) / np.linalg.norm(axis)
        K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
        return np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * K @ K
    R1, t1 = rot([0.1, 1.0, 0.2], 0.3), np.array([0.2, -0.1, 0.3])
    R2, t2 = rot([0.3, 0.9, -0.1], 0.45), np.array([0.2 - baseline, 0.05, 0.25])
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    P = d * rng.uniform(2.0, 10.0, (n, 1))
    def to_cam(R, t): return P @ R.T + t
    c1, c2 = to_cam(R1, t1), to_cam(R2, t2)
    b1 = (c1 / np.linalg.norm(c1, axis=1, keepdims=True)).astype(np.float32)
    b2 = (c2 / np.linalg.norm(c2, axis=1, keepdims=True)).astype(np.float32)
    T1 = np.eye(4); T1[:3, :3] = R1; T1[:3, 3] = t1
    T2 = np.eye(4); T2[:3, :3] = R2; T2[:3, 3] = t2
    return b1, b2, T1.astype(np.float32), T2.astype(np.float32), P
