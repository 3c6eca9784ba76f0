"""Deterministic synthetic scenes and cameras for tests and bench (numpy only).

Camera matrices restate the reference's conventions:
  getWorld2View2      /root/reference/utils/graphics_utils.py:38-49
  getProjectionMatrix /root/reference/utils/graphics_utils.py:51-71
  Camera              /root/reference/scene/cameras.py:54-57 (matrices stored transposed, so the
                      contiguous memory is column-major, which is what the kernels index)
Scene distributions follow SURVEY.md section 8(d) (configs C1..C5 of BASELINE.json).
"""
import math
import numpy as np


def look_at_camera(eye, target=(0.0, 0.0, 0.0), up=(0.0, 1.0, 0.0), fovx_deg=60.0, width=800, height=800,
                   znear=0.01, zfar=100.0):
    eye = np.asarray(eye, np.float64)
    target = np.asarray(target, np.float64)
    fwd = target - eye
    fwd /= np.linalg.norm(fwd)
    right = np.cross(np.asarray(up, np.float64), fwd)
    right /= np.linalg.norm(right)
    dwn = np.cross(fwd, right)
    # world->camera rotation rows: camera looks down +z (COLMAP convention used by the reference)
    Rw2c = np.stack([right, dwn, fwd], 0)
    t = -Rw2c @ eye
    w2v = np.eye(4)
    w2v[:3, :3] = Rw2c
    w2v[:3, 3] = t
    fovx = math.radians(fovx_deg)
    fovy = 2.0 * math.atan(math.tan(fovx * 0.5) * height / width)
    tanx, tany = math.tan(fovx * 0.5), math.tan(fovy * 0.5)
    top, right_ = tany * znear, tanx * znear
    Pm = np.zeros((4, 4))
    Pm[0, 0] = 2.0 * znear / (2 * right_)
    Pm[1, 1] = 2.0 * znear / (2 * top)
    Pm[3, 2] = 1.0
    Pm[2, 2] = zfar / (zfar - znear)
    Pm[2, 3] = -(zfar * znear) / (zfar - znear)
    world_view_transform = w2v.T.astype(np.float32)                      # cameras.py:54
    full_proj = (w2v.T @ Pm.T).astype(np.float32)                        # cameras.py:55-56
    campos = np.linalg.inv(w2v.T.astype(np.float32))[3, :3].astype(np.float32)  # cameras.py:57
    return dict(viewmatrix=np.ascontiguousarray(world_view_transform), projmatrix=np.ascontiguousarray(full_proj),
                campos=np.ascontiguousarray(campos), tanfovx=float(tanx), tanfovy=float(tany),
                image_width=int(width), image_height=int(height))


def ring_cameras(n, radius=3.0, height_y=0.0, **kw):
    """n cameras on a ring around the origin (mirrors Camera.yaw, scene/cameras.py:85-105)."""
    cams = []
    for k in range(n):
        a = 2.0 * math.pi * k / n
        cams.append(look_at_camera((radius * math.sin(a), height_y, -radius * math.cos(a)), **kw))
    return cams


_CONFIGS = {
    # name: (P, W, H, xyz mode, xyz scale, log-scale mean, opacity mean, opacity std)
    "tiny": (2000, 160, 120, "uniform", 1.0, math.log(0.03), 0.0, 2.0),
    "small": (20000, 320, 240, "uniform", 1.0, math.log(0.02), 0.0, 2.0),
    "C1": (100_000, 800, 800, "uniform", 1.0, math.log(0.01), 0.0, 2.0),
    "C2": (1_000_000, 1920, 1080, "normal", 0.5, math.log(0.004), 0.0, 2.0),
    "C3": (3_000_000, 1920, 1080, "normal", 0.5, math.log(0.004), 0.0, 2.0),
    "C4": (6_000_000, 3840, 2160, "slab", 1.0, math.log(0.02), 0.0, 2.0),
    "C5": (2_000_000, 1920, 1080, "normal", 0.3, math.log(0.08), 3.0, 0.5),
}


def make_scene(name="C1", seed=None, P=None, width=None, height=None, sh_degree=3):
    cfgP, W, H, mode, xs, lsm, om, osd = _CONFIGS[name]
    P = cfgP if P is None else P
    W = W if width is None else width
    H = H if height is None else height
    seed = (1234 + list(_CONFIGS).index(name)) if seed is None else seed
    rng = np.random.default_rng(seed)
    f32 = np.float32
    if mode == "uniform":
        xyz = rng.uniform(-xs, xs, (P, 3))
    elif mode == "normal":
        xyz = np.clip(rng.normal(0, xs, (P, 3)), -3 * xs, 3 * xs)
    else:  # street-scale slab 40 x 6 x 40
        xyz = rng.uniform(-1, 1, (P, 3)) * np.array([20.0, 3.0, 20.0])
    scales = np.exp(rng.normal(lsm, 0.5, (P, 3)))
    q = rng.normal(0, 1, (P, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    opac = 1.0 / (1.0 + np.exp(-rng.normal(om, osd, (P, 1))))
    M = 16
    shs = np.concatenate([rng.normal(0, 0.9, (P, 1, 3)), rng.normal(0, 0.1, (P, M - 1, 3))], 1)
    if mode == "slab":
        cam = look_at_camera((0.0, 0.0, -19.0), target=(0.0, 0.0, 0.0), width=W, height=H)
    else:
        cam = look_at_camera((0.0, 0.0, -3.0), width=W, height=H)
    dL = rng.normal(0, 1, (3, H, W)) / (3.0 * H * W)
    return dict(name=name, P=P, sh_degree=sh_degree, means3D=xyz.astype(f32), scales=scales.astype(f32),
                rotations=q.astype(f32), opacities=opac.astype(f32), shs=shs.astype(f32),
                bg=np.zeros(3, f32), cam=cam, dL_dpix=dL.astype(f32))
