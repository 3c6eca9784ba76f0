"""Shared test helpers: run the product (through its C ABI), the CPU oracle and -- when built --
the reference CUDA (oracle/_ref/libdgr_ref.so) on the same synthetic scene."""
import ctypes as C
import os

import numpy as np
import torch

from gsr_synth import make_scene  # noqa: F401  (re-exported)

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libdgr_ref.so")


GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
# scene recipes of the committed fixtures (tools/make_golden.py ran the reference CUDA on them)
GOLDEN_SCENES = {
    "tiny_sh3": dict(name="tiny"),
    "tiny_sh1_bg": dict(name="tiny", seed=77, sh_degree=1, width=150, height=100),
    "tiny_precomp": dict(name="tiny", seed=78, width=200, height=136),
}


def golden_case(name):
    """(scene, kwargs for run_gsr/run_oracle, dict of reference-CUDA outputs) of a committed fixture."""
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    scene = make_scene(**GOLDEN_SCENES[name])
    kw = {k[3:]: z[k] for k in z.files if k.startswith("in_")}
    gold = {k: z[k] for k in z.files if not k.startswith(("in_", "grad_"))}
    gold["grads"] = {k[5:]: z[k] for k in z.files if k.startswith("grad_")}
    gold["num_rendered"] = int(gold["num_rendered"])
    return scene, kw, gold


def settings_for(scene, device="cuda", debug=False, prefiltered=False, bg=None):
    from diff_gaussian_rasterization import GaussianRasterizationSettings
    c = scene["cam"]
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(device)
    return GaussianRasterizationSettings(
        image_height=c["image_height"], image_width=c["image_width"], tanfovx=c["tanfovx"], tanfovy=c["tanfovy"],
        bg=t(scene["bg"] if bg is None else np.asarray(bg, np.float32)), scale_modifier=1.0, viewmatrix=t(c["viewmatrix"]),
        projmatrix=t(c["projmatrix"]), sh_degree=scene["sh_degree"], campos=t(c["campos"]), prefiltered=prefiltered,
        debug=debug)


def scene_tensors(scene, device="cuda", requires_grad=True):
    out = {}
    for k in ("means3D", "scales", "rotations", "opacities", "shs"):
        out[k] = torch.from_numpy(scene[k]).to(device).requires_grad_(requires_grad)
    return out


def gsr_fetch(name, dtype, P, W, H, L, geom, image, scratch, point_list=None):
    """Copy an internal array of the product's last forward out (gsr_debug_fetch)."""
    from diff_gaussian_rasterization import _abi
    lib = _abi.lib()
    p = lambda t: C.c_void_p(t.data_ptr()) if t is not None and t.numel() else None
    n = lib.gsr_debug_fetch(name.encode(), P, W, H, L, p(geom), p(image), p(scratch), p(point_list), None, 0, None)
    assert n >= 0, _abi.last_error()
    buf = torch.empty(max(int(n), 1), dtype=torch.uint8, device="cuda")
    r = lib.gsr_debug_fetch(name.encode(), P, W, H, L, p(geom), p(image), p(scratch), p(point_list), p(buf), n, None)
    assert r >= 0, _abi.last_error()
    torch.cuda.synchronize()
    return buf[:n].cpu().numpy().view(dtype)


def run_gsr(scene, grad=True, fetch=True, obj=None, bg=None, colors_precomp=None, cov3D_precomp=None, dL=None):
    """Forward (+backward) through the public Python API; returns dict of numpy arrays."""
    import diff_gaussian_rasterization as dgr
    rs = settings_for(scene, bg=bg)
    t = scene_tensors(scene, requires_grad=grad or fetch)  # a grad_fn is needed to reach the saved state
    P = scene["P"]
    means2D = torch.zeros((P, 3), device="cuda", requires_grad=grad)
    kw = {}
    extra = {}
    if colors_precomp is not None:
        extra["colors"] = torch.from_numpy(colors_precomp).cuda().requires_grad_(grad)
        kw["colors_precomp"] = extra["colors"]
    else:
        kw["shs"] = t["shs"]
    if cov3D_precomp is not None:
        extra["cov3D"] = torch.from_numpy(cov3D_precomp).cuda().requires_grad_(grad)
        kw["cov3D_precomp"] = extra["cov3D"]
    else:
        kw["scales"], kw["rotations"] = t["scales"], t["rotations"]
    if obj is not None:
        extra["obj"] = torch.from_numpy(obj).cuda().requires_grad_(grad)
        kw["sh_objs"] = extra["obj"]
    outs = dgr.GaussianRasterizer(rs)(means3D=t["means3D"], means2D=means2D, opacities=t["opacities"], **kw)
    res = dict(color=outs[0].detach().cpu().numpy(), radii=outs[1].cpu().numpy())
    if obj is not None:
        res["objects"] = outs[2].detach().cpu().numpy()
    fn = outs[0].grad_fn
    if fn is not None:
        res["num_rendered"] = fn.num_rendered
    if fetch and P:
        W, H = scene["cam"]["image_width"], scene["cam"]["image_height"]
        L = fn.num_rendered
        saved = fn.saved_tensors
        geom, point_list, image = saved[9], saved[10], saved[11]
        scratch = dgr._state(torch.device("cuda", torch.cuda.current_device())).scratch
        res["num_rendered"] = L
        for name, dt in (("depths", np.float32), ("means2D", np.float32), ("cov3D", np.float32),
                         ("conic_opacity", np.float32), ("rgb", np.float32), ("clamped", np.uint8),
                         ("tiles_touched", np.uint32), ("point_offsets", np.uint32), ("final_T", np.float32),
                         ("n_contrib", np.uint32), ("ranges", np.uint32), ("keys_sorted", np.uint64)):
            res[name] = gsr_fetch(name, dt, P, W, H, L, geom, image, scratch, point_list)
        res["rgb"] = res["rgb"].reshape(P, 4)[:, :3].copy()
        res["point_list"] = point_list[:L].cpu().numpy().view(np.uint32)
    if grad:
        dL_t = torch.from_numpy(scene["dL_dpix"] if dL is None else dL).cuda()
        loss = (outs[0] * dL_t).sum()
        if obj is not None and "dL_dobj" in scene:
            loss = loss + (outs[2] * torch.from_numpy(scene["dL_dobj"]).cuda()).sum()
        loss.backward()
        g = dict(means3D=t["means3D"].grad, means2D=means2D.grad, opacity=t["opacities"].grad)
        if colors_precomp is None:
            g["sh"] = t["shs"].grad
        else:
            g["colors"] = extra["colors"].grad
        if cov3D_precomp is None:
            g["scales"], g["rotations"] = t["scales"].grad, t["rotations"].grad
        else:
            g["cov3D"] = extra["cov3D"].grad
        if obj is not None:
            g["obj"] = extra["obj"].grad
        res["grads"] = {k: v.detach().cpu().numpy() for k, v in g.items() if v is not None}
    return res


def run_oracle(scene, grad=True, obj=None, bg=None, colors_precomp=None, cov3D_precomp=None, dL=None):
    from oracle.cpu_ref import Oracle
    c = scene["cam"]
    o = Oracle()
    outs = o.forward(scene["bg"] if bg is None else np.asarray(bg, np.float32), scene["means3D"], scene["opacities"],
                     c["viewmatrix"], c["projmatrix"], c["campos"], c["image_width"], c["image_height"], c["tanfovx"],
                     c["tanfovy"], sh_degree=scene["sh_degree"], shs=None if colors_precomp is not None else scene["shs"],
                     colors_precomp=colors_precomp, scales=None if cov3D_precomp is not None else scene["scales"],
                     rotations=None if cov3D_precomp is not None else scene["rotations"], cov3D_precomp=cov3D_precomp,
                     obj_feats=obj)
    res = dict(color=outs[0], radii=outs[1], num_rendered=o.num_rendered)
    if obj is not None:
        res["objects"] = outs[2]
    for name in ("depths", "means2D", "cov3D", "conic_opacity", "rgb", "clamped", "tiles_touched", "point_offsets",
                 "final_T", "n_contrib", "ranges", "keys_sorted", "point_list"):
        res[name] = o.get(name)
    res["n_pairs"] = int(o.get("n_pairs")[0])
    if grad:
        g = o.backward(scene["dL_dpix"] if dL is None else dL, scene.get("dL_dobj") if obj is not None else None)
        res["grads"] = g
    res["t_stage"] = o.get("t_stage")
    return res


# ---------------------------------------------------------------- reference CUDA (oracle/_ref)
_ref = None


def ref_available():
    return os.path.exists(REF_SO) and torch.cuda.is_available()


def ref_lib():
    global _ref
    if _ref is None:
        _ref = C.CDLL(REF_SO)
        _ref.ref_fetch.restype = C.c_longlong
    return _ref


def run_ref(scene, grad=True, bg=None, colors_precomp=None, cov3D_precomp=None, dL=None, fetch=True):
    """Reference CUDA rasterizer on the same inputs (default stream)."""
    L = ref_lib()
    c = scene["cam"]
    P, W, H = scene["P"], c["image_width"], c["image_height"]
    cu = lambda a: None if a is None else torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).cuda()
    p = lambda t: None if t is None else C.c_void_p(t.data_ptr())
    T = dict(bg=cu(scene["bg"] if bg is None else np.asarray(bg, np.float32)), means3D=cu(scene["means3D"]),
             shs=None if colors_precomp is not None else cu(scene["shs"]), colors=cu(colors_precomp),
             opac=cu(scene["opacities"]), scales=None if cov3D_precomp is not None else cu(scene["scales"]),
             rot=None if cov3D_precomp is not None else cu(scene["rotations"]), cov=cu(cov3D_precomp),
             view=cu(c["viewmatrix"]), proj=cu(c["projmatrix"]), campos=cu(c["campos"]))
    M = 0 if T["shs"] is None else T["shs"].shape[1]
    out_color = torch.zeros((3, H, W), device="cuda")
    radii = torch.zeros((P,), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    R = L.ref_forward(P, scene["sh_degree"], M, p(T["bg"]), W, H, p(T["means3D"]), p(T["shs"]), p(T["colors"]),
                      p(T["opac"]), p(T["scales"]), C.c_float(1.0), p(T["rot"]), p(T["cov"]), p(T["view"]), p(T["proj"]),
                      p(T["campos"]), C.c_float(c["tanfovx"]), C.c_float(c["tanfovy"]), 0, p(out_color), p(radii), 0)
    torch.cuda.synchronize()
    assert R >= 0
    res = dict(color=out_color.cpu().numpy(), radii=radii.cpu().numpy(), num_rendered=R)
    if fetch:
        for name, dt in (("depths", np.float32), ("means2D", np.float32), ("cov3D", np.float32),
                         ("conic_opacity", np.float32), ("rgb", np.float32), ("clamped", np.uint8),
                         ("tiles_touched", np.uint32), ("point_offsets", np.uint32), ("final_T", np.float32),
                         ("n_contrib", np.uint32), ("ranges", np.uint32), ("keys_sorted", np.uint64),
                         ("point_list", np.uint32)):
            n = L.ref_fetch(name.encode(), None, C.c_longlong(0))
            buf = torch.empty(max(int(n), 1), dtype=torch.uint8, device="cuda")
            L.ref_fetch(name.encode(), p(buf), C.c_longlong(n))
            torch.cuda.synchronize()
            res[name] = buf[:n].cpu().numpy().view(dt)
    if grad:
        z = lambda *s: torch.zeros(s, device="cuda")
        g = dict(means2D=z(P, 3), conic=z(P, 4), opacity=z(P, 1), colors=z(P, 3), means3D=z(P, 3), cov3D=z(P, 6),
                 sh=z(P, max(M, 1), 3), scales=z(P, 3), rotations=z(P, 4))
        dLt = cu(scene["dL_dpix"] if dL is None else dL)
        r = L.ref_backward(P, scene["sh_degree"], M, R, p(T["bg"]), W, H, p(T["means3D"]), p(T["shs"]), p(T["colors"]),
                           p(T["scales"]), C.c_float(1.0), p(T["rot"]), p(T["cov"]), p(T["view"]), p(T["proj"]),
                           p(T["campos"]), C.c_float(c["tanfovx"]), C.c_float(c["tanfovy"]), p(radii), p(dLt),
                           p(g["means2D"]), p(g["conic"]), p(g["opacity"]), p(g["colors"]), p(g["means3D"]),
                           p(g["cov3D"]), p(g["sh"]), p(g["scales"]), p(g["rotations"]), 0)
        torch.cuda.synchronize()
        assert r == 0
        res["grads"] = {k: v.cpu().numpy() for k, v in g.items()}
    return res


def grad_close(a, b, rtol=1e-3, atol_frac=1e-4):
    """Gradient tolerance: |a-b| <= rtol*|b| + atol_frac*max|b| elementwise. Returns (ok, worst excess ratio)."""
    a = np.asarray(a, np.float64).reshape(-1)
    b = np.asarray(b, np.float64).reshape(-1)
    scale = np.abs(b).max() if b.size else 0.0
    tol = rtol * np.abs(b) + atol_frac * scale + 1e-30
    err = np.abs(a - b)
    worst = float((err / tol).max()) if b.size else 0.0
    return worst <= 1.0, worst
