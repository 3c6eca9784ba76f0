"""Loads the reference's OWN Python caller of the rasterizer -- unmodified, byte-compiled by
`make -C oracle refpy` from /root/reference into oracle/_ref/refpy/ (sourceless byte code in *.pycode files, git-ignored, travels to
the GPU box like oracle/_ref/libdgr_ref.so):

  gaussian_renderer.render      /root/reference/gaussian_renderer/__init__.py:18-103
  scene.gaussian_model.GaussianModel   /root/reference/scene/gaussian_model.py:22-121 (raw parameters + activations)
  scene.cameras.Camera          /root/reference/scene/cameras.py:17-57 (transposed, non-contiguous CUDA matrices)
  arguments.PipelineParams      /root/reference/arguments/__init__.py:64-69

`import diff_gaussian_rasterization` inside that code resolves to THIS repo's package (tests/conftest.py puts it
on sys.path).  plyfile and simple_knn (scene/gaussian_model.py:15,17) are absent from the image and unused on this
path: they are stubbed in sys.modules.  scene/__init__.py (dataset readers, PIL) is not taken; `scene` is a bare
namespace over the compiled files.  Test infrastructure only.
"""
import importlib
import importlib.abc
import importlib.machinery
import importlib.util
import math
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFPY = os.path.join(ROOT, "oracle", "_ref", "refpy")

_loaded = None


_TOP = ("gaussian_renderer", "scene", "utils", "arguments")


def available():
    return os.path.exists(os.path.join(REFPY, "gaussian_renderer", "__init__.pycode"))


class _RefFinder(importlib.abc.MetaPathFinder):
    """Resolves the reference's module names to the byte-compiled files under oracle/_ref/refpy (*.pycode)."""

    def find_spec(self, name, path=None, target=None):
        parts = name.split(".")
        if parts[0] not in _TOP:
            return None
        base = os.path.join(REFPY, *parts)
        pkg, mod = os.path.join(base, "__init__.pycode"), base + ".pycode"
        if os.path.exists(pkg):
            return importlib.util.spec_from_file_location(
                name, pkg, loader=importlib.machinery.SourcelessFileLoader(name, pkg), submodule_search_locations=[base])
        if os.path.exists(mod):
            return importlib.util.spec_from_file_location(name, mod,
                                                          loader=importlib.machinery.SourcelessFileLoader(name, mod))
        return None


def load():
    """Returns a namespace with render, GaussianModel, Camera, PipelineParams, eval_sh of the reference."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError("oracle/_ref/refpy is missing: run `make -C oracle refpy` where /root/reference exists")
    if not any(isinstance(f, _RefFinder) for f in sys.meta_path):
        sys.meta_path.insert(0, _RefFinder())
    ply = types.ModuleType("plyfile")
    ply.PlyData = ply.PlyElement = object
    sys.modules.setdefault("plyfile", ply)
    knn = types.ModuleType("simple_knn")
    knn_c = types.ModuleType("simple_knn._C")
    knn_c.distCUDA2 = lambda *a, **k: (_ for _ in ()).throw(RuntimeError("simple_knn is stubbed"))
    knn._C = knn_c
    sys.modules.setdefault("simple_knn", knn)
    sys.modules.setdefault("simple_knn._C", knn_c)
    scene = types.ModuleType("scene")
    scene.__path__ = [os.path.join(REFPY, "scene")]
    sys.modules.setdefault("scene", scene)
    ns = types.SimpleNamespace()
    ns.render = importlib.import_module("gaussian_renderer").render
    ns.GaussianModel = importlib.import_module("scene.gaussian_model").GaussianModel
    ns.Camera = importlib.import_module("scene.cameras").Camera
    ns.eval_sh = importlib.import_module("utils.sh_utils").eval_sh
    ns.PipelineParams = importlib.import_module("arguments").PipelineParams
    import diff_gaussian_rasterization as dgr
    assert importlib.import_module("gaussian_renderer").GaussianRasterizer is dgr.GaussianRasterizer
    _loaded = ns
    return ns


def pipe(convert_SHs_python=False, compute_cov3D_python=False, debug=False):
    """The `pipe` argument of render(): arguments.PipelineParams fields (arguments/__init__.py:66-68)."""
    return types.SimpleNamespace(convert_SHs_python=convert_SHs_python, compute_cov3D_python=compute_cov3D_python,
                                 debug=debug)


def inverse_activated_model(ns, scene, obj=None, device="cuda"):
    """A reference GaussianModel whose raw parameters reproduce the synthetic scene's activated attributes
    (inverse of scene/gaussian_model.py:30-40: log scale, logit opacity, unnormalised quaternion x 1.7)."""
    P = scene["P"]
    gm = ns.GaussianModel(3)
    gm.active_sh_degree = scene["sh_degree"]
    t = lambda a: torch.nn.Parameter(torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).to(device))
    op = np.clip(scene["opacities"].astype(np.float64), 1e-6, 1 - 1e-6)
    gm._xyz = t(scene["means3D"])
    gm._features_dc = t(scene["shs"][:, :1, :])
    gm._features_rest = t(scene["shs"][:, 1:, :])
    gm._scaling = t(np.log(scene["scales"].astype(np.float64)))
    gm._rotation = t(scene["rotations"] * 1.7)
    gm._opacity = t(np.log(op / (1 - op)))
    if obj is None:
        obj = np.random.default_rng(5).uniform(0, 1, (P, 1, 16)).astype(np.float32)
    gm._objects_dc = t(obj)
    return gm


def camera_from(ns, cam_dict, eye=(0.0, 0.0, -3.0), target=(0.0, 0.0, 0.0), up=(0.0, 1.0, 0.0), uid=0):
    """A reference Camera (scene/cameras.py:17-57) looking from `eye` at `target` with the FoV / size of a
    gsr_synth camera.  R is camera-to-world (getWorld2View2 transposes it), T the world-to-camera translation."""
    eye = np.asarray(eye, np.float64)
    fwd = np.asarray(target, np.float64) - eye
    fwd /= np.linalg.norm(fwd)
    right = np.cross(np.asarray(up, np.float64), fwd)
    right /= np.linalg.norm(right)
    dwn = np.cross(fwd, right)
    Rw2c = np.stack([right, dwn, fwd], 0)
    T = -Rw2c @ eye
    W, H = cam_dict["image_width"], cam_dict["image_height"]
    fovx, fovy = 2 * math.atan(cam_dict["tanfovx"]), 2 * math.atan(cam_dict["tanfovy"])
    return ns.Camera(colmap_id=uid, R=Rw2c.T.copy(), T=T, FoVx=fovx, FoVy=fovy, image=torch.zeros(3, H, W),
                     gt_alpha_mask=None, image_name=f"synthetic_{uid}", uid=uid)


def camera_arrays(cam):
    """The Camera's matrices as the contiguous float32 arrays the reference's _C call would read
    (tensor.contiguous().data_ptr() of world_view_transform / full_proj_transform / camera_center)."""
    f = lambda x: np.ascontiguousarray(x.detach().float().cpu().numpy())
    return dict(viewmatrix=f(cam.world_view_transform), projmatrix=f(cam.full_proj_transform),
                campos=f(cam.camera_center), tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5),
                image_width=int(cam.image_width), image_height=int(cam.image_height))
