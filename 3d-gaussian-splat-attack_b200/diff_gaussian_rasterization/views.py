"""View-level forward / accumulate-backward API for multi-camera gradient steps (new).

The reference sums a loss over the cameras of a batch (attack.py:481-494): each camera goes through
_RasterizeGaussians, whose backward returns nine freshly zero-filled tensors (rasterize_points.cu:151-159)
that autograd then ADDS to the leaves' .grad, one elementwise kernel per attribute per camera.  Here a view's
backward adds its gradients directly into caller-owned buffers (gsr_backward_accumulate), touching only the
rows of Gaussians that view rendered; with a FlatGradBucket (dist.py) those buffers are the slices of the
all-reduce bucket, so a step is: zero bucket -> for each view {forward_view, loss, backward_view} -> all_reduce.

No autograd graph is built through the rasterizer: the image returned by forward_view is a plain tensor; the
caller differentiates its loss w.r.t. that image (autograd or by hand) and passes dL/dimage to backward_view.
There is no CPU path.
"""
import ctypes as C
from typing import Dict, Optional

import torch

from . import _abi
from . import _f32c, _make_inputs, _ptr, _run_forward, _state

__all__ = ["ViewState", "forward_view", "backward_view", "GRAD_KEYS"]

# gradient buffers backward_view() accepts, named after GaussianRasterizer.forward's arguments
GRAD_KEYS = ("means3D", "means2D", "opacities", "shs", "colors_precomp", "scales", "rotations", "cov3D_precomp")


class ViewState:
    """What one forwarded view keeps until its backward (the role of ctx.save_for_backward,
    DGR/diff_gaussian_rasterization/__init__.py:97)."""
    __slots__ = ("rs", "P", "M", "L", "tensors", "radii", "geom", "point_list", "image")

    def __init__(self, rs, P, M, L, tensors, radii, geom, point_list, image):
        self.rs, self.P, self.M, self.L = rs, P, M, L
        self.tensors, self.radii, self.geom, self.point_list, self.image = tensors, radii, geom, point_list, image

    @property
    def num_rendered(self):
        return self.L


def forward_view(raster_settings, means3D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                 cov3D_precomp=None):
    """Render one view.  Same argument meaning and exclusivity rules as GaussianRasterizer.forward
    (DGR/diff_gaussian_rasterization/__init__.py:187-220).  Returns (color (3,H,W), radii (P,) int32, ViewState)."""
    if (shs is None) == (colors_precomp is None):
        raise Exception('Please provide excatly one of either SHs or precomputed colors!')
    if ((scales is None or rotations is None) and cov3D_precomp is None) or \
            ((scales is not None or rotations is not None) and cov3D_precomp is not None):
        raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
    if means3D.ndimension() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise RuntimeError("means3D must be a CUDA tensor (this rasterizer has no CPU path)")
    P = means3D.size(0)
    with torch.no_grad():
        m = _f32c(means3D, "means3D") if P else means3D
        t = dict(shs=_f32c(shs, "shs"), colors_precomp=_f32c(colors_precomp, "colors_precomp"),
                 opacities=_f32c(opacities, "opacities"), scales=_f32c(scales, "scales"),
                 rotations=_f32c(rotations, "rotations"), cov3D_precomp=_f32c(cov3D_precomp, "cov3D_precomp"))
        M = t["shs"].size(1) if t["shs"] is not None else 0
        color, _, radii, L, geom, point_list, image = _run_forward(
            raster_settings, P, M, 0, False, m, t["shs"], t["colors_precomp"], t["opacities"], t["scales"],
            t["rotations"], t["cov3D_precomp"], None)
    t["means3D"] = m
    return color, radii, ViewState(raster_settings, P, M, L, t, radii, geom, point_list, image)


def backward_view(state: ViewState, dL_dimage: torch.Tensor, grads: Dict[str, Optional[torch.Tensor]]) -> None:
    """ADD the gradients of one forwarded view to `grads` (keys from GRAD_KEYS; float32 CUDA contiguous tensors
    shaped like the matching input; `means2D` (P,3) is optional).  Rows of Gaussians the view did not render are
    left as they are.  Arithmetic per view is that of Rasterizer::backward (rasterizer_impl.cu:340-433)."""
    rs, P, t = state.rs, state.P, state.tensors
    if P == 0:
        return
    H, W = int(rs.image_height), int(rs.image_width)
    dev = t["means3D"].device
    for k in grads:
        if k not in GRAD_KEYS:
            raise Exception(f"unknown gradient buffer '{k}' (expected one of {GRAD_KEYS})")

    def buf(name, need, shape_numel):
        g = grads.get(name)
        if g is None:
            if need:
                raise Exception(f"grads['{name}'] is required for the inputs this view was rendered with")
            return None
        if not g.is_cuda or g.dtype != torch.float32 or not g.is_contiguous() or g.numel() != shape_numel:
            raise RuntimeError(f"grads['{name}'] must be a contiguous float32 CUDA tensor with {shape_numel} elements")
        return g

    g_m3 = buf("means3D", True, P * 3)
    g_m2 = buf("means2D", False, P * 3)
    g_op = buf("opacities", True, P)
    g_sh = buf("shs", t["shs"] is not None, P * state.M * 3)
    g_col = buf("colors_precomp", t["colors_precomp"] is not None, P * 3)
    g_sc = buf("scales", t["scales"] is not None, P * 3)
    g_rot = buf("rotations", t["rotations"] is not None, P * 4)
    g_cov = buf("cov3D_precomp", t["cov3D_precomp"] is not None, P * 6)
    dpix = _f32c(dL_dimage, "dL_dimage")
    if dpix is None or dpix.numel() != 3 * H * W:
        raise RuntimeError(f"dL_dimage must have 3*{H}*{W} elements")
    with torch.cuda.device(dev):
        lib = _abi.lib()
        st = _state(dev)
        stream = torch.cuda.current_stream(dev)
        keep = []
        inp = _make_inputs(rs, P, state.M, 0, t["means3D"], t["shs"], t["colors_precomp"], t["opacities"], t["scales"],
                           t["rotations"], t["cov3D_precomp"], None, keep)
        scratch = st.ensure_scratch(lib.gsr_scratch_bytes(P, 0, W, H))
        _abi.check(lib.gsr_backward_accumulate(
            C.byref(inp), _ptr(state.geom), _ptr(state.radii), state.L, _ptr(state.point_list), _ptr(state.image),
            _ptr(scratch), scratch.numel(), _ptr(dpix), _ptr(g_m2), _ptr(g_op), _ptr(g_col), _ptr(g_m3), _ptr(g_cov),
            _ptr(g_sh), _ptr(g_sc), _ptr(g_rot), C.c_void_p(stream.cuda_stream)), "backward_accumulate")
        if rs.debug:
            stream.synchronize()
