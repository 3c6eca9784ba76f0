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

__all__ = ["ViewState", "forward_view", "forward_views", "backward_view", "backward_views", "ViewPipeline", "GRAD_KEYS"]

# gradient buffers backward_view() accepts, named after GaussianRasterizer.forward's arguments
GRAD_KEYS = ("means3D", "means2D", "opacities", "shs", "shs_rest", "colors_precomp", "scales", "rotations", "cov3D_precomp")


class ViewState:
    """What one forwarded view keeps until its backward (the role of ctx.save_for_backward,
    DGR/diff_gaussian_rasterization/__init__.py:97)."""
    __slots__ = ("rs", "P", "M", "L", "tensors", "radii", "geom", "point_list", "image", "raw_params")

    def __init__(self, rs, P, M, L, tensors, radii, geom, point_list, image, raw_params=False):
        self.rs, self.P, self.M, self.L, self.raw_params = rs, P, M, L, raw_params
        self.tensors, self.radii, self.geom, self.point_list, self.image = tensors, radii, geom, point_list, image

    @property
    def num_rendered(self):
        return self.L


def forward_view(raster_settings, means3D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                 cov3D_precomp=None, shs_rest=None, raw_params=False):
    """Render one view.  Same argument meaning and exclusivity rules as GaussianRasterizer.forward
    (DGR/diff_gaussian_rasterization/__init__.py:187-220).  Returns (color (3,H,W), radii (P,) int32, ViewState).

    Fused model activations (render() materialises them per call, gaussian_renderer/__init__.py:53-81):
      shs_rest     pass the model's `_features_dc` (P,1,3) as `shs` and `_features_rest` (P,M-1,3) here: no torch.cat
      raw_params   `opacities`, `scales`, `rotations` are the RAW `_opacity`, `_scaling`, `_rotation`; sigmoid / exp /
                   normalize (scene/gaussian_model.py:97-121) run inside the preprocess kernels and backward_view
                   returns gradients w.r.t. the raw parameters."""
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
                 rotations=_f32c(rotations, "rotations"), cov3D_precomp=_f32c(cov3D_precomp, "cov3D_precomp"),
                 shs_rest=_f32c(shs_rest, "shs_rest"))
        M = _sh_count(t["shs"], t["shs_rest"])
        color, _, radii, L, geom, point_list, image = _run_forward(
            raster_settings, P, M, 0, False, m, t["shs"], t["colors_precomp"], t["opacities"], t["scales"],
            t["rotations"], t["cov3D_precomp"], None, shs_rest=t["shs_rest"], raw_params=raw_params)
    t["means3D"] = m
    return color, radii, ViewState(raster_settings, P, M, L, t, radii, geom, point_list, image, bool(raw_params))


def _same_scene_settings(settings):
    """The grouped kernels take the SH degree and the scale modifier once per group: every view must agree on them."""
    if not settings:
        return True
    d0, m0 = int(settings[0].sh_degree), float(settings[0].scale_modifier)
    return all(int(rs.sh_degree) == d0 and float(rs.scale_modifier) == m0 for rs in settings)


def _view_fwd(inp, geom, radii):
    r = _abi.GsrViewFwd()
    r.viewmatrix, r.projmatrix, r.campos = inp.viewmatrix, inp.projmatrix, inp.campos
    r.tanfovx, r.tanfovy, r.W, r.H = inp.tanfovx, inp.tanfovy, inp.W, inp.H
    r.geom, r.radii = _ptr(geom), _ptr(radii)
    return r


def forward_views(settings, means3D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                  cov3D_precomp=None):
    """Render SEVERAL views of the same Gaussians: the per-Gaussian preprocess runs once per group of up to 16 views
    (gsr_forward_preprocess_views: every Gaussian's attributes are read once instead of once per view), then each
    view is ordered, binned and blended as forward_view does.  Per view the results are bit-identical to
    forward_view's.  Returns a list of (color, radii, ViewState), usable with backward_view / backward_views."""
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
    if P == 0 or any(rs.prefiltered for rs in settings) or not _same_scene_settings(settings):
        return [forward_view(rs, means3D, opacities, shs=shs, colors_precomp=colors_precomp, scales=scales,
                             rotations=rotations, cov3D_precomp=cov3D_precomp) for rs in settings]
    dev = means3D.device
    out = []
    with torch.no_grad(), torch.cuda.device(dev):
        lib = _abi.lib()
        st = _state(dev)
        stream = torch.cuda.current_stream(dev)
        sp = C.c_void_p(stream.cuda_stream)
        t = dict(means3D=_f32c(means3D, "means3D"), shs=_f32c(shs, "shs"),
                 colors_precomp=_f32c(colors_precomp, "colors_precomp"), opacities=_f32c(opacities, "opacities"),
                 scales=_f32c(scales, "scales"), rotations=_f32c(rotations, "rotations"),
                 cov3D_precomp=_f32c(cov3D_precomp, "cov3D_precomp"), shs_rest=None)
        M = _sh_count(t["shs"], None)
        F = _abi.MAX_FUSED_VIEWS
        for c0 in range(0, len(settings), F):
            group, keep = [], []
            for rs in settings[c0:c0 + F]:
                inp = _make_inputs(rs, P, M, 0, t["means3D"], t["shs"], t["colors_precomp"], t["opacities"], t["scales"],
                                   t["rotations"], t["cov3D_precomp"], None, keep)
                geom = torch.empty(lib.gsr_geom_bytes(P), dtype=torch.uint8, device=dev)
                radii = torch.empty((P,), dtype=torch.int32, device=dev)
                group.append((rs, inp, geom, radii))
            arr = (_abi.GsrViewFwd * len(group))(*[_view_fwd(inp, geom, radii) for _, inp, geom, radii in group])
            _abi.check(lib.gsr_forward_preprocess_views(C.byref(group[0][1]), len(group), arr, sp),
                       "forward_preprocess_views")
            for rs, inp, geom, radii in group:
                H, W = int(rs.image_height), int(rs.image_width)
                image = torch.empty(lib.gsr_image_bytes(W, H), dtype=torch.uint8, device=dev)
                color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
                scratch = st.ensure_scratch(lib.gsr_scratch_bytes(P, 0, W, H))
                _abi.check(lib.gsr_forward_order(C.byref(inp), _ptr(geom), _ptr(scratch), scratch.numel(),
                                                 _ptr(st.count_dev), C.c_void_p(st.count_host.data_ptr()), sp),
                           "forward_order")
                stream.synchronize()  # the duplicate count (reference: rasterizer_impl.cu:281)
                L = int(st.count_host.item()) & 0xFFFFFFFF
                point_list = torch.empty((max(L, 1),), dtype=torch.int32, device=dev)
                scratch = st.ensure_scratch(lib.gsr_scratch_bytes(P, L, W, H))
                _abi.check(lib.gsr_forward_render(C.byref(inp), _ptr(geom), _ptr(radii), L, _ptr(point_list),
                                                  _ptr(image), _ptr(scratch), scratch.numel(), _ptr(color), None, sp),
                           "forward_render")
                out.append((color, radii, ViewState(rs, P, M, L, t, radii, geom, point_list, image, False)))
    return out


def _check_aligned(name, g):
    """shs / rotations gradients are accumulated with 16-byte vector reductions (include/gsr_b200.h)."""
    if name in ("shs", "rotations") and g.data_ptr() & 15:
        raise RuntimeError(f"grads['{name}'] must be 16-byte aligned: lay flat buffers out with dist.flat_layout "
                           "(FlatGradBucket does)")


def _sh_count(shs, shs_rest):
    """Total SH coefficients M from either layout: shs (P,M,3), or shs (P,1,3) + shs_rest (P,M-1,3)."""
    if shs is None:
        if shs_rest is not None:
            raise Exception("shs_rest needs shs (the DC block)")
        return 0
    if shs_rest is None:
        return shs.size(1)
    if shs.size(1) != 1:
        raise Exception("with shs_rest, shs must be the DC block (P,1,3)")
    return 1 + shs_rest.size(1)


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
        _check_aligned(name, g)
        return g

    g_m3 = buf("means3D", True, P * 3)
    g_m2 = buf("means2D", False, P * 3)
    g_op = buf("opacities", True, P)
    split = t.get("shs_rest") is not None
    g_sh = buf("shs", t["shs"] is not None, P * (1 if split else state.M) * 3)
    g_rest = buf("shs_rest", split, P * (state.M - 1) * 3)
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
                           t["rotations"], t["cov3D_precomp"], None, keep, shs_rest=t.get("shs_rest"),
                           raw_params=state.raw_params, dL_dshs_rest=g_rest)
        scratch = st.ensure_scratch(lib.gsr_scratch_bytes(P, 0, W, H))
        _abi.check(lib.gsr_backward_accumulate(
            C.byref(inp), _ptr(state.geom), _ptr(state.radii), state.L, _ptr(state.point_list), _ptr(state.image),
            _ptr(scratch), scratch.numel(), _ptr(dpix), _ptr(g_m2), _ptr(g_op), _ptr(g_col), _ptr(g_m3), _ptr(g_cov),
            _ptr(g_sh), _ptr(g_sc), _ptr(g_rot), C.c_void_p(stream.cuda_stream)), "backward_accumulate")
        if rs.debug:
            stream.synchronize()


def _view_ref(inp, geom, radii, accum, mean2D=None):
    """gsr_view_ref of one forwarded view (camera fields from the gsr_inputs its forward used)."""
    r = _abi.GsrViewRef()
    r.viewmatrix, r.projmatrix, r.campos = inp.viewmatrix, inp.projmatrix, inp.campos
    r.tanfovx, r.tanfovy, r.W, r.H = inp.tanfovx, inp.tanfovy, inp.W, inp.H
    r.geom, r.radii, r.accum, r.dL_dmean2D = _ptr(geom), _ptr(radii), _ptr(accum), _ptr(mean2D)
    return r


def _preprocess_views(lib, inp, refs, accumulate, grads, stream):
    """One gsr_backward_preprocess_views call over up to MAX_FUSED_VIEWS gsr_view_ref's."""
    arr = (_abi.GsrViewRef * len(refs))(*refs)
    g = lambda n: _ptr(grads.get(n))
    _abi.check(lib.gsr_backward_preprocess_views(
        C.byref(inp), len(refs), arr, int(bool(accumulate)), g("opacities"), g("colors_precomp"), g("means3D"),
        g("cov3D_precomp"), g("shs"), g("scales"), g("rotations"), C.c_void_p(stream.cuda_stream)),
        "backward_preprocess_views")


def backward_views(states, dL_dimages, grads: Dict[str, Optional[torch.Tensor]], accumulate: bool = True,
                   means2D_grads=None) -> None:
    """Backward of SEVERAL forwarded views of the same Gaussians with the per-Gaussian half done once for all of
    them (gsr_backward_blend_view per view, then gsr_backward_preprocess_views per group of up to 16 views): every
    Gaussian's attributes are read once and its gradient rows written once per group instead of once per view.
    `states` / `dL_dimages`: ViewStates of forward_view calls on the SAME input tensors and dL/dimage of each.
    accumulate=True ADDS the summed gradients to `grads` (as repeated backward_view calls would, up to summation
    order); accumulate=False overwrites `grads` (rows of Gaussians no view rendered become zero), so the caller
    need not zero them.  `means2D_grads`: optional list of per-view (P,3) tensors for the screen-space gradients
    (they are per view, not summed).  Fused activations (raw_params / shs_rest) are not supported here."""
    if len(states) != len(dL_dimages):
        raise Exception("one dL_dimage per view state")
    if not states:
        return
    s0 = states[0]
    P, t = s0.P, s0.tensors
    if P == 0:
        return
    for st_ in states:
        if st_.P != P or st_.M != s0.M or st_.raw_params or st_.tensors.get("shs_rest") is not None or \
                any(st_.tensors[k] is not t[k] and (st_.tensors[k] is None or t[k] is None or
                                                   st_.tensors[k].data_ptr() != t[k].data_ptr())
                    for k in ("means3D", "shs", "colors_precomp", "opacities", "scales", "rotations", "cov3D_precomp")):
            raise Exception("backward_views: all views must come from forward_view on the same input tensors, "
                            "without raw_params / shs_rest")
    if not _same_scene_settings([st_.rs for st_ in states]):
        raise Exception("backward_views: the views differ in sh_degree / scale_modifier; use backward_view per view")
    for k in grads:
        if k not in GRAD_KEYS:
            raise Exception(f"unknown gradient buffer '{k}' (expected one of {GRAD_KEYS})")
    for name, key in (("means3D", "means3D"), ("opacities", "opacities"), ("shs", "shs"), ("colors_precomp", "colors_precomp"),
                      ("scales", "scales"), ("rotations", "rotations"), ("cov3D_precomp", "cov3D_precomp")):
        if t[key] is None:
            continue
        gbuf = grads.get(name)
        if gbuf is None:
            raise Exception(f"grads['{name}'] is required for the inputs these views were rendered with")
        if not gbuf.is_cuda or gbuf.dtype != torch.float32 or not gbuf.is_contiguous() or gbuf.numel() != t[key].numel():
            raise RuntimeError(f"grads['{name}'] must be a contiguous float32 CUDA tensor shaped like the input")
        _check_aligned(name, gbuf)
    m2_sum = grads.get("means2D")  # summed over the views, as repeated backward_view calls leave it
    if m2_sum is not None:
        if means2D_grads is not None:
            raise Exception("pass either grads['means2D'] (summed over the views) or means2D_grads (one per view)")
        if not m2_sum.is_cuda or m2_sum.dtype != torch.float32 or not m2_sum.is_contiguous() or m2_sum.numel() != 3 * P:
            raise RuntimeError("grads['means2D'] must be a contiguous float32 CUDA tensor with 3*P elements")
        if not accumulate:
            m2_sum.zero_()
    dev = t["means3D"].device
    with torch.no_grad(), torch.cuda.device(dev):
        lib = _abi.lib()
        stream = torch.cuda.current_stream(dev)
        sp = C.c_void_p(stream.cuda_stream)
        F = _abi.MAX_FUSED_VIEWS
        first = m2_sum is None  # with a summed means2D buffer every group runs in adding mode
        if not accumulate and not first:
            for k, gbuf in grads.items():
                if gbuf is not None:
                    gbuf.zero_()
        for c0 in range(0, len(states), F):
            refs, keep = [], []
            for i in range(c0, min(c0 + F, len(states))):
                st_ = states[i]
                rs = st_.rs
                H, W = int(rs.image_height), int(rs.image_width)
                dpix = _f32c(dL_dimages[i], "dL_dimage")
                if dpix is None or dpix.numel() != 3 * H * W:
                    raise RuntimeError(f"dL_dimage must have 3*{H}*{W} elements")
                inp = _make_inputs(rs, P, st_.M, 0, t["means3D"], t["shs"], t["colors_precomp"], t["opacities"],
                                   t["scales"], t["rotations"], t["cov3D_precomp"], None, keep)
                accum = torch.empty(lib.gsr_view_accum_bytes(P) // 4, dtype=torch.float32, device=dev)
                _abi.check(lib.gsr_backward_blend_view(C.byref(inp), _ptr(st_.geom), st_.L, _ptr(st_.point_list),
                                                       _ptr(st_.image), _ptr(dpix), _ptr(accum), sp), "backward_blend_view")
                m2 = means2D_grads[i] if means2D_grads is not None else None
                if m2 is not None and (not m2.is_cuda or m2.dtype != torch.float32 or not m2.is_contiguous() or m2.numel() != 3 * P):
                    raise RuntimeError("means2D_grads[i] must be a contiguous float32 CUDA tensor with 3*P elements")
                if m2 is not None and not accumulate and not first:
                    m2.zero_()  # later groups run in adding mode (for the summed buffers); per-view buffers start at 0
                refs.append(_view_ref(inp, st_.geom, st_.radii, accum, m2 if m2_sum is None else m2_sum))
                keep.extend([accum, dpix, inp])
            _preprocess_views(lib, keep[-1], refs, accumulate or not first, grads, stream)
            first = False
        if s0.rs.debug:
            stream.synchronize()


class _PipeSlot:
    def __init__(self, dev):
        self.dev = dev
        self.geom = self.image = self.radii = self.accum = None
        self.point_list = torch.empty(0, dtype=torch.int32, device=dev)
        self.count_dev = torch.zeros(1, dtype=torch.int32, device=dev)
        self.count_host = torch.zeros(1, dtype=torch.int32).pin_memory()
        self.ev_geom, self.ev_bin, self.ev_done = torch.cuda.Event(), torch.cuda.Event(), torch.cuda.Event()
        self.busy = False     # a backward on the render stream may still be reading the slot
        self.keep = None      # camera / input tensors referenced by kernels in flight
        self.inp = None
        self.L = 0
        self.sig = None


class ViewPipeline:
    """Multi-camera forward+backward with the memory- and latency-bound half of the NEXT camera (preprocess, depth
    order, scan, duplicate, tile sort, tile ranges) running on a high-priority stream underneath the issue-bound
    blending kernels of the CURRENT camera (new; the reference runs everything on the legacy default stream).

    Per camera k:   G stream:  geometry(k) -> [host reads L] -> binning(k)
                    R stream (the caller's current stream):  blend_fwd(k) -> dL_fn -> blend_bwd(k)
                    B stream:  preprocess_bwd(k), adding into `grads` (HBM-bound, under blend_fwd(k+1))
    geometry(k+1) is queued before blend_fwd(k), binning(k+1) right after backward(k) has been queued.  Per-view
    state lives in a ring of `n_slots` slots owned by the pipeline; all scratch used on G is disjoint from the
    backward's accumulators.  Results are those of forward_view / backward_view (same kernels, same order of
    accumulation into `grads`)."""

    def __init__(self, device=None, n_slots: int = 3, overlap: bool = True, fuse_backward: int = _abi.MAX_FUSED_VIEWS,
                 fuse_forward: bool = True, parallel_order: bool = False):
        """fuse_backward = F > 1: the per-Gaussian half of the backward (preprocess_bwd) is deferred and done ONCE per
        group of up to F cameras (gsr_backward_preprocess_views: attributes read once, gradient rows written once per
        group instead of once per camera); each slot then keeps its own 48-byte-per-Gaussian accumulator and the ring
        grows to F + 2 slots.  fuse_backward <= 1: one preprocess_bwd per camera, as backward_view does.
        fuse_forward (with fuse_backward > 1): the per-Gaussian preprocess of a whole group of F cameras is one launch
        (gsr_forward_preprocess_views: attributes read once per Gaussian instead of once per camera; per-camera results
        bit-identical), followed per camera by gsr_forward_order; the ring then holds 2 F + 2 slots.
        parallel_order (experiment, off): the depth orders of a group's cameras side by side on F extra streams.
        Measured on C2: 967 -> 969 views/s, but with eight more streams the copy streams of pipeline.HostStepPipeline
        end up sharing hardware queues with compute streams (CUDA_DEVICE_MAX_CONNECTIONS = 8) and `e2e` halves."""
        if n_slots < 2:
            raise ValueError("n_slots must be >= 2")
        import os
        if os.environ.get("GSR_FUSE_VIEWS"):  # experiment knob: group size of the deferred / grouped kernels
            fuse_backward = int(os.environ["GSR_FUSE_VIEWS"])
        self.fuse = max(1, min(int(fuse_backward), _abi.MAX_FUSED_VIEWS))
        self.fuse_fwd = bool(fuse_forward) and self.fuse > 1
        if self.fuse > 1:
            # a group's slots stay in use until its deferred backward has run; with the grouped preprocess the NEXT
            # group's geometry is written while the current group is still pending
            n_slots = max(n_slots, (2 * self.fuse if self.fuse_fwd else self.fuse) + 2)
        self.pending = []     # slots whose blend_bwd has been queued and whose per-Gaussian backward is deferred
        self.overwrite_next = False  # the next deferred pass SETS the gradient rows instead of adding to them
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.dev = dev
        self.overlap = overlap
        self.slots = [_PipeSlot(dev) for _ in range(n_slots)]
        import os
        pg, pb = {"none": (0, 0), "g": (-1, 0), "b": (0, -1)}.get(os.environ.get("GSR_PIPE_PRIO", ""), (-1, -1))
        self.G = torch.cuda.Stream(dev, priority=pg) if overlap else None  # measured: 710 views/s vs 680 at equal priority
        self.B = torch.cuda.Stream(dev, priority=pb) if overlap else None
        self.scratch = torch.empty(0, dtype=torch.uint8, device=dev)
        n_order = self.fuse if (overlap and self.fuse_fwd and parallel_order) else 0
        self.O = [torch.cuda.Stream(dev, priority=-1) for _ in range(n_order)]  # ordering streams (grouped preprocess)
        self.o_scratch = [None] * n_order
        self.ev_pre = torch.cuda.Event()
        self.ev_start, self.ev_bb, self.ev_pb = torch.cuda.Event(), torch.cuda.Event(), torch.cuda.Event()
        self.pb_pending = False

    # ---- helpers ------------------------------------------------------------------------------------------
    def _gstream(self):
        return self.G if self.overlap else torch.cuda.current_stream(self.dev)

    def _ensure(self, cur, nbytes, dtype=torch.uint8):
        """Grow a pipeline-owned buffer; both streams are drained first because the old block goes back to the
        caching allocator, which only orders reuse on the allocating stream."""
        if cur is not None and cur.numel() * cur.element_size() >= nbytes:
            return cur
        torch.cuda.synchronize(self.dev)
        n = int(nbytes * 1.25) + 256
        return torch.empty(n // torch.empty(0, dtype=dtype).element_size() + 1, dtype=dtype, device=self.dev)

    def _geometry(self, slot, rs, t, P, M, grads=None, raw_params=False):
        lib = _abi.lib()
        G = self._gstream()
        self._prepare_slot(slot, rs, t, P, M, grads, raw_params)
        _abi.check(lib.gsr_forward_geometry(C.byref(slot.inp), _ptr(slot.geom), _ptr(slot.radii), _ptr(self.scratch),
                                            self.scratch.numel(), _ptr(slot.count_dev),
                                            C.c_void_p(slot.count_host.data_ptr()), C.c_void_p(G.cuda_stream)),
                   "forward_geometry")
        slot.ev_geom.record(G)

    def _preprocess_group(self, slots, settings, t, P, M, grads):
        """Phase 1a of a whole group of cameras in one launch on G (gsr_forward_preprocess_views): attributes read once."""
        lib = _abi.lib()
        G = self._gstream()
        for slot, rs in zip(slots, settings):
            self._prepare_slot(slot, rs, t, P, M, grads, False)
        arr = (_abi.GsrViewFwd * len(slots))(*[_view_fwd(sl.inp, sl.geom, sl.radii) for sl in slots])
        _abi.check(lib.gsr_forward_preprocess_views(C.byref(slots[0].inp), len(slots), arr, C.c_void_p(G.cuda_stream)),
                   "forward_preprocess_views")
        self.last_fwd_group = list(slots)  # measurement hook (replay_last_preprocess)
        if self.overlap:
            self.ev_pre.record(G)

    def _order(self, slot, lane=None):
        """Phase 1b of one camera: depth order, prefix sum, duplicate count (the rest of gsr_forward_geometry).
        lane = None: on G with the pipeline's scratch.  lane = i: on the i-th ordering stream with that stream's own
        scratch -- the depth sorts of a group's cameras are independent and each is latency-bound (one wave of
        tiles, four dependent launches), so queued side by side their tails and launch gaps fill each other."""
        lib = _abi.lib()
        if lane is None or not self.overlap:
            stream, scratch = self._gstream(), self.scratch
        else:
            stream = self.O[lane]
            need = lib.gsr_scratch_bytes(slot.inp.P, 0, slot.inp.W, slot.inp.H)
            self.o_scratch[lane] = scratch = self._ensure(self.o_scratch[lane], need)
            stream.wait_event(self.ev_pre)
        _abi.check(lib.gsr_forward_order(C.byref(slot.inp), _ptr(slot.geom), _ptr(scratch), scratch.numel(),
                                         _ptr(slot.count_dev), C.c_void_p(slot.count_host.data_ptr()),
                                         C.c_void_p(stream.cuda_stream)), "forward_order")
        slot.ev_geom.record(stream)

    def _prepare_slot(self, slot, rs, t, P, M, grads=None, raw_params=False):
        """Buffers and the gsr_inputs of one camera; G waits until the slot's previous user is done with it."""
        lib = _abi.lib()
        W, H = int(rs.image_width), int(rs.image_height)
        G = self._gstream()
        if any(sl is slot for sl in self.pending):  # (many cameras without a backward since this one: finish its group first)
            self._flush_pending(self._step_grads)
        if slot.busy and self.overlap:
            G.wait_event(slot.ev_done)
        sig = (P, W, H)
        if slot.sig != sig:
            slot.geom = self._ensure(None, lib.gsr_geom_bytes(P))
            slot.image = self._ensure(None, lib.gsr_image_bytes(W, H))
            slot.radii = torch.empty((P,), dtype=torch.int32, device=self.dev)
            slot.accum = None
            slot.sig = sig
        self.scratch = self._ensure(self.scratch, lib.gsr_scratch_bytes(P, max(slot.point_list.numel(), 1), W, H))
        keep = []
        inp = _make_inputs(rs, P, M, 0, t["means3D"], t["shs"], t["colors_precomp"], t["opacities"], t["scales"],
                           t["rotations"], t["cov3D_precomp"], None, keep, shs_rest=t.get("shs_rest"),
                           raw_params=raw_params, dL_dshs_rest=(grads or {}).get("shs_rest"))
        slot.keep, slot.inp = keep, inp

    def _binning(self, slot, rs, P):
        lib = _abi.lib()
        W, H = int(rs.image_width), int(rs.image_height)
        G = self._gstream()
        slot.ev_geom.synchronize()  # the one host sync per view (reference: rasterizer_impl.cu:281)
        L = slot.L = int(slot.count_host.item()) & 0xFFFFFFFF
        if slot.point_list.numel() < max(L, 1):
            slot.point_list = self._ensure(slot.point_list, 4 * max(L, 1), torch.int32)
        self.scratch = self._ensure(self.scratch, lib.gsr_scratch_bytes(P, L, W, H))
        _abi.check(lib.gsr_forward_binning(C.byref(slot.inp), _ptr(slot.geom), _ptr(slot.radii), L, _ptr(slot.point_list),
                                           _ptr(slot.image), _ptr(self.scratch), self.scratch.numel(),
                                           C.c_void_p(G.cuda_stream)), "forward_binning")
        slot.ev_bin.record(G)

    def _blend(self, slot, rs):
        lib = _abi.lib()
        R = torch.cuda.current_stream(self.dev)
        if self.overlap:
            R.wait_event(slot.ev_bin)
        color = torch.empty((3, int(rs.image_height), int(rs.image_width)), dtype=torch.float32, device=self.dev)
        _abi.check(lib.gsr_forward_blend(C.byref(slot.inp), _ptr(slot.geom), _ptr(slot.point_list), _ptr(slot.image),
                                         _ptr(color), None, C.c_void_p(R.cuda_stream)), "forward_blend")
        return color

    def _backward(self, slot, rs, t, P, M, dL_dimage, grads, fused=False):
        lib = _abi.lib()
        R = torch.cuda.current_stream(self.dev)
        H, W = int(rs.image_height), int(rs.image_width)
        dpix = _f32c(dL_dimage, "dL_dimage")
        if dpix is None or dpix.numel() != 3 * H * W:
            raise RuntimeError(f"dL_dimage must have 3*{H}*{W} elements")
        g = lambda n: _ptr(grads.get(n))

        def phase(ph, stream):
            _abi.check(lib.gsr_backward_accumulate_phase(
                C.byref(slot.inp), _ptr(slot.geom), _ptr(slot.radii), slot.L, _ptr(slot.point_list), _ptr(slot.image),
                _ptr(self.scratch), self.scratch.numel(), _ptr(dpix), g("means2D"), g("opacities"), g("colors_precomp"),
                g("means3D"), g("cov3D_precomp"), g("shs"), g("scales"), g("rotations"), ph,
                C.c_void_p(stream.cuda_stream)), "backward_accumulate")

        if fused:
            # per-tile half now, into the slot's own accumulator; the per-Gaussian half of the whole group later
            if slot.accum is None:
                slot.accum = torch.empty(lib.gsr_view_accum_bytes(P) // 4, dtype=torch.float32, device=self.dev)
            _abi.check(lib.gsr_backward_blend_view(C.byref(slot.inp), _ptr(slot.geom), slot.L, _ptr(slot.point_list),
                                                   _ptr(slot.image), _ptr(dpix), _ptr(slot.accum),
                                                   C.c_void_p(R.cuda_stream)), "backward_blend_view")
            self.pending.append(slot)
            if len(self.pending) >= self.fuse:
                self._flush_pending(grads)
            return
        if not self.overlap:
            phase(3, R)
            slot.ev_done.record(R)
        else:
            # the issue-bound per-tile half stays on R; the HBM-bound per-Gaussian half goes to stream B, where it
            # runs underneath the next camera's blend_fwd.  The accumulators are shared: the next blend_bwd waits.
            if self.pb_pending:
                R.wait_event(self.ev_pb)
            phase(1, R)
            self.ev_bb.record(R)
            self.B.wait_event(self.ev_bb)
            phase(2, self.B)
            self.ev_pb.record(self.B)
            slot.ev_done.record(self.B)
            self.pb_pending = True
        slot.busy = True

    def _flush_pending(self, grads):
        """The deferred per-Gaussian backward of the cameras in self.pending, once for all of them, on stream B (under
        the next camera's blending) -- or on the caller's stream without overlap."""
        if not self.pending:
            return
        lib = _abi.lib()
        R = torch.cuda.current_stream(self.dev)
        refs = [_view_ref(sl.inp, sl.geom, sl.radii, sl.accum, grads.get("means2D")) for sl in self.pending]
        if self.overlap:
            self.ev_bb.record(R)
            self.B.wait_event(self.ev_bb)
            stream = self.B
        else:
            stream = R
        _preprocess_views(lib, self.pending[0].inp, refs, not self.overwrite_next, grads, stream)
        self.overwrite_next = False
        self.last_group = list(self.pending)  # measurement hook (replay_last_group)
        for sl in self.pending:
            sl.ev_done.record(stream)
            sl.busy = True
        if self.overlap:
            self.ev_pb.record(self.B)
            self.pb_pending = True
        self.pending = []

    def replay_last_preprocess(self):
        """Measurement hook: the most recent grouped preprocess launch again, on the current stream, after the step
        has finished (it rewrites the same geometry state; the slots' depth keys are not needed any more)."""
        group = getattr(self, "last_fwd_group", None)
        if not group:
            raise RuntimeError("no grouped preprocess to replay (fuse_forward off, or no step yet)")
        arr = (_abi.GsrViewFwd * len(group))(*[_view_fwd(sl.inp, sl.geom, sl.radii) for sl in group])
        _abi.check(_abi.lib().gsr_forward_preprocess_views(
            C.byref(group[0].inp), len(group), arr, C.c_void_p(torch.cuda.current_stream(self.dev).cuda_stream)),
            "forward_preprocess_views")
        return [sl.radii for sl in group]

    def replay_last_group(self, grads, accumulate=False):
        """Measurement hook: run the deferred per-Gaussian backward of the most recent group again on the current
        stream (its per-view accumulators and geometry are still in the slots until the next step).  With
        accumulate=False `grads` ends up as a one-group step left it.  Returns the radii tensors of the group."""
        group = getattr(self, "last_group", None)
        if not group:
            raise RuntimeError("no deferred group to replay (fuse_backward <= 1, or no step with a backward yet)")
        refs = [_view_ref(sl.inp, sl.geom, sl.radii, sl.accum, grads.get("means2D")) for sl in group]
        _preprocess_views(_abi.lib(), group[0].inp, refs, accumulate, grads, torch.cuda.current_stream(self.dev))
        return [sl.radii for sl in group]

    # ---- public -------------------------------------------------------------------------------------------
    def step(self, settings, dL_fn, grads, means3D, opacities, shs=None, colors_precomp=None, scales=None,
             rotations=None, cov3D_precomp=None, shs_rest=None, raw_params=False, overwrite=False):
        """Render every camera in `settings` (GaussianRasterizationSettings) and ADD its gradients to `grads` (as
        backward_view does).  `dL_fn(image, k)` receives the plain (3,H,W) render of camera k on the current stream
        and returns dL/dimage -- or None to skip that camera's backward (forward-only use).  Returns the list of
        duplicate counts (num_rendered) per camera.  `shs_rest` / `raw_params` as in forward_view (fused model
        activations; gradients then refer to the raw parameters, `grads["shs"]` to the DC block and
        `grads["shs_rest"]` to the rest block).  overwrite=True SETS `grads` to the sum over this step's cameras
        instead of adding to it: the caller need not zero the buffers, and with the deferred per-Gaussian backward
        (fuse_backward > 1) the first group's pass writes each row without reading it."""
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        if means3D.ndimension() != 2 or means3D.size(1) != 3:
            raise RuntimeError("means3D must have dimensions (num_points, 3)")
        if not means3D.is_cuda:
            raise RuntimeError("means3D must be a CUDA tensor (this rasterizer has no CPU path)")
        for k in grads:
            if k not in GRAD_KEYS:
                raise Exception(f"unknown gradient buffer '{k}' (expected one of {GRAD_KEYS})")
        P, n = means3D.size(0), len(settings)
        counts = []
        if P == 0 or n == 0:
            if overwrite:
                for gbuf in grads.values():
                    if gbuf is not None:
                        gbuf.zero_()
            return counts
        with torch.no_grad(), torch.cuda.device(self.dev):
            t = dict(means3D=_f32c(means3D, "means3D"), shs=_f32c(shs, "shs"),
                     colors_precomp=_f32c(colors_precomp, "colors_precomp"), opacities=_f32c(opacities, "opacities"),
                     scales=_f32c(scales, "scales"), rotations=_f32c(rotations, "rotations"),
                     cov3D_precomp=_f32c(cov3D_precomp, "cov3D_precomp"), shs_rest=_f32c(shs_rest, "shs_rest"))
            M = _sh_count(t["shs"], t["shs_rest"])
            for name in ("means3D", "opacities", "shs", "shs_rest", "colors_precomp", "scales", "rotations", "cov3D_precomp"):
                need = t[name] is not None
                gbuf = grads.get(name)
                if gbuf is None and need and any(grads.get(x) is not None for x in grads):
                    raise Exception(f"grads['{name}'] is required for the inputs these views are rendered with")
                if gbuf is not None and (not gbuf.is_cuda or gbuf.dtype != torch.float32 or not gbuf.is_contiguous()
                                         or (need and gbuf.numel() != t[name].numel())):
                    raise RuntimeError(f"grads['{name}'] must be a contiguous float32 CUDA tensor shaped like the input")
                if gbuf is not None:
                    _check_aligned(name, gbuf)
            R = torch.cuda.current_stream(self.dev)
            # Camera tensors of EVERY view are brought to contiguous float32 device memory here, on the caller's
            # stream and before ev_start: Camera.world_view_transform is a transposed (non-contiguous) CUDA tensor
            # (scene/cameras.py:54), so .contiguous() launches a copy that the G and B streams must not overtake.
            # The list keeps them alive until the end of the step, where R has waited for G and B.
            settings = [rs._replace(bg=_f32c(rs.bg.to(self.dev), "bg"),
                                    viewmatrix=_f32c(rs.viewmatrix.to(self.dev), "viewmatrix"),
                                    projmatrix=_f32c(rs.projmatrix.to(self.dev), "projmatrix"),
                                    campos=_f32c(rs.campos.to(self.dev), "campos")) for rs in settings]
            if self.overlap:  # whatever produced the inputs on the caller's stream must be visible to G
                self.ev_start.record(R)
                self.G.wait_event(self.ev_start)
            # (a single camera gains nothing from the grouped kernels: they run at lower occupancy than the per-view ones)
            fused = self.fuse > 1 and n > 1 and not raw_params and t["shs_rest"] is None and (M <= 16) and \
                _same_scene_settings(settings)  # (SH degree and scale modifier are per-group inputs of the grouped kernels)
            self._step_grads = grads
            self.pending = []  # (a previous step that raised part-way may have left deferred cameras behind)
            self.overwrite_next = bool(overwrite) and fused and grads.get("means2D") is None and \
                any(g is not None for g in grads.values())
            if overwrite and not self.overwrite_next:
                for gbuf in grads.values():
                    if gbuf is not None:
                        gbuf.zero_()
            ns = len(self.slots)
            fused_fwd = fused and self.fuse_fwd and not any(rs.prefiltered for rs in settings)
            F = self.fuse

            def geometry(k):  # phase 1 of camera k on G
                if not fused_fwd:
                    self._geometry(self.slots[k % ns], settings[k], t, P, M, grads, raw_params)
                    return
                if k % F == 0:  # first camera of a group: the whole group's preprocess in one launch
                    ks = range(k, min(k + F, n))
                    self._preprocess_group([self.slots[j % ns] for j in ks], [settings[j] for j in ks], t, P, M, grads)
                if not self.O:
                    self._order(self.slots[k % ns])
                elif k % F == 0:  # the group's first camera is ordered at once on G ...
                    self._order(self.slots[k % ns])
                elif k % F == 1:  # ... the others side by side on the ordering streams once its binning is queued
                    for j in range(k, min(k - 1 + F, n)):
                        self._order(self.slots[j % ns], lane=j % F)

            geometry(0)
            self._binning(self.slots[0], settings[0], P)
            for k in range(n):
                slot = self.slots[k % ns]
                if k + 1 < n:
                    geometry(k + 1)
                color = self._blend(slot, settings[k])
                counts.append(slot.L)
                with torch.enable_grad():
                    dL = dL_fn(color, k)
                if dL is not None:
                    self._backward(slot, settings[k], t, P, M, dL, grads, fused)
                else:
                    slot.ev_done.record(R)
                    slot.busy = True
                if k + 1 < n:
                    self._binning(self.slots[(k + 1) % ns], settings[k + 1], P)
            self._flush_pending(grads)
            if self.overwrite_next:  # no camera had a backward: the sum over no cameras is zero
                for gbuf in grads.values():
                    if gbuf is not None:
                        gbuf.zero_()
                self.overwrite_next = False
            if self.overlap:  # later work on the caller's stream (and frees of the inputs) must follow G's and B's
                ev = torch.cuda.Event()
                ev.record(self.G)
                R.wait_event(ev)
                for o in self.O:  # (every ordering launch was followed by a host wait on its event: already complete)
                    R.wait_stream(o)
                if self.pb_pending:
                    R.wait_event(self.ev_pb)
                    self.pb_pending = False
        return counts
