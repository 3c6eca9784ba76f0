"""Drop-in `diff_gaussian_rasterization` package backed by the B200-native C-ABI library.

Same public names and behaviour as the reference package
(DGR/diff_gaussian_rasterization/__init__.py:21-221; DGR = submodules/gaussian-splatting/
submodules/diff-gaussian-rasterization of poloclub/3d-gaussian-splat-attack):
  GaussianRasterizationSettings  (:157-169)  same 12 fields
  GaussianRasterizer             (:171-221)  markVisible(), forward()
  rasterize_gaussians            (:21-42)
  _RasterizeGaussians            (:44-155)   autograd Function, same argument / gradient order
plus the superset that poloclub's own gaussian_renderer/__init__.py:86-95 relies on: an optional
`sh_objs` (P,1,16) input and a third output `rendered_objects` (16,H,W) when it is given
(Gaussian-Grouping's fork; not vendored and not version-pinned in the reference, so that channel is EXPERIMENTAL: its parity
is unpinned -- 16 extra features blended with the colour's alpha*T weights, no background term, checked against the
oracle's restatement only).  A backward in which no gradient reaches `rendered_objects` (attack.py:485-494 reads
render_pkg["render"] only) runs the colour-only kernels and returns None for `sh_objs`.

Host logic is Python/PyTorch; all compute goes through include/gsr_b200.h via ctypes.
There is no CPU path: non-CUDA tensors raise.
"""
from typing import NamedTuple
import ctypes as C
import os

import torch
import torch.nn as nn

from . import _abi
from ._abi import GsrInputs

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians", "_RasterizeGaussians"]


def cpu_deep_copy_tuple(input_tuple):
    copied_tensors = [item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple]
    return tuple(copied_tensors)


class _DeviceState:
    """Reusable scratch and pinned read-back word of one (device, stream): calls issued on different streams of a
    device (or from different host threads using different streams) never share sort buffers, backward
    accumulators or the duplicate-count word (the reference allocates these per call)."""

    def __init__(self, device):
        self.device = device
        self.scratch = torch.empty(0, dtype=torch.uint8, device=device)
        self.count_dev = torch.zeros(1, dtype=torch.int32, device=device)
        self.count_host = torch.zeros(1, dtype=torch.int32).pin_memory()
        self.ev_count = torch.cuda.Event()
        self.last_L = {}  # (P, W, H) -> duplicate count of the previous view of that shape (sizes the next launch)

    def ensure_scratch(self, nbytes):
        if self.scratch.numel() < nbytes:
            # grow geometrically; the old block is released to the caching allocator (stream-ordered)
            self.scratch = torch.empty(int(nbytes * 1.25) + 256, dtype=torch.uint8, device=self.device)
        return self.scratch


_states = {}
SPECULATE = os.environ.get("GSR_SPECULATE", "1") != "0"  # gsr_forward_render_speculative in steady state


def _state(device):
    index = device.index if device.index is not None else torch.cuda.current_device()
    key = (index, torch.cuda.current_stream(index).cuda_stream)
    st = _states.get(key)
    if st is None:
        st = _states[key] = _DeviceState(torch.device("cuda", index))
    return st


def _ptr(t):
    return None if t is None or t.numel() == 0 else C.c_void_p(t.data_ptr())


def _f32c(t, name):
    if t is None or t.numel() == 0:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (this rasterizer has no CPU path)")
    if t.dtype != torch.float32:
        t = t.float()
    t = t.contiguous()
    if t.data_ptr() & 15:  # a view into a packed buffer: the kernels move shs / rotations / sh_objs as 16-byte words
        t = t.clone()
    return t


def _make_inputs(rs, P, M, K, means3D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp, sh_objs, keep,
                 shs_rest=None, raw_params=False, dL_dshs_rest=None):
    dev = means3D.device
    bg = _f32c(rs.bg.to(dev), "bg")
    view = _f32c(rs.viewmatrix.to(dev), "viewmatrix")
    proj = _f32c(rs.projmatrix.to(dev), "projmatrix")
    campos = _f32c(rs.campos.to(dev), "campos")
    if bg.numel() < 3 or view.numel() != 16 or proj.numel() != 16 or campos.numel() < 3:
        raise RuntimeError("bg needs >=3 elements, viewmatrix/projmatrix 16, campos 3")
    keep.extend([bg, view, proj, campos])
    inp = GsrInputs()
    inp.P, inp.D, inp.M, inp.W, inp.H, inp.K = P, int(rs.sh_degree), M, int(rs.image_width), int(rs.image_height), K
    inp.prefiltered, inp.debug = int(bool(rs.prefiltered)), int(bool(rs.debug))
    inp.scale_modifier, inp.tanfovx, inp.tanfovy = float(rs.scale_modifier), float(rs.tanfovx), float(rs.tanfovy)
    inp.bg, inp.means3D, inp.shs, inp.colors_precomp = _ptr(bg), _ptr(means3D), _ptr(sh), _ptr(colors_precomp)
    inp.opacities, inp.scales, inp.rotations = _ptr(opacities), _ptr(scales), _ptr(rotations)
    inp.cov3D_precomp, inp.viewmatrix, inp.projmatrix = _ptr(cov3Ds_precomp), _ptr(view), _ptr(proj)
    inp.campos, inp.obj_feats = _ptr(campos), _ptr(sh_objs)
    inp.shs_rest, inp.raw_params, inp.dL_dshs_rest = _ptr(shs_rest), int(bool(raw_params)), _ptr(dL_dshs_rest)
    return inp


def _run_forward(rs, P, M, K, want_obj, means3D_c, sh_c, col_c, op_c, sc_c, rot_c, cov_c, obj_c, shs_rest=None,
                 raw_params=False):
    """Both forward phases of one view through the C ABI (Rasterizer::forward, rasterizer_impl.cu:198-336).
    Returns (color, objects, radii, L, geom, point_list, image); the last three are what backward needs."""
    dev = means3D_c.device
    H, W = int(rs.image_height), int(rs.image_width)
    color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
    objects = torch.empty((K, H, W), dtype=torch.float32, device=dev) if want_obj else None
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    L = 0
    geom = image = point_list = None
    with torch.cuda.device(dev):
        lib = _abi.lib()
        st = _state(dev)
        stream = torch.cuda.current_stream(dev)
        sp = C.c_void_p(stream.cuda_stream)
        keep = []
        inp = _make_inputs(rs, P, M, K if want_obj else 0, means3D_c, sh_c, col_c, op_c, sc_c, rot_c, cov_c,
                           obj_c, keep, shs_rest=shs_rest, raw_params=raw_params)
        if P == 0:
            color.zero_()
            if want_obj:
                objects.zero_()
        else:
            geom = torch.empty(lib.gsr_geom_bytes(P), dtype=torch.uint8, device=dev)
            image = torch.empty(lib.gsr_image_bytes(W, H), dtype=torch.uint8, device=dev)
            scratch = st.ensure_scratch(lib.gsr_scratch_bytes(P, 0, W, H))
            # Steady state without a GPU bubble at the read-back of the duplicate count (rasterizer_impl.cu:281):
            # when a previous view of this shape is known, point_list / scratch are sized for 1.25 x its count and
            # phase 2 is queued at once, reading the actual count on the device; the host then waits only for
            # phase 1 (while phase 2 executes) and repeats phase 2 with the exact size in the rare overflow.
            key = (P, W, H)
            guess = st.last_L.get(key, 0) if SPECULATE and not rs.debug and not rs.prefiltered else 0
            L_cap = max(int(guess * 1.25) + 4096, 1) if guess else 0
            scratch = st.ensure_scratch(lib.gsr_scratch_bytes(P, L_cap, W, H))
            _abi.check(lib.gsr_forward_geometry(C.byref(inp), _ptr(geom), _ptr(radii), _ptr(scratch),
                                                scratch.numel(), _ptr(st.count_dev),
                                                C.c_void_p(st.count_host.data_ptr()), sp), "forward_geometry")
            if L_cap:
                st.ev_count.record(stream)
                point_list = torch.empty((L_cap,), dtype=torch.int32, device=dev)
                _abi.check(lib.gsr_forward_render_speculative(
                    C.byref(inp), _ptr(geom), _ptr(radii), L_cap, _ptr(st.count_dev), _ptr(point_list), _ptr(image),
                    _ptr(scratch), scratch.numel(), _ptr(color), _ptr(objects), sp), "forward_render_speculative")
                st.ev_count.synchronize()  # phase 1 only; phase 2 is running
            else:
                stream.synchronize()  # first view of this shape: the one host sync (reference: rasterizer_impl.cu:281)
            L = int(st.count_host.item()) & 0xFFFFFFFF
            st.last_L[key] = L
            if L > L_cap:  # not speculated, or the guess was too small: phase 2 with the exact size
                point_list = torch.empty((max(L, 1),), dtype=torch.int32, device=dev)
                scratch = st.ensure_scratch(lib.gsr_scratch_bytes(P, L, W, H))
                _abi.check(lib.gsr_forward_render(C.byref(inp), _ptr(geom), _ptr(radii), L, _ptr(point_list),
                                                  _ptr(image), _ptr(scratch), scratch.numel(), _ptr(color),
                                                  _ptr(objects), sp), "forward_render")
            if rs.debug:
                stream.synchronize()
    return color, objects, radii, L, geom, point_list, image


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings, sh_objs=None):
    out = _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                    cov3Ds_precomp, raster_settings, sh_objs)
    return out if sh_objs is not None else out[:2]


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings, sh_objs=None):
        rs = raster_settings
        if means3D.ndimension() != 2 or means3D.size(1) != 3:
            raise RuntimeError("means3D must have dimensions (num_points, 3)")  # rasterize_points.cu:57-59
        if not means3D.is_cuda:
            raise RuntimeError("means3D must be a CUDA tensor (this rasterizer has no CPU path)")
        dev = means3D.device
        P, H, W = means3D.size(0), int(rs.image_height), int(rs.image_width)
        means3D_c = _f32c(means3D, "means3D") if P else means3D
        sh_c, col_c, op_c = _f32c(sh, "sh"), _f32c(colors_precomp, "colors_precomp"), _f32c(opacities, "opacities")
        sc_c, rot_c, cov_c = _f32c(scales, "scales"), _f32c(rotations, "rotations"), _f32c(cov3Ds_precomp, "cov3Ds_precomp")
        obj_c = _f32c(sh_objs, "sh_objs")
        M = sh_c.size(1) if sh_c is not None else 0
        K = obj_c.numel() // P if (obj_c is not None and P) else (16 if sh_objs is not None else 0)
        want_obj = sh_objs is not None

        debug_args = None
        if rs.debug:  # snapshot in the reference's _C argument order (__init__.py:60-90)
            debug_args = cpu_deep_copy_tuple((rs.bg, means3D, colors_precomp, opacities, scales, rotations,
                                              rs.scale_modifier, cov3Ds_precomp, rs.viewmatrix, rs.projmatrix,
                                              rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh,
                                              rs.sh_degree, rs.campos, rs.prefiltered, rs.debug))
        try:
            color, objects, radii, L, geom, point_list, image = _run_forward(
                rs, P, M, K, want_obj, means3D_c, sh_c, col_c, op_c, sc_c, rot_c, cov_c, obj_c)
        except Exception as ex:
            if rs.debug:
                torch.save(debug_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
            raise ex

        # undefined output gradients arrive as None instead of zero tensors: a loss that never reads
        # `rendered_objects` (attack.py:485-494 uses render_pkg["render"] only) then runs the K = 0 backward
        ctx.set_materialize_grads(False)
        ctx.raster_settings = rs
        ctx.num_rendered = L
        ctx.dims = (P, M, K if want_obj else 0)
        ctx.want_obj = want_obj
        ctx.obj_shape = tuple(sh_objs.shape) if want_obj else None
        ctx.dev = dev
        empty = torch.empty(0, device=dev)
        ctx.save_for_backward(*[t if t is not None else empty for t in
                                (col_c, means3D_c, sc_c, rot_c, cov_c, radii, sh_c, op_c, obj_c, geom, point_list, image)])
        ctx.mark_non_differentiable(radii)
        if want_obj:
            return color, radii, objects
        return color, radii

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii=None, grad_out_objects=None):
        rs = ctx.raster_settings
        P, M, K = ctx.dims
        L = ctx.num_rendered
        (col_c, means3D_c, sc_c, rot_c, cov_c, radii, sh_c, op_c, obj_c, geom, point_list, image) = \
            [t if t.numel() else None for t in ctx.saved_tensors]
        dev = ctx.dev
        H, W = int(rs.image_height), int(rs.image_width)
        if grad_out_color is None and grad_out_objects is None:
            return (None,) * 10
        if grad_out_color is None:
            grad_out_color = torch.zeros((3, H, W), dtype=torch.float32, device=dev)
        if grad_out_objects is None:  # no gradient reaches the object channels: colour-only backward
            K, obj_c = 0, None

        def new(*shape):
            return torch.empty(shape, dtype=torch.float32, device=dev)

        g_means2D, g_opac, g_means3D = new(P, 3), new(P, 1), new(P, 3)
        g_colors = new(P, 3) if col_c is not None else None
        g_cov = new(P, 6) if cov_c is not None else None
        g_sh = new(P, M, 3) if sh_c is not None else None
        g_scales = new(P, 3) if sc_c is not None else None
        g_rot = new(P, 4) if rot_c is not None else None
        g_obj = new(P, K) if K else None
        if P:
            dpix = _f32c(grad_out_color, "grad_out_color")
            dobj = _f32c(grad_out_objects, "grad_out_objects") if K else None
            debug_args = None
            if rs.debug:  # reference _C backward argument order (__init__.py:109-139)
                debug_args = cpu_deep_copy_tuple((rs.bg, means3D_c, radii, col_c, sc_c, rot_c, rs.scale_modifier, cov_c,
                                                  rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color,
                                                  sh_c, rs.sh_degree, rs.campos, geom, L, point_list, image, rs.debug))
            try:
                with torch.cuda.device(dev):
                    lib = _abi.lib()
                    st = _state(dev)
                    stream = torch.cuda.current_stream(dev)
                    keep = []
                    inp = _make_inputs(rs, P, M, K, means3D_c, sh_c, col_c, op_c, sc_c, rot_c, cov_c, obj_c, keep)
                    scratch = st.ensure_scratch(lib.gsr_scratch_bytes(P, 0, W, H))
                    _abi.check(lib.gsr_backward(C.byref(inp), _ptr(geom), _ptr(radii), L, _ptr(point_list), _ptr(image),
                                                _ptr(scratch), scratch.numel(), _ptr(dpix), _ptr(dobj),
                                                _ptr(g_means2D), _ptr(g_opac), _ptr(g_colors), _ptr(g_means3D),
                                                _ptr(g_cov), _ptr(g_sh), _ptr(g_scales), _ptr(g_rot), _ptr(g_obj),
                                                C.c_void_p(stream.cuda_stream)), "backward")
                    if rs.debug:
                        stream.synchronize()
            except Exception as ex:
                if rs.debug:
                    torch.save(debug_args, "snapshot_bw.dump")
                    print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        if g_obj is not None:
            g_obj = g_obj.view(ctx.obj_shape)
        # same order as the forward inputs (reference __init__.py:143-153), sh_objs appended
        return (g_means3D, g_means2D, g_sh, g_colors, g_opac, g_scales, g_rot, g_cov, None, g_obj)


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """Boolean mask of points passing the near-plane test (rasterizer_impl.cu:54-66,141-153)."""
        with torch.no_grad():
            rs = self.raster_settings
            if not positions.is_cuda:
                raise RuntimeError("positions must be a CUDA tensor (this rasterizer has no CPU path)")
            P = positions.size(0)
            pos = positions.float().contiguous()
            present = torch.zeros((P,), dtype=torch.bool, device=positions.device)
            if P:
                with torch.cuda.device(positions.device):
                    view = rs.viewmatrix.to(positions.device).float().contiguous()
                    proj = rs.projmatrix.to(positions.device).float().contiguous()
                    stream = torch.cuda.current_stream(positions.device)
                    _abi.check(_abi.lib().gsr_mark_visible(P, _ptr(pos), _ptr(view), _ptr(proj), _ptr(present),
                                                           C.c_void_p(stream.cuda_stream)), "mark_visible")
        return present

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None, sh_objs=None):
        raster_settings = self.raster_settings

        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')

        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')

        if shs is None:
            shs = torch.Tensor([])
        if colors_precomp is None:
            colors_precomp = torch.Tensor([])
        if scales is None:
            scales = torch.Tensor([])
        if rotations is None:
            rotations = torch.Tensor([])
        if cov3D_precomp is None:
            cov3D_precomp = torch.Tensor([])

        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                                   raster_settings, sh_objs)
