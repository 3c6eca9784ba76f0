"""ctypes binding of oracle/cpu_ref.cpp -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package (diff_gaussian_rasterization) never does.

Mirrors CudaRasterizer::Rasterizer::{forward,backward,markVisible}
(DGR/cuda_rasterizer/rasterizer.h:20-85) on numpy float32 arrays.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "cpu_ref.cpp")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.orc_create.restype = C.c_void_p
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_forward.restype = C.c_int64
        L.orc_get.restype = C.c_int64
        L.orc_higher_msb.restype = C.c_uint32
        _LIB = L
    return _LIB


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, dtype=np.float32)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Oracle:
    """One rasterization state (geometry / binning / image buffers of the reference)."""

    _DT = dict(depths=np.float32, means2D=np.float32, cov3D=np.float32, rgb=np.float32, conic_opacity=np.float32,
               final_T=np.float32, radii=np.int32, tiles_touched=np.uint32, point_offsets=np.uint32,
               point_list=np.uint32, vals_unsorted=np.uint32, n_contrib=np.uint32, ranges=np.uint32,
               clamped=np.uint8, keys_unsorted=np.uint64, keys_sorted=np.uint64, t_stage=np.float64,
               n_pairs=np.int64)

    def __init__(self):
        self.h = C.c_void_p(lib().orc_create())
        self._keep = None

    def __del__(self):
        try:
            lib().orc_destroy(self.h)
        except Exception:
            pass

    def get(self, name):
        n = lib().orc_get(self.h, name.encode(), None, C.c_int64(0))
        out = np.empty(max(n, 0), dtype=self._DT[name])
        if n > 0:
            r = lib().orc_get(self.h, name.encode(), _p(out), C.c_int64(out.nbytes))
            assert r == n
        return out

    def forward(self, bg, means3D, opacities, view, proj, campos, W, H, tanfovx, tanfovy, sh_degree=0, shs=None,
                colors_precomp=None, scales=None, rotations=None, cov3D_precomp=None, scale_modifier=1.0,
                prefiltered=False, obj_feats=None):
        means3D = _f32(means3D)
        P = means3D.shape[0]
        shs, colors_precomp, scales, rotations, cov3D_precomp = map(_f32, (shs, colors_precomp, scales, rotations, cov3D_precomp))
        opacities, bg, view, proj, campos, obj_feats = map(_f32, (opacities, bg, view, proj, campos, obj_feats))
        M = 0 if shs is None else shs.shape[1]
        K = 0 if obj_feats is None else obj_feats.reshape(P, -1).shape[1]
        out_color = np.zeros((3, H, W), np.float32)
        radii = np.zeros(P, np.int32)
        out_obj = np.zeros((K, H, W), np.float32) if K else None
        self._keep = dict(P=P, M=M, K=K, D=sh_degree, W=W, H=H, bg=bg, means3D=means3D, shs=shs, colors_precomp=colors_precomp,
                          scales=scales, rotations=rotations, cov3D_precomp=cov3D_precomp, view=view, proj=proj,
                          campos=campos, tanfovx=tanfovx, tanfovy=tanfovy, scale_modifier=scale_modifier, radii=radii,
                          obj_feats=obj_feats, opacities=opacities)
        n = 0
        if P:
            n = lib().orc_forward(self.h, P, sh_degree, M, _p(bg), W, H, _p(means3D), _p(shs), _p(colors_precomp),
                                  _p(opacities), _p(scales), C.c_float(scale_modifier), _p(rotations), _p(cov3D_precomp),
                                  _p(view), _p(proj), _p(campos), C.c_float(tanfovx), C.c_float(tanfovy),
                                  int(prefiltered), _p(out_color), _p(radii), K, _p(obj_feats), _p(out_obj))
            if n < 0:
                raise RuntimeError("Point is filtered although prefiltered is set. This shouldn't happen!")
        else:
            out_color[:] = 0
        self.num_rendered = int(n)
        return (out_color, radii, out_obj) if K else (out_color, radii)

    def backward(self, dL_dpix, dL_dobjpix=None):
        k = self._keep
        P, M, K = k["P"], k["M"], k["K"]
        dL_dpix = _f32(dL_dpix)
        dL_dobjpix = _f32(dL_dobjpix)
        g = dict(means2D=np.zeros((P, 3), np.float32), conic=np.zeros((P, 2, 2), np.float32),
                 opacity=np.zeros((P, 1), np.float32), colors=np.zeros((P, 3), np.float32),
                 means3D=np.zeros((P, 3), np.float32), cov3D=np.zeros((P, 6), np.float32),
                 sh=np.zeros((P, M, 3), np.float32), scales=np.zeros((P, 3), np.float32),
                 rotations=np.zeros((P, 4), np.float32), obj=np.zeros((P, max(K, 1)), np.float32))
        if P:
            lib().orc_backward(self.h, P, k["D"], M, _p(k["bg"]), k["W"], k["H"], _p(k["means3D"]), _p(k["shs"]),
                               _p(k["colors_precomp"]), _p(k["scales"]), C.c_float(k["scale_modifier"]),
                               _p(k["rotations"]), _p(k["cov3D_precomp"]), _p(k["view"]), _p(k["proj"]),
                               _p(k["campos"]), C.c_float(k["tanfovx"]), C.c_float(k["tanfovy"]), _p(k["radii"]),
                               _p(dL_dpix), K, _p(k["obj_feats"]), _p(dL_dobjpix),
                               _p(g["means2D"]), _p(g["conic"]), _p(g["opacity"]), _p(g["colors"]), _p(g["means3D"]),
                               _p(g["cov3D"]), _p(g["sh"]), _p(g["scales"]), _p(g["rotations"]), _p(g["obj"]))
        if not K:
            g.pop("obj")
        return g


def mark_visible(means3D, view, proj):
    means3D = _f32(means3D)
    out = np.zeros(means3D.shape[0], np.uint8)
    if means3D.shape[0]:
        lib().orc_mark_visible(means3D.shape[0], _p(means3D), _p(_f32(view)), _p(_f32(proj)), _p(out))
    return out.astype(bool)


def higher_msb(n):
    return int(lib().orc_higher_msb(C.c_uint32(n)))


def num_threads():
    return int(lib().orc_num_threads())


def set_num_threads(n):
    lib().orc_set_num_threads(int(n))
