"""Fused elementwise steps of an attack iteration, through the C ABI.

PGD updates mirror attack.py:25-173 -- the targeted (descending) L-inf and L2 steps the reference applies to
`_xyz`, `_rotation`, `_opacity`, `_scaling`, `_features_dc`, `_features_rest` -- as one in-place kernel per
attribute instead of 6-9 torch elementwise kernels.  Image helpers mirror attack.py:438-443 / :538-543 (uint8 HWC
render for the detector) and detectors/detectron2_detector.py:78 (clip(x*255 + 0.5, 0, 255), differentiable).
CUDA float32 tensors only; there is no CPU path.
"""
import ctypes as C

import torch

from . import _abi

__all__ = ["pgd_linf_step_", "pgd_l2_step_", "pgd_step_", "image_to_uint8_hwc", "detector_input"]


def _chk(t, name, like=None):
    if not isinstance(t, torch.Tensor) or not t.is_cuda or t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be a float32 CUDA tensor (there is no CPU path)")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    if like is not None and t.shape != like.shape:
        raise RuntimeError(f"{name} has shape {tuple(t.shape)}, expected {tuple(like.shape)}")
    return t


def _p(t):
    return None if t is None or t.numel() == 0 else C.c_void_p(t.data_ptr())


def _stream(t):
    return C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def pgd_linf_step_(param, grad, orig, alpha, epsilon):
    """param <- clamp(param - alpha*sign(grad) - orig, -epsilon, epsilon) + orig, in place
    (gaussian_position_linf_attack and siblings, attack.py:25-52,117-133)."""
    _chk(param, "param"); _chk(grad, "grad", param); _chk(orig, "orig", param)
    with torch.no_grad(), torch.cuda.device(param.device):
        _abi.check(_abi.lib().gsr_pgd_linf_update(_p(param), _p(grad), _p(orig), param.numel(), float(alpha),
                                                  float(epsilon), _stream(param)), "pgd_linf_update")
    return param


_scratch = {}


def pgd_l2_step_(param, grad, orig, alpha, epsilon):
    """param <- param - alpha*grad/||grad||_2, then the per-Gaussian rows of (param - orig) are clipped to length
    epsilon (renorm over dim 0) and added back to orig, in place (gaussian_position_l2_attack and siblings,
    attack.py:54-115,135-173)."""
    _chk(param, "param"); _chk(grad, "grad", param); _chk(orig, "orig", param)
    rows = param.shape[0] if param.dim() else 1
    width = param.numel() // rows if rows else 0
    dev = param.device
    s = _scratch.get(dev)
    if s is None:
        s = _scratch[dev] = torch.zeros(1, dtype=torch.float64, device=dev)
    with torch.no_grad(), torch.cuda.device(dev):
        _abi.check(_abi.lib().gsr_pgd_l2_update(_p(param), _p(grad), _p(orig), rows, width, float(alpha), float(epsilon),
                                                _p(s), _stream(param)), "pgd_l2_update")
    return param


def pgd_step_(params, grads, origs, alpha, epsilon, norm="linf"):
    """Apply the same step to several attributes (dicts keyed alike), as attack.py:497-511 does per iteration."""
    fn = {"linf": pgd_linf_step_, "l2": pgd_l2_step_}[norm]
    for k, p in params.items():
        fn(p, grads[k], origs[k], alpha, epsilon)


def image_to_uint8_hwc(image):
    """(3,H,W) float render -> (H,W,3) uint8: (clamp(x,0,1)*255).byte().permute(1,2,0) (attack.py:438-443)."""
    _chk(image, "image")
    if image.dim() != 3 or image.shape[0] != 3:
        raise RuntimeError("image must be (3, H, W)")
    H, W = int(image.shape[1]), int(image.shape[2])
    out = torch.empty((H, W, 3), dtype=torch.uint8, device=image.device)
    with torch.cuda.device(image.device):
        _abi.check(_abi.lib().gsr_image_to_u8_hwc(_p(image), _p(out), W, H, _stream(image)), "image_to_u8_hwc")
    return out


class _DetectorInput(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        xc = _chk(x.contiguous(), "x")
        y = torch.empty_like(xc)
        with torch.cuda.device(xc.device):
            _abi.check(_abi.lib().gsr_detector_scale(_p(xc), _p(y), None, None, xc.numel(), _stream(xc)), "detector_scale")
        ctx.save_for_backward(xc)
        return y

    @staticmethod
    def backward(ctx, dy):
        (xc,) = ctx.saved_tensors
        dyc = _chk(dy.contiguous(), "dy", xc)
        dx = torch.empty_like(xc)
        with torch.cuda.device(xc.device):
            _abi.check(_abi.lib().gsr_detector_scale(_p(xc), None, _p(dyc), _p(dx), xc.numel(), _stream(xc)), "detector_scale")
        return dx


def detector_input(x):
    """clip(x*255 + 0.5, 0, 255) with its gradient (detectors/detectron2_detector.py:78), one kernel each way."""
    return _DetectorInput.apply(x)
