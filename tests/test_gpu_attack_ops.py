"""GPU tests of the fused attack-iteration kernels (attack_ops) against the torch op sequences of the reference
(attack.py:25-173, :438-443; detectors/detectron2_detector.py:78), restated here with plain torch."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref_linf(p, g, p0, alpha, eps):  # attack.py:25-31
    p = p.clone()
    eta = alpha * torch.sign(g)
    eta.mul_(-1)
    p.add_(eta)
    p.sub_(p0).clamp_(-eps, eps).add_(p0)
    return p


def _ref_l2(p, g, p0, alpha, eps):  # attack.py:54-71
    p = p.clone()
    norm = torch.norm(g.view(-1), p=2)
    eta = alpha * (g / norm) if norm > 0 else torch.zeros_like(g)
    eta.mul_(-1)
    p.add_(eta)
    delta = (p - p0).renorm(p=2, dim=0, maxnorm=eps)
    p.copy_(p0 + delta)
    return p


@pytest.mark.parametrize("shape", [(1000, 3), (777, 4), (5001, 1), (300, 15, 3), (129, 1, 3)])
def test_pgd_linf_is_bit_exact_vs_torch_sequence(shape):
    from diff_gaussian_rasterization import attack_ops as ao
    g0 = torch.Generator(device="cuda").manual_seed(1)
    p0 = torch.randn(shape, device="cuda", generator=g0)
    p = p0 + 0.02 * torch.randn(shape, device="cuda", generator=g0)
    g = torch.randn(shape, device="cuda", generator=g0)
    g.view(-1)[::7] = 0.0  # sign(0) = 0: no step
    want = _ref_linf(p, g, p0, 0.01, 0.03)
    got = ao.pgd_linf_step_(p.clone(), g, p0, 0.01, 0.03)
    assert torch.equal(got, want)
    assert float((got - p0).abs().max()) <= 0.03 + 1e-6


@pytest.mark.parametrize("shape", [(1000, 3), (777, 4), (5001, 1), (300, 15, 3), (129, 1, 3)])
def test_pgd_l2_matches_torch_sequence(shape):
    from diff_gaussian_rasterization import attack_ops as ao
    g0 = torch.Generator(device="cuda").manual_seed(2)
    p0 = torch.randn(shape, device="cuda", generator=g0)
    p = p0 + 0.05 * torch.randn(shape, device="cuda", generator=g0)
    g = torch.randn(shape, device="cuda", generator=g0)
    for alpha, eps in ((0.5, 0.08), (5.0, 0.01)):
        want = _ref_l2(p, g, p0, alpha, eps)
        got = ao.pgd_l2_step_(p.clone(), g, p0, alpha, eps)
        assert torch.allclose(got, want, rtol=1e-5, atol=1e-6)
        rows = (got - p0).reshape(shape[0], -1).norm(dim=1)
        assert float(rows.max()) <= eps * (1 + 1e-4)
    # zero gradient: no step, only the projection
    z = torch.zeros_like(g)
    assert torch.allclose(ao.pgd_l2_step_(p.clone(), z, p0, 0.5, 0.08), _ref_l2(p, z, p0, 0.5, 0.08), rtol=1e-5, atol=1e-6)


def test_image_to_uint8_hwc_and_detector_input():
    from diff_gaussian_rasterization import attack_ops as ao
    g0 = torch.Generator(device="cuda").manual_seed(3)
    img = torch.rand((3, 67, 93), device="cuda", generator=g0) * 1.4 - 0.2  # values outside [0,1] too
    want = (torch.clamp(img, min=0, max=1.0) * 255).byte().permute(1, 2, 0).contiguous()
    got = ao.image_to_uint8_hwc(img)
    assert got.dtype == torch.uint8 and got.shape == (67, 93, 3) and torch.equal(got, want)
    x = img.clone().requires_grad_(True)
    xr = img.clone().requires_grad_(True)
    w = torch.randn_like(img)
    y = ao.detector_input(x)
    yr = torch.clip(xr * 255 + 0.5, 0, 255)
    assert torch.equal(y, yr)
    (y * w).sum().backward()
    (yr * w).sum().backward()
    assert torch.equal(x.grad, xr.grad)


def test_attack_ops_reject_cpu_and_mismatched_tensors():
    from diff_gaussian_rasterization import attack_ops as ao
    a = torch.zeros(4, 3)
    with pytest.raises(RuntimeError):
        ao.pgd_linf_step_(a, a, a, 0.1, 0.1)
    b = torch.zeros(4, 3, device="cuda")
    with pytest.raises(RuntimeError):
        ao.pgd_l2_step_(b, torch.zeros(4, 4, device="cuda"), b, 0.1, 0.1)
    with pytest.raises(RuntimeError):
        ao.image_to_uint8_hwc(torch.zeros(4, 5, 5, device="cuda"))
