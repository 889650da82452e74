"""texture_mapping / mask_iou on the B200: the reference tests'
known-answer vectors exactly, and random inputs against the torch restatement of the
reference functions (values 1e-6, gradients 1e-5 relative)."""
import pytest
import torch

import kaolin_windows_b200 as kal
from oracle import mesh_oracle as mo
import epilogue_kats as kats

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("mode", ["nearest", "bilinear"])
def test_texture_mapping_kats(mode):
    t1, t3 = (t.to(DEV) for t in kats.texture_maps())
    tm = kal.render.mesh.texture_mapping
    out = tm(kats.sparse_coords().to(DEV), t1, mode)
    assert out.shape == (2, 4, 1) and torch.equal(out.cpu(), kats.sparse_expected(mode, 1))
    assert torch.equal(tm(kats.sparse_coords().to(DEV), t3, mode).cpu(), kats.sparse_expected(mode, 3))
    assert torch.equal(tm(kats.dense_coords().to(DEV), t3, mode).cpu(), kats.dense_expected(mode))


@pytest.mark.parametrize("mode", ["nearest", "bilinear"])
@pytest.mark.parametrize("shape", [(2, 37, 29, 3, 16, 24), (1, 256, 256, 3, 128, 128),
                                   (3, 1000, None, 5, 7, 9), (1, 5, None, 1, 1, 1)])
def test_texture_mapping_random(mode, shape):
    B, h, w, C, Ht, Wt = shape
    g = torch.Generator(device=DEV).manual_seed(h + C)
    cshape = (B, h, w, 2) if w is not None else (B, h, 2)
    coords = torch.rand(cshape, device=DEV, generator=g) * 1.3 - 0.15     # some outside [0, 1]
    tex = torch.randn((B, C, Ht, Wt), device=DEV, generator=g)
    c1, t1 = coords.clone().requires_grad_(True), tex.clone().requires_grad_(True)
    c2, t2 = coords.double().requires_grad_(True), tex.double().requires_grad_(True)
    out = kal.render.mesh.texture_mapping(c1, t1, mode)
    want = mo.texture_mapping(c2, t2, mode)
    assert out.shape == want.shape and out.is_contiguous()
    if mode == 'nearest':
        # a coordinate within float rounding of a texel boundary may pick the neighbour
        close = (out.double() - want).abs().amax(-1) <= 1e-5
        assert float(close.float().mean()) > 0.999
        cot = torch.randn(out.shape, device=DEV, generator=g) * close.unsqueeze(-1)
    else:
        assert float((out.detach().double() - want.detach()).abs().max()) <= 1e-5 * float(want.detach().abs().max())
        cot = torch.randn(out.shape, device=DEV, generator=g)
    out.backward(cot)
    want.backward(cot.double())
    assert float((t1.grad.double() - t2.grad).abs().max()) <= 1e-4 * float(t2.grad.abs().max()) + 1e-6
    if mode == 'bilinear':
        # d/d coordinate jumps at texel boundaries: compare away from them
        err = (c1.grad.double() - c2.grad).abs().amax(-1)
        ok = err <= 1e-4 * float(c2.grad.abs().max()) + 1e-6
        assert float(ok.float().mean()) > 0.999
    else:
        assert float(c1.grad.abs().max()) == 0.0


def test_mask_iou_kat_and_random():
    lhs, rhs = (t.to(DEV) for t in kats.iou_masks())
    assert torch.allclose(kal.metrics.render.mask_iou(lhs, rhs).cpu(), torch.tensor([0.3105]))
    g = torch.Generator(device=DEV).manual_seed(1)
    for B, H, W in ((1, 5, 7), (8, 512, 512), (3, 33, 1000)):
        l = torch.rand((B, H, W), device=DEV, generator=g)
        r = (torch.rand((B, H, W), device=DEV, generator=g) > 0.5).float()
        l1, r1 = l.clone().requires_grad_(True), r.clone().requires_grad_(True)
        l2, r2 = l.double().requires_grad_(True), r.double().requires_grad_(True)
        loss = kal.metrics.render.mask_iou(l1, r1)
        want = mo.mask_iou(l2, r2)
        assert abs(float(loss) - float(want)) <= 1e-5
        (loss * 3.0).backward()
        (want * 3.0).backward()
        for a, b in ((l1.grad, l2.grad), (r1.grad, r2.grad)):
            assert float((a.double() - b).abs().max()) <= 1e-4 * float(b.abs().max())


def test_epilogue_errors():
    with pytest.raises(ValueError):
        kal.render.mesh.texture_mapping(torch.zeros(1, 2, 2, device=DEV),
                                        torch.zeros(1, 1, 2, 2, device=DEV), mode="bicubic")
    with pytest.raises(RuntimeError):
        kal.render.mesh.texture_mapping(torch.zeros(1, 2, 2), torch.zeros(1, 1, 2, 2))
    with pytest.raises(AssertionError):
        kal.metrics.render.mask_iou(torch.zeros(1, 2, 2, device=DEV), torch.zeros(1, 2, 3, device=DEV))
