"""One process, two GPUs: the kernels that need more than 48 KB of dynamic shared memory
(64-pixel blocks: images with more than 1024 32-pixel bins, e.g. 1100 x 1090) opt in per
DEVICE (csrc/dibr_api.cu smem_optin; round 1 kept one flag per process, so the second GPU of
a process failed to launch).  Needs two GPUs; skipped otherwise."""
import pytest
import torch

import workloads as wl

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_render_on_second_device_with_64px_blocks():
    import kaolin_windows_b200 as kal
    g = wl.fixture_geometry(batch_size=2)
    H, W = 1100, 1090
    res = []
    for dev in ("cuda:0", "cuda:1", "cuda:0"):
        z, img, feat, nz = (g[k].to(dev) for k in ("face_vertices_z", "face_vertices_image",
                                                   "face_uvs", "face_normals_z"))
        img = img.clone().requires_grad_(True)
        out, mask, idx = kal.render.mesh.dibr_rasterization(H, W, z, img, feat, nz)
        (out.sum() + mask.sum()).backward()
        torch.cuda.synchronize(dev)
        assert idx.device == torch.device(dev) and int((idx >= 0).sum()) > 100000
        res.append((out.cpu(), mask.cpu(), idx.cpu(), img.grad.cpu()))
        # the soft-mask-only and rasterize-only entry points on that device as well
        o2, i2 = kal.render.mesh.rasterize(H, W, z, img.detach(), feat, nz >= 0)
        m2 = kal.render.mesh.dibr_soft_mask(img.detach(), i2)
        assert torch.equal(i2.cpu(), res[-1][2]) and torch.equal(m2.cpu(), res[-1][1])
    for a, b in zip(res[0][:3], res[1][:3]):
        assert torch.equal(a, b)
    # gradients are sums of float atomics: equal up to the order of the additions
    for k in (1, 2):
        assert float((res[0][3] - res[k][3]).abs().max()) <= 1e-5 * float(res[0][3].abs().max())
