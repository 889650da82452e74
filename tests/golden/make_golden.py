"""Generates the committed golden fixtures under tests/golden/ from /root/reference.

Run ONCE in the build container (the reference tree does not exist on the GPU box):

    python tests/golden/make_golden.py

What it writes (all small, compressed .npz):

* dibr_simple.npz / dibr_sphere.npz -- the reference's own golden tensors for the soft
  mask (tests/samples/dibr/{simple,sphere}/*.pt, "From Kaolin V0.10.0",
  consumed by tests/python/kaolin/render/mesh/test_dibr.py:77-107,158-165,277-307,359-366),
  re-packed to compact dtypes (close_face_idx converted to 0-based as the test does at
  test_dibr.py:91, stored int16; dist_type uint8).  Values are unchanged.
* model_obj.npz -- tests/samples/model.obj (960 triangles), the mesh every hot-path test
  of the reference renders (test_rasterization.py:38-41, test_dibr.py:203-206).
* pizza_obj.npz -- sample_data/meshes/pizza.obj, BASELINE.json config 1.
* rasterize_naive.npz -- outputs of the reference's pure-torch oracle
  ``_naive_deftet_sparse_render`` (kaolin/render/mesh/deftet.py:25-267), imported from
  /root/reference and run here on CPU on the fixture geometry of
  test_rasterization.py:50-134, forward (face_idx, features) and backward (grads of
  image coordinates and features for a fixed cotangent) -- the same function the
  reference's tests use as ground truth for rasterize (test_rasterization.py:148-152,216-236).
"""
import glob
import math
import os
import sys
import types

import numpy as np
import torch

REF = os.environ.get("KAOLIN_REFERENCE", "/root/reference")
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, os.pardir, os.pardir))
from workloads import read_obj, fixture_geometry  # noqa: E402


def pack_dibr(subdir):
    out = {}
    for f in sorted(glob.glob(os.path.join(REF, "tests/samples/dibr", subdir, "*.pt"))):
        name = os.path.basename(f)[:-3]
        t = torch.load(f, map_location="cpu")
        if name.startswith("close_face_idx"):
            a = (t.long() - 1).numpy().astype(np.int16)
            assert np.array_equal(a.astype(np.float64) + 1, t.double().numpy())
        elif name.startswith("close_face_dist_type"):
            a = t.numpy().astype(np.uint8)
            assert np.array_equal(a.astype(np.float64), t.double().numpy())
        elif name.startswith("new_face_idx"):
            a = t.numpy().astype(np.int16)
        else:
            a = t.double().numpy()
            a32 = a.astype(np.float32)
            if np.array_equal(a32.astype(np.float64), a):
                a = a32  # stored doubles were exactly float32 values
        out[name] = a
    np.savez_compressed(os.path.join(HERE, "dibr_%s.npz" % subdir), **out)
    print(subdir, len(out), "tensors")


def pack_obj(path, name):
    v, f, vt, ft = read_obj(path)
    d = dict(vertices=v.astype(np.float32), faces=f.astype(np.int32))
    if vt is not None and ft is not None:
        d["uvs"] = vt.astype(np.float32)
        d["face_uvs_idx"] = ft.astype(np.int32)
    np.savez_compressed(os.path.join(HERE, name), **d)
    print(name, v.shape, f.shape)


def load_naive_oracle():
    """Import _naive_deftet_sparse_render from the reference tree without the compiled
    kaolin._C (the function itself is pure torch)."""
    pkg = types.ModuleType("kaolin")
    pkg.__path__ = []
    pkg._C = types.ModuleType("kaolin._C")
    sys.modules["kaolin"] = pkg
    sys.modules["kaolin._C"] = pkg._C
    src = open(os.path.join(REF, "kaolin/render/mesh/deftet.py")).read()
    mod = types.ModuleType("kaolin_ref_deftet")
    exec(compile(src, "deftet.py", "exec"), mod.__dict__)
    return mod._naive_deftet_sparse_render


def naive_golden():
    naive = load_naive_oracle()
    H, W = 35, 31
    out = {}
    g = torch.Generator().manual_seed(0)
    for dtype, flip in [(d, f) for d in (torch.float32, torch.float64) for f in (False, True)]:
        geo = fixture_geometry(os.path.join(HERE, "model_obj.npz"), batch_size=3, flip=flip,
                               dtype=dtype)
        z, img, uvs = geo["face_vertices_z"], geo["face_vertices_image"], geo["face_uvs"]
        B = z.shape[0]
        x = (2 * torch.arange(W, dtype=z.dtype) + 1 - W) / W
        y = (H - 2 * torch.arange(H, dtype=z.dtype) - 1.) / H
        pixel_coords = torch.stack([x.reshape(1, 1, -1).repeat(B, H, 1),
                                    y.reshape(1, -1, 1).repeat(B, 1, W)],
                                   dim=-1).reshape(B, -1, 2)
        vc = geo["vertices_camera"]
        rr = torch.stack([vc[:, :, -1].min(dim=1)[0] - 1e-2,
                          vc[:, :, -1].max(dim=1)[0] + 1e-2], dim=-1)
        render_ranges = rr.unsqueeze(1).repeat(1, H * W, 1)
        min_z = z.reshape(B, -1).min(dim=1, keepdims=True)[0]
        max_z = z.reshape(B, -1).max(dim=1, keepdims=True)[0]
        valid = torch.all(z < ((min_z + max_z) / 2.).unsqueeze(-1), dim=-1)
        out["%s_flip%d_validmask" % (str(dtype)[-7:], flip)] = valid.numpy()
        feats = torch.cat([uvs, torch.ones_like(uvs[..., 1:])], dim=-1)
        for wv in (False, True):
            kw = {"valid_faces": valid} if wv else {}
            img2 = img.detach().clone().requires_grad_(True)
            f2 = feats.detach().clone().requires_grad_(True)
            gt_feat, gt_idx = naive(pixel_coords, render_ranges, z, img2, f2, 1, **kw)
            gt_feat = gt_feat.reshape(B, H, W, -1)
            grad_out = torch.rand(gt_feat.shape, generator=g, dtype=gt_feat.dtype)
            gt_feat.backward(grad_out)
            key = "%s_flip%d_valid%d" % (str(dtype)[-7:], flip, wv)
            out[key + "_idx"] = gt_idx.reshape(B, H, W).numpy().astype(np.int16)
            out[key + "_feat"] = gt_feat.detach().numpy()
            out[key + "_grad_out"] = grad_out.numpy()
            out[key + "_grad_img"] = img2.grad.numpy()
            out[key + "_grad_feat"] = f2.grad.numpy()
            print(key, "covered px:", int((gt_idx >= 0).sum()))
    np.savez_compressed(os.path.join(HERE, "rasterize_naive.npz"), **out)


if __name__ == "__main__":
    pack_obj(os.path.join(REF, "tests/samples/model.obj"), "model_obj.npz")
    pack_obj(os.path.join(REF, "sample_data/meshes/pizza.obj"), "pizza_obj.npz")
    pack_dibr("simple")
    pack_dibr("sphere")
    naive_golden()
    for f in sorted(glob.glob(os.path.join(HERE, "*.npz"))):
        print(os.path.basename(f), os.path.getsize(f) // 1024, "KiB")
