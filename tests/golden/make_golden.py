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
* prepare_vertices.npz -- outputs and gradients of the reference's own (pure torch)
  ``prepare_vertices`` (kaolin/render/mesh/utils.py:128-175, with
  camera/legacy.py:22-37,120-139, ops/mesh/mesh.py:25-47, ops/mesh/trianglemesh.py:313-337),
  functions extracted from /root/reference and run here on CPU on model.obj with 2 random
  look-at cameras, both camera parametrisations, fixed random cotangents.
* rasterize_naive.npz -- outputs of the reference's pure-torch oracle
  ``_naive_deftet_sparse_render`` (kaolin/render/mesh/deftet.py:25-267), imported from
  /root/reference and run here on CPU on the fixture geometry of
  test_rasterization.py:50-134, forward (face_idx, features) and backward (grads of
  image coordinates and features for a fixed cotangent) -- the same function the
  reference's tests use as ground truth for rasterize (test_rasterization.py:148-152,216-236).
* deftet_naive.npz -- the same function as the reference's deftet tests use it
  (test_deftet.py:340-437), see deftet_golden().
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


def deftet_golden():
    """deftet_naive.npz: the reference's pure-torch _naive_deftet_sparse_render run as the
    reference's own test does (test_deftet.py:340-437): model.obj, the three look-at cameras,
    uniform random sample points in [-1, 1]^2, ranges up to the centre / the far end,
    knum 20 (never exceeded: beyond knum hits the naive function keeps the NEAREST while the
    CUDA op keeps the FIRST in face order -- the reference's tests never compare them there),
    forward and backward."""
    naive = load_naive_oracle()
    out = {}
    g = torch.Generator().manual_seed(5)
    P = 257
    for dtype in (torch.float32, torch.float64):
        geo = fixture_geometry(os.path.join(HERE, "model_obj.npz"), batch_size=3, dtype=dtype)
        z, img, uvs = geo["face_vertices_z"], geo["face_vertices_image"], geo["face_uvs"]
        vc = geo["vertices_camera"]
        B = z.shape[0]
        pix = torch.rand((B, P, 2), generator=g, dtype=dtype) * 2. - 1.
        tag = str(dtype)[-7:]
        out[tag + "_pixel_coords"] = pix.numpy()
        min_z = vc[:, :, -1].min(dim=1)[0]
        max_z = vc[:, :, -1].max(dim=1)[0]
        for centre in (False, True):
            lo = (min_z + max_z) / 2. if centre else min_z
            rr = torch.nn.functional.pad(lo.unsqueeze(-1), (0, 1), value=0.)
            rr = rr.unsqueeze(1).repeat(1, P, 1)
            for knum in (20,):
                img2 = img.detach().clone().requires_grad_(True)
                f2 = uvs.detach().clone().requires_grad_(True)
                feat, idx = naive(pix, rr, z, img2, f2, knum)
                grad_out = torch.rand(feat.shape, generator=g, dtype=dtype)
                feat.backward(grad_out)
                key = "%s_centre%d_k%d" % (tag, centre, knum)
                out[key + "_ranges"] = rr.numpy()
                out[key + "_idx"] = idx.numpy().astype(np.int16)
                out[key + "_feat"] = feat.detach().numpy()
                out[key + "_grad_out"] = grad_out.numpy()
                out[key + "_grad_img"] = img2.grad.numpy()
                out[key + "_grad_feat"] = f2.grad.numpy()
                print(key, "hits:", int((idx >= 0).sum()),
                      "max per point:", int((idx >= 0).sum(-1).max()))
    np.savez_compressed(os.path.join(HERE, "deftet_naive.npz"), **out)


def load_ref_functions(relpath, names, globs):
    """exec the named top-level functions of a reference source file (pure torch code whose
    package cannot be imported without the compiled kaolin._C) in `globs`."""
    import ast
    src = open(os.path.join(REF, relpath)).read()
    tree = ast.parse(src)
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in names:
            code = compile(ast.Module(body=[node], type_ignores=[]), relpath, "exec")
            exec(code, globs)
    return [globs[n] for n in names]


def prepare_vertices_golden():
    cam = types.SimpleNamespace()
    g_cam = {"torch": torch, "tan": math.tan}
    (cam.rotate_translate_points, cam.perspective_camera, gen_rt, gen_proj, gen_tf) = \
        load_ref_functions("kaolin/render/camera/legacy.py",
                           ["rotate_translate_points", "perspective_camera",
                            "generate_rotate_translate_matrices",
                            "generate_perspective_projection",
                            "generate_transformation_matrix"], g_cam)
    mesh = types.SimpleNamespace()
    (mesh.index_vertices_by_faces,) = load_ref_functions(
        "kaolin/ops/mesh/mesh.py", ["index_vertices_by_faces"], {"torch": torch})
    (mesh.face_normals,) = load_ref_functions(
        "kaolin/ops/mesh/trianglemesh.py", ["face_normals"], {"torch": torch})
    (prepare_vertices,) = load_ref_functions(
        "kaolin/render/mesh/utils.py", ["prepare_vertices"],
        {"torch": torch, "camera": cam, "ops": types.SimpleNamespace(mesh=mesh)})
    d = np.load(os.path.join(HERE, "model_obj.npz"))
    g = torch.Generator().manual_seed(11)
    B = 2
    verts = torch.from_numpy(d["vertices"]).float()
    verts = (verts - verts.mean(0)) / verts.abs().max()
    vertices = verts.unsqueeze(0).repeat(B, 1, 1)
    vertices = vertices + 0.01 * torch.randn(vertices.shape, generator=g)
    faces = torch.from_numpy(d["faces"]).long()
    pos = torch.tensor([[0.3, 0.5, 2.5], [-1.8, 0.2, 1.9]])
    look = torch.zeros(B, 3)
    up = torch.tensor([[0., 1., 0.]]).repeat(B, 1)
    rot, trans = gen_rt(pos, look, up)
    transform = gen_tf(pos, look, up)
    proj = gen_proj(math.pi / 4, 1.0)
    out = dict(vertices=vertices.numpy(), faces=faces.numpy().astype(np.int32),
               camera_rot=rot.numpy(), camera_trans=trans.numpy(),
               camera_transform=transform.numpy(), camera_proj=proj.numpy())
    F = faces.shape[0]
    cot = [torch.randn((B, F, 3, 3), generator=g), torch.randn((B, F, 3, 2), generator=g),
           torch.randn((B, F, 3), generator=g)]
    for i, c in enumerate(cot):
        out["cot%d" % i] = c.numpy()
    for mode in ("rot_trans", "transform"):
        v = vertices.clone().requires_grad_(True)
        pj = proj.clone().requires_grad_(True)
        if mode == "rot_trans":
            r, t = rot.clone().requires_grad_(True), trans.clone().requires_grad_(True)
            res = prepare_vertices(v, faces, pj, camera_rot=r, camera_trans=t)
            leaves = {"vertices": v, "camera_proj": pj, "camera_rot": r, "camera_trans": t}
        else:
            tf = transform.clone().requires_grad_(True)
            res = prepare_vertices(v, faces, pj, camera_transform=tf)
            leaves = {"vertices": v, "camera_proj": pj, "camera_transform": tf}
        sum((o * c).sum() for o, c in zip(res, cot)).backward()
        for name, o in zip(("fvc", "fvi", "fn"), res):
            out["%s_%s" % (mode, name)] = o.detach().numpy()
        for name, leaf in leaves.items():
            out["%s_grad_%s" % (mode, name)] = leaf.grad.numpy()
    np.savez_compressed(os.path.join(HERE, "prepare_vertices.npz"), **out)
    print("prepare_vertices", {k: v.shape for k, v in out.items() if k.startswith("rot_trans")})


if __name__ == "__main__":
    if sys.argv[1:] == ["prepare_vertices"]:
        prepare_vertices_golden()
        sys.exit(0)
    if sys.argv[1:] == ["deftet"]:
        deftet_golden()
        sys.exit(0)
    pack_obj(os.path.join(REF, "tests/samples/model.obj"), "model_obj.npz")
    pack_obj(os.path.join(REF, "sample_data/meshes/pizza.obj"), "pizza_obj.npz")
    pack_dibr("simple")
    pack_dibr("sphere")
    naive_golden()
    prepare_vertices_golden()
    deftet_golden()
    for f in sorted(glob.glob(os.path.join(HERE, "*.npz"))):
        print(os.path.basename(f), os.path.getsize(f) // 1024, "KiB")
