"""Synthetic workloads and fixture geometry for the DIB-R hot path (tests + bench).

Pure torch/numpy, device-agnostic; nothing here is on the product path.  The camera and
mesh helpers restate the few lines of pure-torch caller code that sit upstream of the
hot path in the reference (cited per function) so tests and bench.py can build
``face_vertices_z / face_vertices_image / face_features / face_normals_z`` of the shapes
BASELINE.json names without importing kaolin.
"""
import math
import os

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(ROOT, "tests", "golden")


# ----------------------------------------------------------------------------- OBJ
def read_obj(path):
    """Minimal OBJ reader: v / vt / f (triangles; polygons are fan-triangulated).
    Returns (vertices (V,3) f64, faces (F,3) i64, uvs (T,2) or None, face_uvs_idx or None)."""
    vs, vts, fs, fts = [], [], [], []
    with open(path) as fh:
        for line in fh:
            p = line.split()
            if not p:
                continue
            if p[0] == "v":
                vs.append([float(a) for a in p[1:4]])
            elif p[0] == "vt":
                vts.append([float(a) for a in p[1:3]])
            elif p[0] == "f":
                vi, ti = [], []
                for tok in p[1:]:
                    q = tok.split("/")
                    vi.append(int(q[0]) - 1)
                    ti.append(int(q[1]) - 1 if len(q) > 1 and q[1] else -1)
                for k in range(1, len(vi) - 1):
                    fs.append([vi[0], vi[k], vi[k + 1]])
                    fts.append([ti[0], ti[k], ti[k + 1]])
    v = np.asarray(vs, dtype=np.float64)
    f = np.asarray(fs, dtype=np.int64)
    vt = np.asarray(vts, dtype=np.float64) if vts else None
    ft = np.asarray(fts, dtype=np.int64) if vts and (np.asarray(fts) >= 0).all() else None
    return v, f, vt, ft


# -------------------------------------------------------------------------- cameras
def generate_perspective_projection(fovyangle, ratio=1.0, dtype=torch.float):
    """kaolin/render/camera/legacy.py:142-159."""
    tanfov = math.tan(fovyangle / 2.0)
    return torch.tensor([[1.0 / (ratio * tanfov)], [1.0 / tanfov], [-1]], dtype=dtype)


def generate_rotate_translate_matrices(camera_position, look_at, camera_up_direction):
    """kaolin/render/camera/legacy.py:39-80."""
    camz = look_at - camera_position
    camz = camz / (camz.norm(dim=1, keepdim=True) + 1e-10)
    if camera_up_direction.shape[0] < camz.shape[0]:
        camera_up_direction = camera_up_direction.repeat(camz.shape[0], 1)
    camx = torch.cross(camz, camera_up_direction, dim=1)
    camx = camx / (camx.norm(dim=1, keepdim=True) + 1e-10)
    camy = torch.cross(camx, camz, dim=1)
    camy = camy / (camy.norm(dim=1, keepdim=True) + 1e-10)
    return torch.stack([camx, camy, -camz], dim=1), camera_position


def rotate_translate_points(points, camera_rot, camera_trans):
    """kaolin/render/camera/legacy.py:22-37."""
    return torch.matmul(points - camera_trans.view(-1, 1, 3), camera_rot.permute(0, 2, 1))


def perspective_camera(points, camera_proj):
    """kaolin/render/camera/legacy.py:120-139."""
    projected = points * camera_proj.view(-1, 1, 3)
    return projected[:, :, :2] / projected[:, :, 2:3]


def index_vertices_by_faces(vertices_features, faces):
    """kaolin/ops/mesh/mesh.py:25-47 (a gather)."""
    return vertices_features[:, faces.long()]


def face_normals(face_vertices, unit=False):
    """kaolin/ops/mesh/trianglemesh.py:313-337."""
    e0 = face_vertices[:, :, 1] - face_vertices[:, :, 0]
    e1 = face_vertices[:, :, 2] - face_vertices[:, :, 0]
    n = torch.cross(e0, e1, dim=2)
    if unit:
        n = n / (n.norm(dim=2, keepdim=True) + 1e-10)
    return n


def prepare_vertices(vertices, faces, camera_proj, camera_rot, camera_trans):
    """kaolin/render/mesh/utils.py:128-175."""
    vertices_camera = rotate_translate_points(vertices, camera_rot, camera_trans)
    vertices_image = perspective_camera(vertices_camera, camera_proj)
    face_vertices_camera = index_vertices_by_faces(vertices_camera, faces)
    face_vertices_image = index_vertices_by_faces(vertices_image, faces)
    return face_vertices_camera, face_vertices_image, face_normals(face_vertices_camera, unit=True)


# ------------------------------------------------- the reference's test-fixture recipe
def fixture_geometry(npz_path=None, batch_size=3, flip=False, dtype=torch.float32):
    """Geometry of the reference's hot-path tests (test_rasterization.py:38-103,
    test_dibr.py:203-261): model.obj min-max normalised to [0,1]^3, look-at cameras at
    (0.5,0.5,3), (2,2,-2), (3,0.5,0.5) towards (0.5,0.5,0.5), up +Y, fovy pi/4."""
    d = np.load(npz_path or os.path.join(GOLDEN, "model_obj.npz"))
    vertices = torch.from_numpy(d["vertices"]).to(dtype).unsqueeze(0)
    faces = torch.from_numpy(d["faces"]).long()
    face_uvs_idx = torch.from_numpy(d["face_uvs_idx"]).long()
    uvs = torch.from_numpy(d["uvs"]).to(dtype).unsqueeze(0)
    if flip:
        faces = torch.flip(faces, dims=(-1,))
        face_uvs_idx = torch.flip(face_uvs_idx, dims=(-1,))
    camera_pos = torch.tensor([[0.5, 0.5, 3.], [2., 2., -2.], [3., 0.5, 0.5]],
                              dtype=dtype)[:batch_size]
    look_at = torch.full((batch_size, 3), 0.5, dtype=dtype)
    camera_up = torch.tensor([[0., 1., 0.]], dtype=dtype).repeat(batch_size, 1)
    camera_proj = generate_perspective_projection(math.pi / 4., dtype=dtype)
    vmin = vertices.min(dim=1, keepdim=True)[0]
    vmax = vertices.max(dim=1, keepdim=True)[0]
    vertices = (vertices - vmin) / (vmax - vmin)
    rot, trans = generate_rotate_translate_matrices(camera_pos, look_at, camera_up)
    vertices_camera = rotate_translate_points(vertices, rot, trans)
    vertices_image = perspective_camera(vertices_camera, camera_proj)
    face_vertices_camera = index_vertices_by_faces(vertices_camera, faces)
    return dict(
        faces=faces,
        vertices_camera=vertices_camera,
        face_vertices_z=face_vertices_camera[..., 2].contiguous(),
        face_vertices_image=index_vertices_by_faces(vertices_image, faces).contiguous(),
        face_uvs=index_vertices_by_faces(uvs, face_uvs_idx).repeat(batch_size, 1, 1, 1).contiguous(),
        face_normals_z=face_normals(face_vertices_camera, unit=True)[..., 2].contiguous(),
    )


def simple_soft_mask_fixture(dtype=torch.float32):
    """The three hand-made triangles x 2 batches of test_dibr.py:43-62."""
    img = torch.tensor(
        [[[[-0.7, 0.], [0., -0.7], [0., 0.7]],
          [[-0.7, 0.], [0., 0.7], [0., -0.7]],
          [[0., -0.7], [0., 0.7], [0.7, 0.]]],
         [[[-0.7, -0.7], [0.7, -0.7], [-0.7, 0.7]],
          [[-0.7, -0.7], [0.7, -0.7], [-0.7, 0.7]],
          [[-0.7, -0.7], [0.7, -0.7], [-0.7, 0.7]]]], dtype=dtype)
    z = torch.tensor(
        [[[-2., -1., -1.], [-2.5, -3., -3.], [-2., -2., -2.]],
         [[-2., -1., -3.], [-2., -2., -2.], [-2., -3., -1.]]], dtype=dtype)
    return img, z


# ------------------------------------------------------------------ synthetic meshes
def grid_sphere(nu, nv, noise=0.0, seed=0):
    """Lat-long sphere with nu x nv quads -> 2*nu*nv - 2*nu triangles (the two pole rows
    are fans), radius 1 with optional radial noise N(0, noise).  Returns (V,3) f32, (F,3) i64."""
    rng = np.random.RandomState(seed)
    theta = np.linspace(0.0, np.pi, nv + 1)[1:-1]          # nv-1 interior rings
    phi = np.linspace(0.0, 2 * np.pi, nu, endpoint=False)
    ring = np.stack([np.outer(np.sin(theta), np.cos(phi)),
                     np.outer(np.cos(theta), np.ones_like(phi)),
                     np.outer(np.sin(theta), np.sin(phi))], axis=-1).reshape(-1, 3)
    verts = np.concatenate([[[0., 1., 0.]], ring, [[0., -1., 0.]]], axis=0)
    if noise > 0:
        verts = verts * (1.0 + rng.normal(0.0, noise, size=(verts.shape[0], 1)))
    faces = []
    top, bot = 0, verts.shape[0] - 1
    idx = lambda r, c: 1 + r * nu + (c % nu)
    c = np.arange(nu)
    faces.append(np.stack([np.full(nu, top), idx(0, c + 1), idx(0, c)], axis=1))
    for r in range(nv - 2):
        a, b, cc, d = idx(r, c), idx(r, c + 1), idx(r + 1, c), idx(r + 1, c + 1)
        faces.append(np.stack([a, b, cc], axis=1))
        faces.append(np.stack([b, d, cc], axis=1))
    faces.append(np.stack([np.full(nu, bot), idx(nv - 2, c), idx(nv - 2, c + 1)], axis=1))
    return verts.astype(np.float32), np.concatenate(faces, axis=0).astype(np.int64)


def circle_cameras(n, radius=2.5, seed=0, dtype=torch.float32):
    """look-at cameras on a circle of the given radius around the origin,
    azimuth evenly spaced, elevation ~U(-30deg, 30deg), up +Y, fovy pi/4."""
    rng = np.random.RandomState(seed)
    az = np.arange(n) * (2 * np.pi / n) + 0.1
    el = np.deg2rad(rng.uniform(-30.0, 30.0, size=n))
    pos = np.stack([radius * np.cos(el) * np.sin(az), radius * np.sin(el),
                    radius * np.cos(el) * np.cos(az)], axis=1)
    pos = torch.tensor(pos, dtype=dtype)
    rot, trans = generate_rotate_translate_matrices(
        pos, torch.zeros(n, 3, dtype=dtype), torch.tensor([[0., 1., 0.]], dtype=dtype))
    return rot, trans, generate_perspective_projection(math.pi / 4., dtype=dtype)


def render_inputs(vertices, faces, n_views, vertex_features=None, seed=0, device="cpu",
                  dtype=torch.float32, view_slice=None):
    """vertices (V,3) or (B,V,3) -> dict of hot-path inputs for n_views circle cameras.
    view_slice=(lo, hi) keeps only those views (rank sharding)."""
    v = torch.as_tensor(vertices, dtype=dtype)
    if v.dim() == 2:
        v = v.unsqueeze(0).expand(n_views, -1, -1)
    f = torch.as_tensor(faces).long()
    rot, trans, proj = circle_cameras(n_views, seed=seed, dtype=dtype)
    if vertex_features is None:
        g = torch.Generator().manual_seed(seed)
        vertex_features = torch.rand(v.shape[1], 3, generator=g, dtype=dtype)
    vf = torch.as_tensor(vertex_features, dtype=dtype)
    if vf.dim() == 2:
        vf = vf.unsqueeze(0).expand(n_views, -1, -1)
    if view_slice is not None:
        lo, hi = view_slice
        v, rot, trans, vf = v[lo:hi], rot[lo:hi], trans[lo:hi], vf[lo:hi]
    v, f, rot, trans, proj, vf = (t.to(device) for t in (v, f, rot, trans, proj, vf))
    fvc, fvi, fn = prepare_vertices(v, f, proj, rot, trans)
    vcam = rotate_translate_points(v, rot, trans)
    return dict(face_vertices_z=fvc[..., 2].contiguous(),
                face_vertices_image=fvi.contiguous(),
                face_features=index_vertices_by_faces(vf, f).contiguous(),
                face_normals_z=fn[..., 2].contiguous(),
                faces=f, vertices=v, vertex_features=vf,
                # mesh-level inputs (dibr_rasterization_mesh): the same camera pipeline
                vertices_camera=vcam.contiguous(),
                vertices_image=perspective_camera(vcam, proj).contiguous(),
                camera=(rot, trans, proj))


# BASELINE.json configs.  F is what the generator actually produces.
CONFIGS = {
    # name: (views, mesh spec, H, W)
    "cfg1_pizza_256": dict(views=1, mesh=("obj", "pizza_obj.npz"), H=256, W=256),
    "cfg2_70k_8x512": dict(views=8, mesh=("sphere", 187, 188, 0.02), H=512, W=512),
    "cfg3_200k_64x1024": dict(views=64, mesh=("sphere", 316, 317, 0.0), H=1024, W=1024),
    "cfg4_1m_16x2048": dict(views=16, mesh=("sphere", 1000, 501, 0.0), H=2048, W=2048),
    "cfg5_256x2k_128": dict(views=256, mesh=("sphere_batch", 32, 32, 0.05), H=128, W=128),
}


def make_config(name, device="cpu", views=None, view_slice=None, seed=0):
    """Build the hot-path inputs of one BASELINE.json config.  `views` overrides the
    number of views (parity tests use fewer); view_slice shards them across ranks."""
    c = CONFIGS[name]
    n = views or c["views"]
    spec = c["mesh"]
    if spec[0] == "obj":
        d = np.load(os.path.join(GOLDEN, spec[1]))
        v = d["vertices"].astype(np.float64)
        v = v - (v.max(0) + v.min(0)) / 2
        v = (v / np.linalg.norm(v, axis=1).max()).astype(np.float32)
        out = render_inputs(v, d["faces"], n, seed=seed, device=device, view_slice=view_slice)
    elif spec[0] == "sphere":
        v, f = grid_sphere(spec[1], spec[2], noise=spec[3], seed=seed)
        out = render_inputs(v, f, n, seed=seed, device=device, view_slice=view_slice)
    else:  # one distinct mesh per view
        v0, f = grid_sphere(spec[1], spec[2], noise=0.0, seed=seed)
        rng = np.random.RandomState(seed)
        disp = 1.0 + rng.normal(0.0, spec[3], size=(n, v0.shape[0], 1)).astype(np.float32)
        out = render_inputs(v0[None] * disp, f, n, seed=seed, device=device,
                            view_slice=view_slice)
    out["H"], out["W"] = c["H"], c["W"]
    return out


def raster_pair_count(H, W, face_vertices_image, valid_faces, multiplier=1000.0):
    """N_pair of SURVEY 8(d): the number of (pixel centre, valid face) pairs that pass the
    reference's half-open box test (rasterization_cuda.cu:115) -- tiling independent, the
    unit of the secondary (FP32) roofline.  Closed form per face: the pixel-centre
    coordinates are formed exactly as the kernel forms them (:85-86) and counted against the
    float box of the x multiplier coordinates (rasterization.py:337-344) with searchsorted.
    Checked against the oracle's own pair counter in tests/test_oracle_golden.py."""
    img = face_vertices_image.detach().float() * multiplier
    dev = img.device
    # (the step as an IEEE float division, with numpy: torch divides a tensor by a host
    # scalar as a multiplication by the rounded reciprocal, 1 ulp off when W is not 2^k)
    xstep = float(np.float32(multiplier) / np.float32(W))
    ystep = float(np.float32(multiplier) / np.float32(H))
    xs = torch.arange(1 - W, W + 1, 2, device=dev, dtype=torch.int32).float() * xstep
    ys = torch.arange(1 - H, H + 1, 2, device=dev, dtype=torch.int32).float() * ystep  # ascending
    lo = img.min(dim=2)[0]
    hi = img.max(dim=2)[0]
    # pixels with lo <= c < hi  ->  searchsorted(c, hi, left) - searchsorted(c, lo, left)
    nx = torch.searchsorted(xs, hi[..., 0].contiguous()) - torch.searchsorted(xs, lo[..., 0].contiguous())
    ny = torch.searchsorted(ys, hi[..., 1].contiguous()) - torch.searchsorted(ys, lo[..., 1].contiguous())
    n = nx.clamp_min(0).long() * ny.clamp_min(0).long()
    if valid_faces is not None:
        n = n * valid_faces.to(n.dtype)
    return int(n.sum().item())
