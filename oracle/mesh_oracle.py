"""TEST INFRASTRUCTURE ONLY -- CPU restatement (plain torch, any float dtype) of the
reference steps next to the DIB-R path.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline may import this; the product never does.

Parity: PINNED by tests/golden/prepare_vertices.npz (outputs and gradients of the
reference's own functions, extracted from /root/reference and run by
tests/golden/make_golden.py) -- see tests/test_oracle_golden.py.
"""
import torch


def rotate_translate_points(points, camera_rot, camera_trans):
    """kaolin/render/camera/legacy.py:22-37 -- P_new = R (P_old - T)."""
    return torch.matmul(points - camera_trans.view(-1, 1, 3), camera_rot.permute(0, 2, 1))


def perspective_camera(points, camera_proj):
    """kaolin/render/camera/legacy.py:120-139."""
    p = points * camera_proj.view(-1, 1, 3)
    return p[:, :, :2] / p[:, :, 2:3]


def index_vertices_by_faces(vertices_features, faces):
    """kaolin/ops/mesh/mesh.py:25-47 -- (B,V,C), (F,3) -> (B,F,3,C) gather."""
    B, V, C = vertices_features.shape
    src = vertices_features.unsqueeze(2).expand(-1, -1, faces.shape[-1], -1)
    idx = faces[None, ..., None].expand(B, -1, -1, C)
    return torch.gather(src, 1, idx)


def face_normals(face_vertices, unit=False):
    """kaolin/ops/mesh/trianglemesh.py:313-337."""
    e0 = face_vertices[:, :, 1] - face_vertices[:, :, 0]
    e1 = face_vertices[:, :, 2] - face_vertices[:, :, 0]
    n = torch.cross(e0, e1, dim=2)
    if unit:
        n = n / (n.norm(dim=2, keepdim=True) + 1e-10)
    return n


def prepare_vertices(vertices, faces, camera_proj, camera_rot=None, camera_trans=None,
                     camera_transform=None):
    """kaolin/render/mesh/utils.py:128-175."""
    if camera_transform is None:
        vertices_camera = rotate_translate_points(vertices, camera_rot, camera_trans)
    else:
        padded = torch.nn.functional.pad(vertices, (0, 1), mode='constant', value=1.)
        vertices_camera = padded @ camera_transform
    vertices_image = perspective_camera(vertices_camera, camera_proj)
    fvc = index_vertices_by_faces(vertices_camera, faces)
    fvi = index_vertices_by_faces(vertices_image, faces)
    return fvc, fvi, face_normals(fvc, unit=True)


def texture_mapping(texture_coordinates, texture_maps, mode='nearest'):
    """kaolin/render/mesh/utils.py:23-76 (any device, any float dtype)."""
    batch_size = texture_coordinates.shape[0]
    num_channels = texture_maps.shape[1]
    c = texture_coordinates.reshape(batch_size, -1, 1, 2)
    c = torch.clamp(c, 0., 1.) * 2 - 1                       # [0, 1] -> [-1, 1]
    c = torch.stack([c[..., 0], -c[..., 1]], dim=-1)         # y axis bottom-to-top -> top-to-bottom
    out = torch.nn.functional.grid_sample(texture_maps, c, mode=mode, align_corners=False,
                                          padding_mode='border')
    out = out.permute(0, 2, 3, 1)
    return out.reshape(batch_size, *texture_coordinates.shape[1:-1], num_channels)


def mask_iou(lhs_mask, rhs_mask):
    """kaolin/metrics/render.py:18-40."""
    batch_size = lhs_mask.shape[0]
    sil_mul = lhs_mask * rhs_mask
    sil_add = lhs_mask + rhs_mask
    iou_up = torch.sum(sil_mul.reshape(batch_size, -1), dim=1)
    iou_down = torch.sum((sil_add - sil_mul).reshape(batch_size, -1), dim=1)
    return 1.0 - torch.mean(iou_up / (iou_down + 1e-10))
