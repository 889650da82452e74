"""Known-answer vectors of the reference's own deftet tests
(tests/python/kaolin/render/mesh/test_deftet.py:36-238, TestSimpleDeftetSparseRender),
restated as data: two meshes of three hand-made triangles, seven sample points each."""
import torch


def inputs(dtype=torch.float32):
    img = torch.tensor(
        [[[[-1., 0.], [0., -1.], [0., 1.]],
          [[-1., 0.], [0., 1.], [0., -1.]],
          [[0., -1.], [0., 1.], [1., 0.]]],
         [[[-1., -1.], [1., -1.], [-1., 1.]],
          [[-1., -1.], [1., -1.], [-1., 1.]],
          [[-1., -1.], [1., -1.], [-1., 1.]]]], dtype=dtype)
    z = torch.tensor(
        [[[-2., -1., -1.], [-2.5, -3., -3.], [-2., -2., -2.]],
         [[-2., -1., -3.], [-2., -2., -2.], [-2., -3., -1.]]], dtype=dtype)
    per_face = torch.arange(6, dtype=dtype).reshape(2, 3, 1, 1).repeat(1, 1, 3, 1)
    per_vertex = torch.arange(18, dtype=dtype).reshape(2, 3, 3, 1)
    pix = torch.tensor(
        [[[-0.999, 0.], [-0.001, -0.998], [0.001, 0.998], [0.999, 0.],
          [-0.45, 0.], [0.45, 0.], [-0.999, -0.999]],
         [[-0.998, -0.999], [0.998, -0.999], [-0.999, 0.998],
          [-0.001, -0.], [0., -0.999], [-0.999, 0.], [0.001, 0.001]]], dtype=dtype)
    return pix, z, img, [per_face, per_vertex]


def ranges(lo, dtype=torch.float32):
    return torch.tensor([[[lo, 0.]]], dtype=dtype).repeat(2, 7, 1)


def _pad(rows, k=5, fill=-1):
    return [r + [fill] * (k - len(r)) for r in rows]


def expected(case, dtype=torch.float32):
    """(face_idx (2,7,5), features of the per-vertex list entry (2,7,5,1)); the per-face
    entry is face_idx + 3 * mesh (0 where void).  case: 'full' (ranges [-4, 0), :104-165) or
    'restricted' ([-2.1, 0), :167-229)."""
    if case == "full":
        idx = [_pad([[0, 1], [0, 1], [2], [2], [0, 1], [2], []]),
               _pad([[0, 1, 2], [0, 1, 2], [2, 1, 0], [2, 1, 0], [0, 1, 2], [2, 1, 0], []])]
        f1 = [_pad([[0., 3.], [1., 5.], [7.], [8.], [0.825, 3.825], [7.175], []], fill=0.),
              _pad([[9., 12., 15.], [10., 13., 16.], [17., 14., 11.], [16.5, 13.5, 10.5],
                    [9.5, 12.5, 15.5], [16., 13., 10.], []], fill=0.)]
    else:
        idx = [_pad([[0], [0], [2], [2], [0], [2], []]),
               _pad([[0, 1, 2], [0, 1], [2, 1], [2, 1, 0], [0, 1], [2, 1], []])]
        f1 = [_pad([[0.], [1.], [7.], [8.], [0.825], [7.175], []], fill=0.),
              _pad([[9., 12., 15.], [10., 13.], [17., 14.], [16.5, 13.5, 10.5], [9.5, 12.5],
                    [16., 13.], []], fill=0.)]
    idx = torch.tensor(idx, dtype=torch.long)
    f0 = (idx + torch.arange(2).view(2, 1, 1) * 3).to(dtype).unsqueeze(-1)
    f0[idx == -1] = 0.
    return idx, f0, torch.tensor(f1, dtype=dtype).unsqueeze(-1)
