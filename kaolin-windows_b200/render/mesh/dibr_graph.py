"""CUDA-graph replay of dibr_rasterization (opt-in: ``dibr_rasterization(..., cuda_graph=True)``).

The eager call costs ~25 CUDA launches and ~0.25 ms of host time per forward + backward,
as much as the kernels themselves take on a B200, so small batches (and slow hosts) are
launch-bound.  Here the forward (memsets, binning, rasterizer, soft mask, the 16-byte size
report) and the backward (gradient memsets, both backward kernels) are each captured once
per problem -- same shapes, parameters AND input addresses -- into a CUDA graph working on
buffers owned by a cache entry, and replayed afterwards: one launch each way.

Semantics that differ from the eager call (hence opt-in):
  * the returned tensors are views of the entry's static buffers: the next graphed call on
    the same inputs overwrites them (as with torch.cuda.make_graphed_callables);
  * backward must run before the next forward of the same entry (checked: raises);
  * gradients come back as copies of the static buffers (a view could be adopted as a leaf's
    ``.grad`` and then be rewritten by the next replay).
Correctness is kept as in the eager path: the list-size report of every replay is checked
before the call returns; an overflow grows the workspace, re-issues the call eagerly and
re-captures.  A first sighting of a problem runs eagerly; it is captured when seen again.
"""
import collections
import time

import torch
from torch.autograd import Function

from ... import _lib, _workspace
from .rasterization import _prep, _ptr, _stream

MAX_ENTRIES = 8          # problems kept (each owns outputs, gradients and a workspace)
MAX_BWD_GRAPHS = 4       # cotangent address pairs captured per entry

_entries = collections.OrderedDict()
_sightings = collections.OrderedDict()


def clear():
    """Drop every captured graph and its buffers."""
    _entries.clear()
    _sightings.clear()


class _Entry:
    def __init__(self, H, W, z, img, feat, nz, params):
        self.H, self.W = H, W
        # input ADDRESSES only: holding the tensors would keep the allocator from handing
        # the same blocks to the next iteration's inputs, i.e. from ever hitting this entry
        # again.  The graph is replayed only for a call whose inputs live at these addresses.
        self.zp, self.imgp, self.featp, self.nzp = (t.data_ptr() for t in (z, img, feat, nz))
        self.params = params                      # sigmainv, margin, knum, multiplier, eps
        self.B, self.F = z.shape[:2]
        self.D = feat.shape[-1]
        dev = self.dev = z.device
        self._devctx = torch.cuda.device(dev)   # reused: constructing one costs ~10 us of Python
        B, D = self.B, self.D
        self.idx = torch.empty((B, H, W), dtype=torch.long, device=dev)
        self.out = torch.empty((B, H, W, D), dtype=torch.float32, device=dev)
        self.mask = torch.empty((B, H, W), dtype=torch.float32, device=dev)
        # both gradients in ONE buffer, exactly adjacent: the backward call zero-fills them with
        # one memset node and the result leaves as one clone (fewer nodes in the launch chain)
        self._n_img = img.numel()
        self._gpool = torch.empty(self._n_img + feat.numel(), dtype=torch.float32, device=dev)
        self.grad_img = self._gpool[:self._n_img].view(tuple(img.shape))
        self.grad_feat = self._gpool[self._n_img:].view(tuple(feat.shape))
        self.caps = _workspace.initial_caps(B, self.F, H, W, params[2])
        if not (_workspace.USE_PAIRS and 0 < params[2] <= _workspace.PAIR_MAX_K):
            self.caps[2] = 0
        self.slot = torch.full((4,), -1, dtype=torch.int64).pin_memory()
        self.slot_np = self.slot.numpy()      # same memory: the host polls without torch overhead
        self.ws = None
        self.graph = None
        self.bwd_graphs = collections.OrderedDict()
        self.bwd_seen = set()
        self.generation = 0
        self._alloc_ws()

    # ------------------------------------------------------------------ workspace
    def _alloc_ws(self):
        lib = _lib.lib()
        cap_r, cap_s, cap_p = self.caps
        n = lib.dibr_workspace_bytes_for(self.B, self.F, self.H, self.W, cap_r, cap_s, cap_p,
                                         self.params[2])
        if n < 0:
            _lib.check(int(n), "dibr_workspace_bytes_for")
        self.ws_bytes = int(n)
        self.ws = torch.empty(self.ws_bytes, dtype=torch.uint8, device=self.dev)
        self.graph = None                 # captured graphs point into the old workspace
        self.bwd_graphs.clear()

    # ------------------------------------------------------------------ forward
    def _issue_forward(self):
        sigmainv, margin, knum, multiplier, eps = self.params
        cap_r, cap_s, cap_p = self.caps
        rc = _lib.lib().dibr_render_fwd(
            self.B, self.F, self.H, self.W, self.D, self.zp, self.imgp, self.featp, self.nzp,
            sigmainv, margin, knum, multiplier, eps,
            _ptr(self.idx), _ptr(self.out), _ptr(self.mask), self.ws.data_ptr(), self.ws_bytes,
            cap_r, cap_s, cap_p, self.slot.data_ptr(), _stream(self.dev))
        _lib.check(rc, "dibr_render_fwd")

    def _poll(self):
        slot = self.slot_np
        spins = 0
        t0 = None
        while True:
            nr, ns = int(slot[0]), int(slot[1])
            if nr >= 0 and ns >= 0:
                return nr, ns
            spins += 1
            if spins & 0xfff == 0:
                if t0 is None:
                    t0 = time.perf_counter()
                elif time.perf_counter() - t0 > 120.0:
                    torch.cuda.synchronize(self.dev)
                    nr, ns = int(slot[0]), int(slot[1])
                    if nr < 0 or ns < 0:
                        raise RuntimeError("libdibr_b200: list-size report never arrived")
                    return nr, ns

    def forward(self, capture_ok):
        # pair buffer: what the previous call of this entry wanted (long arrived)
        want = max(int(self.slot_np[2]), int(self.slot_np[3]))
        if self.caps[2] > 0 and want > self.caps[2]:
            self.caps[2] = min(int(want * 1.25) + 1024, 1 << 23)
            self._alloc_ws()
        with self._devctx:
            for _attempt in range(4):
                self.slot_np[:] = -1
                if self.graph is not None:
                    self.graph.replay()
                else:
                    self._issue_forward()
                nr, ns = self._poll()
                if nr <= self.caps[0] and ns <= self.caps[1]:
                    break
                if nr > self.caps[0]:
                    self.caps[0] = int(nr * 1.25) + 4096
                if ns > self.caps[1]:
                    self.caps[1] = int(ns * 1.25) + 4096
                self._alloc_ws()
            else:
                raise RuntimeError("libdibr_b200: workspace kept overflowing")
            if self.graph is None and capture_ok:
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, capture_error_mode="thread_local"):
                    self._issue_forward()
                self.graph = g
        self.generation += 1

    # ------------------------------------------------------------------ backward
    def _issue_backward(self, g_out, g_mask):
        sigmainv, margin, knum, multiplier, eps = self.params
        cap_r, cap_s, cap_p = self.caps
        rc = _lib.lib().dibr_render_bwd(
            self.B, self.F, self.H, self.W, self.D, _ptr(g_out), _ptr(g_mask), _ptr(self.idx),
            self.imgp, self.featp, _ptr(self.mask), sigmainv, margin, knum, multiplier, eps,
            _ptr(self.grad_img), _ptr(self.grad_feat), self.ws.data_ptr(), self.ws_bytes,
            cap_r, cap_s, cap_p, _stream(self.dev))
        _lib.check(rc, "dibr_render_bwd")

    def backward(self, g_out, g_mask):
        key = (None if g_out is None else g_out.data_ptr(),
               None if g_mask is None else g_mask.data_ptr())
        with self._devctx:
            g = self.bwd_graphs.get(key)
            if g is not None:
                g.replay()
                return
            self._issue_backward(g_out, g_mask)
            if key in self.bwd_seen:      # same cotangent buffers again: worth a graph
                if len(self.bwd_graphs) >= MAX_BWD_GRAPHS:
                    self.bwd_graphs.popitem(last=False)
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, capture_error_mode="thread_local"):
                    self._issue_backward(g_out, g_mask)
                self.bwd_graphs[key] = graph     # addresses only, as for the inputs
            else:
                self.bwd_seen.add(key)
                if len(self.bwd_seen) > 64:
                    self.bwd_seen.clear()


def _entry_for(H, W, z, img, feat, nz, params):
    key = (z.device.index, H, W, tuple(z.shape), feat.shape[-1], params,
           z.data_ptr(), img.data_ptr(), feat.data_ptr(), nz.data_ptr())
    e = _entries.get(key)
    if e is not None:
        _entries.move_to_end(key)
        return e, True
    n = _sightings.get(key, 0) + 1
    _sightings[key] = n
    _sightings.move_to_end(key)
    while len(_sightings) > 64:
        _sightings.popitem(last=False)
    if n < 2:
        return None, False               # first sighting: the caller runs eagerly
    while len(_entries) >= MAX_ENTRIES:
        _entries.popitem(last=False)
    e = _entries[key] = _Entry(H, W, z, img, feat, nz, params)
    del _sightings[key]
    return e, True


class DibrRasterizationGraphed(Function):
    """DibrRasterizationB200 (dibr.py) replayed from CUDA graphs; see the module docstring."""

    @staticmethod
    def forward(ctx, entry, img, feat, face_vertices_image, face_features):
        entry.forward(capture_ok=True)
        ctx.entry = entry
        ctx.save_for_backward(img, feat)   # the backward reads them: keep them alive
        ctx.generation = entry.generation
        ctx.set_materialize_grads(False)
        idx = entry.idx.view_as(entry.idx)
        ctx.mark_non_differentiable(idx)
        return entry.out.view_as(entry.out), entry.mask.view_as(entry.mask), idx

    @staticmethod
    def backward(ctx, grad_features, grad_soft_mask, grad_face_idx):
        entry = ctx.entry
        if ctx.generation != entry.generation:
            raise RuntimeError("dibr_rasterization(cuda_graph=True): backward of a call whose "
                               "static buffers were overwritten by a later forward")
        if grad_features is None and grad_soft_mask is None:
            return None, None, None, None, None
        g = None if grad_features is None else _prep(grad_features, "grad_interpolated_features")
        gm = None if grad_soft_mask is None else _prep(grad_soft_mask, "grad_soft_mask")
        entry.backward(g, gm)
        # clones: AccumulateGrad may adopt a returned tensor as a leaf's .grad without copying
        # it (it looks at the tensor's use count, not at storage sharing); a view of the
        # entry's static buffer would then be rewritten by the next replay
        flat = entry._gpool.clone()
        return (None, None, None, flat[:entry._n_img].view(entry.grad_img.shape),
                flat[entry._n_img:].view(entry.grad_feat.shape))


def graphed_dibr_rasterization(H, W, face_vertices_z, face_vertices_image, face_features,
                               face_normals_z, params):
    """Returns (features, soft_mask, face_idx) from a graph entry, or None when this problem
    is seen for the first time (the caller then takes the eager path)."""
    z = _prep(face_vertices_z, "face_vertices_z")
    img = _prep(face_vertices_image, "face_vertices_image")
    feat = _prep(face_features, "face_features")
    nz = _prep(face_normals_z, "face_normals_z")
    B, F = z.shape[:2]
    D = feat.shape[-1]
    if (img.shape != (B, F, 3, 2) or z.shape != (B, F, 3) or feat.shape != (B, F, 3, D)
            or nz.shape != (B, F)):
        raise RuntimeError("dibr_rasterization: inconsistent input shapes")
    if any(t.device != z.device for t in (img, feat, nz)):
        raise RuntimeError("dibr_rasterization: all tensors must be on the same GPU")
    if B == 0 or F == 0:
        return None
    entry, ok = _entry_for(int(H), int(W), z, img, feat, nz, params)
    if not ok:
        return None
    return DibrRasterizationGraphed.apply(entry, img.detach(), feat.detach(),
                                          face_vertices_image, face_features)


# --------------------------------------------------------------------------- mesh level
class _MeshEntry(_Entry):
    """Graph entry of dibr_rasterization_mesh: same replay / overflow / re-capture logic, the
    mesh-level C calls (dibr_render_mesh_fwd / _bwd) and vertex-level gradient buffers."""

    def __init__(self, H, W, vcam, vimg, faces32, vfeat, valid, params):
        self.H, self.W = H, W
        self.vcamp, self.vimgp, self.facesp, self.vfeatp = (
            t.data_ptr() for t in (vcam, vimg, faces32, vfeat))
        self.validp = None if valid is None else valid.data_ptr()
        self.params = params
        self.B, self.V = vcam.shape[:2]
        self.F = faces32.shape[0]
        self.D = vfeat.shape[-1]
        self.shared = int(vfeat.shape[0] == 1)
        dev = self.dev = vcam.device
        self._devctx = torch.cuda.device(dev)   # reused: constructing one costs ~10 us of Python
        B, D = self.B, self.D
        self.idx = torch.empty((B, H, W), dtype=torch.long, device=dev)
        self.out = torch.empty((B, H, W, D), dtype=torch.float32, device=dev)
        self.mask = torch.empty((B, H, W), dtype=torch.float32, device=dev)
        # [grad_img | grad_feat | tail] in one buffer (see _Entry); the vertex entry keeps its
        # world-space vertex gradient in the tail so that both results leave as one clone
        self._n_img, self._n_feat = vimg.numel(), vfeat.numel()
        self._gpool = torch.empty(self._n_img + self._n_feat + getattr(self, "_pool_tail", 0),
                                  dtype=torch.float32, device=dev)
        self.grad_img = self._gpool[:self._n_img].view(tuple(vimg.shape))
        self.grad_feat = self._gpool[self._n_img:self._n_img + self._n_feat].view(tuple(vfeat.shape))
        self.caps = _workspace.initial_caps(B, self.F, H, W, params[2])
        if not (_workspace.USE_PAIRS and 0 < params[2] <= _workspace.PAIR_MAX_K):
            self.caps[2] = 0
        self.slot = torch.full((4,), -1, dtype=torch.int64).pin_memory()
        self.slot_np = self.slot.numpy()
        self.ws = None
        self.graph = None
        self.bwd_graphs = collections.OrderedDict()
        self.bwd_seen = set()
        self.generation = 0
        self._alloc_ws()

    def _issue_forward(self):
        sigmainv, margin, knum, multiplier, eps = self.params
        cap_r, cap_s, cap_p = self.caps
        rc = _lib.lib().dibr_render_mesh_fwd(
            self.B, self.V, self.F, self.H, self.W, self.D, self.vcamp, self.vimgp, self.facesp,
            self.vfeatp, self.shared, self.validp, sigmainv, margin, knum, multiplier, eps,
            _ptr(self.idx), _ptr(self.out), _ptr(self.mask), self.ws.data_ptr(), self.ws_bytes,
            cap_r, cap_s, cap_p, self.slot.data_ptr(), _stream(self.dev))
        _lib.check(rc, "dibr_render_mesh_fwd")

    def _issue_backward(self, g_out, g_mask):
        sigmainv, margin, knum, multiplier, eps = self.params
        cap_r, cap_s, cap_p = self.caps
        rc = _lib.lib().dibr_render_mesh_bwd(
            self.B, self.V, self.F, self.H, self.W, self.D, _ptr(g_out), _ptr(g_mask),
            _ptr(self.idx), self.vimgp, self.facesp, self.vfeatp, self.shared, _ptr(self.mask),
            sigmainv, margin, knum, multiplier, eps, _ptr(self.grad_img), _ptr(self.grad_feat),
            self.ws.data_ptr(), self.ws_bytes, cap_r, cap_s, cap_p,
            _stream(self.dev))
        _lib.check(rc, "dibr_render_mesh_bwd")


class DibrRasterizationMeshGraphed(Function):
    @staticmethod
    def forward(ctx, entry, keep, vertices_image, vertex_features):
        entry.forward(capture_ok=True)
        ctx.entry = entry
        ctx.keep = keep                    # inputs the backward reads: keep them alive
        ctx.generation = entry.generation
        ctx.set_materialize_grads(False)
        idx = entry.idx.view_as(entry.idx)
        ctx.mark_non_differentiable(idx)
        return entry.out.view_as(entry.out), entry.mask.view_as(entry.mask), idx

    @staticmethod
    def backward(ctx, grad_features, grad_soft_mask, grad_face_idx):
        entry = ctx.entry
        if ctx.generation != entry.generation:
            raise RuntimeError("dibr_rasterization_mesh(cuda_graph=True): backward of a call "
                               "whose static buffers were overwritten by a later forward")
        if grad_features is None and grad_soft_mask is None:
            return None, None, None, None
        g = None if grad_features is None else _prep(grad_features, "grad_interpolated_features")
        gm = None if grad_soft_mask is None else _prep(grad_soft_mask, "grad_soft_mask")
        entry.backward(g, gm)
        # clones: autograd may adopt a returned tensor as .grad without copying it, and the
        # next replay rewrites the entry's buffers (V-sized: the copy is cheap)
        n_img, n_feat = entry._n_img, entry._n_feat
        flat = entry._gpool[:n_img + n_feat].clone()
        return (None, None, flat[:n_img].view(entry.grad_img.shape),
                flat[n_img:].view(entry.grad_feat.shape))


def graphed_dibr_rasterization_mesh(H, W, vertices_camera, vertices_image, faces32,
                                    vertex_features, valid_faces, params):
    """(features, soft_mask, face_idx) from a graph entry, or None on the first sighting of
    this problem (the caller then takes the eager path)."""
    vcam = _prep(vertices_camera, "vertices_camera")
    vimg = _prep(vertices_image, "vertices_image")
    vfeat = _prep(vertex_features, "vertex_features")
    B, V = vcam.shape[:2]
    if (vcam.shape != (B, V, 3) or vimg.shape != (B, V, 2) or vfeat.dim() != 3
            or vfeat.shape[1] != V or vfeat.shape[0] not in (1, B)):
        raise RuntimeError("dibr_rasterization_mesh: expected vertices_camera (B,V,3), "
                           "vertices_image (B,V,2), vertex_features (B or 1,V,D)")
    F = faces32.shape[0]
    valid = None
    if valid_faces is not None:
        if valid_faces.dtype != torch.bool or not valid_faces.is_contiguous():
            return None                      # needs a converted copy: eager path
        valid = valid_faces.view(torch.uint8)
    if B == 0 or F == 0 or V == 0:
        return None
    key = ("mesh", vcam.device.index, int(H), int(W), tuple(vcam.shape), F, tuple(vfeat.shape),
           params, vcam.data_ptr(), vimg.data_ptr(), faces32.data_ptr(), vfeat.data_ptr(),
           None if valid is None else valid.data_ptr())
    e = _entries.get(key)
    if e is None:
        n = _sightings.get(key, 0) + 1
        _sightings[key] = n
        _sightings.move_to_end(key)
        while len(_sightings) > 64:
            _sightings.popitem(last=False)
        if n < 2:
            return None
        while len(_entries) >= MAX_ENTRIES:
            _entries.popitem(last=False)
        e = _entries[key] = _MeshEntry(int(H), int(W), vcam, vimg, faces32, vfeat, valid, params)
        del _sightings[key]
    else:
        _entries.move_to_end(key)
    return DibrRasterizationMeshGraphed.apply(e, (vcam, vimg, vfeat, faces32, valid),
                                              vertices_image, vertex_features)


# ----------------------------------------------------------------- vertices + cameras in
class _VertexEntry(_MeshEntry):
    """Graph entry of dibr_rasterization_vertices: the camera transform / projection kernel
    in front of the mesh-level forward, its backward behind the mesh-level backward -- one
    graph launch each way for the whole step."""

    def __init__(self, H, W, vertices, faces32, vfeat, proj, rot, trans, tf, valid, batch,
                 params, exchange=None):
        dev = vertices.device
        self.exchange = exchange      # VertexGradExchange in peer mode, or None
        V = vertices.shape[1]
        self.vertsp, self.projp = vertices.data_ptr(), proj.data_ptr()
        self.rotp = None if rot is None else rot.data_ptr()
        self.transp = None if trans is None else trans.data_ptr()
        self.tfp = None if tf is None else tf.data_ptr()
        self.verts_shared = int(vertices.shape[0] == 1)
        self.vcam = torch.empty((batch, V, 3), dtype=torch.float32, device=dev)
        self.vimg = torch.empty((batch, V, 2), dtype=torch.float32, device=dev)
        self._pool_tail = vertices.numel()
        super().__init__(H, W, self.vcam, self.vimg, faces32, vfeat, valid, params)
        self.grad_verts = self._gpool[self._n_img + self._n_feat:].view(tuple(vertices.shape))
        if exchange is not None:
            # the gradients are written straight into this rank's buffer of the exchange's
            # symmetric allocation; the all-reduce is the last node of the backward graph
            nv, nf = self.grad_verts.numel(), self.grad_feat.numel()
            self.nv, self.nf = nv, nf
            self.off_f = (nv + 63) // 64 * 64
            self.grad_verts = exchange.flat[:nv].view(self.grad_verts.shape)
            self.grad_feat = exchange.flat[self.off_f:self.off_f + nf].view(self.grad_feat.shape)

    def _issue_forward(self):
        _lib.check(_lib.lib().dibr_vertex_transform_fwd(
            self.B, self.V, self.verts_shared, self.vertsp, self.rotp, self.transp, self.tfp,
            self.projp, _ptr(self.vcam), _ptr(self.vimg), _stream(self.dev)),
            "dibr_vertex_transform_fwd")
        super()._issue_forward()

    def _issue_backward(self, g_out, g_mask):
        if self.exchange is not None:
            self.exchange.issue_wait()        # peers have read the previous step's gradients
        super()._issue_backward(g_out, g_mask)
        _lib.check(_lib.lib().dibr_vertex_transform_bwd(
            self.B, self.V, self.verts_shared, _ptr(self.vcam), self.rotp, self.tfp, self.projp,
            None, _ptr(self.grad_img), _ptr(self.grad_verts), _stream(self.dev)),
            "dibr_vertex_transform_bwd")
        if self.exchange is not None:
            self.exchange.issue_allreduce()   # result = sum over ranks

    def result_grads(self):
        """Copies of the two gradients (summed over ranks when an exchange is folded in)."""
        if self.exchange is None:
            flat = self._gpool[self._n_img:].clone()     # [grad_feat | grad_verts], one copy
            return (flat[self._n_feat:].view(self.grad_verts.shape),
                    flat[:self._n_feat].view(self.grad_feat.shape))
        flat = self.exchange.result.clone()    # one copy, both gradients are views of it
        return (flat[:self.nv].view(self.grad_verts.shape),
                flat[self.off_f:self.off_f + self.nf].view(self.grad_feat.shape))


class DibrRasterizationVerticesGraphed(Function):
    @staticmethod
    def forward(ctx, entry, keep, vertices, vertex_features):
        entry.forward(capture_ok=True)
        ctx.entry = entry
        ctx.keep = keep
        ctx.generation = entry.generation
        ctx.set_materialize_grads(False)
        idx = entry.idx.view_as(entry.idx)
        ctx.mark_non_differentiable(idx)
        return entry.out.view_as(entry.out), entry.mask.view_as(entry.mask), idx

    @staticmethod
    def backward(ctx, grad_features, grad_soft_mask, grad_face_idx):
        entry = ctx.entry
        if ctx.generation != entry.generation:
            raise RuntimeError("dibr_rasterization_vertices(cuda_graph=True): backward of a call "
                               "whose static buffers were overwritten by a later forward")
        if grad_features is None and grad_soft_mask is None:
            return None, None, None, None
        g = None if grad_features is None else _prep(grad_features, "grad_interpolated_features")
        gm = None if grad_soft_mask is None else _prep(grad_soft_mask, "grad_soft_mask")
        entry.backward(g, gm)
        gv, gf = entry.result_grads()
        return None, None, gv, gf


def graphed_dibr_rasterization_vertices(H, W, vertices, faces32, vertex_features, camera_proj,
                                        camera_rot, camera_trans, camera_transform, valid_faces,
                                        batch, params, exchange=None):
    """(features, soft_mask, face_idx) from a graph entry, or None on the first sighting."""
    v = _prep(vertices, "vertices")
    vfeat = _prep(vertex_features, "vertex_features")
    proj = _prep(camera_proj.reshape(-1), "camera_proj")
    rot = None if camera_rot is None else _prep(camera_rot, "camera_rot")
    trans = None if camera_trans is None else _prep(camera_trans.reshape(-1, 3), "camera_trans")
    tf = None if camera_transform is None else _prep(camera_transform, "camera_transform")
    V = v.shape[1] if v.dim() == 3 else -1
    if (v.dim() != 3 or v.shape[-1] != 3 or v.shape[0] not in (1, batch) or proj.numel() != 3
            or vfeat.dim() != 3 or vfeat.shape[1] != V or vfeat.shape[0] not in (1, batch)
            or (rot is not None and rot.shape != (batch, 3, 3))
            or (trans is not None and trans.shape != (batch, 3))
            or (tf is not None and tf.shape != (batch, 4, 3))):
        raise RuntimeError("dibr_rasterization_vertices: inconsistent input shapes")
    F = faces32.shape[0]
    valid = None
    if valid_faces is not None:
        if valid_faces.dtype != torch.bool or not valid_faces.is_contiguous():
            return None
        valid = valid_faces.view(torch.uint8)
    if batch == 0 or F == 0 or V == 0:
        return None
    cams = tuple(None if t is None else t.data_ptr() for t in (proj, rot, trans, tf))
    key = ("vertices", v.device.index, int(H), int(W), batch, tuple(v.shape), F,
           tuple(vfeat.shape), params, v.data_ptr(), faces32.data_ptr(), vfeat.data_ptr(), cams,
           None if valid is None else valid.data_ptr(), None if exchange is None else id(exchange))
    e = _entries.get(key)
    if e is None:
        n = _sightings.get(key, 0) + 1
        _sightings[key] = n
        _sightings.move_to_end(key)
        while len(_sightings) > 64:
            _sightings.popitem(last=False)
        if n < 2:
            return None
        while len(_entries) >= MAX_ENTRIES:
            _entries.popitem(last=False)
        e = _entries[key] = _VertexEntry(int(H), int(W), v, faces32, vfeat, proj, rot, trans, tf,
                                         valid, batch, params, exchange)
        del _sightings[key]
    else:
        _entries.move_to_end(key)
    return DibrRasterizationVerticesGraphed.apply(e, (v, vfeat, faces32, proj, rot, trans, tf,
                                                      valid), vertices, vertex_features)
