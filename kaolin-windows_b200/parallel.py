"""Multi-GPU plumbing of the DIB-R path (one process per GPU, torch.distributed).

The path shards over independent units -- camera views / meshes (every kernel indexes the
batch element only; rasterization_cuda.cu:62,71-72, dibr_soft_mask_cuda.cu:51,83) -- so the
data path needs no collective.  When the views share one mesh, the only exchange is a
SUM all-reduce of that mesh's vertex-position and vertex-feature gradients, issued once
per step on a single flat buffer, after the local face->vertex scatter-add has shrunk the
message from (views, faces) to (vertices).
"""
import torch
import torch.distributed as dist


def shard_range(n_units, rank, world_size):
    """Contiguous [lo, hi) slice of `n_units` views/meshes owned by `rank`."""
    base, rem = divmod(n_units, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def allreduce_mesh_grads(grads, group=None):
    """In-place SUM all-reduce of the shared mesh's gradients (a list of tensors, e.g.
    [vertices.grad (V,3), vertex_features.grad (V,D)]) as ONE flat NCCL call."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return grads
    grads = [g for g in grads if g is not None]
    if not grads:
        return grads
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    off = 0
    for g in grads:
        n = g.numel()
        g.copy_(flat[off:off + n].view_as(g))
        off += n
    return grads


def scatter_mesh_grads(faces, num_vertices, *face_grads):
    """Local face -> vertex scatter-add of this rank's views onto the shared mesh: each
    (views, faces, 3, C_i) face-level gradient is summed over views and corners into the
    columns of ONE flat (num_vertices, sum C_i) buffer (csrc/dibr_mesh.cu, one kernel per
    gradient, no intermediate).  The buffer is the message of `allreduce_mesh_grads`."""
    import ctypes
    from . import _lib
    face_grads = [g.contiguous() for g in face_grads]
    dev = face_grads[0].device
    if not face_grads[0].is_cuda:
        raise RuntimeError("scatter_mesh_grads needs CUDA tensors (there is no CPU path)")
    faces = faces.contiguous()
    width = sum(g.shape[-1] for g in face_grads)
    flat = torch.zeros((num_vertices, width), dtype=torch.float32, device=dev)
    lib = _lib.lib()
    col = 0
    with torch.cuda.device(dev):
        st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        for g in face_grads:
            B, F, _, C = g.shape
            _lib.check(lib.dibr_scatter_face_grads(
                B, num_vertices, F, C, 1, ctypes.c_void_p(g.data_ptr()),
                ctypes.c_void_p(faces.data_ptr()), ctypes.c_void_p(flat.data_ptr()),
                width, col, st), "scatter_mesh_grads")
            col += C
    return flat


def exchange_mesh_grads(faces, num_vertices, *face_grads, group=None):
    """scatter_mesh_grads + the path's only collective: one NCCL SUM all-reduce of the
    flat vertex-level buffer.  Returns the (num_vertices, sum C_i) total over all ranks."""
    flat = scatter_mesh_grads(faces, num_vertices, *face_grads)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return flat


class MeshGradExchange:
    """exchange_mesh_grads with its local part -- the buffer memset and the scatter kernels
    -- as ONE CUDA-graph launch, captured once per set of gradient addresses (a training
    loop hands back the same blocks every iteration) and replayed; the NCCL all-reduce
    follows as a normal call (kept out of the graph: a process group torn down while
    graphs still hold its kernels hangs).  The result is a static (num_vertices, sum C_i)
    buffer, overwritten by the next call.  Addresses seen for the first time run eagerly."""

    MAX_GRAPHS = 4

    def __init__(self, faces, num_vertices, group=None, capture_collective=False):
        # capture_collective: put the all-reduce into the graph as well (one launch for the
        # whole exchange).  The owner must then call release() before the process group is
        # destroyed -- tearing a communicator down under live graphs hangs.
        self.capture_collective = bool(capture_collective)
        self.faces = faces.contiguous()
        self.num_vertices = int(num_vertices)
        self.group = group
        self.flat = None
        self.graphs = {}
        self.seen = set()

    def _issue(self, face_grads):
        import ctypes
        from . import _lib
        lib = _lib.lib()
        dev = self.flat.device
        self.flat.zero_()
        width = self.flat.shape[1]
        col = 0
        st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        for g in face_grads:
            B, F, _, C = g.shape
            _lib.check(lib.dibr_scatter_face_grads(
                B, self.num_vertices, F, C, 1, ctypes.c_void_p(g.data_ptr()),
                ctypes.c_void_p(self.faces.data_ptr()), ctypes.c_void_p(self.flat.data_ptr()),
                width, col, st), "scatter_mesh_grads")
            col += C

    def release(self):
        """Drop the captured graphs (required before destroy_process_group when the
        collective is captured)."""
        self.graphs.clear()
        self.seen.clear()

    def _reduce(self):
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)
        return self.flat

    def __call__(self, *face_grads):
        face_grads = [g.contiguous() for g in face_grads]
        dev = face_grads[0].device
        if not face_grads[0].is_cuda:
            raise RuntimeError("MeshGradExchange needs CUDA tensors (there is no CPU path)")
        width = sum(g.shape[-1] for g in face_grads)
        if self.flat is None or self.flat.shape[1] != width:
            self.flat = torch.empty((self.num_vertices, width), dtype=torch.float32, device=dev)
            self.graphs.clear()
        key = tuple((g.data_ptr(), tuple(g.shape)) for g in face_grads)
        with torch.cuda.device(dev):
            graph = self.graphs.get(key)
            if graph is not None:
                graph.replay()
                return self.flat if self.capture_collective else self._reduce()
            self._issue(face_grads)
            if key in self.seen and len(self.graphs) < self.MAX_GRAPHS:
                graph = torch.cuda.CUDAGraph()
                self._reduce()            # this call's result, eagerly
                with torch.cuda.graph(graph, capture_error_mode="thread_local"):
                    self._issue(face_grads)
                    if self.capture_collective:
                        self._reduce()
                self.graphs[key] = graph
                return self.flat
            else:
                self.seen.add(key)
                if len(self.seen) > 64:
                    self.seen.clear()
            return self._reduce()
