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


class _PeerBuffers:
    """This rank's part of a symmetric allocation for the one-shot peer all-reduce
    (csrc/dibr_peer.cu): a float buffer every rank maps, its signal pad, a local result
    buffer.  torch.distributed._symmetric_memory provides allocation and rendezvous."""

    def __init__(self, numel, dev, group):
        import ctypes
        import torch.distributed._symmetric_memory as symm_mem
        from . import _lib
        self.n = (int(numel) + 3) // 4 * 4
        self.buf = symm_mem.empty(self.n, dtype=torch.float32, device=dev)
        hdl = symm_mem.rendezvous(self.buf, group if group is not None else dist.group.WORLD)
        self.hdl = hdl
        self.rank, self.world = int(hdl.rank), int(hdl.world_size)
        if (9 * self.world + self.world) * 4 > int(hdl.signal_pad_size):
            raise RuntimeError("signal pad too small")
        self.bufs = (ctypes.c_int64 * self.world)(*[int(p) for p in hdl.buffer_ptrs])
        self.pads = (ctypes.c_int64 * self.world)(*[int(p) for p in hdl.signal_pad_ptrs])
        self.out = torch.empty(self.n, dtype=torch.float32, device=dev)
        self.sync = torch.zeros(2, dtype=torch.int32, device=dev)
        self.buf.zero_()
        with torch.cuda.device(dev):
            st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
            _lib.check(_lib.lib().dibr_peer_signal(self.rank, self.world, self.pads, 0, st),
                       "dibr_peer_signal")
        torch.cuda.synchronize(dev)


def _make_peer_buffers(numel, dev, group):
    """_PeerBuffers, or None where symmetric memory is unavailable; all ranks agree."""
    import os
    if os.environ.get("DIBR_B200_EXCHANGE", "peer") == "nccl":
        return None
    peer, ok = None, 1
    try:
        peer = _PeerBuffers(numel, dev, group)
    except Exception as e:      # no P2P mapping in this environment: NCCL does the exchange
        import warnings
        warnings.warn("kaolin-windows_b200: peer-memory exchange unavailable (%r), using NCCL" % (e,))
        ok = 0
    flag = torch.tensor([ok], device=dev, dtype=torch.int32)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
    return peer if int(flag.item()) == 1 else None


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
        self.peer = None          # _PeerBuffers when the exchange runs over peer memory
        self.mode = "local"
        self.graphs = {}
        self.seen = set()

    def _setup(self, width, dev):
        """The (num_vertices, width) buffer the scatter fills: in a symmetric allocation when
        the exchange can run as our own peer-memory kernel, else plain memory + NCCL."""
        self.graphs.clear()
        self.peer = None
        multi = (dist.is_available() and dist.is_initialized()
                 and dist.get_world_size(self.group) > 1)
        if multi:
            self.peer = _make_peer_buffers(self.num_vertices * width, dev, self.group)
        if self.peer is not None:
            self.flat = self.peer.buf[:self.num_vertices * width].view(self.num_vertices, width)
            self.result = self.peer.out[:self.num_vertices * width].view(self.num_vertices, width)
            self.mode = "peer"
        else:
            self.flat = torch.empty((self.num_vertices, width), dtype=torch.float32, device=dev)
            self.result = self.flat
            self.mode = "nccl" if multi else "local"

    def _issue(self, face_grads):
        import ctypes
        from . import _lib
        lib = _lib.lib()
        dev = self.flat.device
        st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        pb = self.peer
        if pb is not None:      # peers may still be reading the previous exchange
            _lib.check(lib.dibr_peer_signal(pb.rank, pb.world, pb.pads, 1, st), "dibr_peer_signal")
            pb.buf.zero_()
        else:
            self.flat.zero_()
        width = self.flat.shape[1]
        col = 0
        for g in face_grads:
            B, F, _, C = g.shape
            _lib.check(lib.dibr_scatter_face_grads(
                B, self.num_vertices, F, C, 1, ctypes.c_void_p(g.data_ptr()),
                ctypes.c_void_p(self.faces.data_ptr()), ctypes.c_void_p(self.flat.data_ptr()),
                width, col, st), "scatter_mesh_grads")
            col += C
        if pb is not None:
            _lib.check(lib.dibr_peer_allreduce(
                pb.rank, pb.world, pb.bufs, pb.pads, pb.n, ctypes.c_void_p(pb.out.data_ptr()),
                ctypes.c_void_p(pb.sync.data_ptr()), st), "dibr_peer_allreduce")

    def release(self):
        """Drop the captured graphs (required before destroy_process_group when the
        collective is captured)."""
        self.graphs.clear()
        self.seen.clear()

    def _reduce(self):
        if self.mode == "nccl":
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)
        return self.result

    def __call__(self, *face_grads):
        face_grads = [g.contiguous() for g in face_grads]
        dev = face_grads[0].device
        if not face_grads[0].is_cuda:
            raise RuntimeError("MeshGradExchange needs CUDA tensors (there is no CPU path)")
        width = sum(g.shape[-1] for g in face_grads)
        if self.flat is None or self.flat.shape[1] != width:
            self._setup(width, dev)
        key = tuple((g.data_ptr(), tuple(g.shape)) for g in face_grads)
        with torch.cuda.device(dev):
            graph = self.graphs.get(key)
            if graph is not None:
                graph.replay()
                return self.result if self.capture_collective else self._reduce()
            self._issue(face_grads)
            if key in self.seen and len(self.graphs) < self.MAX_GRAPHS:
                graph = torch.cuda.CUDAGraph()
                self._reduce()            # this call's result, eagerly
                with torch.cuda.graph(graph, capture_error_mode="thread_local"):
                    self._issue(face_grads)
                    if self.capture_collective:
                        self._reduce()
                self.graphs[key] = graph
                return self.result
            else:
                self.seen.add(key)
                if len(self.seen) > 64:
                    self.seen.clear()
            return self._reduce()


class VertexGradExchange:
    """SUM all-reduce of the shared mesh's VERTEX-level gradients -- what the mesh-level
    backward (render.mesh.dibr_rasterization_vertices / _mesh) produces directly, so there is
    no face -> vertex scatter in front of the exchange any more.  The gradients (e.g.
    vertices.grad (1, V, 3) and vertex_features.grad (1, V, D)) travel as ONE flat message:
    copied into this rank's buffer of a symmetric allocation and reduced by our own one-shot
    kernel over NVLink peer memory (csrc/dibr_peer.cu), NCCL where peer memory cannot be
    mapped (or DIBR_B200_EXCHANGE=nccl).  Copies + kernel replay as one CUDA graph once the
    same gradient addresses are seen again.  Returns views of a static result buffer."""

    MAX_GRAPHS = 16      # gradient tensors come out of the caching allocator: a few address sets

    def __init__(self, group=None):
        self.group = group
        self.peer = None
        self.flat = None
        self.result = None
        self.mode = "local"
        self.n = 0
        self.graphs = {}
        self.seen = set()

    def _setup(self, n, dev):
        self.graphs.clear()
        self.n = n
        multi = (dist.is_available() and dist.is_initialized()
                 and dist.get_world_size(self.group) > 1)
        self.peer = _make_peer_buffers(n, dev, self.group) if multi else None
        if self.peer is not None:
            self.flat, self.result, self.mode = self.peer.buf, self.peer.out, "peer"
        else:
            self.flat = torch.empty(n, dtype=torch.float32, device=dev)
            self.result, self.mode = self.flat, ("nccl" if multi else "local")

    def _issue(self, grads):
        import ctypes
        from . import _lib
        lib = _lib.lib()
        pb = self.peer
        dev = self.flat.device
        st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        if pb is not None:      # peers may still be reading the previous exchange
            _lib.check(lib.dibr_peer_signal(pb.rank, pb.world, pb.pads, 1, st), "dibr_peer_signal")
        off = 0
        for g in grads:
            k = g.numel()
            self.flat[off:off + k].copy_(g.reshape(-1))
            off += k
        if pb is not None:      # (the padding behind `off` was zeroed at set-up and stays so)
            _lib.check(lib.dibr_peer_allreduce(
                pb.rank, pb.world, pb.bufs, pb.pads, pb.n, ctypes.c_void_p(pb.out.data_ptr()),
                ctypes.c_void_p(pb.sync.data_ptr()), st), "dibr_peer_allreduce")

    def _reduce(self):
        if self.mode == "nccl":
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)

    def release(self):
        self.graphs.clear()
        self.seen.clear()

    # ---- folding the exchange into the producer's own CUDA graph (render.mesh.dibr_graph):
    # the backward kernels write the gradients straight into this rank's buffer of the
    # symmetric allocation -- no copy -- and the all-reduce kernel is the graph's last node.
    def prepare(self, n, dev):
        """Set up for messages of n floats; returns True when the exchange can be captured
        (peer-memory mode).  Every rank must call it at the same point."""
        n = (int(n) + 3) // 4 * 4
        if self.flat is None or self.n < n:
            self._setup(n, dev)
        return self.mode == "peer"

    def issue_wait(self):
        """(capturable) wait until every peer has read this rank's buffer of the previous
        exchange; the buffer may be overwritten afterwards."""
        import ctypes
        from . import _lib
        pb = self.peer
        st = ctypes.c_void_p(torch.cuda.current_stream(self.flat.device).cuda_stream)
        _lib.check(_lib.lib().dibr_peer_signal(pb.rank, pb.world, pb.pads, 1, st), "dibr_peer_signal")

    def issue_allreduce(self):
        """(capturable) result = sum over ranks of their buffers."""
        import ctypes
        from . import _lib
        pb = self.peer
        st = ctypes.c_void_p(torch.cuda.current_stream(self.flat.device).cuda_stream)
        _lib.check(_lib.lib().dibr_peer_allreduce(
            pb.rank, pb.world, pb.bufs, pb.pads, pb.n, ctypes.c_void_p(pb.out.data_ptr()),
            ctypes.c_void_p(pb.sync.data_ptr()), st), "dibr_peer_allreduce")

    def __call__(self, *grads):
        grads = [g.contiguous() for g in grads]
        if not grads[0].is_cuda:
            raise RuntimeError("VertexGradExchange needs CUDA tensors (there is no CPU path)")
        dev = grads[0].device
        n = sum(g.numel() for g in grads)
        if self.flat is None or self.n < n:      # (a larger buffer is reused: the padding is
            self._setup(n, dev)                  #  reduced along, nobody reads it)
        key = tuple((g.data_ptr(), tuple(g.shape)) for g in grads)
        with torch.cuda.device(dev):
            graph = self.graphs.get(key)
            if graph is not None:
                graph.replay()
            else:
                self._issue(grads)
                if key in self.seen and len(self.graphs) < self.MAX_GRAPHS:
                    self._reduce()
                    graph = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(graph, capture_error_mode="thread_local"):
                        self._issue(grads)
                    self.graphs[key] = graph
                    return self._views(grads)
                self.seen.add(key)
                if len(self.seen) > 64:
                    self.seen.clear()
            self._reduce()
        return self._views(grads)

    def _views(self, grads):
        out, off = [], 0
        for g in grads:
            k = g.numel()
            out.append(self.result[off:off + k].view_as(g))
            off += k
        return out


class SumGradsAcrossRanks(torch.autograd.Function):
    """Identity on (vertices, vertex_features) whose backward SUM-all-reduces both gradients
    through a VertexGradExchange: the shared mesh's gradients come back already summed over
    all ranks."""

    @staticmethod
    def forward(ctx, exchange, vertices, vertex_features):
        ctx.exchange = exchange
        return vertices.view_as(vertices), vertex_features.view_as(vertex_features)

    @staticmethod
    def backward(ctx, gv, gf):
        if gv is None or gf is None:
            raise RuntimeError("SumGradsAcrossRanks: both gradients are exchanged together")
        rv, rf = ctx.exchange(gv, gf)
        return None, rv.clone(), rf.clone()
