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
