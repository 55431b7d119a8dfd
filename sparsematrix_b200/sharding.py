"""Row-slab sharding of W = S^T across the GPUs of one box.

The product has no exchange step: output rows of Y = W X (= column panels of S, which the
reference already treats independently, sparse-matrix.cc:164-190) are split into contiguous,
panel-aligned, nnz-balanced slabs (``sblas_b200_partition``); slab g lives on GPU g, X is
replicated, every GPU writes its own Y slab.  A collective is needed only when the output must
be reassembled on every rank: ``gather_output`` is an all-gather of the (padded) C slabs.

One process per GPU; ``torch.distributed`` (NCCL on GPUs, gloo on CPU for the tests) is the
plumbing.  Nothing here computes: the product is ``DeviceMatrix.addmatmat_dev`` on the local slab.
"""
import numpy as np

from .matrix import DeviceMatrix, SparseMatrix, partition


def slab_bounds(host_matrix: SparseMatrix, world_size: int):
    """Column splits [world_size + 1] of S (= row splits of W), nnz-balanced on panel boundaries."""
    _, _, _, _, bb, bib = host_matrix.members()
    return partition(bb, bib, host_matrix.NumCols(), world_size)


def upload_local_slab(host_matrix: SparseMatrix, rank: int, world_size: int, device: int = 0):
    """-> (DeviceMatrix of this rank's slab or None when the slab is empty, (col_begin, col_end))."""
    splits = slab_bounds(host_matrix, world_size)
    lo, hi = int(splits[rank]), int(splits[rank + 1])
    if hi == lo:
        return None, (lo, hi)
    return host_matrix.upload_slab(lo, hi, device), (lo, hi)


def gather_output(c_local, m: int, splits, group=None):
    """All-gather the per-rank C slabs (m x n_g, row-major) into the full m x n matrix on every rank.

    Slabs are uneven (nnz-balanced, not row-balanced): each rank contributes its slab padded to the
    widest one, in [n_g][m] (slab-major) order so that a rank's piece is one contiguous message.
    Works on CPU tensors (gloo) and CUDA tensors (NCCL)."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    widths = [int(splits[g + 1] - splits[g]) for g in range(world)]
    wmax = max(widths)
    n = int(splits[-1])
    send = c_local.new_zeros((wmax, m))
    rank = dist.get_rank(group)
    if widths[rank]:
        send[: widths[rank]] = c_local.reshape(m, widths[rank]).t()
    recv = c_local.new_empty((world, wmax, m))
    dist.all_gather_into_tensor(recv, send, group=group) if c_local.is_cuda else \
        dist.all_gather(list(recv.unbind(0)), send, group=group)
    out = c_local.new_empty((m, n))
    for g in range(world):
        if widths[g]:
            out[:, int(splits[g]): int(splits[g + 1])] = recv[g, : widths[g]].t()
    return out


def model_scaling_ceiling(nnz, n, k, m, world_size, bytes_per_nnz=5):
    """Bandwidth-model speed-up ceiling of row sharding with X replicated:
    per-GPU bytes fall with the slab, the replicated X does not."""
    one = bytes_per_nnz * nnz + 4 * (n + 1) + 4 * k * m + 8 * n * m
    per = bytes_per_nnz * nnz / world_size + 4 * (n / world_size + 1) + 4 * k * m + 8 * n * m / world_size
    return one / per
