"""Row-slab sharding of W = S^T across the GPUs of one box.

The product has no exchange step: output rows of Y = W X (= column panels of S, which the
reference already treats independently, sparse-matrix.cc:164-190) are split into contiguous,
panel-aligned, nnz-balanced slabs (``sblas_b200_partition``); slab g lives on GPU g, X is
replicated, every GPU writes its own Y slab.  A collective is needed only when the output must
be reassembled on every rank: ``gather_output`` is an all-gather of the (padded) C slabs.
When the whole C is wanted on ONE GPU there is a way without any collective: ``PeerOutput``
holds C on the root GPU and maps it into every rank over NVLink (CUDA IPC), and each rank's
product kernels store their slab through that pointer -- product and "gather" are one launch.

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


class _CudaBuffer:
    """__cuda_array_interface__ holder so torch can view a raw device allocation."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False),
                                         "version": 2, "strides": None}


class PeerOutput:
    """C (m x n fp32, row-major) resident on ``root``'s GPU and mapped into every rank of ``group``.

    ``write_slab(mat, d_a, ...)`` runs this rank's product with ``d_c`` pointing into the ROOT's
    memory at the slab's first column (``ldc = n``): the kernels' epilogues read-modify-write C
    across NVLink / NVSwitch, so no separate gather follows ("fused alternative").
    ``finish()`` makes the result visible to the root: every rank synchronises the stream it used,
    then a barrier.  The root views C as a torch tensor with ``tensor()``.

    One process per GPU; the 64-byte CUDA IPC handle travels through ``broadcast_object_list``."""

    def __init__(self, m, n, root=0, device=0, group=None):
        import ctypes as C

        import torch.distributed as dist

        from . import _capi
        self.m, self.n, self.root, self.device, self.group = int(m), int(n), int(root), int(device), group
        self.rank = dist.get_rank(group)
        self._lib = _capi.lib()
        self._owned = self._mapped = None
        ptr = C.c_void_p()
        box, err = [None], None
        if self.rank == self.root:
            handle = (C.c_uint8 * 64)()
            try:
                _capi.check(self._lib.sblas_b200_peer_alloc(self.device, self.m * self.n * 4, C.byref(ptr), handle))
                self._owned = ptr.value
                box[0] = bytes(handle)
            except _capi.SblasError as e:
                err = str(e)
        dist.broadcast_object_list(box, src=dist.get_global_rank(group, self.root) if group is not None else self.root,
                                   group=group)
        if self.rank != self.root and box[0] is not None:
            handle = (C.c_uint8 * 64).from_buffer_copy(box[0])
            try:
                _capi.check(self._lib.sblas_b200_peer_open(self.device, handle, self.m * self.n * 4, C.byref(ptr)))
                self._mapped = ptr.value
            except _capi.SblasError as e:
                err = str(e)
        # success or failure is collective: a rank must not walk on into barriers its peers never reach
        errs = [None] * dist.get_world_size(group)
        dist.all_gather_object(errs, err, group=group)
        if box[0] is None or any(errs):
            self.base = None
            self.close()
            raise RuntimeError("PeerOutput: " + "; ".join(f"rank {r}: {e}" for r, e in enumerate(errs) if e))
        self.base = ptr.value

    def slab_pointer(self, col_begin):
        """Device address of C[0][col_begin] as seen from this rank."""
        return self.base + 4 * int(col_begin)

    def write_slab(self, mat: DeviceMatrix, d_a, lda, col_begin, alpha, beta, stream=0):
        """C[:, col_begin : col_begin + mat.n] = beta * (same) + alpha * A * S_slab, stored on the root's GPU."""
        mat.addmatmat_dev(d_a, self.m, lda, self.slab_pointer(col_begin), self.n, alpha, beta, stream=stream)

    def finish(self):
        import torch
        import torch.distributed as dist
        torch.cuda.synchronize(self.device)
        dist.barrier(group=self.group)

    def tensor(self):
        """The whole C as a torch CUDA tensor (root only; it aliases the allocation, keep ``self`` alive)."""
        import torch
        if self.rank != self.root:
            raise RuntimeError("only the root rank holds C")
        return torch.as_tensor(_CudaBuffer(self.base, (self.m, self.n)), device=torch.device("cuda", self.device))

    def close(self):
        """Collective: peers unmap first, then the root frees."""
        import torch.distributed as dist
        if self._mapped is not None:
            self._lib.sblas_b200_peer_close(self._mapped)
            self._mapped = None
        dist.barrier(group=self.group)
        if self._owned is not None:
            self._lib.sblas_b200_peer_free(self._owned)
            self._owned = None
        self.base = None


def model_scaling_ceiling(nnz, n, k, m, world_size, bytes_per_nnz=5):
    """Bandwidth-model speed-up ceiling of row sharding with X replicated:
    per-GPU bytes fall with the slab, the replicated X does not."""
    one = bytes_per_nnz * nnz + 4 * (n + 1) + 4 * k * m + 8 * n * m
    per = bytes_per_nnz * nnz / world_size + 4 * (n / world_size + 1) + 4 * k * m + 8 * n * m / world_size
    return one / per


def csr_slab(row_ptr, col, code, lo, hi):
    """Rows [lo, hi) of a CSR matrix as their own CSR (row_ptr rebased to 0): what ``sblas_b200_upload_csr`` takes for one slab."""
    row_ptr = np.asarray(row_ptr)
    e0, e1 = int(row_ptr[lo]), int(row_ptr[hi])
    rp = (row_ptr[lo:hi + 1].astype(np.int64) - e0).astype(np.int32)
    return np.ascontiguousarray(rp), np.ascontiguousarray(col[e0:e1]), np.ascontiguousarray(code[e0:e1])


class ShardedMatrix:
    """One matrix cut into nnz-balanced row slabs of W, slab r resident on rank r's GPU, X replicated.

    Built on every rank from the same host data -- a ``SparseMatrix`` (panel streams: ``sblas_b200_partition`` +
    ``sblas_b200_upload_slab``) or a CSR triple (``sblas_b200_partition_csr`` + ``sblas_b200_upload_csr``).
    ``product_into(out, d_a, ...)`` is the strong-scaled step: this rank's product stored straight into the
    ``PeerOutput`` on the root GPU.  ``addmatmat_host`` is the host-buffer entry point of a rank: every rank owns a
    full copy of A in host memory, uploads only its 1/world share over its own PCIe link and receives the rest over
    NVLink (all-gather), multiplies its slab and copies its C slab back."""

    def __init__(self, source, rank, world, device, table=None, T=255, k=None, align=256):
        self.rank, self.world, self.device = rank, world, device
        if isinstance(source, SparseMatrix):
            self.n, self.k = source.NumCols(), source.NumRows()
            self.splits = slab_bounds(source, world)
            self.lo, self.hi = int(self.splits[rank]), int(self.splits[rank + 1])
            self.mat = source.upload_slab(self.lo, self.hi, device) if self.hi > self.lo else None
        else:
            from .matrix import partition_csr
            row_ptr, col, code = source
            self.n, self.k = int(len(row_ptr) - 1), int(k)
            self.splits = partition_csr(row_ptr, world, align)
            self.lo, self.hi = int(self.splits[rank]), int(self.splits[rank + 1])
            self.mat = None
            if self.hi > self.lo:
                rp, c, cd = csr_slab(row_ptr, col, code, self.lo, self.hi)
                self.mat = DeviceMatrix.upload_csr(rp, c, cd, table, T, self.hi - self.lo, self.k, device)
        self.nnz_local = int(self.mat.nnz) if self.mat is not None else 0
        self._stage = {}

    def product_into(self, out, d_a, m, alpha, beta, stream=0):
        if self.mat is not None:
            out.write_slab(self.mat, d_a, self.k, self.lo, alpha, beta, stream=stream)

    def product_local(self, d_a, m, d_c_slab, alpha, beta, stream=0):
        if self.mat is not None:
            self.mat.addmatmat_dev(d_a, m, self.k, d_c_slab, self.hi - self.lo, alpha, beta, stream=stream)

    def addmatmat_host(self, h_a, m, h_c_slab, alpha, beta):
        """h_a: pinned torch tensor m x k (the same on every rank); h_c_slab: pinned m x (hi - lo), updated in place."""
        import torch
        import torch.distributed as dist
        dev = torch.device("cuda", self.device)
        total = m * self.k
        share = (total + self.world - 1) // self.world
        if m not in self._stage:
            self._stage[m] = (torch.empty(share * self.world, dtype=torch.float32, device=dev),
                              torch.empty(share, dtype=torch.float32, device=dev),
                              torch.empty((m, max(1, self.hi - self.lo)), dtype=torch.float32, device=dev))
        d_full, d_part, d_c = self._stage[m]
        flat = h_a.view(-1)
        b, e = min(total, self.rank * share), min(total, (self.rank + 1) * share)
        if e > b:
            d_part[: e - b].copy_(flat[b:e], non_blocking=True)          # this rank's share of A over its own PCIe link
        if self.hi > self.lo:
            d_c.copy_(h_c_slab, non_blocking=True)
        dist.all_gather_into_tensor(d_full, d_part)                      # the rest of A over NVLink
        if self.mat is not None:
            self.mat.addmatmat_dev(d_full, m, self.k, d_c, self.hi - self.lo, alpha, beta,
                                   stream=torch.cuda.current_stream().cuda_stream)
            h_c_slab.copy_(d_c, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return h_c_slab
