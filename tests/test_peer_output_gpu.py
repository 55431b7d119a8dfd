"""Two GPUs, one process each (NCCL): every rank's product kernels store their slab of C straight
into rank 0's memory through an NVLink-mapped pointer (sharding.PeerOutput), and the assembled C
must equal, bit for bit, the all-gather of the slabs computed into local memory -- the kernels
are the same, only the C pointer differs (for beta == 1 the SpMM tile kernel adds into peer memory
with red.global.add.f32 instead of a read-modify-write: same single rounding, cases m = 64 and 8; the
SpMV kernels likewise, m = 1 here and on a skewed CSR matrix with split long rows).
Skipped on a box with one GPU."""
import os
import socket
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ret):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    import sparsematrix_b200 as sb
    from sparsematrix_b200 import gen, sharding

    n, k = 2048 + 300, 700
    codes = gen.codes_bernoulli(n, k, 0.08, seed=5)
    codes[512:768] = 255                                    # an empty panel
    S = sb.SparseMatrix(codes, n, k, k, gen.codebook(), 255, sb.SparseMatrix.Trans)
    mat, (lo, hi) = sharding.upload_local_slab(S, rank, world, device=rank)
    splits = sharding.slab_bounds(S, world)
    dev = torch.device("cuda", rank)
    results = []
    for m, alpha, beta in ((1, 1.0, 1.0), (16, 1.3, 2.0), (128, 1.0, 0.5), (64, 1.0, 1.0), (8, 1.7, 1.0)):
        a = torch.from_numpy(gen.dense_uniform((m, k), 10 + m)).to(dev)
        c0 = torch.from_numpy(gen.dense_uniform((m, n), 20 + m)).to(dev)
        # (1) local slab + NCCL all-gather
        c_local = c0[:, lo:hi].clone()                   # (.contiguous() of a one-row slice is a view of c0)
        if mat is not None:
            mat.addmatmat_dev(a, m, k, c_local, hi - lo, alpha, beta)
        full_gather = sharding.gather_output(c_local, m, splits)
        # (2) stores through the peer mapping, no collective on the data path
        out = sharding.PeerOutput(m, n, root=0, device=rank)
        if rank == 0:
            out.tensor().copy_(c0)
        out.finish()                                        # C initialised before anyone updates it
        if mat is not None:
            out.write_slab(mat, a, k, lo, alpha, beta)
        out.finish()
        if rank == 0:
            got = out.tensor().cpu().numpy().view(np.uint32)
            want = full_gather.cpu().numpy().view(np.uint32)
            results.append(bool(np.array_equal(got, want)))
            # and the columns of the empty panel only got the scale
            ref = (c0[:, 512:768] * beta if beta != 1.0 else c0[:, 512:768]).cpu().numpy().view(np.uint32)
            results.append(bool(np.array_equal(got.reshape(m, n)[:, 512:768], ref)))
        out.close()

    # A skewed CSR matrix (the length-sorted SpMV copy and the split long rows), rows halved by hand: the remote
    # rank forms its row sums locally and adds them out in row order (reductions for beta == 1, product.cu:
    # run_spmv_sorted_to_peer); the unsorted row kernels above add theirs by reductions straight from the kernel.
    n2 = k2 = 1 << 17                                       # >= 65536 rows in the remote part: it gets the sorted copy
    row_ptr, col, code = gen.powerlaw_csr(n2, k2, seed=14, x_min=8, cap=1 << 15)
    half = n2 // 2 + 77
    r0, r1 = (half, n2) if rank == 0 else (0, half)      # the 12436-entry row is in the remote rank's part
    e0, e1 = int(row_ptr[r0]), int(row_ptr[r1])
    part = sb.DeviceMatrix.upload_csr((row_ptr[r0:r1 + 1] - e0).astype(np.int32), col[e0:e1], code[e0:e1],
                                      gen.codebook(), 255, r1 - r0, k2, device=rank)
    inf = part.info()
    a = torch.from_numpy(gen.dense_uniform((1, k2), 3)).to(dev)
    c0 = torch.from_numpy(gen.dense_uniform((1, n2), 4)).to(dev)
    for alpha, beta in ((1.0, 1.0), (0.7, 1.0), (1.3, 2.0)):
        c_local = c0[:, r0:r1].clone()
        part.addmatmat_dev(a, 1, k2, c_local, r1 - r0, alpha, beta)
        pieces = [torch.empty((1, n2 - half), device=dev), torch.empty((1, half), device=dev)]
        dist.all_gather(pieces, c_local)
        full_gather = torch.cat([pieces[1], pieces[0]], dim=1)
        out = sharding.PeerOutput(1, n2, root=0, device=rank)
        if rank == 0:
            out.tensor().copy_(c0)
        out.finish()
        out.write_slab(part, a, k2, r0, alpha, beta)
        out.finish()
        if rank == 0:
            results.append(bool(np.array_equal(out.tensor().cpu().numpy().view(np.uint32),
                                               full_gather.cpu().numpy().view(np.uint32))))
        out.close()
    if rank == 1:
        ret["skewed"] = (int(inf.long_rows), int(inf.max_row_nnz))
    if rank == 0:
        ret["ok"] = results
        ret["splits"] = [int(x) for x in splits]
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_peer_mapped_output_equals_allgather_of_local_slabs():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, 2, port, ret)) for r in range(2)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(300)
        for p in procs:
            if p.is_alive():
                p.kill()
            assert p.exitcode == 0, f"worker exit code {p.exitcode}"
        assert ret["ok"] and all(ret["ok"]), dict(ret)
        assert ret["splits"][0] == 0 and ret["splits"][-1] == 2348
        assert ret["skewed"][0] > 0 and ret["skewed"][1] > 8192, dict(ret)     # the remote part had split rows
