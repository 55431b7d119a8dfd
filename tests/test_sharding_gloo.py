"""The N > 1 path on CPU: two gloo ranks, each owning one nnz-balanced, panel-aligned slab.
No GPU here, so each rank produces its slab's result with the oracle (checker standing in for
the device product); what is under test is the host logic -- partition, slab extraction from the
compressed members, and the all-gather reassembly of uneven slabs."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, ret):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import sparsematrix_b200 as sb
    from sparsematrix_b200 import gen, sharding
    from oracle import oracle as O

    n, k, m = 1500, 400, 5
    codes = gen.codes_bernoulli(n, k, 0.05, seed=3)
    codes[256:512] = 255                     # an empty panel in the middle
    codes[1024:1280, ::3] = 255              # a lighter panel
    tab = gen.codebook()
    S = sb.SparseMatrix(codes, n, k, k, tab, 255, sb.SparseMatrix.Trans)
    splits = sharding.slab_bounds(S, world)
    lo, hi = int(splits[rank]), int(splits[rank + 1])
    # the slab as its own stream: panels [lo, hi) of the compressed members
    pos, code, table, T, bb, bib = S.members()
    keep = (bb[:, 1] >= lo) & (bb[:, 1] < hi)
    e0 = int(bib[keep][0, 0]) if keep.any() else 0
    e1 = int(bib[keep][-1, 1]) if keep.any() else 0
    slab = O.Stream(pos[e0:e1].copy(), code[e0:e1].copy(), table, T, bb[keep] - np.array([0, lo], np.int32),
                    bib[keep] - e0, k, hi - lo)
    a = gen.dense_uniform((m, k), 1)
    c0 = gen.dense_uniform((m, n), 2)
    c_local = np.ascontiguousarray(c0[:, lo:hi])
    O.addmatmat(slab, a, m, k, c_local, hi - lo, 1.3, 2.0)
    full = sharding.gather_output(torch.from_numpy(c_local), m, splits).numpy()
    want = O.addmatmat(O.copyform(codes, n, k, k, tab, 255, True), a, m, k, c0.copy(), n, 1.3, 2.0)
    ok = np.array_equal(full.view(np.uint32), want.view(np.uint32))
    ret[rank] = (ok, [int(x) for x in splits], int(slab.entries))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_two_rank_slabs_reassemble_to_the_unsharded_result(world):
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    assert len(ret) == world
    splits = ret[0][1]
    for r in range(world):
        ok, sp, entries = ret[r]
        assert ok, f"rank {r}: gathered result differs from the unsharded one"
        assert sp == splits
    assert splits[0] == 0 and splits[-1] == 1500 and all(x % 256 == 0 for x in splits[1:-1])
    loads = [ret[r][2] for r in range(world)]
    assert max(loads) <= 1.6 * (sum(loads) / world)        # nnz-balanced within panel granularity


def test_scaling_model_matches_survey_numbers():
    sys.path.insert(0, ROOT)
    from sparsematrix_b200.sharding import model_scaling_ceiling
    # config 4 at 8 GPUs, CSR-8B bytes: ceiling 6.7x (x replicated)
    c = model_scaling_ceiling(1 << 27, 1 << 23, 1 << 23, 1, 8, bytes_per_nnz=8)
    assert 6.5 < c < 6.9
