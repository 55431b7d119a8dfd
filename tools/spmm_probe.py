#!/usr/bin/env python
"""SpMM of one shape with the tensor-core path forced off / on (SBLAS_B200_SPMM_TC, read per call), cold (L2 flushed) and
warm.  Usage: tools/spmm_probe.py n:density:m [n:density:m ...]   e.g. 4096:0.10:128 16384:0.05:512"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparsematrix_b200 as sb  # noqa: E402
from sparsematrix_b200 import gen  # noqa: E402

flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
tab = gen.codebook()
for spec in sys.argv[1:]:
    n, dens, m = spec.split(":")
    n, dens, m = int(n), float(dens), int(m)
    k = n
    codes = gen.codes_bernoulli(n, k, dens, 12)
    S = sb.SparseMatrix(codes, n, k, k, tab, 255, sb.SparseMatrix.Trans)
    mat = S.device_matrix()
    a = torch.from_numpy(gen.dense_uniform((m, k), 2)).cuda()
    ref = None
    for mode in ("0", "1"):
        os.environ["SBLAS_B200_SPMM_TC"] = mode
        c = torch.zeros((m, n), device="cuda")
        for _ in range(3):
            mat.addmatmat_dev(a, m, k, c, n, 1.0, 1.0)
        torch.cuda.synchronize()
        evs = []
        for _ in range(20):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); mat.addmatmat_dev(a, m, k, c, n, 1.0, 1.0); e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        cold = np.array([x.elapsed_time(y) for x, y in evs]) * 1e3
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            mat.addmatmat_dev(a, m, k, c, n, 1.0, 1.0)
        e1.record(); torch.cuda.synchronize()
        warm = e0.elapsed_time(e1) / 20 * 1e3
        mat.profile_begin()
        for _ in range(5):
            flush.zero_(); mat.addmatmat_dev(a, m, k, c, n, 1.0, 1.0)
        kms, kcnt = mat.profile_end()
        c.zero_(); mat.addmatmat_dev(a, m, k, c, n, 1.0, 1.0); torch.cuda.synchronize()
        got = c.cpu().numpy()
        if ref is None:
            ref = got
        diff = float(np.abs(got - ref).max() / max(1e-30, np.abs(ref).max()))
        nnz = int((codes != 255).sum())
        print(f"{n}^2 dens {dens} m {m} TC={mode}: {mat.last_product()['kernel']:24s} cold {cold.mean():8.1f} us (min {cold.min():8.1f}) warm {warm:8.1f} us "
              f"kernel {kms / max(kcnt, 1) * 1e3:8.1f} us  {2.0 * nnz * m / cold.mean() / 1e3:9.0f} GFLOP/s sparse, "
              f"{2.0 * n * k * m / cold.mean() / 1e6:7.1f} TFLOP/s dense  normwise diff vs TC=0 {diff:.1e}", flush=True)
    os.environ.pop("SBLAS_B200_SPMM_TC", None)
    del mat, S
