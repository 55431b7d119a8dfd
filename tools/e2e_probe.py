#!/usr/bin/env python
"""End-to-end time of config 3 through the host entry point (caller buffers from sblas_b200_host_alloc) under the
library's experiment switches, one process per setting (SBLAS_B200_HOST_SLICE_ROWS is read once):
    SBLAS_B200_HOST_SLICE_ROWS=128 SBLAS_B200_SPMM_TC=1 python tools/e2e_probe.py [workload [slices ...]]
where each `slices` is a value of SBLAS_B200_HOST_SLICES (slice heights, e.g. "128,256,128"; read per call)."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import sparsematrix_b200 as sb  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
inp = bench.build_inputs(name)
n, k, m = inp["n"], inp["k"], inp["m"]
S = sb.SparseMatrix(inp["codes"], n, k, k, inp["tab"], 255, sb.SparseMatrix.Trans)
mat = S.device_matrix()
h_a = sb.host_buffer((m, k)); h_a[...] = inp["a"]
h_c = sb.host_buffer((m, n)); h_c[...] = 0.0
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
settings = sys.argv[2:] or [""]                       # SBLAS_B200_HOST_SLICES values (read per call), "" = the library's choice
for setting in settings:
    if setting:
        os.environ["SBLAS_B200_HOST_SLICES"] = setting
    else:
        os.environ.pop("SBLAS_B200_HOST_SLICES", None)
    for _ in range(3):
        mat.addmatmat_host(h_a, m, k, h_c, n, 1.0, 1.0)
    ts = []
    for _ in range(12):
        flush.zero_(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        mat.addmatmat_host(h_a, m, k, h_c, n, 1.0, 1.0)
        ts.append((time.perf_counter() - t0) * 1e6)
    env = {k_: v for k_, v in os.environ.items() if k_.startswith("SBLAS_B200")}
    print(f"{name} {env}: {mat.last_product()['kernel']:22s} e2e mean {np.mean(ts):8.1f} us  min {np.min(ts):8.1f} us", flush=True)
