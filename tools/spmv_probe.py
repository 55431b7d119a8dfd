#!/usr/bin/env python
"""Time the m = 1 product of BASELINE config 4 (power-law SpMV) under several settings of the library's experiment
switches in one process: each variant re-uploads the matrix with its environment, runs 20 flushed steps and prints
the mean device time.  Usage: tools/spmv_probe.py "NAME=VALUE,NAME=VALUE" "..." (one argument per variant; "" = defaults)."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparsematrix_b200 as sb  # noqa: E402
from sparsematrix_b200 import gen  # noqa: E402

n = int(os.environ.get("PROBE_N", 1 << 23))
k = int(os.environ.get("PROBE_K", n))
row_ptr, col, code = gen.powerlaw_csr(n, k, 14)
tab = gen.codebook()
x = torch.from_numpy(gen.dense_uniform((1, k), 2)).cuda()
y = torch.zeros((1, n), device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ref = None
for variant in (sys.argv[1:] or [""]):
    env = dict(kv.split("=", 1) for kv in variant.split(",") if kv)
    old = {key: os.environ.get(key) for key in env}
    os.environ.update(env)
    t0 = time.perf_counter()
    mat = sb.DeviceMatrix.upload_csr(row_ptr, col, code, tab, 255, n, k)
    torch.cuda.synchronize()
    up = time.perf_counter() - t0
    for _ in range(3):
        mat.addmatmat_dev(x, 1, k, y, n, 1.0, 1.0)
    evs = []
    for _ in range(20):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); mat.addmatmat_dev(x, 1, k, y, n, 1.0, 1.0); b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    us = np.array([a.elapsed_time(b) for a, b in evs]) * 1e3
    y.zero_()
    mat.addmatmat_dev(x, 1, k, y, n, 1.0, 1.0)
    got = y.cpu().numpy().copy()
    y.zero_()
    if ref is None:
        ref = got
    dev = float(np.abs(got - ref).max())
    lp = mat.last_product()
    print(f"{variant or 'defaults':60s} {lp['kernel']:22s} mean {us.mean():8.1f} us  min {us.min():8.1f}  upload {up*1e3:7.0f} ms  "
          f"max|y - y_first| {dev:.2e}  {lp['algorithmic_bytes']/us.mean()/1e3:7.0f} GB/s", flush=True)
    mat.free()
    for key, v in old.items():
        if v is None: os.environ.pop(key, None)
        else: os.environ[key] = v
