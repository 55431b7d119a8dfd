#!/usr/bin/env python
"""CopyForm timing: device route vs host scan of this build's class vs the reference library (oracle/_ref),
4096^2 and 16384^2 at 90 % / 95 % sparsity, Trans input (as the benchmark builds its matrices).
Run under gpurun:  python tools/copyform_bench.py > profiles/rNN_copyform.md"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sparsematrix_b200 as sb  # noqa: E402
from sparsematrix_b200 import gen  # noqa: E402
from oracle import oracle as O  # noqa: E402   (CPU baseline only)


def best(fn, reps):
    t = 1e9
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); t = min(t, time.perf_counter() - t0)
    return t * 1e3


def main():
    import torch
    torch.zeros(1, device="cuda")
    tab = gen.codebook(255, 1)
    print("| n = k | sparsity | entries | device route (host codes in, host streams out, device matrix built) ms | device, codes already in HBM ms | host scan of this build ms | reference library CopyForm ms |")
    print("|---|---|---|---|---|---|---|")
    for n, dens in ((4096, 0.10), (16384, 0.05)):
        codes = gen.codes_bernoulli(n, n, dens, 12)
        d_codes = torch.from_numpy(codes).cuda()

        def route(mode):
            os.environ["SBLAS_B200_COPYFORM"] = mode
            S = sb.SparseMatrix(codes, n, n, n, tab, 255, sb.SparseMatrix.Trans)
            return S
        t_dev = best(lambda: route("device"), 3)
        t_host = best(lambda: route("host"), 2 if n <= 4096 else 1)
        t_hbm = best(lambda: sb.copyform_device(d_codes, n, n, n, tab, 255, True), 3)
        entries = sb.copyform_device(d_codes, n, n, n, tab, 255, True, want_matrix=False)[0].shape[0]
        t_ref = float("nan")
        if O.have_ref():
            R = O.RefMatrix()
            t_ref = best(lambda: R.CopyForm(codes, n, n, n, tab, 255, True), 2 if n <= 4096 else 1)
        print(f"| {n} | {100 * (1 - dens):.0f}% | {entries} | {t_dev:.1f} | {t_hbm:.1f} | {t_host:.1f} | {t_ref:.1f} |", flush=True)


if __name__ == "__main__":
    main()
