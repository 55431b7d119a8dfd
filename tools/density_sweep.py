#!/usr/bin/env python
"""Density sweep (BASELINE.json configs[4]): n = k = 4096 (and optionally 16384), sparsity 50 % .. 99.9 %,
m in {1, 32, 256}: GPU product (device pointers, cold = L2 flushed, and warm) against the reference's own
CPU AddMatMat (oracle/_ref, 1 thread -- the library is single threaded) on the same inputs.
Writes a markdown table; run under gpurun:  python tools/density_sweep.py [--big] > profiles/rNN_density_sweep.md"""
import argparse
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import sparsematrix_b200 as sb  # noqa: E402
from sparsematrix_b200 import gen  # noqa: E402
from oracle import oracle as O  # noqa: E402   (CPU baseline leg only)


def gpu_time(mat, d_a, d_c, m, k, n, flush, steps=30):
    st = torch.cuda.current_stream()
    for _ in range(5):
        mat.addmatmat_dev(d_a, m, k, d_c, n, 1.0, 1.0, stream=st.cuda_stream)
    torch.cuda.synchronize()
    cold = 0.0
    for _ in range(steps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        mat.addmatmat_dev(d_a, m, k, d_c, n, 1.0, 1.0, stream=st.cuda_stream)
        b.record(st)
        torch.cuda.synchronize()
        cold += a.elapsed_time(b)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    for _ in range(steps):
        mat.addmatmat_dev(d_a, m, k, d_c, n, 1.0, 1.0, stream=st.cuda_stream)
    b.record(st)
    torch.cuda.synchronize()
    return cold / steps * 1e3, a.elapsed_time(b) / steps * 1e3          # us


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--big", action="store_true", help="also n = k = 16384 (GPU only)")
    ap.add_argument("--m", type=int, nargs="*", default=[1, 32, 256])
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--sparsity", type=float, nargs="*", default=[50.0, 75.0, 90.0, 95.0, 99.0, 99.9])
    ap.add_argument("--sizes", type=int, nargs="*", default=None)
    ap.add_argument("--cpu-big", action="store_true", help="reference CPU and OpenBLAS columns at 16384^2 too (m <= 32 for the reference)")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    tab = gen.codebook(255, 1)
    print("| n=k | sparsity | m | nnz | kernel | GPU cold us | GPU warm us | GPU GFLOP/s (cold) | ref CPU ms (1 thread) | CPU GFLOP/s | GPU/CPU | "
          "dense OpenBLAS ms, 1 thread | dense OpenBLAS ms, all cores |")
    print("|---|---|---|---|---|---|---|---|---|---|---|---|---|")
    try:
        from threadpoolctl import threadpool_limits
    except Exception:
        threadpool_limits = None
    sizes = args.sizes or ([4096] + ([16384] if args.big else []))
    for n in sizes:
        for idx, sp in enumerate([50.0, 75.0, 90.0, 95.0, 99.0, 99.9]):
            if sp not in args.sparsity:
                continue
            codes = gen.codes_bernoulli(n, n, 1.0 - sp / 100.0, 100 + idx)
            nnz = int((codes != 255).sum())
            S = sb.SparseMatrix(codes, n, n, n, tab, 255, sb.SparseMatrix.Trans)
            mat = S.device_matrix()
            R = None
            cpu_ok = O.have_ref() and not args.no_cpu
            if cpu_ok and (n == 4096 or args.cpu_big):
                R = O.RefMatrix()
                R.CopyForm(codes, n, n, n, tab, 255, True)
            w = None
            if cpu_ok and threadpool_limits is not None:     # the reference harness's cross-check: sgemm / sgemv on the dense expansion
                w = np.where(codes < 255, tab[np.minimum(codes, 254)], np.float32(0)).astype(np.float32)
            for m in args.m:
                a = gen.dense_uniform((m, n), 2)
                d_a = torch.from_numpy(a).to(dev)
                d_c = torch.zeros((m, n), dtype=torch.float32, device=dev)
                cold, warm = gpu_time(mat, d_a, d_c, m, n, n, flush)
                kernel = mat.last_product()["kernel"]
                flops = 2.0 * nnz * m
                cpu_ms = cpu_gf = ratio = float("nan")
                if R is not None and (n == 4096 or m <= 32):   # 16384^2 x 256 at 50 % would take a minute per call
                    c = np.zeros((m, n), np.float32)
                    best = 1e9
                    for _ in range(2 if n == 4096 else 1):
                        t0 = time.perf_counter()
                        R.AddMatMat(a, m, n, c, n, 1.0, 1.0)
                        best = min(best, time.perf_counter() - t0)
                    cpu_ms, cpu_gf, ratio = best * 1e3, flops / best / 1e9, best * 1e6 / cold
                dense = [float("nan"), float("nan")]
                if w is not None:
                    for which, threads in enumerate((1, os.cpu_count())):
                        if threads == 1 and n > 4096 and m > 32:
                            continue                          # 137 GFLOP on one thread: skipped
                        with threadpool_limits(limits=threads, user_api="blas"):
                            run = (lambda: np.dot(a, w.T)) if m > 1 else (lambda: np.dot(w, a[0]))
                            run()
                            t0 = time.perf_counter(); run(); dense[which] = (time.perf_counter() - t0) * 1e3
                print(f"| {n} | {sp}% | {m} | {nnz} | {kernel} | {cold:.1f} | {warm:.1f} | {flops / cold / 1e3:.1f} | {cpu_ms:.2f} | {cpu_gf:.2f} | {ratio:.0f}x | "
                      f"{dense[0]:.2f} | {dense[1]:.2f} |", flush=True)
            del S, mat, w


if __name__ == "__main__":
    main()
