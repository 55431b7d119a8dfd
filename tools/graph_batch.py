"""Batched small-m products: eager calls versus one CUDA-graph replay.
Usage (GPU box, repo root): PYTHONPATH=. python tools/graph_batch.py"""
import time

import torch

import sparsematrix_b200 as sb
from sparsematrix_b200 import gen

B = 64
cuda = torch.device("cuda", 0)
print("| n = k | sparsity | m | eager us / call (python loop) | graph replay us / call | speed-up |")
print("|---|---|---|---|---|---|")
for n, dens in ((4096, 0.01), (4096, 0.10), (16384, 0.01)):
    codes = gen.codes_bernoulli(n, n, dens, seed=11)
    S = sb.SparseMatrix(codes, n, n, n, gen.codebook(), 255, 1)
    dev = S.device_matrix()
    for m in (1, 4, 16):
        a = [torch.rand((m, n), device=cuda) for _ in range(B)]
        c = [torch.zeros((m, n), device=cuda) for _ in range(B)]
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            def batch():
                for i in range(B):
                    dev.addmatmat_dev(a[i], m, n, c[i], n, 1.0, 1.0, stream=st.cuda_stream)
            batch(); batch()
            st.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(5):
                batch()
            e1.record(st)
            st.synchronize()
            eager = e0.elapsed_time(e1) / (5 * B) * 1e3
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                batch()
            g.replay(); st.synchronize()
            e0.record(st)
            for _ in range(5):
                g.replay()
            e1.record(st)
            st.synchronize()
            graph = e0.elapsed_time(e1) / (5 * B) * 1e3
        print(f"| {n} | {100 * (1 - dens):.0f} % | {m} | {eager:.2f} | {graph:.2f} | {eager / graph:.2f}x |")
