#!/usr/bin/env python
"""Phase timeline of the tensor-core contraction (experiment build exp_tctl.so, tools/build_variant.sh tctl -DEXP_TC_TIMELINE
src/sparse/spmm_tc.cu): one shape per process, the library prints the phases of the LAST product at exit.
Usage: SBLAS_B200_LIB=$PWD/sparsematrix_b200/lib/exp_tctl.so python tools/tc_timeline.py n:density:m [cold|warm]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import sparsematrix_b200 as sb  # noqa: E402
from sparsematrix_b200 import gen  # noqa: E402

n, dens, m = sys.argv[1].split(":")
n, dens, m = int(n), float(dens), int(m)
cold = len(sys.argv) < 3 or sys.argv[2] == "cold"
os.environ["SBLAS_B200_SPMM_TC"] = "1"
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
S = sb.SparseMatrix(gen.codes_bernoulli(n, n, dens, 12), n, n, n, gen.codebook(), 255, sb.SparseMatrix.Trans)
mat = S.device_matrix()
a = torch.from_numpy(gen.dense_uniform((m, n), 2)).cuda()
c = torch.zeros((m, n), device="cuda")
for _ in range(4):
    if cold:
        flush.zero_()
    mat.addmatmat_dev(a, m, n, c, n, 1.0, 1.0)
torch.cuda.synchronize()
print(sys.argv[1], "cold" if cold else "warm", mat.last_product()["kernel"], flush=True)
