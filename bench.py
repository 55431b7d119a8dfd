#!/usr/bin/env python
"""bench.py -- the hot path (SparseMatrix::AddMatMat on the GPU) on BASELINE.json's configs.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload all|cfg1..cfg5] [--impl ours|reference]

A step = one pass of the product over one batch of synthetic input:
  cfg1  SpMV  4096 x 4096, 1 % dense, m = 1                             (BASELINE.json configs[0])
  cfg2  SpMM  4096 x 4096, 90 % sparse, x dense 4096 x 128              (configs[1])
  cfg3  SpMM  16384 x 16384, 95 % sparse, x dense N = 512               (configs[2]; HEADLINE: the config the
                                                                         metric's "1/2/4/8 B200" clause names)
  cfg4  SpMV  power-law rows (Pareto, ~16 nnz/row), 2^23 rows, direct CSR   (configs[3])
  cfg5  SpMV  16384 x 16384 at 50 % sparsity (large end of the density sweep, configs[4])
alpha = beta = 1 (the reference's own sparse benchmark setting, src/test/blas_test.h:313).

Default (--workload all): the JSON line's top-level keys are cfg3's, `per_workload` carries the same
measurements for cfg1, cfg2, cfg4 and cfg5.  --workload cfgX measures that one only (ncu captures).

N > 1 (torchrun, one rank per GPU): STRONG scaling of the ONE matrix.  Every rank builds the same host
matrix, `sblas_b200_partition` (panel streams) or `sblas_b200_partition_csr` (cfg4) cuts it into nnz-balanced
row slabs of W, rank r uploads slab r (`sblas_b200_upload_slab` / `_upload_csr`), X is replicated, and the
timed step INCLUDES assembling C on rank 0: every rank's product kernels store their slab straight into
rank 0's memory over NVLink (`sharding.PeerOutput`), no collective on the data path.  `value` = flops of the
whole matrix / max-over-ranks device time.  After the timed region rank 0 checks the assembled C (sampled
rows against an fp64 evaluation, and bit-equality with the single-GPU product of the same matrix for m > 1).

One JSON line on stdout (rank 0).  Additions to the contract's keys: roofline (dominant kernel, event-timed
through sblas_b200_profile_*; bound = whichever of HBM / fp32 FMA takes longer for the algorithmic work),
cpu_baseline (the unmodified reference, oracle/_ref, on the host cores), e2e (host buffers through the
reference-facing host entry point, copies inside the timed region), parity_check, upload_ms.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from sparsematrix_b200 import gen  # noqa: E402

WORKLOADS = {
    #        kind    n      k      density  m    seed
    "cfg1": ("spmv", 4096, 4096, 0.01, 1, 11),
    "cfg2": ("spmm", 4096, 4096, 0.10, 128, 12),
    "cfg3": ("spmm", 16384, 16384, 0.05, 512, 13),
    "cfg4": ("spmv_powerlaw", 1 << 23, 1 << 23, None, 1, 14),
    "cfg5": ("spmv", 16384, 16384, 0.50, 1, 100),
}
HEADLINE = "cfg3"
ORDER = ["cfg3", "cfg4", "cfg2", "cfg1", "cfg5"]
DESCR = {
    "cfg1": "SpMV random 4096x4096 1% density (BASELINE.json configs[0])",
    "cfg2": "pruned-weight SpMM 4096x4096 at 90% sparsity x dense 4096x128 (BASELINE.json configs[1])",
    "cfg3": "SpMM 16384x16384 at 95% sparsity x dense N=512 (BASELINE.json configs[2])",
    "cfg4": "power-law SpMV 2^23 rows ~16 nnz/row, direct CSR (BASELINE.json configs[3])",
    "cfg5": "density sweep, large end: SpMV 16384x16384 at 50% sparsity (BASELINE.json configs[4])",
}
FP32_PEAK_NOMINAL_TFLOPS = 2 * 128 * 148 * 1.965e9 / 1e12     # CUDA-core FMA peak at max clock
TF32_PEAK_NOMINAL_TFLOPS = 2250.0 / 2                         # dense TF32 on tcgen05: half the nominal bf16 rate
L2_NOTE = "GPU arm: L2 flushed before every timed step (256 MiB write); per-step CUDA events summed"
METRIC = "SpMV/SpMM GFLOP/s (2*nnz*m / time)"
TOL = 1e-5                                                    # backward error bound of the parity check


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def build_inputs(workload, scale_rows=1.0):
    """Host-side inputs, the same on every rank: the matrix (codes or CSR), A (m x k), C (m x n)."""
    kind, n, k, dens, m, seed = WORKLOADS[workload]
    n = max(256, int(n * scale_rows) // 256 * 256)
    tab = gen.codebook(255, 1)
    a = gen.dense_uniform((m, k), 2)
    c = np.zeros((m, n), np.float32)                       # C = 0 for timing
    if kind == "spmv_powerlaw":
        row_ptr, col, code = gen.powerlaw_csr(n, k, seed)
        return dict(name=workload, kind=kind, n=n, k=k, m=m, tab=tab, a=a, c=c, csr=(row_ptr, col, code), nnz=int(row_ptr[-1]))
    codes = gen.codes_bernoulli(n, k, dens, seed)
    return dict(name=workload, kind=kind, n=n, k=k, m=m, tab=tab, a=a, c=c, codes=codes, nnz=int((codes != 255).sum()))


def config_of(inp):
    """Identical in both arms (the driver compares them)."""
    return {"workload": inp["name"], "descr": DESCR[inp["name"]], "n": inp["n"], "k": inp["k"], "m": inp["m"],
            "nnz": inp["nnz"], "alpha": 1.0, "beta": 1.0, "l2": L2_NOTE}


class ClockSampler:
    """SM clock + throttle reasons of one GPU while the timed region runs (pynvml, ~2 ms period)."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self._stop, self._t = [], set(), None, False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._stop:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.002)

    def __enter__(self):
        if self.nv:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *exc):
        self._stop = True
        if self._t:
            self._t.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


class CpuReference:
    """The reference's own CPU implementation of the step: the unmodified library (oracle/_ref, kind
    "reference") when it was built, else the C port (kind "port").  The 2^23-row power-law shape cannot be
    held by the reference's format at all: there the CSR restatement runs on a bounded
    sample (the first 2^20 rows) and is scaled by nnz."""

    def __init__(self, inp):
        from oracle import oracle as O      # bench.py's only use of oracle/: the CPU baseline legs
        self.inp, self.scale, self.O = inp, 1.0, O
        a, m, n, k = inp["a"], inp["m"], inp["n"], inp["k"]
        if inp["kind"] == "spmv_powerlaw":
            row_ptr, col, code = inp["csr"]
            rows = min(n, 1 << 20)
            rp = np.ascontiguousarray(row_ptr[: rows + 1])
            self.scale = inp["nnz"] / float(rp[-1])
            self.kind, self.rows = "port", rows
            self.sample = f"first {rows} rows of the same matrix ({1 / self.scale:.3f} of the nnz), time scaled by nnz"
            self._run = lambda c: O.addmatmat_csr(rp, col, code, inp["tab"], 255, rows, k, a, 1, k, c, rows, 1.0, 1.0)
            return
        self.rows = n
        self.sample = "full workload, one AddMatMat call per step"
        if O.have_ref():
            R = O.RefMatrix()
            R.CopyForm(inp["codes"], n, k, k, inp["tab"], 255, True)
            self.kind, self._keep = "reference", R
            self._run = lambda c: R.AddMatMat(a, m, k, c, n, 1.0, 1.0)
        else:
            st = O.copyform(inp["codes"], n, k, k, inp["tab"], 255, True)
            self.kind = "port"
            self._run = lambda c: O.addmatmat(st, a, m, k, c, n, 1.0, 1.0)

    def call(self):
        """One step; seconds for the whole workload."""
        c = np.ascontiguousarray(self.inp["c"][:, : self.rows])
        c = c.copy()
        t0 = time.perf_counter()
        self._run(c)
        return (time.perf_counter() - t0) * self.scale

    def omp_csr(self):
        """cfg4 only: the CSR restatement with OpenMP over rows on all host cores -- NOT the reference, which is
        single-threaded.  -> (seconds, threads)"""
        row_ptr, col, code = self.inp["csr"]
        x = np.ascontiguousarray(self.inp["a"].reshape(-1))
        y = np.zeros(self.inp["n"], np.float32)
        self.O.addmatmat_csr_omp(row_ptr, col, code, self.inp["tab"], 255, self.inp["n"], x, y, 1.0, 1.0)
        best, threads = 1e9, 1
        for _ in range(3):
            t0 = time.perf_counter()
            threads = self.O.addmatmat_csr_omp(row_ptr, col, code, self.inp["tab"], 255, self.inp["n"], x, y, 1.0, 1.0)
            best = min(best, time.perf_counter() - t0)
        return best, threads


def dense_openblas_legs(inp, flops):
    """The reference harness's own cross-check (src/test/blas_test.cc:16-19): dense sgemm / sgemv on the
    CopyTo-expanded matrix, with one OpenBLAS thread and with all host cores (numpy's bundled OpenBLAS, thread
    count through threadpoolctl).  Reported in sparse flops per second."""
    try:
        from threadpoolctl import threadpool_limits
    except Exception:
        return None
    w = np.where(inp["codes"] < 255, inp["tab"][np.minimum(inp["codes"], 254)], np.float32(0)).astype(np.float32)
    wt = np.ascontiguousarray(w.T) if inp["m"] > 1 else w      # sgemm: A (m x k) . W^T (k x n); sgemv: W (n x k) . x
    a = inp["a"]
    out = {}
    for label, threads in (("1_thread", 1), ("all_cores", os.cpu_count())):
        with threadpool_limits(limits=threads, user_api="blas"):
            run = (lambda: np.dot(a, wt)) if inp["m"] > 1 else (lambda: np.dot(w, a[0]))
            run()
            best = 1e9
            for _ in range(2 if threads == 1 else 3):
                t0 = time.perf_counter(); run(); best = min(best, time.perf_counter() - t0)
        out[label] = {"value": flops / best / 1e9, "unit": "GFLOP/s (sparse flops)", "ms_per_step": best * 1e3, "threads": threads,
                      "dense_gflops": 2.0 * inp["n"] * inp["k"] * inp["m"] / best / 1e9}
    out["what"] = "numpy.dot = OpenBLAS sgemm/sgemv on the dense expansion of the same matrix (blas_test.cc:16-19)"
    return out


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the same step on the host cores.
    The library is single-threaded, so "all the host threads it can use" is 1."""
    rank, _, _ = dist_env()
    if rank != 0:
        return
    name = HEADLINE if args.workload == "all" else args.workload
    inp = build_inputs(name, args.scale_rows)
    flops = 2.0 * inp["nnz"] * inp["m"]
    ref = CpuReference(inp)
    first = ref.call()                                             # also the first warm-up step
    # bound the run: at most ~90 s of CPU work in total
    budget = max(1, int(90.0 / max(first / ref.scale, 1e-6)))
    warm = max(0, min(args.warmup - 1, budget // 4))
    steps = max(1, min(args.steps, budget - warm))
    for _ in range(warm):
        ref.call()
    t0 = time.perf_counter()
    times = [ref.call() for _ in range(steps)]
    wall = time.perf_counter() - t0
    sec = float(np.mean(times))
    val = flops / sec / 1e9
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "GFLOP/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": warm + 1, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "strong" if args.gpus > 1 else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_of(inp),
        "cpu_baseline": {"value": val, "unit": "GFLOP/s", "cores": 1, "kind": ref.kind, "sample": ref.sample,
                         "host_cores_available": os.cpu_count()},
        "e2e": {"value": val, "unit": "GFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": wall,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------
# the GPU arm
# ---------------------------------------------------------------------------------------------------------
class Ctx:
    pass


def fp64_sample_check(inp, y, rng_seed=77):
    """Sampled output rows of W against an fp64 evaluation: max over samples of
    |y - y64| / (|alpha| sum |a||w|)  (C = 0, alpha = beta = 1).  numpy only."""
    n, k, m = inp["n"], inp["k"], inp["m"]
    rng = np.random.default_rng(rng_seed)
    if "csr" in inp:
        row_ptr, col, code = inp["csr"]
        lens = np.diff(row_ptr)
        rows = np.unique(np.concatenate([rng.integers(0, n, 512), np.argsort(lens)[-4:], [0, n - 1]]))
        x64 = inp["a"][0].astype(np.float64)
        tab64 = inp["tab"].astype(np.float64)
        worst = 0.0
        for j in rows:
            e0, e1 = int(row_ptr[j]), int(row_ptr[j + 1])
            t = x64[col[e0:e1]] * tab64[code[e0:e1]]
            bound = float(np.abs(t).sum())
            err = abs(float(y[0, j]) - float(t.sum()))
            worst = max(worst, err / bound if bound > 0 else (0.0 if err == 0 else np.inf))
        return worst, int(len(rows))
    rows = np.unique(np.concatenate([rng.integers(0, n, 62), [0, n - 1]]))
    cj = inp["codes"][rows]
    w = np.where(cj < 255, inp["tab"][np.minimum(cj, 254)], np.float32(0)).astype(np.float64)     # |J| x k
    a64 = inp["a"].astype(np.float64)
    y64 = a64 @ w.T                                                                             # m x |J|
    bound = np.abs(a64) @ np.abs(w).T
    err = np.abs(y[:, rows].astype(np.float64) - y64)
    with np.errstate(divide="ignore", invalid="ignore"):
        q = np.where(bound > 0, err / bound, np.where(err == 0, 0.0, np.inf))
    return float(q.max()), int(len(rows))


def measure(name, args, cx):
    """All measurements of one workload; returns the result dict on rank 0 (None elsewhere)."""
    torch, dist, sb, sharding = cx.torch, cx.dist, cx.sb, cx.sharding
    rank, world, dev, stream, flush = cx.rank, cx.world, cx.dev, cx.stream, cx.flush
    L = sb.lib()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def maxr(vals):
        t = torch.tensor(vals, dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return [float(v) for v in t.tolist()]

    inp = build_inputs(name, args.scale_rows)
    n, k, m, nnz = inp["n"], inp["k"], inp["m"], inp["nnz"]
    flops = 2.0 * nnz * m

    # ---- upload (once per matrix; not in the timed region, reported as upload_ms) ------------------------
    barrier()
    t0 = time.perf_counter()
    host_obj, copyform_ms = None, None
    if "codes" in inp:
        host_obj = sb.SparseMatrix(inp["codes"], n, k, k, inp["tab"], 255, sb.SparseMatrix.Trans)   # CopyForm (on the device from 2^18 elements)
        torch.cuda.synchronize()
        copyform_ms = (time.perf_counter() - t0) * 1e3
    t1 = time.perf_counter()
    sh = None
    if world == 1:
        if host_obj is None:
            row_ptr, col, code = inp["csr"]
            mat = sb.DeviceMatrix.upload_csr(row_ptr, col, code, inp["tab"], 255, n, k, device=cx.local_rank)
        else:
            mat = host_obj.device_matrix()                      # upload (once); borrowed handle
        lo, hi, splits = 0, n, [0, n]
    else:
        sh = sharding.ShardedMatrix(host_obj if host_obj is not None else inp["csr"], rank, world, cx.local_rank,
                                    table=inp["tab"], T=255, k=k)
        mat, lo, hi, splits = sh.mat, sh.lo, sh.hi, [int(s) for s in sh.splits]
    torch.cuda.synchronize()
    upload_ms = (time.perf_counter() - t1) * 1e3
    nnz_local = int(mat.nnz) if mat is not None else 0
    if world == 1:
        assert nnz_local == nnz, (nnz_local, nnz)

    d_a = torch.from_numpy(inp["a"]).to(dev)
    d_c = torch.zeros((m, max(1, hi - lo)), dtype=torch.float32, device=dev)        # this rank's slab of C, local
    out = None
    if world > 1:
        out = sharding.PeerOutput(m, n, root=0, device=cx.local_rank)              # the whole C on rank 0's GPU
        if rank == 0:
            out.tensor().zero_()
        out.finish()

    def step_local():
        if mat is not None:
            mat.addmatmat_dev(d_a, m, k, d_c, hi - lo, 1.0, 1.0, stream=stream.cuda_stream)

    def step_assembled():                      # the step `value` is quoted on: C ends up whole on rank 0's GPU
        if world == 1:
            step_local()
        else:
            sh.product_into(out, d_a, m, 1.0, 1.0, stream=stream.cuda_stream)

    def timed_pass(step, steps, profile=False):
        """K steps, L2 flushed before each, one event pair per step on the launching stream."""
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        if profile and mat is not None:
            mat.profile_begin()
        barrier()
        for a_ev, b_ev in evs:
            flush.zero_()
            a_ev.record(stream)
            step()
            b_ev.record(stream)
        barrier()
        per_step = [a_ev.elapsed_time(b_ev) for a_ev, b_ev in evs]
        if profile:
            k_ms, k_cnt = mat.profile_end() if mat is not None else (0.0, 0)
            return per_step, k_ms, k_cnt
        return per_step

    warmup = max(args.warmup, 3)
    for _ in range(warmup):
        flush.zero_()
        step_assembled()
    barrier()

    launches0 = L.sblas_b200_launch_count()
    with ClockSampler(cx.local_rank) as clocks:
        per_step = timed_pass(step_assembled, args.steps)
    launches = L.sblas_b200_launch_count() - launches0
    total_ms = sum(per_step)
    srt = sorted(per_step)
    step_median_ms, step_min_ms = srt[len(srt) // 2], srt[0]
    lp = mat.last_product() if mat is not None else {"kernel": "(empty slab)", "algorithmic_bytes": 0, "algorithmic_flops": 0,
                                                     "format_bytes_per_nnz": 5.0}
    # dominant kernel alone (cold), for the roofline: the same protocol as the timed pass, right after it (a pass of
    # back-to-back products in between heats a tensor-core-bound workload into its power cap: cfg3 ran 9 % slower in
    # this pass than in the timed one when it came third), clocks sampled again
    # Its own pass: the library's event pairs around the kernels cost every step about 5 us (cfg1 10 -> 15 us when they
    # were recorded inside the timed pass), so the step time above is measured without them.  At most 20 steps after
    # 0.3 s of idling: config 3 keeps the tensor pipe 90 % busy and runs into the board's power cap after about 50 ms of
    # it (1680 MHz, sw_power_cap: this pass measured 1023 us kernels right after the 50 timed steps of 939 us).
    time.sleep(0.3)
    for _ in range(2):
        flush.zero_()
        step_assembled()
    with ClockSampler(cx.local_rank) as clocks_profile:
        _, kern_ms, kern_cnt = timed_pass(step_assembled, min(args.steps, 20), profile=True)
    kern_us = kern_ms / max(kern_cnt, 1) * 1e3
    # warm (L2-resident) variant, reported beside the cold number
    barrier()
    w0, w1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    w0.record(stream)
    for _ in range(args.steps):
        step_assembled()
    w1.record(stream)
    barrier()
    warm_ms = w0.elapsed_time(w1)
    # N > 1: the same step with C staying sharded (each rank's slab in its own memory), and the NCCL reassembly of that
    local_ms = gather_ms = None
    if world > 1:
        for _ in range(3):
            step_local()
        local_ms = sum(timed_pass(step_local, args.steps))
        for _ in range(2):
            sharding.gather_output(d_c[:, : hi - lo].contiguous() if hi > lo else d_c[:, :0], m, splits)
        barrier()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record(stream)
        for _ in range(5):
            sharding.gather_output(d_c[:, : hi - lo].contiguous() if hi > lo else d_c[:, :0], m, splits)
        g1.record(stream)
        barrier()
        gather_ms = g0.elapsed_time(g1) / 5

    # ---- end to end through the reference-facing host entry point: host buffers, H2D of A and C, kernels, D2H of C
    # inside every timed step.  Caller buffers from the library's allocator (sblas_b200_host_alloc = cudaHostAlloc,
    # mapped; profiles/r01d_e2e.md), and -- N = 1 -- the same call from plain pageable numpy arrays (what an unmodified
    # caller of the reference passes, src/test/blas_test.h:121).
    e2e_steps = max(3, min(args.steps, 30))
    h_a = sb.host_buffer((m, k))
    h_a[...] = inp["a"]
    h_c = sb.host_buffer((m, max(1, hi - lo)))
    h_c[...] = 0.0
    if world == 1:
        def e2e_call():
            mat.addmatmat_host(h_a, m, k, h_c, n, 1.0, 1.0)       # returns when h_c is valid
    else:
        t_a, t_c = torch.from_numpy(h_a), torch.from_numpy(h_c)

        def e2e_call():
            sh.addmatmat_host(t_a, m, t_c, 1.0, 1.0)             # 1/N of A over this rank's PCIe link, the rest over NVLink
    for _ in range(3):
        e2e_call()
    e2e_s = 0.0
    for _ in range(e2e_steps):
        flush.zero_()
        barrier()
        t0 = time.perf_counter()
        e2e_call()
        e2e_s += time.perf_counter() - t0
    checksum = float(h_c.sum())
    pageable_s = registered_s = None
    if world == 1:
        p_a, p_c = inp["a"].copy(), np.zeros((m, n), np.float32)

        def pageable_pass():
            for _ in range(3):
                mat.addmatmat_host(p_a, m, k, p_c, n, 1.0, 1.0)
            sec = 0.0
            for _ in range(e2e_steps):
                flush.zero_()
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                mat.addmatmat_host(p_a, m, k, p_c, n, 1.0, 1.0)
                sec += time.perf_counter() - t0
            return sec
        pageable_s = pageable_pass()
        os.environ["SBLAS_B200_HOST_REGISTER"] = "1"            # opt-in: the library page-locks a buffer it sees a second time
        try:
            registered_s = pageable_pass()
        finally:
            os.environ.pop("SBLAS_B200_HOST_REGISTER", None)
            L.sblas_b200_host_forget(None)

    # ---- parity of what the timed step produces (after the timed region) -----------------------------------
    if world == 1:
        d_c.zero_()
        step_local()
        torch.cuda.synchronize()
        y = d_c.cpu().numpy()
    else:
        if rank == 0:
            out.tensor().zero_()
        out.finish()
        step_assembled()
        out.finish()
        y = out.tensor().cpu().numpy() if rank == 0 else None
    parity = None
    if rank == 0:
        back, nrows = fp64_sample_check(inp, y)
        parity = {"ok": bool(back <= TOL), "backward": back, "tolerance": TOL, "sampled_rows": nrows,
                  "what": "assembled C (one product into C = 0): sampled rows of W against an fp64 evaluation, |y - y64| / sum |a||w|"}
        if world > 1:
            # the single-GPU product of the same matrix on rank 0
            if host_obj is not None:
                full = host_obj.device_matrix()
            else:
                row_ptr, col, code = inp["csr"]
                full = sb.DeviceMatrix.upload_csr(row_ptr, col, code, inp["tab"], 255, n, k, device=cx.local_rank)
            d_full = torch.zeros((m, n), dtype=torch.float32, device=dev)
            full.addmatmat_dev(d_a, m, k, d_full, n, 1.0, 1.0, stream=stream.cuda_stream)
            torch.cuda.synchronize()
            y1 = d_full.cpu().numpy()
            same = bool(np.array_equal(y.view(np.uint32), y1.view(np.uint32)))
            parity["bit_equal_single_gpu"] = same
            parity["max_abs_diff_single_gpu"] = float(np.abs(y - y1).max())
            back1, _ = fp64_sample_check(inp, y1)
            parity["backward_single_gpu"] = back1
            # m > 1: a slab gives the bits of the whole matrix as long as both run the same kernel family (a small slab may
            # take the tile kernel where the whole matrix takes the tensor-core form: different order of additions, both
            # held to the tolerance).  m = 1: a slab's row sums may differ in the last bit
            parity["kernel_single_gpu"] = full.last_product()["kernel"]
            same_family = lp["kernel"].split("<")[0] == parity["kernel_single_gpu"].split("<")[0]
            parity["same_kernel_family"] = same_family
            if m > 1 and same_family:
                parity["ok"] = parity["ok"] and same
            parity["ok"] = parity["ok"] and bool(back1 <= TOL)
            del d_full, full
    ok_flag = maxr([0.0 if (parity is None or parity["ok"]) else 1.0])[0] == 0.0

    total_ms, warm_ms, e2e_ms, kern_us_max = maxr([total_ms, warm_ms, e2e_s * 1e3, kern_us])
    local_ms_max = maxr([local_ms])[0] if local_ms is not None else None
    nnz_all = [nnz_local]
    if world > 1:
        gathered = [None] * world
        dist.all_gather_object(gathered, nnz_local)
        nnz_all = [int(g) for g in gathered]

    res = None
    if rank == 0:
        hbm_peak, peak_src = measured_peaks()
        fp32_peak = cx.fp32_peak_tflops
        ms_step = total_ms / args.steps
        value = flops / (ms_step * 1e-3) / 1e9
        bytes_alg, flops_k = float(lp["algorithmic_bytes"]), float(lp["algorithmic_flops"])      # of rank 0's kernel launch
        t_hbm_us, t_fp32_us = bytes_alg / (hbm_peak * 1e9) * 1e6, flops_k / (fp32_peak * 1e12) * 1e6
        bound = "fp32" if t_fp32_us > t_hbm_us else "hbm"
        ach_gbs, ach_tf = bytes_alg / (kern_us * 1e-6) / 1e9, flops_k / (kern_us * 1e-6) / 1e12
        roof = {"bound": bound,
                "achieved": ach_tf if bound == "fp32" else ach_gbs,
                "peak": fp32_peak if bound == "fp32" else hbm_peak,
                "unit": "TFLOP/s" if bound == "fp32" else "GB/s",
                "frac": (ach_tf / fp32_peak) if bound == "fp32" else (ach_gbs / hbm_peak),
                "traffic": None, "kernel": lp["kernel"], "kernel_us": kern_us, "kernel_us_max_over_ranks": kern_us_max,
                "peak_source": (f"fp32: in-process FMA microbenchmark sblas_b200_measure_fp32_peak ({fp32_peak:.1f} TFLOP/s; "
                                f"nominal 2*128*148*1.965 GHz = {FP32_PEAK_NOMINAL_TFLOPS:.1f}); " if bound == "fp32" else "") + "hbm: " + peak_src,
                "algorithmic_bytes": bytes_alg, "algorithmic_flops": flops_k, "format_bytes_per_nnz": lp["format_bytes_per_nnz"],
                "hbm_frac": ach_gbs / hbm_peak, "hbm_frac_of_8tbs": ach_gbs / 8000.0, "fp32_frac": ach_tf / fp32_peak,
                "t_roof_us": max(t_hbm_us, t_fp32_us), "roof_frac": max(t_hbm_us, t_fp32_us) / kern_us,
                "kernel_share_of_step": kern_us / (ms_step * 1e3)}
        if lp["kernel"].startswith("spmm_tc"):
            # the tensor-core form multiplies the zeros too: what it executes is the dense contraction, three TF32 products per
            # element pair (hi*hi + hi*lo + lo*hi), on W padded to 128 rows / 512 columns and the m tile padded to its width
            n_tile = int(lp["kernel"].split("<")[1].split(">")[0])
            rows_pad, k_pad, m_pad = -(-(hi - lo) // 128) * 128, -(-k // 512) * 512, -(-m // n_tile) * n_tile
            tc_flops = 3.0 * 2.0 * rows_pad * k_pad * m_pad
            tc_peak = TF32_PEAK_NOMINAL_TFLOPS
            tc_ach = tc_flops / (kern_us * 1e-6) / 1e12
            roof.update({"bound": "tensor", "achieved": tc_ach, "peak": tc_peak, "unit": "TFLOP/s", "frac": tc_ach / tc_peak,
                         "executed_tf32_flops": tc_flops,
                         "kernel_us_covers": "tc_split_x_kernel + spmm_tc_kernel (+ tc_fixup_kernel) + tc_fallback_kernel: chained by "
                                             "programmatic dependent launch, timed as one (an event between them undoes the overlap)",
                         "peak_source": "tensor: nominal dense TF32 = half of the 2250 TFLOP/s bf16 figure (B200_PROFILING.md); "
                                        "MEASURED_PEAKS.json has cuBLAS bf16 only: %s TFLOP/s burst, i.e. %.0f for TF32; hbm: %s"
                                        % (cx.bf16_measured, (cx.bf16_measured or 0) / 2, peak_src),
                         "frac_of_half_measured_bf16": (tc_ach / (cx.bf16_measured / 2)) if cx.bf16_measured else None,
                         "useful_tflops": flops_k / (kern_us * 1e-6) / 1e12})
        clk = clocks.summary()
        roof["clocks"] = clocks_profile.summary()                         # during the pass kernel_us comes from
        sm_hz = (clk.get("sm_mhz") or 1965.0) * 1e6
        # what actually bounds the kernel on the chip
        if lp["kernel"].startswith("spmm_tc"):
            roof["onchip"] = {"limiter": "tensor pipe (ncu sm__pipe_tc_cycles_active 90 % on cfg3, 67 % on cfg2: profiles/r02f_cfg{3,2}_full.txt); "
                                         "the CUDA-core form of this product is capped at 25 % of the fp32 peak by the shared-memory pipe"}
        elif lp["kernel"].startswith("spmm_tile"):
            wavefronts = nnz_local * ((m + 31) // 32)                   # 4*m B of X per non-zero through a 128 B/clk/SM pipe
            t_smem_us = wavefronts / (148 * sm_hz) * 1e6
            roof["onchip"] = {"limiter": "shared-memory pipe (one 128 B wavefront per clock per SM)", "x_wavefronts": int(wavefronts),
                              "t_floor_us": t_smem_us, "frac": t_smem_us / kern_us, "fp32_frac_ceiling": 0.25}
        elif lp["kernel"].startswith("spmv_dense"):
            lookups = (hi - lo) * ((k + 511) // 512 * 512) / 32.0 + bytes_alg / 128.0
            roof["onchip"] = {"limiter": "L1/shared data pipe: one conflict-free codebook look-up per element", "wavefronts": int(lookups),
                              "t_floor_us": lookups / (148 * sm_hz) * 1e6, "frac": lookups / (148 * sm_hz) * 1e6 / kern_us}
        elif lp["kernel"].startswith("spmv_blocked"):
            roof["onchip"] = {"limiter": "two streaming passes (x tile of a column block in shared memory; products written, then added per row block)"}
        else:
            roof["onchip"] = {"limiter": "L2 sector requests of the x gathers (one 32 B sector per non-zero)",
                              "l2_to_sm_gbs": (32 * nnz_local + bytes_alg - 4 * k * m) / (kern_us * 1e-6) / 1e9}
        traffic_file = os.path.join(ROOT, "profiles", "traffic.json")     # dram bytes/launch from the ncu --set full capture (N = 1)
        if world == 1 and os.path.exists(traffic_file):
            tr = json.load(open(traffic_file)).get(name)
            if tr:
                roof["traffic"] = tr["dram_bytes_per_launch"]
                roof["traffic_source"] = tr["source"]
        res = {
            "value": value, "unit": "GFLOP/s", "ms_per_step": ms_step, "ms_per_step_median": step_median_ms, "ms_per_step_min": step_min_ms,
            "config": config_of(inp),
            "setup": {"sharding": ("one matrix in %d nnz-balanced row slabs of W (sblas_b200_partition%s + upload), X replicated, C assembled on "
                                   "rank 0 by peer stores inside the product kernels; no data-path collective" % (world, "_csr" if host_obj is None else ""))
                      if world > 1 else "single GPU",
                      "kernel": lp["kernel"], "spmv_lanes": int(mat.info().spmv_lanes) if mat is not None else None,
                      "long_rows": int(mat.info().long_rows) if mat is not None else None},
            "effective_gbs": bytes_alg * (flops / max(flops_k, 1.0)) / (ms_step * 1e-3) / 1e9,
            "warm_l2": {"ms_per_step": warm_ms / args.steps, "value": flops / (warm_ms / args.steps * 1e-3) / 1e9,
                        "note": "products back to back, no flush in between" + (
                            "; the tensor-core form reaches the board's power cap in this pass (sw_power_cap, 1680 MHz measured): "
                            "slower than the flushed steps although L2 is warm" if lp["kernel"].startswith("spmm_tc") else "")},
            "roofline": roof,
            "e2e": {"value": flops / (e2e_ms / e2e_steps * 1e-3) / 1e9, "unit": "GFLOP/s",
                    "h2d_bytes_per_step": int(4 * m * (k + n)), "d2h_bytes_per_step": int(4 * m * n),
                    "ms_per_step": e2e_ms / e2e_steps, "steps": e2e_steps, "checksum": checksum,
                    "api": ("sblas_b200_addmatmat_host (host A, C from sblas_b200_host_alloc: page-locked + mapped; synchronous)" if world == 1 else
                            "ShardedMatrix.addmatmat_host per rank: its 1/N share of A over its own PCIe link, NCCL all-gather of A over NVLink, "
                            "product of its slab, its C slab back to host; wall clock, max over ranks")},
            "gpu_launches": int(launches), "clocks": clk, "upload_ms": upload_ms, "parity_check": parity,
        }
        if copyform_ms is not None:
            res["copyform_ms"] = copyform_ms
        if pageable_s is not None:
            res["e2e_pageable"] = {"value": flops / (pageable_s / e2e_steps) / 1e9, "unit": "GFLOP/s", "ms_per_step": pageable_s / e2e_steps * 1e3,
                                   "what": "the same call with plain numpy (malloc) buffers, as an unmodified caller passes them (driver-staged copies)"}
            res["e2e_pageable_registered"] = {"value": flops / (registered_s / e2e_steps) / 1e9, "unit": "GFLOP/s",
                                              "ms_per_step": registered_s / e2e_steps * 1e3,
                                              "what": "the same malloc buffers with SBLAS_B200_HOST_REGISTER=1: the library page-locks a buffer "
                                                      "it sees a second time (opt-in; for callers that keep their buffers)"}
        if world > 1:
            res["n_gpus"] = world
            res["slab_nnz"] = nnz_all
            res["slab_balance"] = max(nnz_all) / (sum(nnz_all) / world)
            res["sharded_output"] = {"ms_per_step": local_ms_max / args.steps, "value": flops / (local_ms_max / args.steps * 1e-3) / 1e9,
                                     "what": "the same step with C left sharded (each rank's slab in its own memory): the value without reassembly"}
            res["reassembly_allgather_ms"] = gather_ms
            res["model_scaling_ceiling"] = sharding.model_scaling_ceiling(nnz, n, k, m, world, lp["format_bytes_per_nnz"])
        if args.cpu_baseline and world == 1:
            ref = CpuReference(inp)
            ref.call()
            reps = 3 if ref.call() < 5 else 1
            best = min(ref.call() for _ in range(reps))
            res["cpu_baseline"] = {"value": flops / best / 1e9, "unit": "GFLOP/s", "cores": 1, "kind": ref.kind,
                                   "sample": ref.sample + f"; best of {reps}", "ms_per_step": best * 1e3,
                                   "host_cores_available": os.cpu_count()}
            if host_obj is not None:
                legs = dense_openblas_legs(inp, flops)
                if legs:
                    res["cpu_dense_openblas"] = legs
            else:
                sec, threads = ref.omp_csr()
                res["cpu_csr_openmp"] = {"value": flops / sec / 1e9, "unit": "GFLOP/s", "ms_per_step": sec * 1e3, "threads": threads,
                                         "what": "CSR restatement with OpenMP over rows, full matrix, all host cores -- NOT the reference (single-threaded)"}
    # release
    if out is not None:
        out.close()
    if mat is not None and (sh is not None or host_obj is None):
        mat.free()
    del d_a, d_c, host_obj, sh, mat
    torch.cuda.empty_cache()
    return res, ok_flag


def run_ours(args):
    import torch
    import torch.distributed as dist
    import sparsematrix_b200 as sb
    from sparsematrix_b200 import sharding

    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this library has no CPU fallback")
    # stdout carries the one JSON line and nothing else: libraries that print there (NCCL's version
    # banner at N > 1) are sent to stderr for the duration of the run
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    cx = Ctx()
    cx.torch, cx.dist, cx.sb, cx.sharding = torch, dist, sb, sharding
    cx.rank, cx.local_rank, cx.world = rank, local_rank, world
    cx.dev = torch.device("cuda", local_rank)
    cx.flush = torch.empty(256 << 20, dtype=torch.uint8, device=cx.dev)       # > 126 MB L2
    cx.stream = torch.cuda.current_stream()
    cx.bf16_measured = None
    try:
        cx.bf16_measured = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"])
    except Exception:
        pass
    try:
        cx.fp32_peak_tflops = sb.measure_fp32_peak(local_rank)
    except Exception:
        cx.fp32_peak_tflops = FP32_PEAK_NOMINAL_TFLOPS

    names = ORDER if args.workload == "all" else [args.workload]
    head, per, all_ok = None, {}, True
    for i, name in enumerate(names):
        t0 = time.perf_counter()
        if i == 0:
            res, ok = measure(name, args, cx)                    # the headline: a failure here is the run's failure
        else:
            try:
                res, ok = measure(name, args, cx)
            except Exception as e:                              # never lose the headline line over a secondary workload
                import traceback
                traceback.print_exc()
                res, ok = ({"error": f"{type(e).__name__}: {e}"} if rank == 0 else None), False
                if world > 1:
                    raise
        all_ok = all_ok and ok
        if rank == 0:
            res["wall_s"] = time.perf_counter() - t0
            if i == 0:
                head = res
            else:
                per[name] = res

    if rank == 0:
        line = {"metric": METRIC, "value": head["value"], "unit": "GFLOP/s", "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": head["ms_per_step"], "higher_is_better": True,
                "scaling": "strong" if world > 1 else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic"}
        for key, v in head.items():
            if key not in line:
                line[key] = v
        line["fp32_peak_tflops_measured"] = cx.fp32_peak_tflops
        if per:
            line["per_workload"] = per
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if not all_ok:
        print("bench.py: parity check FAILED (see parity_check in the JSON line)", file=sys.stderr)
        raise SystemExit(3)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--workload", choices=["all"] + sorted(WORKLOADS), default="all")
    ap.add_argument("--scale-rows", type=float, default=1.0, help="shrink n for quick checks (not a bench value)")
    ap.add_argument("--no-cpu-baseline", dest="cpu_baseline", action="store_false")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
