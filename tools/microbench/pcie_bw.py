"""PCIe copy times on this box for the sizes of the host path (pinned memory, CUDA events + wall clock)."""
import time
import torch

dev = torch.device("cuda", 0)
for nbytes in (16 << 10, 256 << 10, 1 << 20, 2 << 20, 4 << 20, 32 << 20):
    h = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    d = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    for name, fn in (("h2d", lambda: d.copy_(h, non_blocking=True)), ("d2h", lambda: h.copy_(d, non_blocking=True))):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        a.record()
        for _ in range(20):
            fn()
        b.record()
        torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) / 20 * 1e6
        ev = a.elapsed_time(b) / 20 * 1e3
        # one copy, host-synchronous (what a blocking caller sees)
        t0 = time.perf_counter()
        for _ in range(20):
            fn()
            torch.cuda.synchronize()
        one = (time.perf_counter() - t0) / 20 * 1e6
        print(f"{name} {nbytes >> 10:6d} KiB: {ev:8.1f} us back to back ({nbytes / ev / 1e3:6.1f} GB/s), {one:8.1f} us copy + sync")
