"""Why are some pinned H2D copies slow on this box?  Same size, different allocations / offsets / alignments."""
import time
import torch

dev = torch.device("cuda", 0)


def t_copy(dst, src, reps=20):
    for _ in range(3):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        dst.copy_(src, non_blocking=True)
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3


MB = 1 << 20
d = torch.empty(64 * MB, dtype=torch.uint8, device=dev)
print("-- five separate 2 MiB pinned allocations, h2d / d2h us")
keep = []
for i in range(5):
    h = torch.zeros(2 * MB, dtype=torch.uint8).pin_memory()
    keep.append(h)
    print(f"   alloc {i} ptr%2MiB={h.data_ptr() % (2 * MB):8d}: h2d {t_copy(d[:2 * MB], h):7.1f}  d2h {t_copy(h, d[:2 * MB]):7.1f}")
print("-- one 64 MiB pinned allocation, 2 MiB windows at offsets")
big = torch.zeros(64 * MB, dtype=torch.uint8).pin_memory()
for off in (0, 1 * MB, 2 * MB, 3 * MB, 8 * MB, 8 * MB + 4096, 33 * MB):
    print(f"   off {off / MB:6.2f} MiB: h2d {t_copy(d[:2 * MB], big[off:off + 2 * MB]):7.1f}  d2h {t_copy(big[off:off + 2 * MB], d[:2 * MB]):7.1f}")
print("-- sizes from the big allocation (offset 0)")
for sz in (1 * MB, 1536 << 10, 2 * MB, 2 * MB + 4096, 3 * MB, 4 * MB, 8 * MB, 16 * MB, 32 * MB):
    print(f"   {sz / MB:6.2f} MiB: h2d {t_copy(d[:sz], big[:sz]):7.1f} ({sz / t_copy(d[:sz], big[:sz]) / 1e3:5.1f} GB/s)  d2h {t_copy(big[:sz], d[:sz]):7.1f}")
print("-- device destination offsets (2 MiB h2d from big[0:2MiB])")
for doff in (0, 2 * MB, 5 * MB, 32 * MB):
    print(f"   dst off {doff / MB:5.1f} MiB: {t_copy(d[doff:doff + 2 * MB], big[:2 * MB]):7.1f}")
print("-- two copies in flight on two streams (A and C of the host path), 2 MiB each")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(20):
    with torch.cuda.stream(s1):
        d[:2 * MB].copy_(big[:2 * MB], non_blocking=True)
    with torch.cuda.stream(s2):
        d[4 * MB:6 * MB].copy_(big[4 * MB:6 * MB], non_blocking=True)
    torch.cuda.synchronize()
print(f"   {(time.perf_counter() - t0) / 20 * 1e6:7.1f} us per pair incl. sync")
