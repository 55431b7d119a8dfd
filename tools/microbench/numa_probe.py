"""Where do pinned allocations land (NUMA node) and how fast is H2D from each?"""
import glob
import os
import re
import torch

for p in sorted(glob.glob("/sys/devices/system/node/node*/cpulist")):
    print(p, open(p).read().strip())
print("affinity", sorted(os.sched_getaffinity(0)))
print(open("/proc/self/status").read().split("Cpus_allowed_list")[1].split("\n")[0], "| mems", re.search(r"Mems_allowed_list:\s*(\S+)", open("/proc/self/status").read()).group(1))
try:
    print("gpu numa_node", open(glob.glob("/sys/bus/pci/devices/*/numa_node")[0]).read().strip())
except Exception as e:
    print("no numa_node", e)
dev = torch.device("cuda", 0)
MB = 1 << 20
d = torch.empty(32 * MB, dtype=torch.uint8, device=dev)


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


def node_of(ptr):
    for line in open("/proc/self/numa_maps"):
        f = line.split()
        if int(f[0], 16) <= ptr < int(f[0], 16) + (1 << 40):
            pass
    best = None
    for line in open("/proc/self/numa_maps"):
        f = line.split()
        a = int(f[0], 16)
        if a <= ptr and (best is None or a > best[0]):
            best = (a, " ".join(x for x in f[1:] if x.startswith("N") or x.startswith("kernelpagesize") or x.startswith("huge")))
    return best[1] if best else "?"


keep = []
for i, sz in enumerate((2 * MB, 2 * MB, 32 * MB, 2 * MB, 8 * MB, 32 * MB, 2 * MB, 2 * MB)):
    h = torch.zeros(sz, dtype=torch.uint8).pin_memory()
    keep.append(h)
    t = t_copy(d[:sz], h)
    print(f"alloc {i} {sz // MB:3d} MiB: h2d {t:8.1f} us ({sz / t / 1e3:5.1f} GB/s)  d2h {t_copy(h, d[:sz]):8.1f}  pages: {node_of(h.data_ptr())}")
