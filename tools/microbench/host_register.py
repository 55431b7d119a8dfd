"""Cost of cudaHostRegister / cudaHostUnregister on malloc'ed buffers (would pinning a pageable caller buffer for
the duration of one call pay?)."""
import time
import numpy as np
import torch

rt = torch.cuda.cudart()
torch.cuda.init()
for mb in (0.0625, 2, 32):
    n = int(mb * (1 << 20))
    a = np.ones(n, dtype=np.uint8)
    ptr = a.ctypes.data
    for flags in (0, 2):        # default, mapped
        ts = []
        for _ in range(5):
            t0 = time.perf_counter(); r = rt.cudaHostRegister(ptr, n, flags); t1 = time.perf_counter()
            u = rt.cudaHostUnregister(ptr); t2 = time.perf_counter()
            ts.append(((t1 - t0) * 1e6, (t2 - t1) * 1e6))
        print(f"{mb:8.4f} MiB flags={flags}: register {min(t[0] for t in ts):8.1f} us, unregister {min(t[1] for t in ts):8.1f} us (rc {int(r)}, {int(u)})")
