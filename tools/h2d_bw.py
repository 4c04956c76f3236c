"""Host<->device copy bandwidth of the box (pinned memory), to put the end-to-end number in context."""
import time
import torch
n = 512 << 20
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
for name, fn in (("h2d", lambda: d.copy_(h, non_blocking=True)), ("d2h", lambda: h.copy_(d, non_blocking=True))):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    print(name, "%.1f GB/s" % (5 * n / (time.perf_counter() - t0) / 1e9))
