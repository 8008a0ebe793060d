"""What a pure write stream reaches on this GPU, beside the copy figure MEASURED_PEAKS.json quotes: torch fill_ (write only) and
copy_ (read + write) over 4 GB of int64, CUDA events, best of 10."""
import json

import torch

n = 4 * 1024 ** 3 // 8
x = torch.empty(n, dtype=torch.int64, device="cuda")
y = torch.empty(n, dtype=torch.int64, device="cuda")


def best(fn, reps=10):
    fn()
    torch.cuda.synchronize()
    t = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        t.append(a.elapsed_time(b))
    return min(t)


w = best(lambda: x.fill_(7))
c = best(lambda: y.copy_(x))
print(json.dumps({"bytes": n * 8, "fill_ms": w, "write_only_GBps": n * 8 / w / 1e6, "copy_ms": c, "copy_read_plus_write_GBps": 2 * n * 8 / c / 1e6}))
