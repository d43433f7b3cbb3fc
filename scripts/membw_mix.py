#!/usr/bin/env python
"""What the HBM system delivers for different read:write mixes (torch library kernels, CUDA events, best of 10):
copy (1:1, the MEASURED_PEAKS method), fill (0:1), sum (1:0), u8->f32 cast (1:4, the sample kernel's mix)."""
import json, torch
dev = torch.device("cuda", 0)
def best(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    b = 1e9
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        b = min(b, e0.elapsed_time(e1))
    return b * 1e-3
N = 1 << 30
a = torch.empty(N, dtype=torch.bfloat16, device=dev).normal_()
b = torch.empty_like(a)
out = {}
out["copy_1r1w_GBs"] = 4 * N / best(lambda: b.copy_(a)) / 1e9
out["fill_0r1w_GBs"] = 2 * N / best(lambda: b.fill_(1.0)) / 1e9
out["sum_1r0w_GBs"] = 2 * N / best(lambda: a.sum()) / 1e9
del a, b
u = torch.randint(0, 256, (1 << 29,), dtype=torch.uint8, device=dev)
f = torch.empty(1 << 29, dtype=torch.float32, device=dev)
out["cast_u8_f32_1r4w_GBs"] = 5 * (1 << 29) / best(lambda: f.copy_(u)) / 1e9
print(json.dumps(out))
