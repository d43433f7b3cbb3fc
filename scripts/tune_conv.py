"""Part-count sweep of the stand-alone random_conv (option conv_part: 0 = 12-row band kernel, 1 = auto, >= 2 = forced)."""
import json, sys, torch
from augcl_b200 import augmentations as A, _lib as L
for B in (128, 256, 512, 2048):
    x = torch.rand(B, 9, 84, 84, device="cuda") * 255
    w = torch.randn(B, 3, 3, 3, 3, device="cuda")
    for parts in (0, 1, 2, 3, 4, 6):
        L.set_option("conv_part", str(parts))
        for _ in range(5):
            A.random_conv(x, weights=w)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        it = 200 if B <= 512 else 50
        e0.record()
        for _ in range(it):
            A.random_conv(x, weights=w)
        e1.record(); torch.cuda.synchronize()
        print(json.dumps({"batch": B, "conv_part": parts, "us": round(e0.elapsed_time(e1) * 1e3 / it, 2)}))
L.set_option("conv_part", "1")
