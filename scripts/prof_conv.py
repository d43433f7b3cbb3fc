"""ncu target: a few launches of the stand-alone random_conv at batch 2048 (scripts/ncu_brief.py reads the report)."""
import torch
from augcl_b200 import augmentations as A
x = torch.rand(2048, 9, 84, 84, device="cuda") * 255
w = torch.randn(2048, 3, 3, 3, 3)
for _ in range(3):
    y = A.random_conv(x, weights=w)
torch.cuda.synchronize()
