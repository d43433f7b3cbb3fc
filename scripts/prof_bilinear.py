"""ncu target: a few launches of the bilinear shift at batch 2048 (scripts/ncu_brief.py reads the report)."""
import torch
from augcl_b200 import augmentations as A
x = torch.rand(2048, 9, 84, 84, device="cuda") * 255
off = torch.randint(0, 9, (2048, 2))
for _ in range(3):
    y = A.random_shift(x, 4, offsets=off, mode="bilinear")
torch.cuda.synchronize()
