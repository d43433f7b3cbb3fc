#!/usr/bin/env python
"""ncu target: two launches of each stand-alone augmentation kernel at batch 2048 (9x84x84; crop from 100x100)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from augcl_b200 import augmentations as A
torch.manual_seed(0); np.random.seed(0)
B = 2048
x = torch.randint(0, 256, (B, 9, 84, 84), dtype=torch.uint8, device="cuda").float()
x100 = torch.randint(0, 256, (B, 9, 100, 100), dtype=torch.uint8, device="cuda").float()
A.set_background_bank(torch.randint(0, 256, (4096, 3, 84, 84), dtype=torch.uint8, device="cuda"))
w = torch.randn(B, 3, 3, 3, 3)
for _ in range(2):
    A.random_shift(x, 4); A.random_shift(x, 4, mode="bilinear"); A.random_crop(x100, 84); A.random_conv(x, weights=w)
    A.random_overlay(x); A.splice(x); A.CS_splice(x); A.random_grayscale(x); A.random_flip(x); A.random_rotation(x)
    A.random_cutout_color(x); A.color_jitter(x.clone())
torch.cuda.synchronize()
print("ok")
