#!/usr/bin/env python
"""Small fixed workload for Nsight Compute: three launches of each hot kernel at the BASELINE.json
batch sizes (sample/shift/conv/splice at 128, overlay at 256).  Run plain first, then under ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import __graft_entry__ as entry
entry.build()
from augcl_b200 import augmentations as A, utils as U

torch.manual_seed(0)
B, CAP, PAD = 128, 100_000, 4
buf = U.ReplayBuffer((9, 84, 84), (6,), CAP, B)
buf.fill_synthetic(seed=3)
bank = torch.randint(0, 256, (4096, 3, 84, 84), dtype=torch.uint8, device="cuda")
A.set_background_bank(bank)
idx = torch.randint(0, CAP, (8, B))
for rep in range(3):
    obs, act, rew, nxt, nd = buf.sample_drq(pad=PAD, idxs=idx[rep].numpy())         # ring_sample_band_kernel
    obs2 = buf.sample(idxs=idx[rep + 3].cuda())[0]
    sh = A.random_shift(obs2, PAD)                                                 # window_copy_vec_kernel
    sp = A.splice(obs2)                                                            # splice_vec_kernel
    cv = A.random_conv(sp)                                                         # random_conv_tma_kernel
    x256 = torch.cat([obs2, nxt], 0)
    ov = A.random_overlay(x256)                                                    # overlay_vec_kernel
torch.cuda.synchronize()
print("ok", float(cv.mean()), float(ov.mean()), float(sh.mean()))
