import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import __graft_entry__ as entry
entry.build()
from augcl_b200 import utils as U, augmentations as A
from oracle import augcl_oracle as O
sys.path.insert(0, "tests")
from test_gpu_fused import build_pair
rng = np.random.RandomState(1)
S, k, pad, cap, B = 84, 3, 4, 40, 11
buf, ora = build_pair(U, O, rng, S, k, cap, B)
bank = torch.from_numpy(rng.randint(0, 256, (9, 3, S, S)).astype(np.uint8))
A.set_background_bank(bank.cuda())
for chain in ["conv", "splice-conv", "splice"]:
    idxs = rng.randint(0, cap, size=B)
    off = {"obs": rng.randint(0, 2 * pad + 1, (B, 2)), "next_obs": rng.randint(0, 2 * pad + 1, (B, 2))}
    rows = rng.randint(0, 9, size=B)
    wt = torch.from_numpy(rng.randn(B, 3, 3, 3, 3).astype(np.float32))
    out = buf.sample_fused(idxs=idxs, pad=pad, data_aug=chain, aug_params={}, aug_input="plain",
                           want=("obs", "obs_shift", "obs_aug", "next_obs"), offsets=off, bank_idx=rows, weights=wt)
    o_obs = ora.sample(idxs)[0]
    imgs = O.bank_rows_f32(bank, rows)
    x = o_obs
    for key in chain.split("-"):
        x = O.splice(x, imgs) if key == "splice" else O.random_conv(x, wt)
    err = (out["obs_aug"].cpu() - x).abs() / 255.0
    print(chain, "max err", err.max().item())
    bad = (err > 1e-5)
    print("  bad frac", bad.float().mean().item())
    if bad.any():
        print("  bad by row:", bad.sum(dim=(0, 1, 3)).tolist())
        print("  bad by col:", bad.sum(dim=(0, 1, 2)).tolist())
        print("  bad by ch:", bad.sum(dim=(0, 2, 3)).tolist())
