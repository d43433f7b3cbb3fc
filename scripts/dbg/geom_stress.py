import sys, os
sys.path.insert(0, "/root/repo" if os.path.isdir("/root/repo/augcl_b200") else ".")
import numpy as np, torch
from augcl_b200 import utils as U, augmentations as A
torch.manual_seed(0)
bad = 0
for (H, W, k, pad, B) in [(8, 8, 1, 2, 5), (12, 16, 2, 3, 7), (36, 36, 3, 4, 33), (64, 48, 2, 4, 40), (128, 128, 1, 4, 9), (100, 100, 3, 4, 70), (16, 128, 3, 2, 12), (52, 8, 2, 1, 6)]:
    try:
        buf = U.ReplayBuffer((3 * k, H, W), (4,), 300, B)
    except Exception as e:
        print("ring", (H, W, k), "unsupported:", str(e)[:80]); continue
    buf.fill_synthetic(seed=H * 7 + W)
    bank = torch.randint(0, 256, (13, 3, H, W), dtype=torch.uint8, device="cuda")
    A.set_background_bank(bank)
    g = torch.Generator().manual_seed(H + W)
    idxs = torch.randint(0, 300, (B,), generator=g).numpy()
    off = {"obs": torch.randint(0, 2 * pad + 1, (B, 2), generator=g).numpy(), "next_obs": torch.randint(0, 2 * pad + 1, (B, 2), generator=g).numpy()}
    rows = torch.randint(0, 13, (B,), generator=g).numpy()
    wt = torch.randn((B, 3, 3, 3, 3), generator=g)
    for chain in ["splice-conv", "overlay", "conv", "CS_splice-conv", "splice", "overlay-conv", ""]:
        for aug_input in ["plain", "shift"]:
            try:
                want = ("obs", "obs_shift", "obs_aug", "next_obs", "next_obs_shift") if chain else ("obs", "obs_shift", "next_obs", "next_obs_shift")
                out = buf.sample_fused(idxs=idxs, pad=pad, data_aug=chain, aug_input=aug_input, offsets=off, bank_idx=rows, weights=wt, want=want)
            except NotImplementedError as e:
                print((H, W, k), chain, "unsupported:", str(e)[:60]); break
            obs, _, _, nxt, _ = buf.sample(idxs=idxs)
            ok = torch.equal(out["obs"], obs) and torch.equal(out["next_obs"], nxt)
            sh = A.random_shift(obs, pad, offsets=off["obs"])
            ok = ok and torch.equal(out["obs_shift"], sh) and torch.equal(out["next_obs_shift"], A.random_shift(nxt, pad, offsets=off["next_obs"]))
            err = 0.0
            if chain:
                x = sh if aug_input == "shift" else obs
                for key in chain.split("-"):
                    if key == "splice": x = A.splice(x, bank_idx=rows)
                    elif key == "CS_splice": x = A.CS_splice(x, bank_idx=rows)
                    elif key == "overlay": x = A.random_overlay(x, bank_idx=rows)
                    elif key == "conv": x = A.random_conv(x, weights=wt)
                err = (out["obs_aug"] - x).abs().max().item() / 255.0
            if not ok or err > 1e-5:
                bad += 1
                print("MISMATCH", (H, W, k, pad, B), chain, aug_input, ok, err)
    A.set_background_bank(None)
print("bad", bad)
