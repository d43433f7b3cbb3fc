#!/usr/bin/env python
"""Device-timed throughput of the fused per-algorithm sampler (augcl_ring_sample_fused) against the unfused
product chain (augcl_ring_sample_ex + stand-alone kernels) for BASELINE.json configs[2] (SVEA overlay, batch 256)
and configs[3] (AugCL splice-conv, batch 128), plus a batch sweep.  One JSON line per (case, batch).
Called straight through the C ABI with pre-allocated, rotating outputs (> L2), device-resident draws.
    python scripts/bench_fused.py [--batches 128,256,1024,4096] [--opts fused_ctas_per_sm=6]"""
import argparse, ctypes as C, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import __graft_entry__ as entry
entry.build()
from augcl_b200 import _lib as L, augmentations as A, utils as U

K, S, PAD, CAP = 3, 84, 4, 100_000
CH = 3 * K
S8, S32, F = CH * S * S, 4 * CH * S * S, 3 * S * S
PEAK = 6550.4
try:
    PEAK = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass


def timeit(fn, iters):
    st = torch.cuda.current_stream()
    for t in range(3):
        fn(t)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for t in range(iters):
            fn(t)
        e1.record(st)
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / iters)
    return best * 1e3   # us


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", default="128,256,1024,4096")
    ap.add_argument("--opts", default="")
    ap.add_argument("--cases", default="augcl_splice_conv,svea_overlay,svea_conv,augcl_splice,drq_shift_both")
    a = ap.parse_args()
    for kv in filter(None, a.opts.split(",")):
        k, v = kv.split("=")
        L.set_option(k, v)
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    buf = U.ReplayBuffer((CH, S, S), (6,), CAP, 128)
    buf.fill_synthetic(seed=0)
    A.seed(0)
    A.load_dataloader(128, S)
    bank = A.get_background_bank()
    lib = L.lib()
    st = torch.cuda.current_stream().cuda_stream
    g = torch.Generator().manual_seed(0)
    # case -> (ops, aug_input, outputs, algorithmic bytes per sample, unfused chain)
    cases = {
        "augcl_splice_conv": (L.OP_SPLICE | L.OP_CONV, 0, ("obs_shift", "obs_aug", "next"), 2 * S8 + 3 * F + 3 * S32),
        "augcl_splice": (L.OP_SPLICE, 0, ("obs_shift", "obs_aug", "next"), 2 * S8 + 3 * F + 3 * S32),
        "svea_overlay": (L.OP_OVERLAY, 1, ("obs_shift", "obs_aug", "next"), 2 * S8 + F + 3 * S32),
        "svea_conv": (L.OP_CONV, 1, ("obs_shift", "obs_aug", "next"), 2 * S8 + 3 * S32),
        "drq_shift_both": (0, 0, ("obs_shift", "next_shift"), 2 * S8 + 2 * S32),
    }
    for B in [int(x) for x in a.batches.split(",")]:
        n_sets = max(2, min(8, (800 << 20) // (3 * B * S32)))
        n_idx = 16
        idx = [torch.randint(0, CAP, (B,), generator=g).to(dev) for _ in range(n_idx)]
        off = torch.randint(0, 2 * PAD + 1, (B, 2), generator=g, dtype=torch.int32).to(dev)
        off2 = torch.randint(0, 2 * PAD + 1, (B, 2), generator=g, dtype=torch.int32).to(dev)
        rows = torch.randint(0, bank.shape[0], (B,), generator=g).to(dev)
        wts = torch.randn((B, 3, 3, 3, 3), generator=g).to(dev)
        outs = [torch.empty((4, B, CH, S, S), device=dev) for _ in range(n_sets)]
        act = torch.empty((B, 6), device=dev); rew = torch.empty((B,), device=dev); nd = torch.empty((B,), device=dev)
        iters = max(20, min(400, (1 << 16) // B))
        for name in a.cases.split(","):
            ops, aug_in, wants, abytes = cases[name]
            descs = []
            for i in range(n_sets * n_idx):
                o = outs[i % n_sets]
                d = L.FusedDesc()
                d.struct_size = C.sizeof(L.FusedDesc)
                d.B, d.idx_src, d.off_src, d.idx = B, L.SRC_DEVICE, L.SRC_DEVICE, idx[i % n_idx].data_ptr()
                d.off_obs, d.off_next, d.pad = off.data_ptr(), off2.data_ptr(), PAD
                d.out_obs_shift = o[0].data_ptr() if "obs_shift" in wants else None
                d.out_obs_aug = o[1].data_ptr() if "obs_aug" in wants else None
                d.out_next = o[2].data_ptr() if "next" in wants else None
                d.out_next_shift = o[2].data_ptr() if "next_shift" in wants else None
                d.out_act, d.out_rew, d.out_nd = act.data_ptr(), rew.data_ptr(), nd.data_ptr()
                d.aug_input, d.aug_ops = aug_in, ops
                d.bank, d.bank_len, d.bank_idx = bank.data_ptr(), bank.shape[0], rows.data_ptr()
                d.conv_weights = wts.data_ptr()
                d.hue_thres, d.sat_thres, d.val_thres, d.alpha = 0.0, 0.0, 0.6, 0.5
                descs.append(d)

            def fused(t):
                rc = lib.augcl_ring_sample_fused(buf._ring, C.byref(descs[t % len(descs)]), st)
                assert rc == 0, lib.augcl_last_error()

            def unfused(t):
                o = outs[t % n_sets]
                ix = idx[t % n_idx].data_ptr()
                if name == "drq_shift_both":
                    L.call("augcl_ring_sample", buf._ring, ix, B, o[0].data_ptr(), o[2].data_ptr(), act.data_ptr(), rew.data_ptr(),
                           nd.data_ptr(), L.AUG_SHIFT, off.data_ptr(), off2.data_ptr(), PAD, 0, st)
                    return
                if aug_in == 0:   # AugCL: plain obs -> o[3]; shift -> o[0]; chain on plain -> o[1]
                    L.call("augcl_ring_sample", buf._ring, ix, B, o[3].data_ptr(), o[2].data_ptr(), act.data_ptr(), rew.data_ptr(),
                           nd.data_ptr(), L.AUG_NONE, None, None, 0, 0, st)
                    L.call("augcl_random_shift", o[3].data_ptr(), o[0].data_ptr(), B, CH, S, S, PAD, off.data_ptr(), 0, None, st)
                    src = o[3]
                else:             # SVEA: shifted obs straight from the ring -> o[0]; chain on it -> o[1]
                    L.call("augcl_ring_sample", buf._ring, ix, B, o[0].data_ptr(), o[2].data_ptr(), act.data_ptr(), rew.data_ptr(),
                           nd.data_ptr(), L.AUG_SHIFT_OBS, off.data_ptr(), None, PAD, 0, st)
                    src = o[0]
                if ops & L.OP_SPLICE:
                    dst = o[1] if not (ops & L.OP_CONV) else outs[(t + 1) % n_sets][3]
                    L.call("augcl_splice", src.data_ptr(), dst.data_ptr(), B, CH, S, S, bank.data_ptr(), 0, bank.shape[0],
                           rows.data_ptr(), 0.0, 0.0, 0.6, 0, st)
                    src = dst
                if ops & L.OP_OVERLAY:
                    L.call("augcl_random_overlay", src.data_ptr(), o[1].data_ptr(), B, CH, S, S, bank.data_ptr(), 0, bank.shape[0],
                           rows.data_ptr(), 0.5, st)
                if ops & L.OP_CONV:
                    L.call("augcl_random_conv", src.data_ptr(), o[1].data_ptr(), B, CH, S, S, wts.data_ptr(), st)

            tf, tu = timeit(fused, iters), timeit(unfused, iters)
            gbs = B * abytes / tf / 1e3
            print(json.dumps({"case": name, "batch": B, "fused_us": round(tf, 2), "unfused_us": round(tu, 2),
                              "speedup": round(tu / tf, 2), "samples_per_s": round(B / tf * 1e6),
                              "algorithmic_bytes_per_sample": abytes, "GB/s": round(gbs, 1),
                              "frac_of_measured_hbm": round(gbs / PEAK, 4)}), flush=True)


if __name__ == "__main__":
    main()
