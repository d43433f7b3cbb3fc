#!/usr/bin/env python
"""Device-timed throughput of the stand-alone augmentation kernels against their algorithmic bytes
(DESIGN.md §4): one JSON line per (kernel, batch).  Inputs rotate over buffer sets larger than L2.
    python scripts/bench_augs.py [--batches 128,256,2048]"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import __graft_entry__ as entry
entry.build()
from augcl_b200 import _lib as L, augmentations as A

C, S, PAD = 9, 84, 4
S32, F = 4 * C * S * S, 3 * S * S
PEAK = 6546.2
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
    return best * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", default="128,256,2048")
    ap.add_argument("--only", default="")
    ap.add_argument("--opts", default="", help="augcl_set_option pairs, e.g. conv_part=-1")
    a = ap.parse_args()
    for kv in filter(None, a.opts.split(",")):
        key, val = kv.split("=")
        L.set_option(key, val)
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(0)
    bank = torch.randint(0, 256, (4096, 3, S, S), generator=g, dtype=torch.uint8).to(dev)
    A.set_background_bank(bank)
    lib = L.lib()
    sp = torch.cuda.current_stream().cuda_stream
    for B in [int(x) for x in a.batches.split(",")]:
        n_sets = max(2, min(8, (700 << 20) // (2 * B * S32)))
        xs = [torch.randint(0, 256, (B, C, S, S), generator=g, dtype=torch.uint8).to(dev).float() for _ in range(n_sets)]
        x100 = [torch.randint(0, 256, (B, C, 100, 100), generator=g, dtype=torch.uint8).to(dev).float() for _ in range(n_sets)]
        outs = [torch.empty((B, C, S, S), device=dev) for _ in range(n_sets)]
        off = torch.randint(0, 2 * PAD + 1, (B, 2), generator=g, dtype=torch.int32).to(dev)
        offc = torch.randint(0, 16, (B, 2), generator=g, dtype=torch.int32).to(dev)
        wts = torch.randn((B, 3, 3, 3, 3), generator=g).to(dev)
        rows = torch.randint(0, 4096, (B,), generator=g).to(dev)
        base = A._base_grid(S, PAD, dev)
        iters = max(20, min(500, (1 << 17) // B))
        cases = {
            "random_shift_int": (2 * S32, lambda t: lib.augcl_random_shift(xs[t % n_sets].data_ptr(), outs[t % n_sets].data_ptr(), B, C, S, S, PAD, off.data_ptr(), 0, None, sp)),
            "random_shift_bilinear": (2 * S32, lambda t: lib.augcl_random_shift(xs[t % n_sets].data_ptr(), outs[t % n_sets].data_ptr(), B, C, S, S, PAD, off.data_ptr(), 1, base.data_ptr(), sp)),
            "random_crop_100_84": (2 * S32, lambda t: lib.augcl_random_crop(x100[t % n_sets].data_ptr(), outs[t % n_sets].data_ptr(), B, C, 100, 100, 84, offc.data_ptr(), sp)),
            "random_conv": (2 * S32, lambda t: lib.augcl_random_conv(xs[t % n_sets].data_ptr(), outs[t % n_sets].data_ptr(), B, C, S, S, wts.data_ptr(), sp)),
            "random_overlay_u8bank": (2 * S32 + F, lambda t: lib.augcl_random_overlay(xs[t % n_sets].data_ptr(), outs[t % n_sets].data_ptr(), B, C, S, S, bank.data_ptr(), 0, 4096, rows.data_ptr(), 0.5, sp)),
            "splice_u8bank": (2 * S32 + 3 * F, lambda t: lib.augcl_splice(xs[t % n_sets].data_ptr(), outs[t % n_sets].data_ptr(), B, C, S, S, bank.data_ptr(), 0, 4096, rows.data_ptr(), 0.0, 0.0, 0.6, 0, sp)),
        }
        # the remaining aug_to_func entries (csrc/augment2.cu), through the module functions (device draws; the output tensor
        # comes from torch's caching allocator, so small batches also show the host side of the call)
        def mod(f):
            def run(t):
                f(xs[t % n_sets])
                return 0
            return run
        cases.update({
            "random_grayscale": (2 * S32, mod(lambda x: A.random_grayscale(x, p=0.3))),
            "random_flip": (2 * S32, mod(lambda x: A.random_flip(x, p=0.2))),
            "random_rotation": (2 * S32, mod(lambda x: A.random_rotation(x, p=0.3))),
            "random_cutout_color": (2 * S32, mod(lambda x: A.random_cutout_color(x))),
            "color_jitter": (2 * S32, mod(lambda x: A.color_jitter(x))),
            "CS_splice_u8bank": (2 * S32 + 3 * F, mod(lambda x: A.CS_splice(x))),
        })
        for name, (bytes_per_sample, fn) in cases.items():
            if a.only and a.only not in name:
                continue
            rc = fn(0)
            assert rc == 0, lib.augcl_last_error()
            us = timeit(fn, iters)
            gbs = B * bytes_per_sample / us / 1e3
            print(json.dumps({"kernel": name, "batch": B, "us": round(us, 2), "samples_per_s": round(B / us * 1e6),
                              "algorithmic_bytes_per_sample": bytes_per_sample, "GB/s": round(gbs, 1), "frac_of_measured_hbm": round(gbs / PEAK, 4)}), flush=True)
        del xs, x100, outs


if __name__ == "__main__":
    main()
