#!/usr/bin/env python
"""A/B timing of augcl_ring_sample variants on one B200 (device-timed, fresh indices per step).
    python scripts/tune_sample.py [--batches 128,1024,8192]
Prints one JSON line per (variant, batch).  Not a bench line: evidence for the defaults in state.cu."""
import argparse, itertools, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import __graft_entry__ as entry
entry.build()
from augcl_b200 import _lib as L, utils as U

CAP, C, S, PAD = 100_000, 9, 84, 4
BYTES = 2 * C * S * S * 5

def run(buf, B, iters, dev):
    g = torch.Generator().manual_seed(B)
    n_sets = max(2, min(8, (600 << 20) // (2 * B * C * S * S * 4)))
    outs = [(torch.empty((B, C, S, S), device=dev), torch.empty((B, C, S, S), device=dev)) for _ in range(n_sets)]
    idx = torch.randint(0, CAP, (iters + 5, B), generator=g).to(dev)
    off = torch.randint(0, 2 * PAD + 1, (iters + 5, 2, B, 2), generator=g, dtype=torch.int32).to(dev)
    st = torch.cuda.current_stream(dev)
    fn = L.lib().augcl_ring_sample
    def step(t):
        a, b = outs[t % n_sets]
        rc = fn(buf._ring, idx[t].data_ptr(), B, a.data_ptr(), b.data_ptr(), None, None, None, L.AUG_SHIFT,
                off[t, 0].data_ptr(), off[t, 1].data_ptr(), PAD, 0, st.cuda_stream)
        assert rc == 0, L.lib().augcl_last_error()
    for t in range(5):
        step(t)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for t in range(5, iters + 5):
            step(t)
        e1.record(st)
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / iters)
    return best * 1e3

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", default="128,1024,8192")
    ap.add_argument("--full", action="store_true")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    buf = U.ReplayBuffer((C, S, S), (6,), CAP, 128, device=dev)
    buf.fill_synthetic(seed=1)
    peak = 6546.2
    variants = [dict(sample_band=0, sample_pdl=0, sample_stages=3, sample_ctas_per_sm=6)]
    grid = itertools.product([0, 1], [2, 3, 4], [4, 5, 6, 7]) if a.full else [(0, 3, 6), (1, 3, 6), (1, 4, 6), (1, 2, 6), (1, 3, 5), (1, 3, 4), (0, 4, 6), (1, 4, 5)]
    for pdl, stg, cps in grid:
        variants.append(dict(sample_band=1, sample_pdl=pdl, sample_stages=stg, sample_ctas_per_sm=cps))
    for v in variants:
        for k, val in v.items():
            L.set_option(k, str(val))
        for B in [int(x) for x in a.batches.split(",")]:
            iters = max(30, min(2000, (1 << 20) // B))
            us = run(buf, B, iters, dev)
            gbs = B * BYTES / us / 1e3
            print(json.dumps({**v, "batch": B, "us": round(us, 2), "GB/s": round(gbs, 1), "frac": round(gbs / peak, 4)}), flush=True)

if __name__ == "__main__":
    main()
