#!/usr/bin/env python
"""Host-side cost of one public-API step (ReplayBuffer.sample_drq with host draws): cProfile + wall clock."""
import cProfile, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import __graft_entry__ as entry
entry.build()
from augcl_b200 import utils as U
B, PAD, CAP = 128, 4, 100_000
buf = U.ReplayBuffer((9, 84, 84), (6,), CAP, B)
buf.fill_synthetic(seed=1)
rng = np.random.RandomState(0)
N = 3000
idx = rng.randint(0, CAP, (N, B))
off = rng.randint(0, 2 * PAD + 1, (N, 2, B, 2)).astype(np.int32)
h_rew = torch.empty((B, 1)).pin_memory()
st = torch.cuda.current_stream()
def step(t, sync=True):
    obs, act, rew, nxt, nd = buf.sample_drq(pad=PAD, idxs=idx[t], offsets={"obs": off[t, 0], "next_obs": off[t, 1]})
    if sync:
        h_rew.copy_(rew, non_blocking=True)
        st.synchronize()
for t in range(50): step(t)
for name, sync in (("with D2H+sync", True), ("launch only (async)", False)):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for t in range(N): step(t, sync)
    torch.cuda.synchronize(); el = time.perf_counter() - t0
    print(f"{name}: {1e6 * el / N:.1f} us/step")
pr = cProfile.Profile(); pr.enable()
for t in range(N): step(t)
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
