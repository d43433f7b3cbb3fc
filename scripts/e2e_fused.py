"""Host-side cost of the fused per-algorithm sampler calls: wall time per ReplayBuffer.sample_augcl / sample_svea_aug call
(device draws, stream synchronize every step) beside the kernel's device time."""
import time, json, numpy as np, torch
from augcl_b200 import utils as U, augmentations as A
B = 128
buf = U.ReplayBuffer((9, 84, 84), (6,), 100_000, B)
buf.fill_synthetic(seed=0)
bank = torch.randint(0, 256, (4096, 3, 84, 84), dtype=torch.uint8, device="cuda")
A.set_background_bank(bank)
st = torch.cuda.current_stream()
fa = U.FastFusedSampler(buf, kind="augcl", data_aug="splice-conv")
fv = U.FastFusedSampler(buf, kind="svea_aug", data_aug="overlay")
for name, fn, kw in (("FastFusedSampler augcl splice-conv", fa.sample, {}), ("FastFusedSampler svea_aug overlay", fv.sample, {}),
                     ("sample_augcl splice-conv", buf.sample_augcl, dict(data_aug="splice-conv")),
                     ("sample_svea_aug overlay", buf.sample_svea_aug, dict(data_aug="overlay")),
                     ("sample_drq", buf.sample_drq, {})):
    for _ in range(20):
        fn(**kw)
    st.synchronize()
    t0 = time.perf_counter()
    for _ in range(1000):
        fn(**kw)
        st.synchronize()
    wall = (time.perf_counter() - t0) / 1000
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(1000):
        fn(**kw)
    e1.record()
    st.synchronize()
    print(json.dumps({"call": name, "batch": B, "wall_us_per_step_with_sync": round(wall * 1e6, 2),
                      "device_us_per_step_back_to_back": round(e0.elapsed_time(e1), 2)}))
