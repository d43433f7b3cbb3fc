"""The sample entry points are CUDA-graph capturable when every draw is device-resident: they enqueue one kernel on the
given stream, allocate nothing themselves, never synchronise, and the fused sampler's ticket counters re-arm at kernel end
(SURVEY.md 7.2 "offer CUDA-graph capture").  A captured update replays with new indices / offsets written into the same
device tensors and equals the eager call."""
import numpy as np
import pytest
import torch

from conftest import TOL_NORM

pytestmark = pytest.mark.gpu


def test_captured_samplers_replay_with_new_draws():
    from augcl_b200 import _lib as L, augmentations as A, utils as U
    cap, B, S = 5000, 128, 84
    buf = U.ReplayBuffer((9, S, S), (6,), cap, B)
    buf.fill_synthetic(seed=2)
    A.set_background_bank(None)
    A.seed(3)
    A.load_dataloader(B, S)
    g = torch.Generator().manual_seed(1)
    idx = torch.randint(0, cap, (B,), generator=g).cuda()
    off = torch.randint(0, 9, (B, 2), generator=g, dtype=torch.int32).cuda()
    off2 = torch.randint(0, 9, (B, 2), generator=g, dtype=torch.int32).cuda()
    rows = torch.randint(0, A.SYNTHETIC_BANK_LEN, (B,), generator=g).cuda()
    wt = torch.randn((B, 3, 3, 3, 3), generator=g).cuda()

    def update():
        drq = buf.sample_drq(idxs=idx, offsets={"obs": off, "next_obs": off2})
        aug = buf.sample_augcl(idxs=idx, offsets={"obs": off}, bank_idx=rows, weights=wt)
        return drq, aug

    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    try:
        with torch.cuda.stream(side):
            update()                                   # warm-up outside the capture (lazy init, func attributes)
        torch.cuda.current_stream().wait_stream(side)
        graph = torch.cuda.CUDAGraph()
        n0 = L.launch_count()
        with torch.cuda.graph(graph):
            drq_g, aug_g = update()
        assert L.launch_count() - n0 == 2              # two kernels in the graph
        for trial in range(3):
            idx.copy_(torch.randint(0, cap, (B,), generator=g))
            off.copy_(torch.randint(0, 9, (B, 2), generator=g, dtype=torch.int32))
            graph.replay()
            drq_e, aug_e = update()
            torch.cuda.synchronize()
            for a, b in zip(drq_g, drq_e):
                assert torch.equal(a, b)
            assert torch.equal(aug_g[0], aug_e[0]) and torch.equal(aug_g[4], aug_e[4])
            assert (aug_g[1] - aug_e[1]).abs().max().item() / 255.0 <= TOL_NORM
    finally:
        A.set_background_bank(None)


def test_graph_sampler_equals_eager_calls():
    from augcl_b200 import utils as U
    cap, B, S, pad = 3000, 128, 84, 4
    buf = U.ReplayBuffer((9, S, S), (6,), cap, B)
    buf.fill_synthetic(seed=4)
    rng = np.random.RandomState(0)
    for kind in ("drq", "svea", "plain"):
        gs = U.GraphSampler(buf, kind=kind, pad=pad, host_reward=(kind == "drq"))
        for _ in range(3):
            idxs = rng.randint(0, cap, size=B)
            o1, o2 = rng.randint(0, 2 * pad + 1, (B, 2)), rng.randint(0, 2 * pad + 1, (B, 2))
            if kind == "drq":
                got = gs.sample(idxs, {"obs": o1, "next_obs": o2})
                want = buf.sample_drq(pad=pad, idxs=idxs, offsets={"obs": o1, "next_obs": o2})
            elif kind == "svea":
                got = gs.sample(idxs, o1)
                want = buf.sample_svea(pad=pad, idxs=idxs, offsets=o1)
            else:
                got = gs.sample(idxs)
                want = buf.sample(idxs=idxs)
            torch.cuda.synchronize()
            for a, b in zip(got, want):
                assert torch.equal(a, b), kind
            if kind == "drq":       # the captured D2H: the step's rewards are in pinned host memory after the sync
                assert torch.equal(gs.reward_host, want[2].cpu())
        with pytest.raises(IndexError):
            gs.sample(np.full(B, cap))
    # transitions added after the capture are visible to the next replay (the ring is read at replay time)
    gs = U.GraphSampler(buf, kind="plain")
    frame = np.full((9, S, S), 7, dtype=np.uint8)
    slot = buf.idx
    buf.add(frame, np.zeros(6, np.float32), 1.5, frame, False)
    got = gs.sample(np.full(B, slot))
    torch.cuda.synchronize()
    assert float(got[0].min()) == 7.0 and float(got[0].max()) == 7.0 and float(got[2][0]) == 1.5


def test_fast_sampler_equals_eager_calls():
    """utils.FastSampler: host draws by value in the launch, fixed outputs, rewards written straight to pinned host
    memory: same results as the drop-in calls, draws validated by the library, later adds visible."""
    from augcl_b200 import utils as U
    from augcl_b200._lib import AugclError
    S, cap, B, pad = 84, 300, 32, 4
    buf = U.ReplayBuffer((9, S, S), (6,), cap, B)
    buf.fill_synthetic(seed=3)
    rng = np.random.RandomState(1)
    for kind in ("drq", "svea", "plain"):
        fs = U.FastSampler(buf, kind=kind, pad=pad, host_reward=(kind != "svea"))
        for _ in range(3):
            idxs = rng.randint(0, cap, size=B)
            o1, o2 = rng.randint(0, 2 * pad + 1, (B, 2)), rng.randint(0, 2 * pad + 1, (B, 2))
            if kind == "drq":
                got = fs.sample(idxs, {"obs": o1, "next_obs": o2})
                want = buf.sample_drq(pad=pad, idxs=idxs, offsets={"obs": o1, "next_obs": o2})
            elif kind == "svea":
                got = fs.sample(idxs, o1)
                want = buf.sample_svea(pad=pad, idxs=idxs, offsets=o1)
            else:
                got = fs.sample(idxs)
                want = buf.sample(idxs=idxs)
            torch.cuda.synchronize()
            for i, (a, b) in enumerate(zip(got, want)):
                assert torch.equal(a.cpu(), b.cpu()), (kind, i)
            if kind != "svea":
                assert fs.reward_host is got[2] and not got[2].is_cuda
        with pytest.raises(AugclError):
            fs.sample(np.full(B, cap))
        if kind == "drq":
            with pytest.raises(AugclError):
                fs.sample(np.zeros(B, np.int64), {"obs": np.full((B, 2), 2 * pad + 1), "next_obs": np.zeros((B, 2))})
    fs = U.FastSampler(buf, kind="plain")
    frame = np.full((9, S, S), 9, dtype=np.uint8)
    slot = buf.idx
    buf.add(frame, np.zeros(6, np.float32), 2.5, frame, False)
    got = fs.sample(np.full(B, slot))
    torch.cuda.synchronize()
    assert float(got[0].min()) == 9.0 and float(got[0].max()) == 9.0 and float(got[2][0]) == 2.5
    with pytest.raises(ValueError):
        U.FastSampler(buf, batch=257)


@pytest.mark.parametrize("kind,data_aug", [("augcl", "splice-conv"), ("svea_aug", "overlay"), ("augcl", "conv"),
                                           ("svea_aug", "CS_splice")])
def test_fast_fused_sampler_equals_sample_fused(kind, data_aug):
    """utils.FastFusedSampler (cached descriptor, fixed outputs, host indices by value in the launch) returns what
    ReplayBuffer.sample_augcl / sample_svea_aug return for the same seed, indices and (optional) host offsets."""
    from augcl_b200 import utils as U, augmentations as A
    S, cap, B, pad = 84, 200, 24, 4
    buf = U.ReplayBuffer((9, S, S), (6,), cap, B)
    buf.fill_synthetic(seed=5)
    bank = torch.randint(0, 256, (32, 3, S, S), dtype=torch.uint8, device="cuda")
    A.set_background_bank(bank)
    try:
        fs = U.FastFusedSampler(buf, kind=kind, data_aug=data_aug, pad=pad)
        ref = buf.sample_augcl if kind == "augcl" else buf.sample_svea_aug
        rng = np.random.RandomState(2)
        for step in range(3):
            idxs = rng.randint(0, cap, size=B)
            offs = None if step == 0 else {"obs": rng.randint(0, 2 * pad + 1, (B, 2))}
            A.seed(11 + step)
            want = ref(idxs=idxs, data_aug=data_aug, pad=pad, offsets=offs)
            A.seed(11 + step)
            got = fs.sample(idxs, offs)
            torch.cuda.synchronize()
            for i, (a, b) in enumerate(zip(got, want)):
                assert torch.equal(a, b), (kind, data_aug, step, i)
        np.random.seed(3)
        A.seed(1)
        want = ref(data_aug=data_aug, pad=pad)           # the reference's own index draw (numpy global RNG)
        np.random.seed(3)
        A.seed(1)
        got = fs.sample()
        for a, b in zip(got, want):
            assert torch.equal(a, b)
    finally:
        A.set_background_bank(None)
