"""Drop-in for the replay-buffer half of the reference's ``utils`` module
(/root/reference/src/utils.py:127-322): ``ReplayBuffer`` with the same constructor, attributes
(``capacity, batch_size, idx, full, last_save, actions, rewards, not_dones``) and methods
(``add, sample, sample_obs, sample_curl, sample_drq, save, load``) that ``src/train.py`` and
``src/algorithms/*`` use, and ``LazyFrames``.

The Python list of (LazyFrames, LazyFrames) becomes a device-resident uint8 ring
(include/augcl_b200.h, "v1" full-stack layout); every ``sample*`` is ONE fused CUDA kernel that
gathers obs + next_obs by index, casts u8->f32 and (sample_drq / sample_curl) applies the random
shift / crop on the way out.  No CPU path: the CUDA library is required.
"""
from __future__ import annotations

import collections
import ctypes as C
import os

import numpy as np
import torch

from . import _lib as L
from . import augmentations
from . import persist


try:                                    # raw cudaStream_t of torch's current stream without building a Stream object
    _raw_stream = torch._C._cuda_getCurrentRawStream
except AttributeError:                  # pragma: no cover
    def _raw_stream(index):
        return torch.cuda.current_stream(index).cuda_stream


def k_dev_keep(*srcs):
    """True when a call made device-side copies of its offset sources (they must outlive the launch)."""
    for x in srcs:
        if x is not None and x[2] is not None:
            return True
    return False


class LazyFrames(object):
    """k frames joined along the channel axis on demand: the pickle-compatible counterpart of the reference's class
    (src/utils.py:283-322).  The three attribute names are part of the on-disk chunk format (instances are pickled by
    ReplayBuffer.save and unpickled by the reference, and the other way round), so they are kept; the behaviour is the
    reference's: ``extremely_lazy`` never caches, otherwise the joined array is cached and the frame list dropped."""

    def __init__(self, frames, extremely_lazy=True):
        self._frames, self._extremely_lazy, self._out = frames, extremely_lazy, None

    frames = property(lambda self: self._frames)

    def _force(self):
        if self._out is not None:
            return self._out
        joined = np.concatenate(self._frames, axis=0)
        if not self._extremely_lazy:             # cache, and let go of the per-frame references
            self._out, self._frames = joined, None
        return joined

    def __array__(self, dtype=None, copy=None):
        arr = self._force()
        return arr if dtype is None else arr.astype(dtype)

    def __len__(self):
        return len(self._frames) if self._extremely_lazy else len(self._force())

    def __getitem__(self, i):
        return self._force()[i]

    def count(self):
        return len(self._frames) if self._extremely_lazy else self._force().shape[0] // 3

    def frame(self, i):
        return self._force()[3 * i:3 * i + 3]


# chunks written by save() unpickle under the reference (its class lives in module ``utils``)
LazyFrames.__module__ = "utils"


try:
    _storage_use_count = torch._C._storage_Use_Count
except AttributeError:                  # pragma: no cover
    _storage_use_count = None
_getrefcount = __import__("sys").getrefcount


class _OutputSet(object):
    """The five tensors one sample() returns -- obs, next_obs [B,C,Ho,Wo], actions [B,...], rewards, not_dones [B,1] --
    as views of ONE flat f32 device block, with the pointers the C ABI needs computed once."""
    __slots__ = ("flat", "views", "ptrs", "n_storage", "storage")

    def __init__(self, B, chw, act_shape, device):
        c, h, w = chw
        n_img = B * c * h * w
        n_act = 1
        for d in act_shape:
            n_act *= d
        n_act *= B
        pad = (-n_act) % 4                                   # keep the reward / not_done rows 16-byte aligned
        flat = torch.empty(2 * n_img + n_act + pad + 2 * B, dtype=torch.float32, device=device)
        o0 = 2 * n_img
        r0 = o0 + n_act + pad
        obs, nxt = flat[:n_img].view(B, c, h, w), flat[n_img:o0].view(B, c, h, w)
        act = flat[o0:o0 + n_act].view((B,) + tuple(act_shape))
        rew, nd = flat[r0:r0 + B].view(B, 1), flat[r0 + B:r0 + 2 * B].view(B, 1)
        self.flat = flat
        self.views = (obs, act, rew, nxt, nd)               # the order sample() returns them in
        self.ptrs = (obs.data_ptr(), nxt.data_ptr(), act.data_ptr(), rew.data_ptr(), nd.data_ptr())
        self.storage = flat.untyped_storage()._cdata
        self.n_storage = _storage_use_count(self.storage)   # flat + the five views (+ this query's wrapper, every time)

    def free(self):
        """True when nothing outside the pool can still see this memory: no other tensor shares the storage (a view the
        caller made, a tensor autograd saved), nobody holds one of the five Python objects, nobody holds their
        TensorImpls from C++, and the caller left no autograd state on them."""
        if _storage_use_count(self.storage) != self.n_storage:
            return False
        for v in self.views:          # 3 = the tuple, this loop variable and getrefcount's own argument
            if _getrefcount(v) != 3 or v._use_count() != 1 or v.requires_grad:
                return False
        return True


class ReplayBuffer(object):
    """Buffer to store environment transitions — device-resident (src/utils.py:136-280)."""

    STAGE = 256  # transitions staged in pinned host memory between flushes to the ring

    POOL_SETS = 3             # output sets kept per batch geometry (see _outputs)
    POOL_MAX_BYTES = 1 << 30  # ... as long as the sets of one geometry stay below this

    def __init__(self, obs_shape, action_shape, capacity, batch_size, prefill=True, *, device=None,
                 frame_channels=3, dedup=False, pool_factor=1.25, pool_outputs=True):
        # `prefill` only reserved host memory in the reference (src/utils.py:127-133,143-145); the ring
        # is allocated up front, so the flag is accepted and has nothing left to do.
        if not torch.cuda.is_available():
            raise RuntimeError("augcl_b200.utils.ReplayBuffer needs a CUDA device (no CPU path)")
        self.capacity = int(capacity)
        self.batch_size = batch_size
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self._dev_index = self.device.index
        c, h, w = obs_shape
        if c % frame_channels:
            raise ValueError(f"obs channels {c} not a multiple of frame_channels {frame_channels}")
        self.obs_shape = (int(c), int(h), int(w))
        self._frame_channels = int(frame_channels)
        self.action_dim = int(np.prod(action_shape)) if len(action_shape) else 1
        self.actions = np.empty((self.capacity, *action_shape), dtype=np.float32)
        self.rewards = np.empty((self.capacity, 1), dtype=np.float32)
        self.not_dones = np.empty((self.capacity, 1), dtype=np.float32)
        self.idx = 0
        self.last_save = 0
        self.last_requeue_save = 0
        self.full = False
        self.refilled = False

        h_ring = C.c_void_p()
        k = c // frame_channels
        self._k = k
        self.dedup = bool(dedup)
        n = min(self.STAGE, self.capacity)
        if self.dedup:
            # "v2" layout (include/augcl_b200.h): one copy of every distinct frame + 2k references per transition,
            # the sharing LazyFrames / FrameStack have on the host (src/utils.py:283-322, src/env/wrappers.py:183-203)
            self._pool = int(self.capacity * float(pool_factor)) + 4 * k + 16
            L.call("augcl_ring_create_dedup", C.byref(h_ring), self._dev_index, self.capacity, k, frame_channels, h, w,
                   self.action_dim, self._pool)
            self._pool_cur = 0                 # next pool slot to allocate
            self._adds = 0                     # transitions added so far
            self._slot_last_use = np.full(self._pool, -(1 << 62), dtype=np.int64)   # add number of the newest referrer
            self._recent = collections.OrderedDict()   # id(frame) -> (frame, pool slot): the last few distinct frames
            self._recent_max = 4 * k + 8
            self._sealed = False
            self._st_frames = torch.empty((n * 2 * k, frame_channels, h, w), dtype=torch.uint8).pin_memory()
            self._st_frames_np = self._st_frames.numpy()
            self._st_refs = torch.empty((n, 2, k), dtype=torch.int32).pin_memory()
            self._st_refs_np = self._st_refs.numpy()
            self._st_nframes = 0
            self._st_frame0 = 0                # pool slot of staged frame 0
        else:
            L.call("augcl_ring_create", C.byref(h_ring), self._dev_index, self.capacity, k, frame_channels, h, w,
                   self.action_dim)
            self._st_obs = torch.empty((n, c, h, w), dtype=torch.uint8).pin_memory()
            self._st_next = torch.empty((n, c, h, w), dtype=torch.uint8).pin_memory()
            self._st_obs_np, self._st_next_np = self._st_obs.numpy(), self._st_next.numpy()
        self._st_cap = n
        self._ring = h_ring
        self._fn_sample = L.lib().augcl_ring_sample_ex
        self._stack_bytes = c * h * w
        self._st_start = 0      # ring slot of staged transition 0
        self._st_count = 0
        self._st_event = None   # completion of the last flush (staging must not be overwritten before)
        self._host_blocks = {}  # batch size -> host block of the draws of one call (_host_block)
        self._out_pool = {}     # (B, Ho, Wo) -> [_OutputSet]: recycled output sets (_outputs)
        self._pool_outputs = bool(pool_outputs) and _storage_use_count is not None
        self._held = collections.deque()   # (event, device-side sources of a call) not yet known to be consumed

    def __del__(self):
        ring, self._ring = getattr(self, "_ring", None), None
        if ring:
            try:
                L.lib().augcl_ring_destroy(ring)
            except Exception:
                pass

    # ------------------------------------------------------------------ ingest
    def add(self, obs, action, reward, next_obs, done):
        """src/utils.py:216-227.  obs / next_obs: LazyFrames or (C,H,W) uint8 arrays."""
        if self._st_count == 0:
            if self._st_event is not None:
                self._st_event.synchronize()
                self._st_event = None
            self._st_start = self.idx
        j = self._st_count
        if self.dedup:
            self._stage_frames(j, obs, next_obs)
        else:
            self._st_obs_np[j] = np.asarray(obs)
            self._st_next_np[j] = np.asarray(next_obs)
        np.copyto(self.actions[self.idx], action)
        np.copyto(self.rewards[self.idx], reward)
        np.copyto(self.not_dones[self.idx], not done)
        self._st_count += 1
        self.idx = (self.idx + 1) % self.capacity
        self.full = self.full or self.idx == 0
        if self._st_count == self._st_cap or self.idx == 0:   # staging full, or about to wrap
            self._flush()

    # ---- de-duplicated ingest -------------------------------------------------------------------------
    def _frames_of(self, x):
        """The k frame arrays of an observation.  LazyFrames hands out its (shared) frame objects; anything else is
        split into k views, which are never shared (every add then uploads 2k frames)."""
        fr = getattr(x, "_frames", None)
        if fr is not None and len(fr) == self._k:
            return fr
        a = np.asarray(x)
        cpf = self._frame_channels
        return [a[i * cpf:(i + 1) * cpf] for i in range(self._k)]

    def _stage_frames(self, j, obs, next_obs):
        """Resolve the 2k frames of one transition to pool slots (frames seen in the last few adds are shared BY OBJECT
        IDENTITY, like the reference's LazyFrames), stage the new ones, record the references."""
        if self._sealed:
            raise RuntimeError("ReplayBuffer(dedup=True): add() after fill_synthetic() is not supported")
        k = self._k
        if self._pool_cur + 2 * k > self._pool:      # keep the staged frames one contiguous pool range
            self._flush_frames()
            self._pool_cur = 0
        refs = self._st_refs_np[j]
        recent = self._recent
        for t, x in enumerate((obs, next_obs)):
            for f, fr in enumerate(self._frames_of(x)):
                ent = recent.get(id(fr))
                if ent is not None and ent[0] is fr:
                    slot = ent[1]
                else:
                    slot = self._pool_cur
                    if self._slot_last_use[slot] > self._adds - self.capacity:
                        raise RuntimeError(
                            f"ReplayBuffer(dedup=True): frame pool of {self._pool} exhausted (consecutive adds share too few "
                            "frame objects); raise pool_factor or use dedup=False")
                    if self._st_nframes == 0:
                        self._st_frame0 = slot
                    self._st_frames_np[self._st_nframes] = fr
                    self._st_nframes += 1
                    self._pool_cur += 1
                    recent[id(fr)] = (fr, slot)
                    if len(recent) > self._recent_max:
                        recent.popitem(last=False)
                refs[t, f] = slot
                self._slot_last_use[slot] = self._adds
        self._adds += 1

    def _flush_frames(self):
        n = self._st_nframes
        if n:
            L.call("augcl_ring_put_frames", self._ring, self._st_frame0, n, self._st_frames.data_ptr(), self._stream())
            self._st_nframes = 0
            # frames staged from here on must not overwrite pinned memory the copy above still reads
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(self.device))
            ev.synchronize()

    def _stream(self):
        return torch.cuda.current_stream(self.device).cuda_stream

    def _flush(self):
        """Pending staged transitions -> ring (one async H2D per field)."""
        n = self._st_count
        if n == 0:
            return
        s0 = self._st_start
        a = np.ascontiguousarray(self.actions[s0:s0 + n]).reshape(n, -1)
        if self.dedup:
            if self._st_nframes:
                L.call("augcl_ring_put_frames", self._ring, self._st_frame0, self._st_nframes, self._st_frames.data_ptr(),
                       self._stream())
                self._st_nframes = 0
            L.call("augcl_ring_put_refs", self._ring, s0, n, self._st_refs.data_ptr(), a.ctypes.data,
                   self.rewards[s0:s0 + n].ctypes.data, self.not_dones[s0:s0 + n].ctypes.data, self._stream())
        else:
            L.call("augcl_ring_put", self._ring, s0, n, self._st_obs.data_ptr(), self._st_next.data_ptr(),
                   a.ctypes.data, self.rewards[s0:s0 + n].ctypes.data, self.not_dones[s0:s0 + n].ctypes.data,
                   self._stream())
        # the scalar sources are pageable (copied synchronously by the driver); the frame staging is pinned
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        self._st_event = ev
        self._st_count = 0

    # ------------------------------------------------------------------ indices
    def _get_idxs(self, n=None):
        """src/utils.py:229-232 — same numpy global-RNG draw as the reference."""
        if n is None:
            n = self.batch_size
        return np.random.randint(0, self.capacity if self.full else self.idx, size=n)

    # ------------------------------------------------------------------ the fused sample kernel
    def _host_block(self, B):
        """Per-batch-size host block the HOST draws of a call are copied into (one numpy slice assignment each: dtype
        conversion, contiguity and list -> array in one step, ~0.4 us) so the C ABI always gets the same three
        pre-computed pointers.  HOST sources are consumed before the call returns (by value in the launch for B <= 256,
        a synchronous pageable staging copy otherwise), so the block can be reused by the next call."""
        blk = self._host_blocks.get(B)
        if blk is None:
            idx = np.zeros(B, dtype=np.int64)
            off = np.zeros((3, B, 2), dtype=np.int32)
            blk = self._host_blocks[B] = (idx, off, idx.ctypes.data, off[0].ctypes.data, off[1].ctypes.data,
                                          off[2].ctypes.data)
        return blk

    def _idx_arg(self, idxs):
        """-> (AUGCL_SRC_*, pointer, B, keepalive).  numpy / list / CPU tensor = HOST (the reference's case:
        np.random.randint on the host, src/utils.py:229-232); CUDA tensor = DEVICE."""
        if isinstance(idxs, torch.Tensor):
            t = idxs.reshape(-1)
            if t.is_cuda:
                made = t.dtype != torch.int64 or not t.is_contiguous() or t.device != self.device
                if made:        # a converted copy must outlive the launch (_hold); the caller's own tensor is the caller's
                    t = t.to(device=self.device, dtype=torch.int64).contiguous()
                return L.SRC_DEVICE, t.data_ptr(), t.numel(), (t if made else None)
            idxs = t.numpy()
        elif type(idxs) is not np.ndarray:
            idxs = np.asarray(idxs)
        if idxs.ndim != 1:
            idxs = idxs.reshape(-1)
        B = idxs.shape[0]
        if B == 0:
            raise L.AugclError("augcl_ring_sample_ex", -1, "B must be positive (got 0)")
        if idxs.dtype.kind not in "iu":
            raise TypeError(f"ReplayBuffer: indices must be integers, got {idxs.dtype}")
        blk = self._host_block(B)
        blk[0][:] = idxs
        return L.SRC_HOST, blk[2], B, None

    def _off_arg(self, v, B, slot=0):
        """-> None | (AUGCL_SRC_*, pointer, keepalive) for one (B,2) offsets array (slot 0 obs, 1 next_obs, 2 pos)."""
        if v is None:
            return None
        if isinstance(v, torch.Tensor):
            if v.is_cuda:
                if tuple(v.shape) != (B, 2):
                    raise ValueError(f"ReplayBuffer: offsets must have shape ({B}, 2), got {tuple(v.shape)}")
                made = v.dtype != torch.int32 or not v.is_contiguous() or v.device != self.device
                t = v.to(device=self.device, dtype=torch.int32).contiguous() if made else v
                return L.SRC_DEVICE, t.data_ptr(), (t if made else None)
            v = v.numpy()
        elif type(v) is not np.ndarray:
            v = np.asarray(v)
        if v.shape != (B, 2):
            raise ValueError(f"ReplayBuffer: offsets must have shape ({B}, 2), got {v.shape}")
        blk = self._host_block(B)
        blk[1][slot] = v
        return L.SRC_HOST, blk[3 + slot], None

    def _sample(self, idxs=None, *, n=None, obs=True, next_obs=True, scalars=True, aug=L.AUG_NONE, off_obs=None,
                off_next=None, pad_or_size=0, normalize=False, device_indices=False, pos=False, off_pos=None, rng=True):
        """One augcl_ring_sample_ex launch.  Returns (obs, actions, rewards, next_obs, not_dones); entries
        that were not requested are None.  Host indices / offsets travel inside the launch (no extra H2D);
        offsets that are not supplied are drawn inside the kernel (Philox stream of augmentations.seed()).
        rng=False: the call draws nothing and leaves the Philox call counter alone (internal reads: save())."""
        if self._st_count:
            self._flush()
        sub = augmentations._reserve_subsequences(4 if pos else 3) if rng else 0
        if idxs is None and device_indices:
            B = self.batch_size if n is None else int(n)
            i_src, i_ptr, keep_i = L.SRC_PHILOX, None, None
            idx_hi = self.capacity if self.full else self.idx
        else:
            if idxs is None:
                idxs = self._get_idxs(n)
            i_src, i_ptr, B, keep_i = self._idx_arg(idxs)
            idx_hi = 0
        oo, on = self._off_arg(off_obs, B, 0), self._off_arg(off_next, B, 1)
        op = self._off_arg(off_pos, B, 2) if pos else None
        if aug == L.AUG_NONE:
            o_src = L.SRC_DEVICE
        else:
            if aug == L.AUG_SHIFT_OBS:
                need_draw = obs and oo is None
            else:
                need_draw = (obs and oo is None) or (next_obs and on is None) or (pos and op is None)
            k_host = k_dev = False
            for x in (oo, on, op):
                if x is not None:
                    if x[0] == L.SRC_HOST:
                        k_host = True
                    else:
                        k_dev = True
            if k_host and (need_draw or k_dev):
                # PHILOX draws for the missing tensors (or a mix of sources): host arrays become DEVICE arrays
                def to_dev(x, slot):
                    if x is None or x[0] == L.SRC_DEVICE:
                        return x
                    t = torch.from_numpy(self._host_block(B)[1][slot]).to(self.device, non_blocking=False)
                    return L.SRC_DEVICE, t.data_ptr(), t
                oo, on, op = to_dev(oo, 0), to_dev(on, 1), to_dev(op, 2)
                k_host = False
            o_src = L.SRC_PHILOX if need_draw else (L.SRC_HOST if k_host else L.SRC_DEVICE)
        c, h, w = self.obs_shape
        ho, wo = (pad_or_size, pad_or_size) if aug == L.AUG_CROP else (h, w)
        # torch.empty(device=) and the raw-stream query name the ring's device explicitly and the library switches to the
        # ring's device for the call (DeviceGuard), so no device context manager is needed here
        dev = self.device
        o = no = act = rew = nd = None
        po = pn = pa = pr = pd = None
        if obs and next_obs and scalars and not pos:
            (o, act, rew, no, nd), (po, pn, pa, pr, pd) = self._outputs(B, ho, wo)
            scalars = False                     # allocated above
        elif obs and next_obs:
            o, no = torch.empty((2, B, c, ho, wo), dtype=torch.float32, device=dev).unbind(0)
            po = o.data_ptr()
            pn = po + 4 * B * c * ho * wo
        elif obs:
            o = torch.empty((B, c, ho, wo), dtype=torch.float32, device=dev)
            po = o.data_ptr()
        elif next_obs:
            no = torch.empty((B, c, ho, wo), dtype=torch.float32, device=dev)
            pn = no.data_ptr()
        if scalars:
            act = torch.empty(self._act_shape(B), dtype=torch.float32, device=dev)
            rew, nd = torch.empty((2, B, 1), dtype=torch.float32, device=dev).unbind(0)
            pa, pr = act.data_ptr(), rew.data_ptr()
            pd = pr + 4 * B
        if pos:     # sample_curl: a third, independently cropped view of obs from the same launch
            ps = torch.empty((B, c, ho, wo), dtype=torch.float32, device=dev)
            rc = L.lib().augcl_ring_sample_curl(
                self._ring, i_src, i_ptr, idx_hi, B, po, pn, ps.data_ptr(), pa, pr, pd, o_src,
                None if oo is None else oo[1], None if on is None else on[1], None if op is None else op[1],
                int(pad_or_size), 1 if normalize else 0, augmentations._seed, sub, _raw_stream(self._dev_index))
            self._pos = ps
        else:
            rc = self._fn_sample(self._ring, i_src, i_ptr, idx_hi, B, po, pn, pa, pr, pd, aug, o_src,
                                 None if oo is None else oo[1], None if on is None else on[1], int(pad_or_size),
                                 1 if normalize else 0, augmentations._seed, sub, _raw_stream(self._dev_index))
        if rc != 0:
            raise L.AugclError("augcl_ring_sample_curl" if pos else "augcl_ring_sample_ex", rc,
                               L.lib().augcl_last_error().decode(errors="replace"))
        if keep_i is not None or k_dev_keep(oo, on, op):
            self._hold((keep_i, oo, on, op))
        return o, act, rew, no, nd

    def _act_shape(self, B):
        return (B,) + self.actions.shape[1:]

    def _outputs(self, B, ho, wo):
        """Output tensors for one full sample() (obs, actions, rewards, next_obs, not_dones) + their device pointers.
        The reference returns fresh tensors from every call (src/utils.py:206-214) and so does this: a set is handed out
        again only when nothing outside the pool references it any more (_OutputSet.free: storage use count, Python
        reference counts, TensorImpl use counts) -- the caller cannot tell it from a new allocation, but the call skips
        three allocator round trips and the construction of five tensor objects (~10 us of a ~25 us call).  Like torch's
        caching allocator the pool orders reuse by the launching stream only; a caller that consumes the outputs on another
        stream must keep them alive until that work is done (Tensor.record_stream has no effect on pooled memory).
        Geometries whose POOL_SETS sets would exceed POOL_MAX_BYTES, and calls that find every set busy, allocate."""
        c = self.obs_shape[0]
        if self._pool_outputs:
            key = (B, ho, wo)
            sets = self._out_pool.get(key)
            if sets is None:
                per_set = 4 * (2 * B * c * ho * wo + B * (self.action_dim + 2))
                sets = self._out_pool[key] = [] if per_set * self.POOL_SETS <= self.POOL_MAX_BYTES else None
                if len(self._out_pool) > 6:                  # geometries come and go (tests); keep the table small
                    self._out_pool.pop(next(iter(self._out_pool)))
            if sets is not None:
                for i, st in enumerate(sets):
                    if st.free():
                        sets.append(sets.pop(i))             # most recently used last: the next call tries the oldest first
                        return st.views, st.ptrs
                if len(sets) < self.POOL_SETS:
                    st = _OutputSet(B, (c, ho, wo), self.actions.shape[1:], self.device)
                    sets.append(st)
                    return st.views, st.ptrs
        dev = self.device
        o, no = torch.empty((2, B, c, ho, wo), dtype=torch.float32, device=dev).unbind(0)
        act = torch.empty(self._act_shape(B), dtype=torch.float32, device=dev)
        rew, nd = torch.empty((2, B, 1), dtype=torch.float32, device=dev).unbind(0)
        po, pr = o.data_ptr(), rew.data_ptr()
        return (o, act, rew, no, nd), (po, po + 4 * B * c * ho * wo, act.data_ptr(), pr, pr + 4 * B)

    def _hold(self, keep):
        """Device-side sources made for one call (converted offsets, a moved index tensor) stay referenced until the
        kernel that reads them has run: a short FIFO of (event, objects), released as the events complete.  (torch's
        caching allocator is stream-ordered, so dropping them early would be safe on the launching stream; the FIFO
        also covers callers that hand over tensors produced on another stream.)"""
        if torch.cuda.is_current_stream_capturing():
            self._held.append((None, keep))          # kept for the buffer's lifetime: a graph replays the launch
            return
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        q = self._held
        q.append((ev, keep))
        while len(q) > 1 and q[0][0] is not None and q[0][0].query():
            q.popleft()

    # ------------------------------------------------------------------ reference API
    def tensor_buffer_samples(self, idxs):
        """src/utils.py:206-214 (note the return order differs from sample())."""
        obs, actions, rewards, next_obs, not_dones = self._sample(idxs)
        return obs, next_obs, actions, rewards, not_dones

    def sample(self, n=None, idxs=None, *, device_indices=False):
        """src/utils.py:272-280.  device_indices=True draws the indices inside the kernel instead of with
        numpy's global RNG (no host work at all)."""
        return self._sample(idxs, n=n, device_indices=device_indices)

    def sample_obs(self, n=None, *, idxs=None):
        """src/utils.py:242-245."""
        return self._sample(idxs, n=n, next_obs=False, scalars=False)[0]

    def sample_curl(self, n=None, *, idxs=None, size=84, offsets=None):
        """src/utils.py:247-258: three independent random crops (pos, obs, next_obs), fused into the
        gather.  offsets: optional dict {"pos","obs","next_obs"} -> (B,2) (w1,h1) draws."""
        if idxs is None:
            idxs = self._get_idxs(n)
        c, h, w = self.obs_shape
        crop_max = w - size
        if crop_max <= 0:   # random_crop passthrough (src/augmentations.py:396-399)
            obs, actions, rewards, next_obs, not_dones = self._sample(idxs)
            return obs, actions, rewards, next_obs, not_dones, obs.clone()
        offsets = offsets or {}
        obs, actions, rewards, next_obs, not_dones = self._sample(
            idxs, aug=L.AUG_CROP, off_obs=offsets.get("obs"), off_next=offsets.get("next_obs"), pad_or_size=size,
            pos=True, off_pos=offsets.get("pos"))
        pos, self._pos = self._pos, None
        return obs, actions, rewards, next_obs, not_dones, pos

    def sample_drq(self, n=None, pad=4, *, idxs=None, offsets=None, device_indices=False):
        """src/utils.py:260-270: gather + random_shift of obs and next_obs in one kernel.
        offsets: optional dict {"obs","next_obs"} -> (B,2) (row,col) draws in [0, 2*pad]."""
        offsets = offsets or {}
        return self._sample(idxs, n=n, aug=L.AUG_SHIFT, off_obs=offsets.get("obs"), off_next=offsets.get("next_obs"),
                            pad_or_size=pad, device_indices=device_indices)

    def sample_encoder_input(self, n=None, *, idxs=None, center_crop=None, device_indices=False):
        """sample() with the encoder's first two layers folded into the same kernel:
        obs / next_obs = NormalizeImg(CenterCrop(center_crop)(.)) of src/algorithms/modules.py:69-90 — the centre
        ``center_crop`` x ``center_crop`` window (None or the stored size: no crop) divided by 255.0 (IEEE-rounded, bit-equal
        to torch's ``x / 255.0``).  Saves the f32 round trip of the un-normalised batch per encoder call."""
        c, h, w = self.obs_shape
        size = h if center_crop is None else int(center_crop)
        if size == h and size == w:
            return self._sample(idxs, n=n, normalize=True, device_indices=device_indices)
        if not (0 < size <= min(h, w)) or (h - size) % 2 or (w - size) % 2:
            raise ValueError(f"sample_encoder_input: cannot centre a {size}x{size} window in {h}x{w}")
        B = self.batch_size if n is None else int(n)
        if idxs is not None:
            B = int(idxs.numel()) if isinstance(idxs, torch.Tensor) else int(np.asarray(idxs).size)
        key = (B, size)
        off = self._center_off.get(key) if hasattr(self, "_center_off") else None
        if off is None:
            if not hasattr(self, "_center_off"):
                self._center_off = {}
            off = self._center_off[key] = torch.tensor([[(h - size) // 2, (w - size) // 2]], dtype=torch.int32,
                                                       device=self.device).repeat(B, 1).contiguous()
        return self._sample(idxs, n=n, aug=L.AUG_CROP, off_obs=off, off_next=off, pad_or_size=size, normalize=True,
                            device_indices=device_indices)

    def sample_drq_km(self, n=None, pad=4, *, k=2, m=2, idxs=None, offsets=None, device_indices=False):
        """A DrQ update's data (src/algorithms/drq.py:67-70) from ONE kernel: k independently shifted copies of obs and m
        of next_obs.  Returns (obs_list, action, reward, next_obs_list, not_done).
        offsets: optional {"obs": [k arrays (B,2)], "next_obs": [m arrays (B,2)]}; missing ones are drawn in the kernel."""
        k, m = int(k), int(m)
        if k < 0 or m < 0 or not 1 <= k + m <= 6:
            raise ValueError("sample_drq_km: need 1 <= k + m <= 6")
        self._flush()
        nv = k + m
        sub = augmentations._reserve_subsequences(3 + nv)
        if device_indices and idxs is None:
            B = self.batch_size if n is None else int(n)
            i_src, i_ptr, keep_i = L.SRC_PHILOX, None, None
            idx_hi = self.capacity if self.full else self.idx
        else:
            if idxs is None:
                idxs = self._get_idxs(n)
            i_src, i_ptr, B, keep_i = self._idx_arg(idxs)
            idx_hi = 0
        offsets = offsets or {}
        given = list(offsets.get("obs") or [None] * k) + list(offsets.get("next_obs") or [None] * m)
        if len(given) != nv:
            raise ValueError(f"sample_drq_km: offsets must hold {k} obs and {m} next_obs arrays")
        dev = self.device
        keep = [keep_i]
        with torch.cuda.device(dev):
            offs = []
            for g in given:                       # supplied draws become DEVICE arrays; the rest are PHILOX draws
                if g is None:
                    offs.append(None)
                    continue
                t = torch.as_tensor(np.asarray(g.cpu() if isinstance(g, torch.Tensor) else g)).to(device=dev, dtype=torch.int32).contiguous()
                if tuple(t.shape) != (B, 2):
                    raise ValueError(f"sample_drq_km: offsets must have shape ({B}, 2), got {tuple(t.shape)}")
                offs.append(t)
            keep.append(offs)
            c, h, w = self.obs_shape
            block = torch.empty((nv, B, c, h, w), dtype=torch.float32, device=dev)
            views = list(block.unbind(0))
            act = torch.empty((B,) + self.actions.shape[1:], dtype=torch.float32, device=dev)
            rew, nd = torch.empty((2, B, 1), dtype=torch.float32, device=dev).unbind(0)
            src = (C.c_int * nv)(*([0] * k + [1] * m))
            outs = (C.c_void_p * nv)(*[v.data_ptr() for v in views])
            offp = (C.c_void_p * nv)(*[None if o is None else o.data_ptr() for o in offs])
            L.call("augcl_ring_sample_views", self._ring, i_src, i_ptr, idx_hi, B, nv, src, outs, offp, act.data_ptr(),
                   rew.data_ptr(), nd.data_ptr(), L.AUG_SHIFT, L.SRC_PHILOX, int(pad), 0, augmentations._seed, sub,
                   _raw_stream(self._dev_index))
        # `keep` holds tensors made on the launching stream: torch's allocator is stream-ordered, so letting them go here
        # is safe (their memory is not reused before the kernel that reads them has run)
        del keep
        return views[:k], act, rew, views[k:], nd

    def sample_svea(self, n=None, pad=4, *, idxs=None, offsets=None, device_indices=False):
        """The SVEA / AugCL access pattern (src/algorithms/svea.py:66-70, augcl.py:69-72) as one kernel:
        obs is shifted, next_obs is returned un-augmented.  offsets: optional (B,2) draws for obs."""
        return self._sample(idxs, n=n, aug=L.AUG_SHIFT_OBS, off_obs=offsets, pad_or_size=pad,
                            device_indices=device_indices)

    # ------------------------------------------------------------------ fused per-algorithm samplers
    _FUSED_OPS = {"splice": L.OP_SPLICE, "CS_splice": L.OP_CS_SPLICE, "overlay": L.OP_OVERLAY, "conv": L.OP_CONV}
    _FUSED_CHAINS = {(), ("splice",), ("CS_splice",), ("overlay",), ("conv",), ("splice", "conv"), ("CS_splice", "conv"),
                     ("overlay", "conv")}

    def sample_fused(self, n=None, *, idxs=None, pad=4, data_aug="", aug_params=None, aug_input="plain",
                     want=("obs_shift", "obs_aug", "next_obs"), offsets=None, bank_idx=None, weights=None,
                     normalize=False, device_indices=False):
        """ONE kernel (augcl_ring_sample_fused) for a whole update(): gather + cast + random_shift +
        RAD.apply_aug chain, every frame read once.

        data_aug   the reference's ``--data_aug`` string (src/algorithms/rad.py:9-16): "splice-conv", "overlay",
                   "conv", "splice", "CS_splice", "CS_splice-conv", "overlay-conv" ("" / "identity" = no obs_aug).
        aug_params the reference's ``--aug_params`` dict, e.g. {"splice": {"val_thres": 0.4}}.
        aug_input  "plain": chain on the un-shifted obs (AugCL, src/algorithms/augcl.py:70-71);
                   "shift": chain on random_shift(obs) (SVEA, src/algorithms/svea.py:67 + :46).
        want       subset of {"obs", "obs_shift", "obs_aug", "next_obs", "next_obs_shift"}.
        offsets    optional {"obs": (B,2), "next_obs": (B,2)} shift draws; bank_idx (B,), weights (B,3,3,3,3).
        Returns a dict with the wanted tensors plus "action", "reward", "not_done".
        Raises NotImplementedError for chains the fused kernel does not cover (use sample() + augmentations)."""
        keys = tuple(k for k in str(data_aug).split("-") if k and k != "identity")
        if keys not in self._FUSED_CHAINS:
            raise NotImplementedError(f"sample_fused: chain {data_aug!r} is not fused; use sample() + augmentations.*")
        want = set(want)
        bad = want - {"obs", "obs_shift", "obs_aug", "next_obs", "next_obs_shift"}
        if bad:
            raise ValueError(f"sample_fused: unknown outputs {sorted(bad)}")
        if "obs_aug" in want and not keys:
            raise ValueError("sample_fused: obs_aug wanted but data_aug is empty")
        if aug_input not in ("plain", "shift"):
            raise ValueError(f"sample_fused: aug_input must be 'plain' or 'shift', got {aug_input!r}")
        aug_params = aug_params or {}
        offsets = offsets or {}
        self._flush()
        sub = augmentations._reserve_subsequences(3)
        if device_indices and idxs is None:
            B = self.batch_size if n is None else int(n)
            i_src, i_ptr, keep_i = L.SRC_PHILOX, None, None
            idx_hi = self.capacity if self.full else self.idx
        else:
            if idxs is None:
                idxs = self._get_idxs(n)
            i_src, i_ptr, B, keep_i = self._idx_arg(idxs)
            idx_hi = 0
        oo, on = self._off_arg(offsets.get("obs"), B, 0), self._off_arg(offsets.get("next_obs"), B, 1)
        shift_obs = "obs_shift" in want or ("obs_aug" in want and aug_input == "shift")
        need_draw = (shift_obs and oo is None) or ("next_obs_shift" in want and on is None)
        kinds = {x[0] for x in (oo, on) if x is not None}
        if need_draw or len(kinds) > 1:
            def to_dev(x, slot):
                if x is None or x[0] == L.SRC_DEVICE:
                    return x
                t = torch.from_numpy(self._host_block(B)[1][slot]).to(self.device)
                return L.SRC_DEVICE, t.data_ptr(), t
            oo, on = to_dev(oo, 0), to_dev(on, 1)
            o_src = L.SRC_PHILOX if need_draw else L.SRC_DEVICE
        else:
            o_src = kinds.pop() if kinds else L.SRC_DEVICE
        c, h, w = self.obs_shape
        dev = self.device
        d = L.FusedDesc()
        d.struct_size = C.sizeof(L.FusedDesc)
        d.B, d.idx_src, d.off_src, d.idx, d.idx_hi = B, i_src, o_src, i_ptr, idx_hi
        d.off_obs = None if oo is None else oo[1]
        d.off_next = None if on is None else on[1]
        d.pad, d.normalize, d.seed, d.subsequence = int(pad), 1 if normalize else 0, augmentations._seed, sub
        with torch.cuda.device(dev):
            order = [k for k in ("obs", "obs_shift", "obs_aug", "next_obs", "next_obs_shift") if k in want]
            block = torch.empty((len(order), B, c, h, w), dtype=torch.float32, device=dev)
            out = dict(zip(order, block.unbind(0)))
            ptr = {k: v.data_ptr() for k, v in out.items()}
            d.out_obs, d.out_obs_shift, d.out_obs_aug = ptr.get("obs"), ptr.get("obs_shift"), ptr.get("obs_aug")
            d.out_next, d.out_next_shift = ptr.get("next_obs"), ptr.get("next_obs_shift")
            act = torch.empty((B,) + self.actions.shape[1:], dtype=torch.float32, device=dev)
            rew, nd = torch.empty((2, B, 1), dtype=torch.float32, device=dev).unbind(0)
            d.out_act, d.out_rew, d.out_nd = act.data_ptr(), rew.data_ptr(), nd.data_ptr()
            keep = [keep_i, oo, on]
            if "obs_aug" in want:
                ops = 0
                for k in keys:
                    ops |= self._FUSED_OPS[k]
                d.aug_ops = ops
                d.aug_input = L.AUG_INPUT_SHIFTED if aug_input == "shift" else L.AUG_INPUT_PLAIN
                pre = next((k for k in keys if k != "conv"), None)
                if pre is not None:
                    prm = aug_params.get(pre, {})
                    augmentations.load_dataloader(batch_size=B, image_size=h, dataset=prm.get("dataset", "coco"), device=dev)
                    bank = augmentations.get_background_bank()
                    if bank.dtype != torch.uint8 or bank.device != dev or tuple(bank.shape[2:]) != (h, w):
                        raise NotImplementedError("sample_fused: needs a uint8 background bank of the observation size on the ring's device")
                    d.bank, d.bank_len = bank.data_ptr(), bank.shape[0]
                    if bank_idx is None:      # drawn inside the kernel: the values augcl_rng_fill_indices would write
                        d.bank_idx, d.bank_subsequence = None, augmentations._next_subsequence()
                        keep.append(bank)
                    else:
                        rows = augmentations._bank_rows(B, bank.shape[0], bank_idx, dev)
                        d.bank_idx = rows.data_ptr()
                        keep += [bank, rows]
                    if pre == "overlay":
                        d.alpha = float(prm.get("alpha", 0.5))
                    else:
                        dflt = (0, 0, 0.6) if pre == "splice" else (3.5, 0, 0)
                        d.hue_thres = float(prm.get("hue_thres", dflt[0]))
                        d.sat_thres = float(prm.get("sat_thres", dflt[1]))
                        d.val_thres = float(prm.get("val_thres", dflt[2]))
                if "conv" in keys:
                    if weights is None:
                        # drawn inside the kernel: the values augcl_rng_fill_normal would write for this subsequence
                        d.conv_weights = None
                        d.conv_subsequence = augmentations._next_subsequence()
                    else:
                        wt = torch.as_tensor(weights).to(device=dev, dtype=torch.float32).contiguous()
                        if tuple(wt.shape) != (B, 3, 3, 3, 3):
                            raise ValueError(f"sample_fused: weights must be ({B},3,3,3,3), got {tuple(wt.shape)}")
                        d.conv_weights = wt.data_ptr()
                        keep.append(wt)
            rc = L.lib().augcl_ring_sample_fused(self._ring, C.byref(d), _raw_stream(self._dev_index))
            if rc != 0:
                e = L.lib().augcl_last_error().decode(errors="replace")
                if rc == -5:
                    raise NotImplementedError(e)
                raise L.AugclError("augcl_ring_sample_fused", rc, e)
        del keep        # made on the launching stream: the stream-ordered allocator keeps them valid for the kernel
        out.update(action=act, reward=rew, not_done=nd)
        return out

    def sample_augcl(self, n=None, *, data_aug="splice-conv", aug_params=None, pad=4, **kw):
        """AugCL.update's data (src/algorithms/augcl.py:69-72) from one kernel:
        (obs_shift, obs_aug, action, reward, next_obs, not_done), obs_aug = apply_aug(un-shifted obs)."""
        o = self.sample_fused(n, data_aug=data_aug, aug_params=aug_params, pad=pad, aug_input="plain",
                              want=("obs_shift", "obs_aug", "next_obs"), **kw)
        return o["obs_shift"], o["obs_aug"], o["action"], o["reward"], o["next_obs"], o["not_done"]

    def sample_svea_aug(self, n=None, *, data_aug="overlay", aug_params=None, pad=4, **kw):
        """SVEA.update's data (src/algorithms/svea.py:66-70, :45-52) from one kernel:
        (obs, obs_aug, action, reward, next_obs, not_done) with obs = random_shift(obs), obs_aug = apply_aug(obs)."""
        o = self.sample_fused(n, data_aug=data_aug, aug_params=aug_params, pad=pad, aug_input="shift",
                              want=("obs_shift", "obs_aug", "next_obs"), **kw)
        return o["obs_shift"], o["obs_aug"], o["action"], o["reward"], o["next_obs"], o["not_done"]

    # ------------------------------------------------------------------ persistence (reference chunk format)
    IO_BLOCK = 2048   # transitions moved between the ring and pinned host memory per copy

    def _drain(self, start, end):
        """Ring slots [start, end) -> (obs, next) uint8 host arrays, in IO_BLOCK pieces."""
        c, h, w = self.obs_shape
        n = end - start
        obs = np.empty((n, c, h, w), dtype=np.uint8)
        nxt = np.empty((n, c, h, w), dtype=np.uint8)
        if self.dedup:      # no stacks in HBM to copy out: gather them with the sample kernel (exact u8 -> f32 -> u8)
            for s0 in range(0, n, 256):
                m = min(256, n - s0)
                o, _, _, no, _ = self._sample(np.arange(start + s0, start + s0 + m), scalars=False, rng=False)
                obs[s0:s0 + m] = o.to(torch.uint8).cpu().numpy()
                nxt[s0:s0 + m] = no.to(torch.uint8).cpu().numpy()
            return obs, nxt
        blk = min(self.IO_BLOCK, max(n, 1))
        po = torch.empty((blk, c, h, w), dtype=torch.uint8).pin_memory()
        pn = torch.empty((blk, c, h, w), dtype=torch.uint8).pin_memory()
        stream = torch.cuda.current_stream(self.device)
        for s0 in range(0, n, blk):
            m = min(blk, n - s0)
            L.call("augcl_ring_get", self._ring, start + s0, m, po.data_ptr(), pn.data_ptr(), None, None, None,
                   self._stream())
            stream.synchronize()
            obs[s0:s0 + m] = po.numpy()[:m]
            nxt[s0:s0 + m] = pn.numpy()[:m]
        return obs, nxt

    def __save__(self, save_dir, last_save_idx):
        """src/utils.py:156-167: chunk "<last>_<idx>.tar" with the transitions added since the last save."""
        if self.idx == last_save_idx:
            return self.idx
        self._flush()
        end = max(self.idx, last_save_idx)   # the reference slices [last:idx]: empty once the ring has wrapped
        obs, nxt = self._drain(last_save_idx, end)
        pairs = persist.encode_obses(obs, nxt, self._frame_channels, LazyFrames)
        path = os.path.join(save_dir, persist.chunk_name(last_save_idx, self.idx))
        persist.save_chunk(path, pairs, self.actions[last_save_idx:self.idx], self.rewards[last_save_idx:self.idx],
                           self.not_dones[last_save_idx:self.idx])
        return self.idx

    def save(self, save_dir):
        """src/utils.py:169-170."""
        self.last_save = self.__save__(save_dir=save_dir, last_save_idx=self.last_save)

    def requeue_save(self, save_dir):
        """src/utils.py:172-175."""
        self.last_requeue_save = self.__save__(save_dir=save_dir, last_save_idx=self.last_requeue_save)

    def load(self, save_dir, end_step=None, is_requeue_load=True, is_load=True):
        """src/utils.py:177-204: replay the chunks of save_dir, in order, into the ring."""
        if self.refilled:
            return
        self._flush()
        c, h, w = self.obs_shape
        chunks = persist.list_chunks(save_dir)
        stream = torch.cuda.current_stream(self.device)
        for chunk in chunks:
            start, end = persist.parse_chunk_name(chunk)
            payload = persist.load_chunk(os.path.join(save_dir, chunk), LazyFrames)
            assert self.idx == start, (     # same condition and exception type as src/utils.py:188-190
                f"ReplayBuffer.load: chunk {chunk} of {save_dir} starts at slot {start} but the write cursor is at {self.idx} "
                f"(chunks must be contiguous and loaded in order; found {chunks})")
            pairs = payload[0]
            n = len(pairs)
            if start + n > self.capacity:
                raise ValueError(f"chunk {chunk} holds {n} transitions from slot {start}: beyond capacity {self.capacity}")
            self.actions[start:end] = payload[1]
            self.rewards[start:end] = payload[2]
            self.not_dones[start:end] = payload[3]
            if self.dedup:
                # the chunk's LazyFrames share frame objects exactly as they did when saved (pickle keeps identity):
                # replay them through the ingest path, which de-duplicates by identity
                for i, (o, no) in enumerate(pairs):
                    if self._st_count == 0:
                        if self._st_event is not None:
                            self._st_event.synchronize()
                            self._st_event = None
                        self._st_start = start + i
                    self._stage_frames(self._st_count, o, no)
                    self._st_count += 1
                    if self._st_count == self._st_cap:
                        self._flush()
                self._flush()
                self.idx = end
                if end == end_step:
                    break
                continue
            blk = min(self.IO_BLOCK, max(n, 1))
            po = torch.empty((blk, c, h, w), dtype=torch.uint8).pin_memory()
            pn = torch.empty((blk, c, h, w), dtype=torch.uint8).pin_memory()
            for s0 in range(0, n, blk):
                m = min(blk, n - s0)
                persist.decode_obses(pairs[s0:s0 + m], po.numpy(), pn.numpy())
                a = np.ascontiguousarray(self.actions[start + s0:start + s0 + m], dtype=np.float32).reshape(m, -1)
                r = np.ascontiguousarray(self.rewards[start + s0:start + s0 + m], dtype=np.float32)
                d = np.ascontiguousarray(self.not_dones[start + s0:start + s0 + m], dtype=np.float32)
                L.call("augcl_ring_put", self._ring, start + s0, m, po.data_ptr(), pn.data_ptr(), a.ctypes.data,
                       r.ctypes.data, d.ctypes.data, self._stream())
                stream.synchronize()       # the pinned staging is reused by the next block
            self.idx = end
            if end == end_step:
                break
        if is_requeue_load:
            self.last_requeue_save = self.idx
        if is_load:
            self.last_save = self.idx

    def fill_synthetic(self, count=None, seed=0):
        """Benchmarks/tests: fill slots [0,count) with Philox frames and scalars on the device."""
        count = self.capacity if count is None else int(count)
        self._flush()
        if self.dedup:
            if count + self._k > self._pool:
                raise ValueError("fill_synthetic: the frame pool is smaller than count + k")
            self._sealed = True        # the pool now holds a synthetic sliding window the host bookkeeping does not track
        L.call("augcl_ring_fill_synthetic", self._ring, count, int(seed), self._stream())
        n = count
        a = np.empty((n, self.action_dim), dtype=np.float32)
        L.call("augcl_ring_get", self._ring, 0, n, None, None, a.ctypes.data, self.rewards.ctypes.data,
               self.not_dones.ctypes.data, self._stream())
        torch.cuda.current_stream(self.device).synchronize()
        self.actions[:n] = a.reshape((n,) + self.actions.shape[1:])
        self.idx = count % self.capacity
        self.full = self.full or count == self.capacity

    def ring_bytes(self):
        return int(L.lib().augcl_ring_bytes(self._ring))


class GraphSampler(object):
    """A captured ``sample_drq`` / ``sample_svea`` / ``sample`` step for training loops that want the smallest host cost
    per update: the host draws (the reference's ``np.random.randint`` indices, src/utils.py:229-232, and the shift offsets)
    are written into ONE pinned staging block, copied to the device with one async H2D and consumed by a CUDA graph that
    holds the H2D copy and the sample kernel (and, with ``host_reward=True``, a D2H copy of the step's rewards into the
    pinned ``reward_host`` tensor, readable after a stream synchronize): one graph launch per step.  Outputs are the
    graph's own tensors: valid until the next ``sample`` call.

        gs = GraphSampler(buf, kind="drq", pad=4)            # once
        obs, action, reward, next_obs, not_done = gs.sample(idxs, offsets)   # every update
    """

    def __init__(self, buf, kind="drq", pad=4, batch=None, host_reward=False):
        if kind not in ("drq", "svea", "plain"):
            raise ValueError("GraphSampler: kind must be 'drq', 'svea' or 'plain'")
        self.buf, self.kind, self.pad = buf, kind, int(pad)
        B = self.B = int(batch or buf.batch_size)
        dev = buf.device
        # staging layout (bytes): [idx int64 B][off_obs int32 2B][off_next int32 2B]
        self._pin = torch.empty(24 * B, dtype=torch.uint8).pin_memory()
        self._dev = torch.empty(24 * B, dtype=torch.uint8, device=dev)
        hp = self._pin.numpy()
        self._h_idx = hp[:8 * B].view(np.int64)
        self._h_off = hp[8 * B:].view(np.int32).reshape(2, B, 2)
        self._h_idx_u, self._h_off_u = self._h_idx.view(np.uint64), self._h_off.view(np.uint32)
        self._h_off[:] = self.pad
        self._h_idx[:] = 0
        d_idx = self._dev[:8 * B].view(torch.int64)
        d_off = self._dev[8 * B:].view(torch.int32).view(2, B, 2)
        self._dev.copy_(self._pin)

        def step():
            if kind == "drq":
                return buf.sample_drq(pad=self.pad, idxs=d_idx, offsets={"obs": d_off[0], "next_obs": d_off[1]})
            if kind == "svea":
                return buf.sample_svea(pad=self.pad, idxs=d_idx, offsets=d_off[0])
            return buf.sample(idxs=d_idx)

        buf._flush()
        cur = torch.cuda.current_stream(dev)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            step()                                  # warm-up outside the capture
        cur.wait_stream(side)
        self.reward_host = torch.empty((B, 1), dtype=torch.float32).pin_memory() if host_reward else None
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._dev.copy_(self._pin, non_blocking=True)
            self.out = step()
            if host_reward:
                self.reward_host.copy_(self.out[2], non_blocking=True)
        self._done = torch.cuda.Event()              # the previous replay has read the staging block

    def sample(self, idxs, offsets=None, validate=True):
        """idxs: (B,) host integers; offsets: {"obs": (B,2), "next_obs": (B,2)} host integers in [0, 2*pad] (drq), a (B,2)
        array (svea) or None.  Returns (obs, action, reward, next_obs, not_done)."""
        buf = self.buf
        if buf._st_count:
            buf._flush()
        self._done.synchronize()
        self._h_idx[:] = idxs
        if validate and int(self._h_idx_u.max()) >= buf.capacity:      # unsigned view: one reduction catches negatives too
            raise IndexError(f"GraphSampler: index outside [0, {buf.capacity})")
        if offsets is not None:
            if self.kind == "drq":
                self._h_off[0] = offsets["obs"]
                self._h_off[1] = offsets["next_obs"]
            elif self.kind == "svea":
                self._h_off[0] = offsets
            if validate and int(self._h_off_u.max()) > 2 * self.pad:
                raise ValueError(f"GraphSampler: offset outside [0, {2 * self.pad}]")
        self.graph.replay()
        self._done.record()
        return self.out


class FastSampler(object):
    """The leanest host path for ``sample`` / ``sample_drq`` / ``sample_svea`` at batch <= 256: one ctypes call per step
    and nothing else.  The host draws (the reference's ``np.random.randint`` indices, src/utils.py:229-232, and the shift
    offsets) travel BY VALUE in the kernel's parameter block (augcl_ring_sample_ex, HOST source): no H2D copy, no staging
    block that a later call could overwrite, no event.  Outputs are fixed device tensors, valid until the next ``sample``
    call; with ``host_reward=True`` the kernel writes the step's rewards straight into the pinned, device-mapped
    ``reward_host`` tensor (readable after a stream synchronize) instead of a device tensor followed by a D2H copy.

        fs = FastSampler(buf, kind="drq", pad=4)             # once
        obs, action, reward, next_obs, not_done = fs.sample(idxs, offsets)   # every update
    """

    def __init__(self, buf, kind="drq", pad=4, batch=None, host_reward=False):
        if kind not in ("drq", "svea", "plain"):
            raise ValueError("FastSampler: kind must be 'drq', 'svea' or 'plain'")
        B = self.B = int(batch or buf.batch_size)
        if B > 256:
            raise ValueError("FastSampler: batch > 256 does not fit the kernel's parameter block; use GraphSampler")
        self.buf, self.kind, self.pad = buf, kind, int(pad)
        dev = buf.device
        c, h, w = buf.obs_shape
        self._idx = np.zeros(B, dtype=np.int64)
        self._off = np.full((2, B, 2), self.pad, dtype=np.int32)
        obs, nxt = torch.empty((2, B, c, h, w), dtype=torch.float32, device=dev).unbind(0)
        act = torch.empty((B,) + buf.actions.shape[1:], dtype=torch.float32, device=dev)
        nd = torch.empty((B, 1), dtype=torch.float32, device=dev)
        if host_reward:
            rew = self.reward_host = torch.zeros((B, 1), dtype=torch.float32).pin_memory()
        else:
            rew, self.reward_host = torch.empty((B, 1), dtype=torch.float32, device=dev), None
        self.out = (obs, act, rew, nxt, nd)
        aug = {"drq": L.AUG_SHIFT, "svea": L.AUG_SHIFT_OBS, "plain": L.AUG_NONE}[kind]
        self._args = (buf._ring, L.SRC_HOST, self._idx.ctypes.data, 0, B, obs.data_ptr(), nxt.data_ptr(), act.data_ptr(),
                      rew.data_ptr(), nd.data_ptr(), aug, L.SRC_HOST, self._off[0].ctypes.data, self._off[1].ctypes.data,
                      self.pad if kind != "plain" else 0, 0, 0, 0)
        self._fn = buf._fn_sample
        self._dev_index = buf._dev_index

    def sample(self, idxs, offsets=None):
        """idxs: (B,) host integers; offsets: {"obs": (B,2), "next_obs": (B,2)} host integers in [0, 2*pad] (drq), a (B,2)
        array (svea) or None (keep the previous ones).  The library validates both (AugclError, like the drop-in call)."""
        buf = self.buf
        if buf._st_count:
            buf._flush()
        self._idx[:] = idxs
        if offsets is not None:
            if self.kind == "drq":
                self._off[0] = offsets["obs"]
                self._off[1] = offsets["next_obs"]
            elif self.kind == "svea":
                self._off[0] = offsets
        rc = self._fn(*self._args, _raw_stream(self._dev_index))
        if rc != 0:
            raise L.AugclError("augcl_ring_sample_ex", rc, L.lib().augcl_last_error().decode(errors="replace"))
        return self.out


class FastFusedSampler(object):
    """``sample_augcl`` / ``sample_svea_aug`` with the per-call Python work taken out (batch <= 256): fixed output tensors,
    a cached ``augcl_fused_desc``, the host indices by value in the launch.  Per step: the bank-row and conv-weight draws
    (both inside the kernel; the same Philox subsequences ``sample_fused`` uses) and ONE fused sample kernel -- one launch
    per update.

        fs = FastFusedSampler(buf, kind="augcl", data_aug="splice-conv")          # once
        obs_shift, obs_aug, action, reward, next_obs, not_done = fs.sample()      # every update (src/algorithms/augcl.py:69-72)

    kind="svea_aug" augments the SHIFTED obs (src/algorithms/svea.py:67 + :46).  Outputs are valid until the next call."""

    def __init__(self, buf, kind="augcl", data_aug=None, aug_params=None, pad=4, batch=None):
        if kind not in ("augcl", "svea_aug"):
            raise ValueError("FastFusedSampler: kind must be 'augcl' or 'svea_aug'")
        data_aug = data_aug if data_aug is not None else ("splice-conv" if kind == "augcl" else "overlay")
        keys = tuple(k for k in str(data_aug).split("-") if k and k != "identity")
        if not keys or keys not in ReplayBuffer._FUSED_CHAINS:
            raise NotImplementedError(f"FastFusedSampler: chain {data_aug!r} is not fused")
        B = self.B = int(batch or buf.batch_size)
        if B > 256:
            raise ValueError("FastFusedSampler: batch > 256 does not fit the kernel's parameter block")
        self.buf, self.kind, self.pad = buf, kind, int(pad)
        aug_params = aug_params or {}
        dev = buf.device
        c, h, w = buf.obs_shape
        self._idx = np.zeros(B, dtype=np.int64)
        self._off = np.full((2, B, 2), self.pad, dtype=np.int32)
        obs_shift, obs_aug, nxt = torch.empty((3, B, c, h, w), dtype=torch.float32, device=dev).unbind(0)
        act = torch.empty((B,) + buf.actions.shape[1:], dtype=torch.float32, device=dev)
        rew, nd = torch.empty((2, B, 1), dtype=torch.float32, device=dev).unbind(0)
        self.out = (obs_shift, obs_aug, act, rew, nxt, nd)
        d = self._d = L.FusedDesc()
        d.struct_size = C.sizeof(L.FusedDesc)
        d.B, d.idx_src, d.idx, d.idx_hi = B, L.SRC_HOST, self._idx.ctypes.data, 0
        d.pad, d.normalize, d.seed = self.pad, 0, augmentations._seed
        d.out_obs_shift, d.out_obs_aug, d.out_next = obs_shift.data_ptr(), obs_aug.data_ptr(), nxt.data_ptr()
        d.out_act, d.out_rew, d.out_nd = act.data_ptr(), rew.data_ptr(), nd.data_ptr()
        ops = 0
        for k in keys:
            ops |= ReplayBuffer._FUSED_OPS[k]
        d.aug_ops = ops
        d.aug_input = L.AUG_INPUT_SHIFTED if kind == "svea_aug" else L.AUG_INPUT_PLAIN
        self._rows = None
        pre = next((k for k in keys if k != "conv"), None)
        if pre is not None:
            prm = aug_params.get(pre, {})
            augmentations.load_dataloader(batch_size=B, image_size=h, dataset=prm.get("dataset", "coco"), device=dev)
            bank = self._bank = augmentations.get_background_bank()
            if bank.dtype != torch.uint8 or bank.device != dev or tuple(bank.shape[2:]) != (h, w):
                raise NotImplementedError("FastFusedSampler: needs a uint8 background bank of the observation size on the ring's device")
            self._rows = True                # bank rows are drawn inside the kernel (desc.bank_idx = NULL)
            d.bank, d.bank_len, d.bank_idx = bank.data_ptr(), bank.shape[0], None
            if pre == "overlay":
                d.alpha = float(prm.get("alpha", 0.5))
            else:
                dflt = (0, 0, 0.6) if pre == "splice" else (3.5, 0, 0)
                d.hue_thres = float(prm.get("hue_thres", dflt[0]))
                d.sat_thres = float(prm.get("sat_thres", dflt[1]))
                d.val_thres = float(prm.get("val_thres", dflt[2]))
        self._conv = "conv" in keys   # weights are drawn inside the kernel (conv_weights NULL)
        lib = L.lib()
        self._fill_idx, self._fused = lib.augcl_rng_fill_indices, lib.augcl_ring_sample_fused
        self._dref = C.byref(d)
        self._dev_index = buf._dev_index

    def sample(self, idxs=None, offsets=None):
        """idxs: (B,) host integers (default: the reference's np.random.randint draw, src/utils.py:229-232); offsets:
        {"obs": (B,2), "next_obs": (B,2)} host shift draws, default drawn in the kernel (Philox)."""
        buf, d = self.buf, self._d
        if buf._st_count:
            buf._flush()
        self._idx[:] = buf._get_idxs(self.B) if idxs is None else idxs
        st = _raw_stream(self._dev_index)
        d.seed = augmentations._seed
        d.subsequence = augmentations._reserve_subsequences(3)
        if offsets is None:
            d.off_src, d.off_obs, d.off_next = L.SRC_PHILOX, None, None
        else:
            self._off[0] = offsets["obs"]
            d.off_src, d.off_obs, d.off_next = L.SRC_HOST, self._off[0].ctypes.data, None
        rc = 0
        if self._rows is not None:
            d.bank_subsequence = augmentations._next_subsequence()
        if self._conv:
            d.conv_subsequence = augmentations._next_subsequence()
        rc = self._fused(buf._ring, self._dref, st)
        if rc != 0:
            raise L.AugclError("augcl_ring_sample_fused", rc, L.lib().augcl_last_error().decode(errors="replace"))
        return self.out
