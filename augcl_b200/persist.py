"""Replay-buffer persistence in the reference's chunk format (host side; no CUDA needed here).

The reference writes incremental chunks ``"<start>_<end>.tar"`` =
``torch.save([_obses[start:end], actions[start:end], rewards[start:end], not_dones[start:end]])``
(/root/reference/src/utils.py:156-170) where ``_obses`` entries are ``(LazyFrames, LazyFrames)`` tuples whose
3xHxW uint8 frame arrays are SHARED between consecutive transitions (FrameStack's deque,
src/env/wrappers.py:183-206), so pickle stores roughly one frame per environment step; ``load`` replays the
chunks in order (src/utils.py:177-204).  AugCL's second phase starts from such a directory
(src/train.py:156-185).  This module converts between that payload and the dense uint8 stacks the device ring
holds, keeping the frame sharing so the files stay the size the reference writes.
"""
from __future__ import annotations

import contextlib
import io
import os
import pickle
import sys
import types

import numpy as np
import torch


def chunk_name(start: int, end: int) -> str:
    return "%d_%d.tar" % (start, end)


def parse_chunk_name(name: str):
    """'<start>_<end>.tar' -> (start, end)   (src/utils.py:185)."""
    start, end = [int(x) for x in name.split(".")[0].split("_")]
    return start, end


def list_chunks(save_dir: str):
    """Chunk file names ordered by start index (src/utils.py:181-182)."""
    return sorted(os.listdir(save_dir), key=lambda x: int(x.split("_")[0]))


def encode_obses(obs_u8: np.ndarray, next_u8: np.ndarray, frame_channels: int, lazy_frames_cls):
    """Dense stacks (n, k*fc, H, W) x2 -> the reference's list of (LazyFrames, LazyFrames).

    Frames that are byte-identical to the frame one position later in the previous stack (the sliding
    FrameStack window), or to their left neighbour (episode reset repeats the first frame k times,
    src/env/wrappers.py:193-197), are stored as the SAME ndarray object, so pickle memoises them."""
    n, c = obs_u8.shape[:2]
    k = c // frame_channels
    out = []
    prev_next = None
    for t in range(n):
        o = [obs_u8[t, j * frame_channels:(j + 1) * frame_channels] for j in range(k)]
        x = [next_u8[t, j * frame_channels:(j + 1) * frame_channels] for j in range(k)]
        for j in range(k):
            if prev_next is not None and np.array_equal(o[j], prev_next[j]):
                o[j] = prev_next[j]                       # obs[t] == next_obs[t-1] inside an episode
            elif j > 0 and np.array_equal(o[j], o[j - 1]):
                o[j] = o[j - 1]
            else:
                o[j] = np.ascontiguousarray(o[j]).copy()  # own the memory: the staging buffer is reused
        for j in range(k):
            if j + 1 < k and np.array_equal(x[j], o[j + 1]):
                x[j] = o[j + 1]                           # sliding window
            elif j > 0 and np.array_equal(x[j], x[j - 1]):
                x[j] = x[j - 1]
            elif np.array_equal(x[j], o[j]):
                x[j] = o[j]
            else:
                x[j] = np.ascontiguousarray(x[j]).copy()
        out.append((lazy_frames_cls(o), lazy_frames_cls(x)))
        prev_next = x
    return out


def decode_obses(pairs, obs_out: np.ndarray, next_out: np.ndarray) -> None:
    """The reference's list of (LazyFrames, LazyFrames) -> dense stacks written into obs_out / next_out
    (n, C, H, W) uint8 (what LazyFrames.__array__ yields, src/utils.py:293-305)."""
    for t, (o, x) in enumerate(pairs):
        obs_out[t] = np.asarray(o)
        next_out[t] = np.asarray(x)


class _RefUnpickler(pickle.Unpickler):
    """Chunks written by the reference name its classes as ``utils.LazyFrames``; resolve them to ours
    whichever module is installed under the name ``utils`` in this process."""

    lazy_frames_cls = None

    def find_class(self, module, name):
        if name == "LazyFrames" and module in ("utils", "augcl_b200.utils"):
            return type(self).lazy_frames_cls
        return super().find_class(module, name)


class _PickleModule:
    """The slice of the ``pickle`` module interface torch.load(pickle_module=...) uses."""
    __name__ = "augcl_b200.persist"
    Unpickler = _RefUnpickler
    load = staticmethod(lambda f, **kw: _RefUnpickler(f, **kw).load())
    loads = staticmethod(lambda b, **kw: _RefUnpickler(io.BytesIO(b), **kw).load())
    dump, dumps, Pickler = pickle.dump, pickle.dumps, pickle.Pickler
    PickleError, UnpicklingError, PicklingError = pickle.PickleError, pickle.UnpicklingError, pickle.PicklingError
    HIGHEST_PROTOCOL, DEFAULT_PROTOCOL = pickle.HIGHEST_PROTOCOL, pickle.DEFAULT_PROTOCOL


def load_chunk(path: str, lazy_frames_cls):
    """torch.load of one chunk (pickled numpy + LazyFrames: weights_only must be off, as under torch 1.7)."""
    _RefUnpickler.lazy_frames_cls = lazy_frames_cls
    return torch.load(path, map_location="cpu", pickle_module=_PickleModule, weights_only=False)


@contextlib.contextmanager
def _utils_alias(lazy_frames_cls):
    """The reference unpickles chunks as ``utils.LazyFrames`` (its class lives in src/utils.py), so that is the
    name written; pickle insists the name resolves to the class being saved, whatever ``utils`` module this
    process has (none, our drop-in shim, or something unrelated): alias it for the duration of the save."""
    saved = sys.modules.get("utils")
    stub = types.ModuleType("utils")
    stub.LazyFrames = lazy_frames_cls
    sys.modules["utils"] = stub
    try:
        yield
    finally:
        if saved is None:
            sys.modules.pop("utils", None)
        else:
            sys.modules["utils"] = saved


def save_chunk(path: str, pairs, actions, rewards, not_dones) -> None:
    cls = type(pairs[0][0]) if len(pairs) else None
    if cls is not None and cls.__module__ == "utils":
        with _utils_alias(cls):
            torch.save([pairs, actions, rewards, not_dones], path)
    else:
        torch.save([pairs, actions, rewards, not_dones], path)
