"""Rollout store and batcher on pinned / device memory.

Drop-ins for reference ``libs/utils.py:29-55`` (``create_buffers``) and ``:166-218``
(``get_batch``); the slot-write contract of ``microbeast.py:73-100`` is unchanged
(``buffers[key][index][t, ...] = value``).

What changes underneath:

* every key lives in ONE slab ``[n_buffers, T+1, n_envs, ...]`` in shared memory
  (actor processes keep writing into it) that the learner process page-locks with
  ``cudaHostRegister``, so host->device copies are asynchronous DMA;
* wire format: ``obs`` travels as uint8 (microRTS planes are 0/1; 4x fewer bytes than the
  reference's float32) and is widened on the device; ``action_mask`` stays uint8,
  ``action`` int64 (the dtypes of the C ABI);
* keys the learner never reads (``policy_logits``: 80 KB/frame at 16x16, ``last_action``,
  ``ep_return``, ``ep_step``) are backed by a write-discarding slot when ``lean=True``,
  so an unmodified ``act()`` still runs but nothing is stored or copied;
* ``get_batch`` blocks on the queue instead of spinning on ``qsize()``
  (``libs/utils.py:183-188`` burns a core), copies each buffer straight into its column of a
  device-contiguous ``[T+1, B, n_envs, ...]`` staging tensor on a copy stream (no host
  ``torch.stack``), and recycles the indices only after the DMA has finished.

The flattening order is the reference's ``(t, b, e)`` (``libs/utils.py:197-205``), so
``PPO_learn``'s reshape semantics (quirk Q7) are preserved.
"""
from __future__ import annotations

import queue
import threading
import weakref
from typing import Dict, List, Optional

import torch

Buffers = Dict[str, List[torch.Tensor]]

LEARNER_KEYS = ("obs", "reward", "done", "logprobs", "action", "action_mask", "baseline")
DEAD_KEYS = ("ep_return", "ep_step", "policy_logits", "last_action")


def buffer_specs(n_envs: int, T: int, obs_size: tuple, obs_dtype=torch.float32) -> Dict[str, dict]:
    """Shapes / dtypes of one rollout buffer -- reference ``libs/utils.py:33-46``."""
    h = w = obs_size[0]
    return dict(
        obs=dict(size=(T + 1, n_envs, *obs_size), dtype=obs_dtype),
        reward=dict(size=(T + 1, n_envs), dtype=torch.float32),
        done=dict(size=(T + 1, n_envs), dtype=torch.bool),
        ep_return=dict(size=(T + 1, n_envs), dtype=torch.float32),
        ep_step=dict(size=(T + 1, n_envs), dtype=torch.int32),
        policy_logits=dict(size=(T + 1, n_envs, h * w * 78), dtype=torch.float32),
        baseline=dict(size=(T + 1, n_envs), dtype=torch.float32),
        last_action=dict(size=(T + 1, n_envs, h * w * 7), dtype=torch.int64),
        action=dict(size=(T + 1, n_envs, h * w * 7), dtype=torch.int64),
        action_mask=dict(size=(T + 1, n_envs, 78 * h * w), dtype=torch.uint8),
        logprobs=dict(size=(T + 1, n_envs), dtype=torch.float32),
    )


class DiscardSlot:
    """Stands in for a buffer nobody reads: accepts the actor's slot writes and drops them."""

    def __init__(self, size, dtype):
        self.shape = torch.Size(size)
        self.dtype = dtype

    def size(self):
        return self.shape

    def __setitem__(self, index, value):
        pass

    def __getitem__(self, index):
        return self

    def share_memory_(self):
        return self


def pack_mask(mask: torch.Tensor) -> torch.Tensor:
    """``[..., W]`` u8/bool mask -> ``[..., ceil(W/8)]`` u8, bit i of byte j = entry 8j+i (host side)."""
    import numpy as np
    arr = (mask.cpu().numpy() != 0)
    return torch.from_numpy(np.packbits(arr, axis=-1, bitorder="little"))


def pack_obs(obs: torch.Tensor) -> torch.Tensor:
    """``[..., P]`` 0/1 feature planes (P <= 32) -> ``[...]`` int32 with bit c = plane c (host side)."""
    import numpy as np
    arr = (obs.cpu().numpy() != 0)
    weights = (np.uint32(1) << np.arange(arr.shape[-1], dtype=np.uint32))
    return torch.from_numpy((arr * weights).sum(axis=-1, dtype=np.uint32).astype(np.int32))


def unpack_obs_host(bits: torch.Tensor, planes: int = 27) -> torch.Tensor:
    """Inverse of :func:`pack_obs` on the host (uint8 planes); the device path is ``ops.unpack_obs``."""
    import numpy as np
    b = bits.numpy().astype(np.uint32)[..., None]
    return torch.from_numpy(((b >> np.arange(planes, dtype=np.uint32)) & 1).astype(np.uint8))


def pack_batch(batch: dict) -> dict:
    """Reference-format host batch -> the compact wire format ``PPO_learn`` also accepts:
    ``action_mask`` becomes ``action_mask_bits`` (8x smaller) and ``action`` uint8 (8x smaller;
    indices are < 49).  ``obs`` becomes one int32 per cell (the 27 planes are 0/1: 27x smaller than
    the reference's float32)."""
    out = {k: v for k, v in batch.items() if k not in ("action_mask", "action", "obs")}
    out["action_mask_bits"] = pack_mask(batch["action_mask"])
    out["action"] = batch["action"].to(torch.uint8)
    out["obs"] = pack_obs(batch["obs"])      # one int32 per cell; widened to NCHW fp32 on the device
    return out


class PackedMaskSlot:
    """Rollout-buffer slot that keeps the action mask bit-packed: the actor's slot write
    ``buffers["action_mask"][i][t, ...] = mask`` (``microbeast.py:75,96``) packs on the way in."""

    def __init__(self, packed: torch.Tensor, width: int):
        self.packed = packed               # [T+1, n_envs, ceil(width/8)] u8 view of the slab
        self.width = width
        self.shape = torch.Size(tuple(packed.shape[:-1]) + (width,))
        self.dtype = torch.uint8

    def size(self):
        return self.shape

    def __setitem__(self, index, value):
        value = torch.as_tensor(value)
        self.packed[index] = pack_mask(value.reshape(tuple(value.shape[:-1]) + (self.width,))).reshape(
            self.packed[index].shape)

    def unpacked(self) -> torch.Tensor:
        import numpy as np
        bits = np.unpackbits(self.packed.numpy(), axis=-1, bitorder="little")[..., :self.width]
        return torch.from_numpy(bits)


class PackedObsSlot:
    """Rollout-buffer slot that keeps observations as one int32 per cell (bit c = plane c)."""

    def __init__(self, packed: torch.Tensor, planes: int):
        self.packed = packed               # [T+1, n_envs, h, w] int32 view of the slab
        self.planes = planes
        self.shape = torch.Size(tuple(packed.shape) + (planes,))
        self.dtype = torch.uint8

    def size(self):
        return self.shape

    def __setitem__(self, index, value):
        value = torch.as_tensor(value)
        self.packed[index] = pack_obs(value).reshape(self.packed[index].shape)

    def unpacked(self) -> torch.Tensor:
        return unpack_obs_host(self.packed, self.planes)


class RolloutBuffers(dict):
    """``Dict[str, List[Tensor]]`` like the reference's, plus the slabs behind the lists."""

    def __init__(self):
        super().__init__()
        self.slabs: Dict[str, torch.Tensor] = {}
        self.pinned = False
        self.n_buffers = 0
        self._lock = threading.Lock()
        self._staging = {}           # device -> [set0, set1] of {key: tensor}
        self._turn = {}
        self._pending = []           # (event, indices, free_queue) waiting for their DMA
        self._stream = {}
        self._registered = []
        self.compact = False
        self.cells = 0
        self._finalizer = None

    def live_keys(self):
        return [k for k, v in self.items() if isinstance(v[0], torch.Tensor)]

    def pin(self) -> bool:
        """Page-lock the shared slabs in THIS process (call once in the learner process)."""
        if self.pinned or not torch.cuda.is_available():
            return self.pinned
        rt = torch.cuda.cudart()
        for slab in self.slabs.values():
            err = int(rt.cudaHostRegister(slab.data_ptr(), slab.numel() * slab.element_size(), 0))
            if err == 712:          # cudaErrorHostMemoryAlreadyRegistered: a stale registration of a freed
                rt.cudaHostUnregister(slab.data_ptr())       # mapping at the same address; replace it
                err = int(rt.cudaHostRegister(slab.data_ptr(), slab.numel() * slab.element_size(), 0))
            if err != 0:
                raise RuntimeError(f"cudaHostRegister failed with error {err}")
            self._registered.append(slab.data_ptr())
        self.pinned = True
        self._finalizer = weakref.finalize(self, _unregister, list(self._registered))
        return True

    def unpin(self) -> None:
        """Undo :meth:`pin` (also runs when the buffers are garbage collected)."""
        if self.pinned:
            self._finalizer.detach()
            _unregister(self._registered)
            self._registered = []
            self.pinned = False


def _unregister(ptrs) -> None:
    try:
        rt = torch.cuda.cudart()
        torch.cuda.synchronize()
        for ptr in ptrs:
            rt.cudaHostUnregister(ptr)
    except Exception:
        pass


def create_buffers(n_buffers: int, n_envs: int, B: int, T: int, obs_size: tuple, *, lean: bool = True,
                   obs_dtype=torch.uint8, share: bool = True, compact: bool = False) -> Buffers:
    """Creates ``n_buffers`` rollout buffers shared between processes -- reference
    ``libs/utils.py:29-55``, same positional signature and the same 11 keys.

    ``lean=False, obs_dtype=torch.float32`` reproduces the reference layout exactly.
    ``compact=True`` additionally keeps ``action_mask`` bit-packed, ``action`` as uint8 and ``obs``
    as one int32 per cell in the slabs; ``get_batch`` widens masks and actions on the device
    (``mb_unpack_rollout``) so the learner still sees u8 masks and i64 actions, and hands the
    packed observations to ``Agent``, which expands them straight to NCHW fp32 (``mb_unpack_obs``).
    """
    specs = buffer_specs(n_envs, T, tuple(obs_size), obs_dtype)
    buffers = RolloutBuffers()
    buffers.n_buffers = n_buffers
    for key, spec in specs.items():
        if lean and key in DEAD_KEYS:
            buffers[key] = [DiscardSlot(spec["size"], spec["dtype"]) for _ in range(n_buffers)]
            continue
        size, dtype = tuple(spec["size"]), spec["dtype"]
        if compact and key == "action_mask":            # bits on the wire, widened on the device
            size = size[:-1] + ((size[-1] + 7) // 8,)
        if compact and key == "action":                 # indices < 49 fit a byte
            dtype = torch.uint8
        if compact and key == "obs":                    # 27 binary planes -> one int32 per cell
            size, dtype = size[:-1], torch.int32
        slab = torch.zeros((n_buffers,) + size, dtype=dtype)
        if share:
            slab.share_memory_()
        buffers.slabs[key] = slab
        if compact and key == "action_mask":
            buffers[key] = [PackedMaskSlot(slab[i], spec["size"][-1]) for i in range(n_buffers)]
        elif compact and key == "obs":
            buffers[key] = [PackedObsSlot(slab[i], spec["size"][-1]) for i in range(n_buffers)]
        else:
            buffers[key] = [slab[i] for i in range(n_buffers)]
    buffers.compact = compact
    buffers.cells = obs_size[0] * obs_size[1]
    return buffers


def _recycle(buffers: RolloutBuffers, block: bool) -> None:
    still = []
    for event, indices, free_queue in buffers._pending:
        if event is not None and not block and not event.query():
            still.append((event, indices, free_queue))
            continue
        if event is not None:
            event.synchronize()
        for m in indices:
            free_queue.put(m)                                    # utils.py:211-213
    buffers._pending = still


def get_batch(batch_size: int, shapes: dict, gamma: float, device, free_queue, full_queue, buffers: Buffers,
              lock=threading.Lock(), *, keys: Optional[tuple] = None) -> dict:
    """Returns ``batch_size`` rollouts as one flat batch -- reference ``libs/utils.py:166-218``.

    Output: ``{key: [ (T+1)*batch_size*n_envs, ... ]}`` in ``(t, b, e)`` order, on ``device``.
    ``shapes`` and ``gamma`` are accepted and ignored like in the reference.  With a CUDA
    ``device`` the copies run on a side stream and the current stream is made to wait for them;
    buffer indices return to ``free_queue`` once their DMA is done (at the latest on the next
    call).
    """
    if not isinstance(lock, type(threading.Lock())):
        lock = threading.Lock()
    dev = torch.device(device) if device is not None else torch.device("cpu")
    rb = buffers if isinstance(buffers, RolloutBuffers) else None
    with lock:                                                   # utils.py:191-192
        indices = []
        while len(indices) < batch_size:                         # blocking get; no qsize() spin
            try:
                indices.append(full_queue.get(timeout=0.05))
            except queue.Empty:
                if rb is not None and rb._pending:
                    _recycle(rb, block=True)     # actors may be waiting for buffers whose DMA is pending
    live = [k for k, v in buffers.items() if isinstance(v[0], (torch.Tensor, PackedMaskSlot, PackedObsSlot))]
    if keys is not None:
        live = [k for k in live if k in keys]

    if dev.type != "cuda":
        # host batch (reference semantics: stack dim=1, flatten) -- used by CPU tests and actors
        batch = {}
        for key in live:
            slots = [buffers[key][m] for m in indices]
            slots = [sl.unpacked() if isinstance(sl, (PackedMaskSlot, PackedObsSlot)) else sl for sl in slots]
            if key == "action" and slots[0].dtype == torch.uint8:
                slots = [sl.long() for sl in slots]
            x = torch.stack(slots, dim=1)                                         # utils.py:197-199
            batch[key] = x.reshape((-1,) + tuple(x.shape[3:]))                    # utils.py:203-205
        for m in indices:
            free_queue.put(m)
        return batch

    def raw(key, m):                 # the tensor actually stored in the slab for slot (key, m)
        sl = buffers[key][m]
        return sl.packed if isinstance(sl, (PackedMaskSlot, PackedObsSlot)) else sl

    def alloc():
        return {key: torch.empty((raw(key, 0).shape[0], batch_size) + tuple(raw(key, 0).shape[1:]),
                                 dtype=raw(key, 0).dtype, device=dev) for key in live}

    if rb is not None:
        rb.pin()
        _recycle(rb, block=False)
        sets = rb._staging.get((dev, batch_size))
        if sets is None:
            sets = [alloc(), alloc()]
            rb._staging[(dev, batch_size)] = sets
            rb._turn[(dev, batch_size)] = 0
            rb._stream[dev] = torch.cuda.Stream(device=dev)
        turn = rb._turn[(dev, batch_size)]
        rb._turn[(dev, batch_size)] = turn ^ 1
        stage = sets[turn]
        stream = rb._stream[dev]
    else:
        stage = alloc()
        stream = torch.cuda.Stream(device=dev)

    cur = torch.cuda.current_stream(dev)
    stream.wait_stream(cur)          # the staging set may still be read by the step before last
    with torch.cuda.stream(stream):
        for key in live:
            dst = stage[key]
            for j, m in enumerate(indices):
                dst[:, j].copy_(raw(key, m), non_blocking=True)
        batch = {key: stage[key].reshape((-1,) + tuple(stage[key].shape[3:])) for key in live}
        if rb is not None and rb.compact:
            # widen the wire format on the device: bits -> u8 mask, u8 -> i64 action
            from . import ops
            wide = stage.setdefault("_wide", {})
            n = batch["action"].shape[0]
            if "mask" not in wide:
                wide["mask"] = torch.empty((n, rb.cells * 78), dtype=torch.uint8, device=dev)
                wide["action"] = torch.empty((n, rb.cells * 7), dtype=torch.int64, device=dev)
            ops.unpack_rollout(batch["action_mask"], batch["action"], rb.cells, wide["mask"], wide["action"])
            batch["action_mask"], batch["action"] = wide["mask"], wide["action"]
        done = torch.cuda.Event()
        done.record(stream)
    cur.wait_event(done)
    if rb is not None:
        rb._pending.append((done, indices, free_queue))
    else:
        done.synchronize()
        for m in indices:
            free_queue.put(m)
    return batch


def release(buffers: Buffers) -> None:
    """Wait for outstanding DMA and hand every buffer index back (call before shutdown)."""
    if isinstance(buffers, RolloutBuffers):
        _recycle(buffers, block=True)
        buffers.unpin()
