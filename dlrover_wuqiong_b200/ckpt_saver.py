"""Shared-memory data plane and the agent-side asynchronous saver.

Mirror of the reference module dlrover/python/elastic_agent/torch/ckpt_saver.py
("SAVER" below): same class names, method names, shared-object names and meta
layout, so engines, checkpointers and third-party saver subclasses written
against the reference work unchanged.  What changed is *how* bytes move:

* ``SharedMemoryHandler.save_state_dict`` hands every CUDA tensor of the state
  dict to one pack-kernel launch and a pinned DMA into the shm segment
  (``transfer.DeviceShardEngine`` -> libfcp.so) instead of one pageable
  ``copy_`` per tensor (SAVER:174-206);
* ``SharedMemoryHandler.load_state_dict(into=...)`` scatters the shm image
  straight into live tensors, verifying the per-chunk CRC-32C recorded at save
  time; without ``into`` it returns the reference's zero-copy CPU views
  (SAVER:292-311), which is what the agent-side saver needs.

State dicts that live entirely in host memory (the reference's own CPU tests,
config 1 of BASELINE.json) never touch CUDA: their bytes are host-to-host
copies, exactly the reference's path.
"""
import importlib
import os
import pickle
import signal
import threading
import time
from abc import ABCMeta, abstractmethod
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass
from enum import Enum, auto
from pathlib import Path
from typing import Dict, List, Optional

import torch

from . import plan as _plan
from . import raw_ckpt
from .common import env_utils
from .common.constants import CheckpointConstant, NodeEnv
from .common.log import default_logger as logger
from .common.multi_process import SharedDict, SharedLock, SharedMemory, SharedQueue
from .common.serialize import ClassMeta
from .plan import TensorMeta  # noqa: F401  (re-exported, reference name)

DLROVER_CKPT_CONFIG_KEY = "_DLORVER_CKPT_CONFIG"
FCP_CHECKSUM_KEY = "_FCP_CHECKSUMS"


class CheckpointSharedObjPrefix:
    SAVE_STEP_QNAME = "checkpoint_lock_rank_"
    META_NAME = "checkpoint_meta_"
    SHM_NAME = "checkpoint_shm_"
    SHM_LOCK_NAME = "shm_lock_"


class CheckpointEventType(Enum):
    SAVE = auto()
    UPDATE_SHARD = auto()
    EXIT = auto()


@dataclass
class CheckpointEvent:
    type: CheckpointEventType = CheckpointEventType.SAVE
    step: int = 0
    global_shard_num: int = 0


@dataclass
class CheckpointConfig:
    """Per-shard bookkeeping stored in the meta dict under
    DLROVER_CKPT_CONFIG_KEY (SAVER:73-91).  ``writing_shm`` is True while the
    shm bytes are being replaced; ``paths`` maps state names to storage paths.
    """
    rank: int = 0
    group_rank: int = 0
    world_size: int = 0
    step: int = 0
    writing_shm: bool = False
    paths: Dict[str, str] = None  # type: ignore


def _traverse_state_dict(value, visitor):
    return _plan.traverse_state_dict(value, visitor)


def _read_tensor_from_buf(value, shm_tensor_buffer):
    """TensorMeta -> CPU tensor viewing the shm bytes (SAVER:120-137)."""
    if not isinstance(value, TensorMeta):
        return value
    if value.numel == 0:
        return torch.tensor([], dtype=value.dtype)
    flat = torch.frombuffer(buffer=shm_tensor_buffer.buf, dtype=value.dtype,
                            offset=value.offset, count=value.numel)
    return flat.reshape(value.shape)


def _read_state_dict_from_shm(meta_dict, tensor_shm):
    return _traverse_state_dict(meta_dict, lambda m: _read_tensor_from_buf(m, tensor_shm))


def _create_shared_memory(name, create, size=0):
    """Attach to, or (re)create with exactly `size` bytes, the named segment
    (SAVER:140-171)."""
    if not create:
        try:
            return SharedMemory(name=name)
        except FileNotFoundError:
            return None
    if size == 0:
        logger.warning("Cannot create the shared memory with size = 0.")
        return None
    try:
        return SharedMemory(name=name, create=True, size=size)
    except FileExistsError:
        shm = SharedMemory(name=name)
        if shm.size == size:
            return shm
        logger.info("shm %s has %d bytes; re-creating it with %d", name, shm.size, size)
        shm.close()
        shm.unlink()
        return SharedMemory(name=name, create=True, size=size)


def _host_copy_leaves(leaves, buffer):
    """Host-resident leaves -> shm, one frombuffer view + copy_ each; this is
    the reference's own path for CPU tensors (SAVER:197-206)."""
    for lf in leaves:
        if lf.nbytes == 0:
            continue
        view = torch.frombuffer(buffer, dtype=lf.meta.dtype, count=lf.meta.numel,
                                offset=lf.meta.offset).reshape(lf.meta.shape)
        view.copy_(lf.tensor)


def _write_shared_memory(value, meta, buffer):
    """One host tensor -> its slot in `buffer` (SAVER:197-206); kept under the
    reference's name for code written against that module."""
    if value.numel():
        _host_copy_leaves([_plan.Leaf(value, meta)], buffer)


def _traverse_copy_to_shm(value, meta, buffer):
    """Copy the tensors of `value` into `buffer` at the offsets of the meta
    tree `meta` and refresh its non-tensor leaves (SAVER:174-194).  Host path
    only; the handler itself pairs leaves once (plan.collect_leaves) and moves
    CUDA tensors through the pack kernel."""
    leaves = _plan.collect_leaves(value, meta)
    if leaves is None:
        raise KeyError("the state dict no longer has the structure of its meta tree")
    _host_copy_leaves(leaves, buffer)


class PendingSave(object):
    """Handle of a save whose bytes may still be in flight."""

    def __init__(self, handler, meta_dict, conf, engine, host_leaves=None):
        self._handler = handler
        self._meta = meta_dict
        self._conf = conf
        self._engine = engine
        self._host_leaves = host_leaves  # layout indices of host-resident leaves
        self._done = False
        self._transferred = False
        self._sums = None
        self.window = None  # (chunk_begin, chunk_end) of a striped save

    def wait_packed(self):
        """Return once the live tensors may be modified again."""
        if self._engine is not None and not self._done:
            self._engine.save_packed_wait()

    def wait(self):
        """Return once the shm holds the complete shard and the meta says so."""
        if self._done:
            return
        self.publish(self.wait_transfer())

    def wait_transfer(self):
        """Return once this rank's bytes are in the shm; -> its chunk sums (a
        value per chunk of the image; for a striped save only those of
        ``window`` mean anything), None if there are none."""
        if self._transferred:
            return self._sums
        self._transferred = True
        if self._engine is not None:
            self._sums = self._engine.save_finish()
            self._handler.last_stats = dict(self._engine.stats)
        return self._sums

    def publish(self, sums):
        """Record `sums` (of the WHOLE image) in the meta and tell the agent
        the image is complete (writing_shm=False).  Owner of the segment only."""
        if self._done:
            return
        self._done = True
        h = self._handler
        if sums is not None:
            from .transfer import checksums_to_meta
            self._meta[FCP_CHECKSUM_KEY] = checksums_to_meta(
                sums, self._engine.chunk_bytes, self._engine.stats.get("xxh64"),
                host_leaves=self._host_leaves)
        else:
            self._meta.pop(FCP_CHECKSUM_KEY, None)
        self._conf.writing_shm = False
        h._publish(self._meta, self._conf)
        h._note_staged(self._conf.step if self._engine is not None else 0, self.window)

    def abandon(self, step):
        """A rank that wrote one stripe of another rank's segment: nothing to
        publish; remember what its HBM staging now holds."""
        self._done = True
        self._handler._note_staged(step if self._engine is not None else 0, self.window)


class SharedMemoryHandler(object):
    """Writes / reads one shard's state dict to / from its shm segment.

    Args:
        local_rank (int): index of the shard on this node; names the segment
            ``[<run_id>_]checkpoint_shm_<i>`` and the dict ``checkpoint_meta_<i>``.
        host (bool): True in the agent (serves the meta dict), False in the
            training process (client).
    """

    def __init__(self, local_rank, host=True):
        self._buffer_size = 0
        run_id = os.getenv(NodeEnv.TORCHELASTIC_RUN_ID, "")
        self._shm_name = CheckpointSharedObjPrefix.SHM_NAME + str(local_rank)
        if run_id:
            self._shm_name = run_id + "_" + self._shm_name
        self.shared_memory: Optional[SharedMemory] = None
        self.metadata = SharedDict(
            name=CheckpointSharedObjPrefix.META_NAME + str(local_rank), create=host)
        self._need_creation = True
        self._engine = None
        self._n_metas = -1
        self._delta_ok = False  # the agent holds the meta tree of the current layout
        self._changes = []      # non-tensor leaves refreshed by the current save
        self.last_stats = {}
        # what the engine's HBM staging holds: (step, byte_lo, byte_hi) of a
        # complete, checksummed image (or stripe of one); step 0 = nothing
        self._staged = (0, 0, 0)
        self._view_alias = None
        self._published_step = 0  # step of the meta this process last published in full

    # -- lifecycle ----------------------------------------------------------
    def _note_staged(self, step, window=None):
        eng = self._engine
        if eng is None or eng.staging is None or not step:
            self._staged = (0, 0, 0)
            return
        total = self._buffer_size
        if window is None:
            self._staged = (step, 0, total)
        else:
            self._staged = (step, window[0] * eng.chunk_bytes,
                            min(total, window[1] * eng.chunk_bytes))

    def staged_image(self):
        """(step, byte_lo, byte_hi): the part of the checkpoint image of `step`
        that this process still holds, complete, in HBM staging — the source
        of peer recovery over NVLink.  (0, 0, 0) if none."""
        if self._engine is None or self._engine.staging is None:
            return (0, 0, 0)
        return self._staged

    def _drop_engine(self):
        self._staged = (0, 0, 0)
        if self._engine is not None:
            self._engine.close()  # also drops the pinned mapping of the segment
            self._engine = None

    def close(self):
        self._drop_engine()
        if self.shared_memory:
            self.shared_memory.close()

    def unlink(self):
        if not self.shared_memory:
            # the segment may have been created by another process
            self.init_shared_memory()
        if self.shared_memory:
            self.shared_memory.unlink()
        if self.metadata:
            self.metadata.unlink()

    def reset(self):
        self._need_creation = True

    def _wait_populated(self):
        """Host writes into a segment this process just created wait for the
        background threads that make it resident."""
        wait = getattr(self.shared_memory, "wait_populated", None)
        if wait is not None:
            wait()

    def init_shared_memory(self, create=False, size=0):
        self._drop_engine()
        self._view_alias = None  # (unmapped once the views over it are gone)
        self.shared_memory = _create_shared_memory(self._shm_name, create=create, size=size)
        self._need_creation = False

    # -- save -----------------------------------------------------------------
    def _create_tensor_meta(self, value):
        """Kept for API parity (SAVER:255-270); the planner does the same walk
        for the whole dict at once."""
        if not torch.is_tensor(value):
            return value
        meta = TensorMeta(shape=tuple(value.shape), dtype=value.dtype,
                          element_size=value.element_size(), numel=value.numel(),
                          offset=self._buffer_size)
        self._buffer_size += value.numel() * value.element_size()
        return meta

    def _announce(self, meta_dict, conf):
        """Tell the agent the image is being rewritten (writing_shm=True)."""
        if self._delta_ok:
            try:
                if self.metadata.update([((DLROVER_CKPT_CONFIG_KEY,), conf)]):
                    return
            except Exception:  # noqa: BLE001
                pass
            self._delta_ok = False
        self.metadata.set(meta_dict)

    def _publish(self, meta_dict, conf):
        """Make the agent's copy of the meta equal `meta_dict`: by paths when
        it already holds this layout (config, checksums and the refreshed
        non-tensor leaves are all that changed), else as a whole."""
        if self._delta_ok:
            updates = list(self._changes)
            updates.append(((DLROVER_CKPT_CONFIG_KEY,), conf))
            if FCP_CHECKSUM_KEY in meta_dict:
                updates.append(((FCP_CHECKSUM_KEY,), meta_dict[FCP_CHECKSUM_KEY]))
            try:
                if self.metadata.update(updates):
                    self._published_step = conf.step
                    return
            except Exception as e:  # noqa: BLE001 - fall back to the full set
                logger.debug("meta update failed (%s); sending the whole meta", e)
        self.metadata.set(meta_dict)
        self._published_step = conf.step
        self._delta_ok = os.getenv("FCP_META_DELTA", "1") == "1"

    def _layout_for(self, state_dict, create=True):
        """(meta_dict, leaves) for this save: the cached layout if the dict
        still has the same structure, a fresh one (and a fresh segment of the
        exact size) otherwise.  create=False: this rank writes a stripe of a
        segment another rank owns — attach to it, never create."""
        if self.shared_memory is not None:
            meta_dict = self.metadata.get(local=True)
            if meta_dict:
                self._changes = []
                leaves = _plan.collect_leaves(state_dict, meta_dict, changes=self._changes)
                if leaves is not None and len(leaves) == self._n_metas:
                    return meta_dict, leaves
                logger.info("state dict structure changed; rebuilding the shm layout")
        self._delta_ok = False  # new layout: the agent needs the whole tree
        self._changes = []
        meta_dict, leaves, total = _plan.build_layout(state_dict)
        self._buffer_size = total
        self._n_metas = len(leaves)
        if not create:
            if self.shared_memory is None or self.shared_memory.size != total \
                    or self._segment_replaced():
                if self.shared_memory is not None:
                    self._drop_engine()
                    self.shared_memory.close()
                self.init_shared_memory(create=False)
            if self.shared_memory is None or self.shared_memory.size != total:
                raise RuntimeError("the segment %s of the stripe owner has %s bytes, this "
                                   "rank's state needs %d: replicas differ"
                                   % (self._shm_name, getattr(self.shared_memory, "size", None),
                                      total))
            # the owner's meta is the one the agent gets; keep ours locally
            self.metadata._dict = meta_dict
            return meta_dict, leaves
        if self.shared_memory is None or self.shared_memory.size != total:
            if self.shared_memory is not None:
                self._drop_engine()
                self.shared_memory.close()
            self.init_shared_memory(create=True, size=total)
        return meta_dict, leaves

    def _engine_for(self, leaves):
        """The device engine if any leaf with bytes lives on a GPU."""
        for lf in leaves:
            if lf.is_cuda and lf.nbytes:
                dev = lf.tensor.device
                if self._engine is None or self._engine.device != dev:
                    self._drop_engine()
                    from .transfer import DeviceShardEngine
                    self._engine = DeviceShardEngine(dev)
                return self._engine
        return None

    def save_state_dict_async(self, state_dict, stripe=None, attach_barrier=None):
        """`stripe` = (index, count, owner): this rank moves stripe `index` of
        `count` of an image that every rank of a replicated group holds
        identically, into ONE segment; `owner` (exactly one rank) creates the
        segment and talks to the agent.  `attach_barrier()`: called by every
        rank between the owner's (re)creation of the segment + its
        writing_shm=True announcement and the first write of any rank.
        CUDA tensors only."""
        if stripe is not None:
            return self._save_stripe_async(state_dict, stripe, attach_barrier)
        self._staged = (0, 0, 0)
        meta_dict, leaves = self._layout_for(state_dict)
        conf = meta_dict.get(DLROVER_CKPT_CONFIG_KEY)
        if conf is None:
            conf = CheckpointConfig()
            meta_dict[DLROVER_CKPT_CONFIG_KEY] = conf
        conf.writing_shm = True
        engine = self._engine_for(leaves)
        # writing_shm=True reaches the agent before the first byte of the
        # image changes (SAVER:283-288): a trainer that dies mid-save then
        # leaves a meta that says so, and neither the agent nor a restarted
        # trainer's plain load_state_dict() (CPU views, no CRC) trusts the
        # image.  With the agent holding this layout the message is the config
        # alone (a path-wise update, ~0.2 ms).
        self._announce(meta_dict, conf)
        if engine is not None and self.shared_memory is not None:
            engine.save_begin(leaves, self._buffer_size, self.shared_memory.buf,
                              ready=getattr(self.shared_memory, "wait_populated", None))
            host_leaves = [i for i, lf in enumerate(leaves) if not lf.is_cuda]
            return PendingSave(self, meta_dict, conf, engine, host_leaves)
        if self.shared_memory is None:
            # nothing but empty tensors / scalars: the meta carries it all
            return PendingSave(self, meta_dict, conf, None)
        self._wait_populated()
        _host_copy_leaves(leaves, self.shared_memory.buf)
        return PendingSave(self, meta_dict, conf, None)

    def _save_stripe_async(self, state_dict, stripe, attach_barrier):
        index, count, owner = stripe
        self._staged = (0, 0, 0)
        meta_dict = leaves = conf = None
        if owner:
            meta_dict, leaves = self._layout_for(state_dict)
            conf = meta_dict.get(DLROVER_CKPT_CONFIG_KEY)
            if conf is None:
                conf = meta_dict[DLROVER_CKPT_CONFIG_KEY] = CheckpointConfig()
            conf.writing_shm = True
            self._announce(meta_dict, conf)
        if attach_barrier is not None:
            attach_barrier()
        if not owner:
            meta_dict, leaves = self._layout_for(state_dict, create=False)
            conf = meta_dict.get(DLROVER_CKPT_CONFIG_KEY) or CheckpointConfig()
        engine = self._engine_for(leaves)
        if engine is None or self.shared_memory is None:
            raise ValueError("striped saves need a state dict with CUDA tensors")
        window = engine.stripe_of(self._buffer_size, index, count)
        engine.save_begin(leaves, self._buffer_size, self.shared_memory.buf,
                          ready=getattr(self.shared_memory, "wait_populated", None)
                          if owner else None, window=window)
        host_leaves = [i for i, lf in enumerate(leaves) if not lf.is_cuda]
        pending = PendingSave(self, meta_dict, conf, engine, host_leaves)
        pending.window = window
        return pending

    def save_state_dict(self, state_dict):
        """Copy the state dict into the shm segment; complete on return
        (SAVER:272-290)."""
        self.save_state_dict_async(state_dict).wait()

    # -- load -----------------------------------------------------------------
    def load_state_dict(self, into=None, device=None):
        """Without arguments: nested dict of CPU tensors that alias the shm
        (SAVER:292-311) — consume it before the next save.

        ``into``: a state dict of the same structure whose tensors receive the
        bytes in place (CUDA tensors through the scatter kernel).  ``device``:
        allocate fresh tensors there instead (leaves that were host-resident
        at save time stay on the host; the device tensors are views of one
        allocation, so keeping any of them keeps all of it).  Given both, `into` may cover only
        part of the structure (say, the optimizer shards) and the rest is
        allocated on `device`.  All variants verify the CRC-32C recorded at
        save time.  Returns {} when the shm holds no complete checkpoint."""
        meta_dict = self._current_meta()
        config = meta_dict.get(DLROVER_CKPT_CONFIG_KEY, CheckpointConfig())
        if not meta_dict or config.writing_shm:
            return {}
        if self.shared_memory is None or self._need_creation or self._segment_replaced():
            self.init_shared_memory(create=False)
        if not self.shared_memory:
            return {}
        if into is None and device is None:
            return _read_state_dict_from_shm(_strip_private(meta_dict), self._view_source())
        return self._load_into(meta_dict, into, device)

    def _current_meta(self):
        """The agent's meta dict.  A process that published the layout itself
        asks the agent for the config alone and, if step and state agree with
        its own copy, uses that copy: the tree of a few thousand TensorMeta
        pickles to hundreds of KB, milliseconds on the restore path."""
        local = self.metadata.get(local=True)
        mine = local.get(DLROVER_CKPT_CONFIG_KEY) if local else None
        if mine is not None and self._published_step and mine.step == self._published_step:
            theirs = self.metadata.get_keys([DLROVER_CKPT_CONFIG_KEY]).get(DLROVER_CKPT_CONFIG_KEY)
            if theirs is not None and theirs.step == mine.step \
                    and theirs.writing_shm == mine.writing_shm and theirs.rank == mine.rank:
                return local
        return self.metadata.get()

    def _view_source(self):
        """What the zero-copy CPU views of load_state_dict() alias.  In a
        process whose engine pinned the segment piece by piece, views over
        that mapping look like pinned memory to CUDA, and a later
        ``cuda_tensor.copy_(view)`` of a tensor that straddles two pieces is
        rejected (a transfer may not span two registrations).  Such a process
        hands out views over a second, unregistered mapping of the same
        segment: plain pageable memory, exactly what the reference's views
        are."""
        if self._engine is None or getattr(self.shared_memory, "_fd", -1) < 0:
            return self.shared_memory
        alias = self._view_alias
        if alias is None or alias.size != self.shared_memory.size:
            import mmap
            alias = _ShmAlias(mmap.mmap(self.shared_memory._fd, self.shared_memory.size,
                                        flags=mmap.MAP_SHARED))
            self._view_alias = alias
        return alias

    def load_raw_images(self):
        """{state name: RawImage} over the shm, for the saver's raw persist
        format (raw_ckpt.py); {} when the shm holds no complete checkpoint."""
        meta_dict = self.metadata.get()
        config = meta_dict.get(DLROVER_CKPT_CONFIG_KEY, CheckpointConfig())
        if not meta_dict or config.writing_shm:
            return {}
        if self.shared_memory is None or self._need_creation or self._segment_replaced():
            self.init_shared_memory(create=False)
        sums = meta_dict.get(FCP_CHECKSUM_KEY)
        buf = self.shared_memory.buf if self.shared_memory else b""
        return {name: raw_ckpt.RawImage(sub, buf, config.step, sums, total=len(buf))
                for name, sub in _strip_private(meta_dict).items()
                if name != DLROVER_CKPT_CONFIG_KEY}

    def _load_into(self, meta_dict, into, device):
        total = self.shared_memory.size
        public = _strip_private(meta_dict)
        # leaves that lived in host memory when the image was packed (rng
        # states, a CPU `step`) come back as host tensors: their consumers
        # (torch.set_rng_state ...) want them there
        host_idx = (meta_dict.get(FCP_CHECKSUM_KEY) or {}).get("host_leaves") or ()
        alloc = _plan.arena_allocator(device, host_idx)
        leaves = _plan.pair_or_alloc(public, into, alloc)
        engine = self._engine_for(leaves)
        if engine is not None:
            from .transfer import checksums_from_meta
            entry = meta_dict.get(FCP_CHECKSUM_KEY)
            expected = checksums_from_meta(entry, total, engine.chunk_bytes)
            engine.restore(leaves, total, self.shared_memory.buf, expected,
                           expected_xxh64=(entry or {}).get("xxh64") if expected is not None
                           else None)
            self.last_stats = dict(engine.stats)
            # the staging now holds the verified image: this rank can serve it
            self._buffer_size = total
            conf = meta_dict.get(DLROVER_CKPT_CONFIG_KEY)
            self._note_staged(conf.step if conf is not None and expected is not None else 0)
        else:
            for lf in leaves:
                if lf.nbytes:
                    with torch.no_grad():
                        lf.tensor.copy_(_read_tensor_from_buf(lf.meta, self.shared_memory))
        it = iter(leaves)

        def fill(m):
            return next(it).tensor if isinstance(m, TensorMeta) else m

        return _traverse_state_dict(public, fill)

    def restore_from_image(self, meta_dict, image, into):
        """Unpack a packed image that is not in this handler's shm — e.g. one
        received from a surviving replica over NCCL — into the tensors of
        `into`.  `image` is a uint8 tensor (CUDA: scatter kernel with CRC
        verification; CPU: host copies)."""
        public = _strip_private(meta_dict)
        leaves = _plan.collect_leaves(into, public, refresh=False)
        if leaves is None or len(leaves) != _plan.count_tensor_metas(public):
            raise ValueError("`into` does not match the structure of the received checkpoint")
        total = image.numel()
        if image.is_cuda:
            engine = self._engine_for(leaves)
            if engine is None:
                raise ValueError("device image but no CUDA tensor to restore into")
            from .transfer import checksums_from_meta
            expected = checksums_from_meta(meta_dict.get(FCP_CHECKSUM_KEY), total,
                                           engine.chunk_bytes)
            engine.restore_from_device(leaves, total, image, expected)
        else:
            for lf in leaves:
                if lf.nbytes:
                    src = image[lf.meta.offset:lf.meta.offset + lf.nbytes]
                    with torch.no_grad():
                        lf.tensor.copy_(src.view(lf.meta.dtype).reshape(lf.meta.shape)
                                        if lf.meta.offset % lf.meta.element_size == 0
                                        else src.clone().view(lf.meta.dtype).reshape(lf.meta.shape))

    def restore_from_file(self, path, into):
        """Fill the tensors of `into` (one state name's dict) from a raw image
        file written by the saver (raw_ckpt.py).  CUDA tensors: file mapping ->
        pinned ring -> HBM -> scatter kernel, CRC-verified when the file holds
        the whole image; host tensors: copies.  Returns `into`'s structure
        with the file's non-tensor leaves."""
        header, data_off = raw_ckpt.read_header(path)
        meta = raw_ckpt.decode_meta(header)
        leaves = _plan.collect_leaves(into, meta, refresh=False)
        if leaves is None or len(leaves) != _plan.count_tensor_metas(meta):
            raise ValueError("`into` does not match the structure of the checkpoint in %s" % path)
        total = header["nbytes"]
        mm = raw_ckpt.map_payload(path, header, data_off)
        engine = self._engine_for(leaves)
        if engine is not None and mm is not None:
            from .transfer import checksums_from_meta
            # sums of the payload's own windows (files written before they
            # existed carry the whole image's sums, usable only when the
            # payload IS the whole image)
            expected = checksums_from_meta(header.get("payload_checksums"), total,
                                           engine.chunk_bytes)
            image = header.get("image") or {}
            if expected is None and image.get("offset") == 0 \
                    and image.get("total", total) == total:
                expected = checksums_from_meta(image.get("checksums"), total, engine.chunk_bytes)
            engine.restore(leaves, total, mm, expected, transient=True)
            self.last_stats = dict(engine.stats)
        elif mm is not None:
            for lf in leaves:
                if lf.nbytes:
                    src = torch.frombuffer(mm, dtype=lf.meta.dtype, count=lf.meta.numel,
                                           offset=lf.meta.offset).reshape(lf.meta.shape)
                    with torch.no_grad():
                        lf.tensor.copy_(src)
        it = iter(leaves)

        def fill(m):
            return next(it).tensor if isinstance(m, TensorMeta) else m

        return _traverse_state_dict(meta, fill)

    def verify_image(self):
        """Recompute the CRC-32C of every chunk of the shm image on the host
        and compare with the sums the pack kernel recorded in the meta
        (FCP_CHECKSUM_KEY).  The agent calls this before it persists a shard:
        whatever the lock state was, an image that was torn by a trainer dying
        mid-save, or that does not belong to the meta next to it, never
        reaches storage.  No-op for images without recorded sums (host-only
        state dicts).  Host code only (libfcp.so's SSE4.2 routine, no CUDA).
        Returns the number of chunks checked."""
        meta_dict = self.metadata.get()
        entry = meta_dict.get(FCP_CHECKSUM_KEY) if meta_dict else None
        if not entry or entry.get("algo") != "crc32c":
            return 0
        if self.shared_memory is None or self._need_creation or self._segment_replaced():
            self.init_shared_memory(create=False)
        if not self.shared_memory:
            return 0
        from . import _native
        chunk = int(entry["chunk_bytes"])
        expected = entry["sums"]
        total = self.shared_memory.size
        t0 = time.time()
        got = _native.host_chunk_crc32c(self.shared_memory.buf, total, chunk)
        if len(got) != len(expected):
            raise _native.ChecksumMismatch(
                -74, "the shm image has %d chunks, its meta records %d"
                % (len(got), len(expected)))
        for i, (a, b) in enumerate(zip(got.tolist(), expected)):
            if a != int(b):
                raise _native.ChecksumMismatch(
                    -74, "shm image is corrupt: CRC-32C of chunk %d is %08x, recorded %08x"
                    % (i, a, int(b)))
        logger.info("verified %d chunk checksums of the shm image (%.2f GB) in %.2fs",
                    len(got), total / 1e9, time.time() - t0)
        return len(got)

    def _segment_replaced(self):
        """The writer re-creates the segment when the layout changes; a reader
        still mapped to the old one must re-attach."""
        try:
            now = os.stat("/dev/shm/" + self._shm_name)
            if now.st_size != self.shared_memory.size:
                return True
            fd = getattr(self.shared_memory, "_fd", -1)
            return fd >= 0 and os.fstat(fd).st_ino != now.st_ino
        except OSError:
            return False

    def no_checkpint_state(self):
        """True if no shard has ever been saved (handlers attach lazily, so
        this asks the meta dict, not the segment; SAVER:313-323)."""
        meta_dict = self.metadata.get_keys([DLROVER_CKPT_CONFIG_KEY])
        config = meta_dict.get(DLROVER_CKPT_CONFIG_KEY, None)
        return config is None or config.step == 0

    def get_checkpoint_config(self, default_config):
        meta_dict = self.metadata.get_keys([DLROVER_CKPT_CONFIG_KEY])
        return meta_dict.get(DLROVER_CKPT_CONFIG_KEY, default_config)


class _ShmAlias(object):
    """A second mapping of a segment; quacks like SharedMemory for the view
    builders (``.buf``)."""

    def __init__(self, mm):
        self._mmap = mm
        self.buf = memoryview(mm)
        self.size = len(mm)


def _strip_private(meta_dict):
    """The meta dict without the keys this build adds on top of the
    reference's layout."""
    if FCP_CHECKSUM_KEY in meta_dict:
        return {k: v for k, v in meta_dict.items() if k != FCP_CHECKSUM_KEY}
    return meta_dict


# =============================================================== agent side
class AsyncCheckpointSaver(metaclass=ABCMeta):
    """Runs in the agent (or a daemon forked by local rank 0): owns the lock,
    event queue and meta-dict servers of every local shard and persists the
    shm contents to storage when told to (SAVER:344-770).

    Args:
        checkpoint_dir (str): root directory of the checkpoints.
        storage_meta (ClassMeta): how to build the CheckpointStorage.
        local_shard_num (int): shards (= shm segments) on this node.
        global_shard_num (int): shards over all nodes.
        save_timeout (int): seconds to wait for all shards when committing.
    """

    _saver_instance = None
    _STAGE_DIR = "._dlrover_ckpt_stage"

    def __init__(self, checkpoint_dir, storage_meta: ClassMeta, local_shard_num=1,
                 global_shard_num=1, save_timeout=CheckpointConstant.SAVE_TIMEOUT):
        self.checkpoint_dir = checkpoint_dir
        self.local_shard_num = local_shard_num
        self.global_shard_num = global_shard_num
        self._node_rank = env_utils.get_node_rank()
        self._is_agent_rank_0 = self._node_rank == 0
        self._stop_commit = False
        self._save_timeout = save_timeout
        storage_cls = getattr(importlib.import_module(storage_meta.module_path),
                              storage_meta.class_name)
        self.storage = storage_cls(**storage_meta.kwargs)
        self._writing_storage = False
        self._latest_step = 0
        self._event_queue = SharedQueue(
            name=CheckpointSharedObjPrefix.SAVE_STEP_QNAME + "0", create=True)
        self._shm_handlers: List[SharedMemoryHandler] = []
        self._shm_locks: List[SharedLock] = []
        for i in range(local_shard_num):
            self._shm_handlers.append(SharedMemoryHandler(i))
            self._shm_locks.append(SharedLock(
                name=CheckpointSharedObjPrefix.SHM_LOCK_NAME + str(i), create=True))
        self._executor = ThreadPoolExecutor(
            max_workers=local_shard_num, thread_name_prefix="ckpt_saver-")
        logger.info("AsyncSaver: checkpoint_dir=%s local_shard_num=%s global_shard_num=%s",
                    checkpoint_dir, local_shard_num, global_shard_num)

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001 - interpreter shutdown
            pass

    # -- life cycle ----------------------------------------------------------
    @classmethod
    def start_async_saving_ckpt(cls):
        """Start the factory thread: it waits on the queue "factory" for the
        ClassMeta of the saver the training process wants, builds it and runs
        its event loop in a second thread (SAVER:409-465)."""
        factory_queue = SharedQueue(name="factory", create=True)
        creation_guard = threading.Lock()

        def run_saver(class_meta):
            if cls._saver_instance is not None:
                cls._saver_instance.close()
                cls._saver_instance = None
            saver_cls = getattr(importlib.import_module(class_meta.module_path),
                                class_meta.class_name)
            cls._saver_instance = saver_cls(**class_meta.kwargs)
            cls._saver_instance._sync_shm_to_storage()

        def factory():
            logger.info("Start the checkpoint saver factory.")
            worker = None
            while True:
                class_meta = factory_queue.get()
                with creation_guard:
                    if cls._saver_instance and worker and worker.is_alive():
                        logger.info("The saver is already created, skip creating the saver.")
                        continue
                    worker = threading.Thread(target=run_saver, args=(class_meta,),
                                              name="checkpoint-saver", daemon=True)
                    worker.start()

        threading.Thread(target=factory, name="checkpoint-saver-factory", daemon=True).start()

    @classmethod
    def get_ckpt_saver(cls):
        return cls._saver_instance

    @classmethod
    def register_signal_handler(cls):
        """SIGINT: drop the shm; SIGTERM: persist it first (SAVER:471-494)."""
        prev_int = signal.getsignal(signal.SIGINT)
        prev_term = signal.getsignal(signal.SIGTERM)

        def on_sigint(signum, frame):
            if cls._saver_instance:
                cls._saver_instance.close()
            if callable(prev_int):
                # (called by hand with signum=None, Python >= 3.12's default
                # handler raises TypeError instead of KeyboardInterrupt)
                prev_int(signal.SIGINT if signum is None else signum, frame)

        def on_sigterm(signum, frame):
            if cls._saver_instance:
                cls._saver_instance.save_shm_to_storage()
                cls._saver_instance.close()
            if callable(prev_term):
                prev_term(signal.SIGTERM if signum is None else signum, frame)

        signal.signal(signal.SIGINT, on_sigint)
        signal.signal(signal.SIGTERM, on_sigterm)

    @classmethod
    def reset(cls):
        """Called by the agent when it restarts the workers."""
        if cls._saver_instance is None:
            return
        cls._saver_instance.reset_shared_memory()
        logger.info("Reset all shared memory of shards.")

    def reset_shared_memory(self):
        self._stop_commit = True
        for h in self._shm_handlers:
            h.reset()

    def wait_saving_checkpoint(self):
        return self._writing_storage

    def close(self):
        """Stop the event loop and remove every shared object."""
        if not self._event_queue.empty():
            self._event_queue._queue.get()
        self._event_queue.put(CheckpointEvent(type=CheckpointEventType.EXIT))
        for i in range(self.local_shard_num):
            if self._shm_handlers[i]:
                self._shm_handlers[i].close()
                self._shm_handlers[i].unlink()
            self._shm_locks[i].unlink()
        self._event_queue.unlink()
        self._executor.shutdown(wait=False)

    def _sync_shm_to_storage(self):
        """Event loop of the saver thread."""
        logger.info("Async flash checkpoint saver starts!")
        while True:
            event = self._event_queue.get()
            if event.type == CheckpointEventType.UPDATE_SHARD:
                logger.info("global shard number is now %s", event.global_shard_num)
                self.global_shard_num = event.global_shard_num
            elif event.type == CheckpointEventType.SAVE:
                logger.info("persist checkpoint, event %s", event)
                self.save_step_checkpoint(event.step)
            elif event.type == CheckpointEventType.EXIT:
                break

    # -- helpers ----------------------------------------------------------------
    def _shard_config(self, i):
        return self._shm_handlers[i].get_checkpoint_config(CheckpointConfig())

    def _save_shard(self, step, local_shard_id, ckpt_config, step_done_dir):
        """Persist one shard under its lock and drop a done-file."""
        lock = self._shm_locks[local_shard_id]
        try:
            handler = self._shm_handlers[local_shard_id]
            if handler.shared_memory is None:
                handler.init_shared_memory(create=False)
            lock.acquire()
            in_memory = self._shard_config(local_shard_id)
            if in_memory.step != step:
                logger.error("event step %s != step %s in memory", step, in_memory.step)
                return
            logger.info("persisting shard of rank %s: %s", ckpt_config.rank, ckpt_config)
            self.persist_to_storage(local_shard_id, ckpt_config)
            lock.release()
            self.storage.write("done", os.path.join(step_done_dir, str(ckpt_config.rank)))
            logger.info("shard of rank %s persisted", ckpt_config.rank)
            return True
        except Exception as e:  # noqa: BLE001
            logger.error("failed to persist shard of rank %s: %s", ckpt_config.rank, e,
                         exc_info=True)
            return False
        finally:
            lock.release()

    def _dist_make_dir(self, path, timeout=30):
        if self._node_rank == 0:
            self.storage.safe_rmtree(path)
            self.storage.safe_makedirs(path)
        else:
            for _ in range(timeout):
                if self.storage.exists(path):
                    break
                time.sleep(1)

    def _any_rank_locked(self):
        return any(lock.locked() for lock in self._shm_locks)

    def _get_checkpoint_done_dir(self, step):
        return os.path.join(self.checkpoint_dir, self._STAGE_DIR, str(step) + ".done")

    def _check_shard_step_consistence(self, step, timeout=15):
        """Poll until every shard that holds a checkpoint holds `step`."""
        start = time.time()
        while True:
            steps = [s for s in (self._shard_config(i).step
                                 for i in range(self.local_shard_num)) if s > 0]
            if all(s == step for s in steps):
                return True
            time.sleep(1)
            if time.time() - start > timeout:
                logger.info("The cached steps are %s", steps)
                return False

    def _persist_all_shards(self, step, step_done_dir, skip_empty):
        futures = []
        for i in range(self.local_shard_num):
            conf = self._shard_config(i)
            if skip_empty and conf.step == 0:
                continue
            futures.append(self._executor.submit(
                self._save_shard, step, i, conf, step_done_dir))
        ok = 0
        for i, fut in enumerate(futures):
            if fut.result():
                ok += 1
            else:
                logger.error("failed to persist shard %s of step %s", i, step)
        return ok == self.local_shard_num

    def save_shm_to_storage(self, timeout=60, master_client=None):
        """Breakpoint save: persist whatever complete checkpoint the shm holds
        (agent calls this when workers fail or before restarting them;
        SAVER:634-693)."""
        if any(h.no_checkpint_state() for h in self._shm_handlers):
            logger.info("Skip because no any memory buffer with the state dict.")
            return
        steps = [self._shard_config(i).step for i in range(self.local_shard_num)]
        if len(set(steps)) > 1:
            logger.error("Skip because steps in shards are inconsistent: %s", steps)
            return
        step = steps[0]
        if master_client and not self._sync_node_checkpoint(master_client, step, timeout):
            logger.info("Skip saving the checkpoint from the memory to the storage.")
            self._stop_commit = True
            return
        # a held lock means a trainer died mid-copy: the shm may be torn
        if self._writing_storage or self._any_rank_locked():
            logger.info("Saver is writing the checkpoint to storage "
                        "and skips saving at the breakpoint.")
            return
        if step > self._latest_step:
            self.save_step_checkpoint(step)
            logger.info("Saved the checkpoint of step %s from shm to storage.", step)
        else:
            logger.info("The checkpoint of step %s has been saved.", step)

    def _sync_node_checkpoint(self, master_client, step, timeout):
        """All nodes must agree they can persist `step` (some may have died)."""
        start = time.time()
        while True:
            if master_client.sync_checkpoint(step):
                return True
            time.sleep(3)
            if time.time() - start > timeout:
                logger.info("timeout syncing the checkpoint step; some nodes may have failed")
                return False

    def _persist_states(self, local_shard_id, ckpt_config):
        """Write every state name that has a path: ``torch.save`` of views
        over the shm (SAVER:914-921), or with DLROVER_FCP_PERSIST_FORMAT=raw
        the image bytes themselves (raw_ckpt.py)."""
        handler = self._shm_handlers[local_shard_id]
        if os.getenv("FCP_AGENT_VERIFY", "1") == "1":
            handler.verify_image()  # raises ChecksumMismatch: a torn image is never persisted
        if raw_ckpt.persist_format() == "raw":
            states, write_func = handler.load_raw_images(), raw_ckpt.write_raw
        else:
            states, write_func = handler.load_state_dict(), torch.save
        for state_name, sd in states.items():
            if sd and state_name in ckpt_config.paths:
                self.storage.write_state_dict(sd, ckpt_config.paths[state_name], write_func)

    @abstractmethod
    def save_step_checkpoint(self, step: int):
        """Persist the checkpoint of `step` from shm to storage."""

    @abstractmethod
    def persist_to_storage(self, local_shard_id, ckpt_config):
        """Write one shard to the paths in `ckpt_config`."""

    @abstractmethod
    def commit_checkpoint(self, step: int, step_done_dir: str, timeout=600):
        """Publish the checkpoint once every shard reported done."""

    @abstractmethod
    def update_tracker_file(self, step: int):
        """Record `step` as the latest complete checkpoint."""


class CommonDirCheckpointSaver(AsyncCheckpointSaver):
    """Each shard goes straight to the path its framework chose; the tracker
    file is updated when the number of done-files equals the number of global
    shards (SAVER:772-921)."""

    def update_tracker_file(self, step):
        self.storage.write(str(step), os.path.join(
            self.checkpoint_dir, CheckpointConstant.TRACER_FILE_NAME))

    def save_step_checkpoint(self, step: int):
        if not self._check_shard_step_consistence(step):
            logger.warning("Skip persisting step %s: cached steps are not consistent.", step)
            return
        self._writing_storage = True
        try:
            step_done_dir = self._get_checkpoint_done_dir(step)
            self._dist_make_dir(step_done_dir)
            if not self._persist_all_shards(step, step_done_dir, skip_empty=True):
                logger.error("node %s failed to save step %s", self._node_rank, step)
                return
            self._latest_step = step
            if self._is_agent_rank_0:
                self._stop_commit = False
                self.commit_checkpoint(step, step_done_dir, timeout=self._save_timeout)
        finally:
            # (the reference leaves the flag set after a failed save, which
            # disables every later breakpoint save)
            self._writing_storage = False

    def commit_checkpoint(self, step: int, step_done_dir: str, timeout=600):
        start = time.time()
        success = False
        while True:
            if self._stop_commit:
                logger.info("Stop committing: the training processes restarted.")
                break
            done_files = self.storage.listdir(step_done_dir)
            if len(done_files) == self.global_shard_num:
                logger.info("All agents finish saving checkpoint for step %s", step)
                self.update_tracker_file(step)
                self.storage.safe_rmtree(step_done_dir)
                success = True
                break
            logger.info("ready shards %s != %s", len(done_files), self.global_shard_num)
            if round(time.time() - start, 2) > timeout:
                logger.error("commit of step %s timed out; done files: %s", step, done_files)
                self.storage.safe_rmtree(step_done_dir)
                break
            time.sleep(5)
        self.storage.commit(step, success)

    def persist_to_storage(self, local_shard_id, ckpt_config):
        self._persist_states(local_shard_id, ckpt_config)


class TempDirCheckpointSaver(AsyncCheckpointSaver):
    """Shards are first written below ``<dir>/._dlrover_ckpt_stage/<step>/``
    and the folder is moved into place when all are done (SAVER:924-1114)."""

    def update_tracker_file(self, step):
        self.storage.write(str(step), os.path.join(
            self.checkpoint_dir, CheckpointConstant.TRACER_FILE_NAME))

    def save_step_checkpoint(self, step):
        logger.info("node %s persists step %s", self._node_rank, step)
        if not self._check_shard_step_consistence(step):
            logger.warning("Skip persisting step %s: cached steps are not consistent.", step)
            return
        self._writing_storage = True
        try:
            temp_dir = self._get_tmp_ckpt_dir(step)
            self._dist_make_dir(temp_dir)
            step_done_dir = self._get_checkpoint_done_dir(step)
            self._dist_make_dir(step_done_dir)
            futures, ckpt_dir = [], ""
            for i in range(self.local_shard_num):
                conf = self._shard_config(i)
                ckpt_dir = self._replace_path_dir(conf, temp_dir)
                futures.append(self._executor.submit(
                    self._save_shard, step, i, conf, step_done_dir))
            ok = sum(1 for fut in futures if fut.result())
            if ok != self.local_shard_num:
                logger.error("node %s failed to save step %s", self._node_rank, step)
                return
            if self._is_agent_rank_0:
                self.commit_checkpoint(step, step_done_dir=step_done_dir, tmp_path=temp_dir,
                                       target_path=ckpt_dir, timeout=self._save_timeout)
        finally:
            self._writing_storage = False

    def _replace_path_dir(self, ckpt_config, temp_dir):
        """Point the shard's paths into `temp_dir`; returns the directory they
        pointed to (all paths of a shard must share one)."""
        ckpt_dir = ""
        if not ckpt_config.paths:
            return ckpt_dir
        staged = {}
        for name, path in ckpt_config.paths.items():
            path = str(path)
            path_dir = os.path.dirname(path)
            staged[name] = path.replace(path_dir, temp_dir)
            if not ckpt_dir:
                ckpt_dir = path_dir
            elif ckpt_dir != path_dir:
                raise ValueError("The directories must be same. The latest dir is %s and "
                                 "the current dir of %s is %s" % (ckpt_dir, name, path_dir))
        ckpt_config.paths = staged
        return ckpt_dir

    def _get_tmp_ckpt_dir(self, step):
        return os.path.join(self.checkpoint_dir, self._STAGE_DIR, str(step))

    def commit_checkpoint(self, step, step_done_dir, tmp_path, target_path,  # type: ignore
                          timeout=600):
        logger.info("Start commit checkpoint tmp_path: %s, path: %s", tmp_path, target_path)
        start = time.time()
        success = False
        while True:
            done_files = self.storage.listdir(step_done_dir)
            if len(done_files) == self.global_shard_num:
                logger.info("All agents finish saving checkpoint for step %s", step)
                if os.path.exists(target_path):
                    if os.path.isdir(target_path):
                        self.storage.safe_rmtree(target_path)
                    else:
                        self.storage.safe_remove(target_path)
                self.storage.safe_move(tmp_path, target_path)
                self.storage.safe_rmtree(step_done_dir)
                self.update_tracker_file(step)
                success = True
                break
            logger.info("ready shards %s != %s", len(done_files), self.global_shard_num)
            if time.time() - start > timeout:
                logger.error("commit of step %s timed out; done files: %s", step, done_files)
                self.storage.safe_rmtree(tmp_path)
                self.storage.safe_rmtree(step_done_dir)
                break
            time.sleep(5)
        self.storage.commit(step, success)

    def persist_to_storage(self, local_shard_id, ckpt_config):
        self._persist_states(local_shard_id, ckpt_config)


class DdpCheckpointSaver(CommonDirCheckpointSaver):
    """DDP replicas are identical: only node 0 writes."""

    def persist_to_storage(self, local_shard_id, ckpt_config):
        if self._node_rank != 0:
            logger.info("Skip and only rank 0 saves checkpoint in a DDP job.")
            return
        super().persist_to_storage(local_shard_id, ckpt_config)


class _ExtraTrackerSaver(CommonDirCheckpointSaver):
    TRACER_FILE = ""

    def update_tracker_file(self, step):
        super().update_tracker_file(step)
        self.storage.write(str(step), os.path.join(self.checkpoint_dir, self.TRACER_FILE))


class MegatronCheckpointSaver(_ExtraTrackerSaver):
    TRACER_FILE = "latest_checkpointed_iteration.txt"


class DeepSpeedCheckpointSaver(_ExtraTrackerSaver):
    TRACER_FILE = "latest"


class FsdpDcpSaver(CommonDirCheckpointSaver):
    """FSDP distributed checkpoint: the shm image *is* the ``.distcp`` file, so
    it is written out raw; node 0 / shard 0 also writes the pickled DCP
    ``.metadata`` and the tracker (SAVER:1165-1197)."""

    def persist_to_storage(self, local_shard_id, ckpt_config):
        handler = self._shm_handlers[local_shard_id]
        path = ckpt_config.paths[CheckpointConstant.MODEL_STATES_NAME]
        checkpoint_dir = os.path.dirname(path)
        # (the reference calls _dist_make_dir here, whose rmtree lets the
        # shards of one node delete each other's files; SAVER:1181-1185)
        self.storage.safe_makedirs(checkpoint_dir)
        if handler.shared_memory is None or handler._segment_replaced():
            # the trainer re-creates the segment when its plan changes size;
            # a mapping made before that still shows the old image
            handler.init_shared_memory(create=False)
        assert handler.shared_memory is not None
        if os.getenv("FCP_AGENT_VERIFY", "1") == "1":
            handler.verify_image()
        self.storage.write(handler.shared_memory.buf, path)
        if self._is_agent_rank_0 and local_shard_id == 0:
            meta_dict = handler.metadata.get()
            dcp_metadata = meta_dict.get("dcp_metadata", {})
            if dcp_metadata:
                self.storage.write(pickle.dumps(dcp_metadata),
                                   Path(checkpoint_dir) / ".metadata")
            self.storage.write(str(ckpt_config.step), os.path.join(
                self.checkpoint_dir, CheckpointConstant.TRACER_FILE_NAME))
