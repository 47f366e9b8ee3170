"""Checkpoint engine of the training process (reference:
dlrover/trainer/torch/flash_checkpoint/engine.py:136-435, "ENGINE" below).

Same policy as the reference — who saves, non-blocking lock + all-rank vote
(skip, never wait), SAVE events to the agent, memory first / storage second on
load — and the same overridable hooks and attribute names, so subclasses
written against the reference keep working.  The byte movement behind
``save_state_dict_to_memory`` / ``get_state_dict_from_memory`` is the CUDA
path of ``ckpt_saver.SharedMemoryHandler``.

Additions (all optional, defaults keep reference behaviour):
* ``async_save=True`` (or FCP_ASYNC_SAVE=1): ``save_state_dict_to_memory``
  returns as soon as the pack kernel has snapshotted the tensors into HBM
  staging; a finisher thread waits for the DMA into shm, publishes the meta
  (writing_shm=False) and only then releases the shm lock, so the agent can
  never persist a torn image.
* ``get_state_dict_from_memory(into=..., device=...)``: restore straight into
  live tensors through the scatter kernel.
"""
import os
import threading
import time
from abc import ABCMeta, abstractmethod
from datetime import timedelta
from multiprocessing import Process
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from ..ckpt_saver import (
    DLROVER_CKPT_CONFIG_KEY,
    AsyncCheckpointSaver,
    CheckpointConfig,
    CheckpointEvent,
    CheckpointEventType,
    CheckpointSharedObjPrefix,
    SharedMemoryHandler,
)
from ..common import env_utils
from ..common.constants import CheckpointConstant
from ..common.log import default_logger as logger
from ..common.multi_process import SharedLock, SharedQueue
from ..common.serialize import ClassMeta
from ..common.singleton import Singleton
from ..common.storage import CheckpointStorage
from .replica import CkptReplicaManger


def _local_rank0_log(local_rank, message):
    if local_rank == 0:
        logger.info(message)


def _sync_device(group):
    """Device on which the tiny sync collectives of `group` must live."""
    if dist.get_backend(group) == "gloo":
        return "cpu"
    return "cuda:%d" % env_utils.get_local_rank()


class ReadyTensor(Singleton):
    def __init__(self, device) -> None:
        self.tensor = torch.tensor([0], dtype=torch.int32).to(device)


def check_all_rank_ready(group, ready):
    """True iff every rank of `group` passed ready=True (one 4-byte
    all_reduce; ENGINE:53-67)."""
    if not group and not dist.is_initialized():
        return ready
    rt = ReadyTensor.singleton_instance(_sync_device(group))
    rt.tensor[0] = 0 if ready else 1
    dist.all_reduce(rt.tensor, group=group)
    return bool(rt.tensor.item() == 0)


def verify_all_rank_step_consistent(group, step):
    """True iff every rank holds the same checkpoint step (ENGINE:70-91)."""
    if not group and not dist.is_initialized():
        return True
    device = _sync_device(group)
    mine = torch.tensor([float(step)]).to(device)
    world = group.size() if group else dist.get_world_size()
    gathered = [torch.zeros(1).to(device) for _ in range(world)]
    dist.all_gather(gathered, mine, group=group)
    return all(torch.equal(s, gathered[0]) for s in gathered)


def timer(func):
    def wrapper(*args, **kwargs):
        start = time.time()
        result = func(*args, **kwargs)
        logger.info("Local rank %s execute %s in %ss.", env_utils.get_local_rank(),
                    func.__name__, round(time.time() - start, 3))
        return result

    return wrapper


def start_async_save():
    AsyncCheckpointSaver.start_async_saving_ckpt()
    while True:
        time.sleep(60)


def start_saver_process():
    """Outside dlrover-run (ROLE_NAME != "dlrover-trainer") nobody hosts the
    saver, so local rank 0 forks a daemon for it (ENGINE:108-133).  Such a
    daemon dies with the trainer; under dlrover-run the agent hosts the saver
    and the shm survives worker failures."""
    if os.getenv("ROLE_NAME", "") != "dlrover-trainer" and env_utils.get_local_rank() == 0:
        p = Process(target=start_async_save, daemon=True)
        p.start()
        logger.info("Start a process to asynchronously save checkpoint.")
        return p
    return None


class CheckpointEngine(metaclass=ABCMeta):
    """
    Args:
        checkpoint_dir (str): root directory of the checkpoints.
        storage (CheckpointStorage): storage used for persisted checkpoints.
        comm_backend (str): backend of the sync group; default: that of the
            main process group.
        save_timeout (int): seconds the agent waits for all shards.
        replica_count (int): in-memory replicas on other nodes (0 = none).
        async_save (bool): see module docstring.
    """

    saver_proc = None

    def __init__(self, checkpoint_dir: str, storage: CheckpointStorage, comm_backend: str = "",
                 save_timeout: int = CheckpointConstant.SAVE_TIMEOUT, replica_count=0,
                 async_save=None, stripe_replicas=None):
        if not self.saver_proc:
            self.saver_proc = start_saver_process()
        self.checkpoint_dir = checkpoint_dir
        self.storage = storage
        self._save_timeout = save_timeout
        self._local_rank = env_utils.get_local_rank()
        self._cached_step = 0
        self._restart_count = env_utils.get_torch_restart_count()
        if async_save is None:
            async_save = os.getenv("FCP_ASYNC_SAVE", "0") == "1"
        self._async_save = async_save
        if stripe_replicas is None:
            stripe_replicas = os.getenv("FCP_STRIPE_REPLICAS", "0") == "1"
        self._stripe_replicas = bool(stripe_replicas)
        self._node_group = None   # gloo group of this node's ranks (striped saves)
        self._node_ranks: List[int] = []
        self._finisher: Optional[threading.Thread] = None
        # only local rank 0 talks to the agent's event queue
        self._event_queue = SharedQueue(
            name=CheckpointSharedObjPrefix.SAVE_STEP_QNAME + "0", create=False
        ) if self._local_rank == 0 else None
        self.local_shard_id = self._local_rank % self.get_local_shard_num()
        self._shm_lock = SharedLock(
            name=CheckpointSharedObjPrefix.SHM_LOCK_NAME + str(self.local_shard_id),
            create=False)
        self._shm_handler = SharedMemoryHandler(self.local_shard_id, host=False)
        self._rank = 0
        self._group_rank = 0
        self._world_size = 1
        self._loader_group = None
        self._saver_group = None
        self._saving_ranks: Optional[List[int]] = None
        self._init_sync_group(comm_backend)
        self._init_node_group()
        self._notify_agent_to_create_saver()
        self._update_saver_config()
        self._replica_manager = CkptReplicaManger.create_replica_manager(
            self.get_global_shard_num(), replica_count)

    def _init_sync_group(self, comm_backend):
        if not dist.is_initialized():
            self._saving_ranks = [0]
            return
        self._rank = dist.get_rank()
        self._group_rank = env_utils.get_group_rank()
        self._world_size = dist.get_world_size()
        default_backend = dist.get_backend()
        backend = comm_backend or default_backend
        self._loader_group = None if backend == default_backend else dist.new_group(
            backend=backend, timeout=timedelta(seconds=60))
        self._saving_ranks = self.get_saving_ranks()
        everyone = (self._saving_ranks is None
                    or len(self._saving_ranks) == dist.get_world_size())
        if backend == default_backend and everyone:
            self._saver_group = None
            _local_rank0_log(self._local_rank,
                             "Use the default process group to sync when saving checkpoint.")
        else:
            self._saver_group = dist.new_group(
                ranks=self._saving_ranks, backend=backend, timeout=timedelta(seconds=60))
            _local_rank0_log(
                self._local_rank,
                "Create a %s communication group to save checkpoint. Saving ranks are %s."
                % (backend, self._saving_ranks or "all ranks"))

    def _striping(self):
        """Striped saves: replicated state (one global shard), several ranks
        per node, CUDA.  Every local rank then writes 1/n of the ONE image of
        its node over its own host link."""
        return (self._stripe_replicas and self._node_group is not None
                and len(self._node_ranks) > 1 and torch.cuda.is_available())

    def _init_node_group(self):
        if not (self._stripe_replicas and dist.is_initialized()
                and self.get_global_shard_num() == 1):
            return
        per_node = env_utils.get_local_world_size()
        world = dist.get_world_size()
        if per_node < 2 or world % per_node:
            return
        # small CPU collectives (chunk sums, the pre-write barrier), also from
        # the finisher thread of an async save: gloo, apart from the NCCL group
        for node in range(world // per_node):
            ranks = list(range(node * per_node, (node + 1) * per_node))
            group = dist.new_group(ranks=ranks, backend="gloo", timeout=timedelta(seconds=300))
            if self._rank in ranks:
                self._node_group, self._node_ranks = group, ranks

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def close(self):
        """Finish any in-flight save and release the shm handler."""
        self.wait_async_save()
        self._shm_handler.close()

    def wait_async_save(self):
        """Block until the last async save is complete in shm."""
        t = self._finisher
        if t is not None:
            t.join()
            self._finisher = None

    def _notify_agent_to_create_saver(self):
        """Local rank 0 tells the agent which saver class to host."""
        if self._local_rank != 0:
            return
        saver_cls = self.get_saver_class()
        SharedQueue(name="factory").put(ClassMeta(
            module_path=saver_cls.__module__,
            class_name=saver_cls.__name__,
            kwargs={
                "checkpoint_dir": self.checkpoint_dir,
                "storage_meta": self.storage.get_class_meta(),
                "local_shard_num": self.get_local_shard_num(),
                "global_shard_num": self.get_global_shard_num(),
                "save_timeout": self._save_timeout,
            }))

    def _update_saver_config(self):
        if self._local_rank != 0:
            return
        if self._event_queue is None:
            raise ValueError("The event queue cannot be None on local rank 0.")
        self._event_queue.put(CheckpointEvent(
            type=CheckpointEventType.UPDATE_SHARD,
            global_shard_num=self.get_global_shard_num()))

    # -- save ------------------------------------------------------------------
    def save_state_dict_to_memory(self, state_dict, conf: CheckpointConfig):
        """Copy the state dict into this shard's shm.  Returns False (and does
        nothing) if this rank does not save, or if any saving rank could not
        get its shm lock because the agent is still persisting the previous
        checkpoint — training is never blocked on the saver (ENGINE:297-330).
        """
        if self._striping():
            return self._save_striped(state_dict, conf)
        if self._local_rank != self.local_shard_id:
            return False
        if self._saving_ranks and self._rank not in self._saving_ranks:
            return False
        self.wait_async_save()
        conf.rank = self._rank
        conf.group_rank = self._group_rank
        conf.world_size = self._world_size
        acquired = self._shm_lock.acquire(blocking=False)
        logger.info("%s acquired the lock of shared memory: %s.", self._local_rank, acquired)
        all_rank_ready = check_all_rank_ready(self._saver_group, acquired)
        if not all_rank_ready or not state_dict:
            logger.info("Rank %s skips saving to memory: the saver is still persisting "
                        "the previous checkpoint.", self._rank)
            if acquired:
                self._shm_lock.release()
            return False
        state_dict[DLROVER_CKPT_CONFIG_KEY] = conf
        if self._async_save:
            pending = self._shm_handler.save_state_dict_async(state_dict)
            pending.wait_packed()  # tensors are snapshotted: training may go on
            self._cached_step = conf.step

            def finish():
                try:
                    pending.wait()
                    self._replica_manager.backup(self._shm_handler)
                finally:
                    if acquired:
                        self._shm_lock.release()

            self._finisher = threading.Thread(target=finish, name="fcp-finisher", daemon=True)
            self._finisher.start()
            return True
        self._shm_handler.save_state_dict(state_dict)
        if acquired:
            self._shm_lock.release()
        self._cached_step = conf.step
        self._replica_manager.backup(self._shm_handler)
        return True

    def _save_striped(self, state_dict, conf: CheckpointConfig):
        """Replicated state, every local rank takes part (all of them must call
        save with the same state): the owner — the rank that would have saved
        alone (full_ckpt_engine.py:75-88) — takes the lock, owns the segment
        and the meta; every rank packs and copies only its stripe of the image
        into that one segment; the chunk sums meet on the owner, which
        publishes the complete meta and only then releases the lock.  The
        bytes in shm are identical to a single-writer save."""
        import numpy as np
        owner = (self._local_rank == self.local_shard_id
                 and (not self._saving_ranks or self._rank in self._saving_ranks))
        self.wait_async_save()
        owner_rank = self._node_ranks[self.local_shard_id]
        conf.rank = owner_rank
        conf.group_rank = self._group_rank
        conf.world_size = self._world_size
        acquired = self._shm_lock.acquire(blocking=False) if owner else True
        if not check_all_rank_ready(None, acquired) or not state_dict:
            logger.info("Rank %s skips saving to memory: the saver is still persisting "
                        "the previous checkpoint.", self._rank)
            if owner and acquired:
                self._shm_lock.release()
            return False
        state_dict[DLROVER_CKPT_CONFIG_KEY] = conf
        index, count = self._node_ranks.index(self._rank), len(self._node_ranks)
        group = self._node_group
        try:
            pending = self._shm_handler.save_state_dict_async(
                state_dict, stripe=(index, count, owner),
                attach_barrier=lambda: dist.barrier(group=group))
        except BaseException:
            if owner:
                self._shm_lock.release()
            raise

        def finish():
            try:
                sums = pending.wait_transfer()
                w0, w1 = pending.window
                part = (w0, w1, None if sums is None else np.asarray(sums[w0:w1]))
                parts = [None] * count if owner else None
                dist.gather_object(part, parts, dst=owner_rank, group=group)
                if owner:
                    merged = np.zeros(len(sums), dtype=np.uint32)
                    for a, b, values in parts:
                        merged[a:b] = values
                    pending.publish(merged)
                else:
                    pending.abandon(conf.step)
            finally:
                if owner:
                    self._shm_lock.release()

        pending.wait_packed()  # tensors are snapshotted: training may go on
        self._cached_step = conf.step
        if self._async_save:
            self._finisher = threading.Thread(target=finish, name="fcp-finisher", daemon=True)
            self._finisher.start()
        else:
            finish()
        return True

    # -- load ------------------------------------------------------------------
    def _load_from_peers(self, into, device):
        """Replicated state on a GPU job: stripes of the image come from the
        peers' HBM over NVLink, each stripe crossing the host link at most once
        per node (peer_restore.py).  -> (step, state_dict), or None when that
        path does not apply and the reference behaviour follows."""
        from . import peer_restore
        if not peer_restore.applicable(self, into, device):
            return None
        if into is not None and DLROVER_CKPT_CONFIG_KEY not in into:
            into = dict(into)
        try:
            got = peer_restore.restore_replicated(self._shm_handler, into, device)
        except Exception as e:  # noqa: BLE001
            from .._native import ChecksumMismatch
            if not isinstance(e, ChecksumMismatch):
                raise
            logger.error("checkpoint in memory is corrupt (%s); falling back to storage", e)
            self._memory_restore_wrecked_into = into is not None
            return 0, {}
        if got is None:
            return None
        step, state_dict = got
        state_dict.pop(DLROVER_CKPT_CONFIG_KEY, None)
        logger.info("Load checkpoint at step %s from memory (stripes over NVLink).", step)
        return step, state_dict

    def get_state_dict_from_memory(self, into=None, device=None):
        """(step, state_dict) from shm, ({} if shm holds nothing valid or the
        ranks disagree on the step; ENGINE:332-347).  `into` / `device` select
        the in-place GPU restore of SharedMemoryHandler.load_state_dict; a
        checksum mismatch counts as "nothing valid"."""
        self.wait_async_save()
        self._memory_restore_wrecked_into = False
        got = self._load_from_peers(into, device)
        if got is not None:
            return got
        self._restore_memory_from_replica()
        state_dict = {}
        config = self._shm_handler.get_checkpoint_config(CheckpointConfig())
        passed = verify_all_rank_step_consistent(self._loader_group, config.step)
        if passed and config.step > 0:
            if into is not None and DLROVER_CKPT_CONFIG_KEY not in into:
                into = dict(into)
            try:
                state_dict = self._shm_handler.load_state_dict(into=into, device=device)
            except Exception as e:  # noqa: BLE001
                from .._native import ChecksumMismatch
                if not isinstance(e, ChecksumMismatch):
                    raise
                logger.error("checkpoint in memory is corrupt (%s); falling back to storage", e)
                # the scatter kernel verifies while it writes: the tensors of
                # `into` now hold bytes of the corrupt image
                self._memory_restore_wrecked_into = into is not None
                return config.step, {}
            state_dict.pop(DLROVER_CKPT_CONFIG_KEY, None)
            logger.info("Load checkpoint at step %s from memory.", config.step)
        return config.step, state_dict

    def restore_device(self):
        """Where the framework adapters (Megatron, DeepSpeed) ask for restored
        tensors.  The reference hands them CPU views of the shm and the
        framework's ``load_state_dict`` then issues one pageable host->device
        copy per tensor (megatron.py:92-108, megatron_dist_ckpt.py:685-686,
        deepspeed_engine.py:148-158).  When the image in memory was written
        from CUDA tensors (it carries chunk checksums) and this process has a
        GPU, the adapters instead get fresh tensors on the current CUDA
        device, filled by one verified scatter launch, and the framework's
        copies become device-to-device.  None = reference behaviour (CPU
        views).  FCP_RESTORE_DEVICE=cpu|cuda overrides the choice."""
        choice = os.getenv("FCP_RESTORE_DEVICE", "auto").lower()
        if choice == "cpu" or not torch.cuda.is_available():
            return None
        device = torch.device("cuda", torch.cuda.current_device())
        if choice == "cuda":
            return device
        from ..ckpt_saver import FCP_CHECKSUM_KEY
        try:
            gpu_image = bool(self._shm_handler.metadata.get_keys([FCP_CHECKSUM_KEY]))
        except Exception:  # noqa: BLE001 - no agent, nothing in memory
            gpu_image = False
        return device if gpu_image else None

    def _raise_if_into_wrecked(self, restored):
        """After the storage fallback: if the in-place restore from memory
        failed its checksums (so `into` was partly overwritten) and storage had
        nothing either, the caller's live tensors are garbage — say so instead
        of returning an empty dict that reads as "no checkpoint, keep going"."""
        if getattr(self, "_memory_restore_wrecked_into", False) and not restored:
            raise RuntimeError(
                "the checkpoint in shared memory failed its CRC-32C check while it was being "
                "restored in place, and no checkpoint in storage could replace it: the tensors "
                "passed as into= were partly overwritten and must be re-initialised")

    def _restore_memory_from_replica(self):
        """If this node lost its shm (it was replaced), fetch the shard from a
        peer's memory before loading (ENGINE:349-367)."""
        if not self._replica_manager.has_replica():
            return
        self._shm_handler.init_shared_memory()
        byte_tensor, meta = self._replica_manager.gather(self._shm_handler)
        if byte_tensor is not None and meta and not self._shm_handler.shared_memory:
            shm_size = byte_tensor.size()[0]
            self._shm_handler.init_shared_memory(create=True, size=shm_size)
            self._shm_handler._wait_populated()
            torch.frombuffer(self._shm_handler.shared_memory.buf,
                             dtype=torch.uint8).copy_(byte_tensor)
            self._shm_handler.metadata.set(meta)
            logger.info("Restored the checkpoint shard (%s bytes) from the memory of "
                        "an alive node.", shm_size)
        dist.barrier()

    # -- hooks -----------------------------------------------------------------
    @abstractmethod
    def get_saving_ranks(self):
        pass

    @abstractmethod
    def get_saver_class(self):
        """The AsyncCheckpointSaver subclass the agent must host."""

    @abstractmethod
    def get_local_shard_num(self):
        """Number of shards (shm segments) on this node."""

    @abstractmethod
    def get_global_shard_num(self):
        """Number of shards over all nodes."""

    @abstractmethod
    def save_to_memory(self, step, state_dict, paths: Dict[str, str]):
        """Save into shm only.  `paths`: state name -> storage path the agent
        uses if it has to persist this checkpoint after a failure."""

    @abstractmethod
    def save_to_storage(self, step, state_dict, paths: Dict[str, str]):
        """Save into shm, then ask the agent to persist it asynchronously."""

    @abstractmethod
    def load(self, resume_path=""):
        """State dict from shm if valid, else from storage."""


class LocalShardsEngine(CheckpointEngine):
    """Common shape of the DeepSpeed and Megatron engines: there are
    ``get_global_shard_num()`` shards in the job, the first
    ``min(local_world_size, global_shard_num)`` local ranks of every node each
    own one, and after a successful save one rank per node tells its agent to
    persist (deepspeed_engine.py:61-135, megatron_engine.py:73-146)."""

    def get_saving_ranks(self):
        per_node = env_utils.get_local_world_size()
        shards = self.get_local_shard_num()
        return [r for r in range(dist.get_world_size()) if r % per_node < shards]

    def get_local_shard_num(self):
        return min(env_utils.get_local_world_size(), self.get_global_shard_num())

    @timer
    def save_to_memory(self, step, state_dict, paths):
        conf = CheckpointConfig(step=step, paths=paths)
        return self.save_state_dict_to_memory(state_dict, conf)

    def _notifies_agent(self):
        """Whether this rank puts the SAVE event after a save."""
        return self._local_rank == 0

    @timer
    def save_to_storage(self, step, state_dict, paths):
        success = True
        if step > self._cached_step:
            success = self.save_to_memory(step, state_dict, paths)
        if dist.is_initialized():
            dist.barrier()
        if success and self._notifies_agent():
            self._event_queue.put(CheckpointEvent(type=CheckpointEventType.SAVE, step=step))
        return success
