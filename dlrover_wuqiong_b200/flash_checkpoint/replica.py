"""In-memory replicas of checkpoint shards and recovery from a peer
(reference: dlrover/trainer/torch/flash_checkpoint/replica.py:28-350, "REPL").

Two transports:

* host transport (reference behaviour, kept for multi-node backup groups):
  the shm bytes travel as CPU uint8 tensors over a gloo group;
* device transport (new): the *packed image that the last save left in HBM
  staging* travels over the NCCL group — NVLink / NVSwitch inside a node —
  straight from the surviving replica's HBM into the replaced rank's HBM,
  where the scatter kernel verifies the CRC-32C of every chunk and unpacks it
  into the live tensors (``FullCkptReplicaManager.recover_on_device``).  Host
  memory is not touched.

``replica_count == 0`` means "no replicas": no group is created and every
method is a no-op (the reference divides by zero for sharded engines,
REPL:82,107).
"""
from abc import ABCMeta, abstractmethod
from typing import Dict, List

import torch
import torch.distributed as dist

from ..ckpt_saver import (
    DLROVER_CKPT_CONFIG_KEY,
    FCP_CHECKSUM_KEY,
    CheckpointConfig,
    SharedMemoryHandler,
)
from ..common import env_utils
from ..common.log import default_logger as logger


def _shm_bytes(shm_handler):
    """The handler's shm as a CPU uint8 tensor (empty if it has none)."""
    if shm_handler is not None and shm_handler.shared_memory:
        return torch.frombuffer(shm_handler.shared_memory.buf, dtype=torch.uint8)
    return torch.empty(0, dtype=torch.uint8)


class CkptReplicaManger(metaclass=ABCMeta):
    def __init__(self, replica_count) -> None:
        self.replica_count = replica_count
        self.local_rank = env_utils.get_local_rank()
        self.local_world_size = env_utils.get_local_world_size()
        self.node_rank = env_utils.get_node_rank()
        self.node_num = env_utils.get_node_num()
        self.current_device = torch.device("cpu")
        self._rank_shms: Dict[int, SharedMemoryHandler] = {}
        self.backup_ranks: List[int] = []
        self._backup_group = None
        self.rank = dist.get_rank() if dist.is_initialized() else env_utils.get_rank()

    @staticmethod
    def create_replica_manager(shard_num, replica_count):
        if shard_num == 1:
            return FullCkptReplicaManager(replica_count)
        return ShardCkptReplicaManager(replica_count)

    def has_replica(self):
        return self.replica_count > 0

    def _new_host_group(self):
        if dist.is_initialized() and self.replica_count > 0:
            self._backup_group = dist.new_group(backend="gloo", ranks=self.backup_ranks)

    def _get_max_tensor_size(self, byte_tensor):
        """Largest shm among the ranks of the backup group."""
        mine = torch.tensor([byte_tensor.numel()], dtype=torch.long)
        sizes = [torch.zeros(1, dtype=torch.long) for _ in self.backup_ranks]
        dist.all_gather(sizes, mine, group=self._backup_group)
        return int(max(int(s) for s in sizes))

    @abstractmethod
    def backup(self, shm_handler: SharedMemoryHandler):
        """Give the peers of the backup group a copy of this rank's shard."""

    @abstractmethod
    def gather(self, shm_handler: SharedMemoryHandler):
        """Fetch this rank's shard from a peer: (uint8 CPU tensor, meta)."""


class ShardCkptReplicaManager(CkptReplicaManger):
    """Sharded state (FSDP / TP / ZeRO): the same local rank on
    `replica_count` consecutive nodes form a backup group and keep each
    other's shards in extra local shm segments."""

    def __init__(self, replica_count=0) -> None:
        super().__init__(replica_count)
        self.backup_ranks = self._get_backup_ranks(replica_count)
        self._new_host_group()

    def _get_backup_ranks(self, replica_count):
        """E.g. 2 ranks per node, groups of 3 nodes: local rank 0 -> {0, 2, 4},
        local rank 1 -> {1, 3, 5}."""
        if replica_count <= 0:
            return []
        first_node = (self.node_rank // replica_count) * replica_count
        return [(first_node + i) * self.local_world_size + self.local_rank
                for i in range(replica_count)]

    def backup(self, shm_handler: SharedMemoryHandler):
        if self.replica_count == 0:
            return
        assert shm_handler.shared_memory is not None
        meta = shm_handler.metadata.get()
        tensors, metas = self._gather_peer_ckpt(shm_handler.shared_memory.buf, meta)
        self._rank_shms[self.rank] = shm_handler
        self._write_peer_ckpt_to_shm(tensors, metas)

    def _gather_peer_ckpt(self, buffer, meta_data):
        mine = torch.frombuffer(buffer, dtype=torch.uint8) if len(buffer) \
            else torch.empty(0, dtype=torch.uint8)
        max_size = self._get_max_tensor_size(mine)
        padded = torch.zeros(max_size, dtype=torch.uint8)
        padded[:mine.numel()] = mine
        group_size = dist.get_world_size(group=self._backup_group)
        outputs = [torch.empty(max_size, dtype=torch.uint8) for _ in range(group_size)]
        dist.all_gather(outputs, padded, group=self._backup_group)
        metas = [None] * group_size
        dist.all_gather_object(metas, meta_data, group=self._backup_group)
        return outputs, metas

    def _write_peer_ckpt_to_shm(self, shm_tensors, ckpt_metas):
        for shm_tensor, meta in zip(shm_tensors, ckpt_metas):
            if not meta:
                continue
            config: CheckpointConfig = meta[DLROVER_CKPT_CONFIG_KEY]
            if config.rank == self.rank:
                continue
            handler = self._rank_shms.get(config.rank)
            if handler is None:
                handler = SharedMemoryHandler(local_rank=config.rank)
                handler.init_shared_memory(create=True, size=shm_tensor.numel())
                handler._wait_populated()
                self._rank_shms[config.rank] = handler
            local = torch.frombuffer(handler.shared_memory.buf, dtype=torch.uint8)
            local.copy_(shm_tensor[:local.numel()])
            handler.metadata.set(meta)

    def gather(self, shm_handler: SharedMemoryHandler):
        """One all_gather round per member of the group; in its own round a
        rank picks the first copy of its shard that any peer still holds."""
        if self.replica_count == 0:
            return None, None
        handlers = {}
        for rank in self.backup_ranks:
            if rank == self.rank:
                handlers[rank] = shm_handler
            else:
                h = SharedMemoryHandler(local_rank=rank)
                h.init_shared_memory()
                handlers[rank] = h
        return self._gather_owner_checkpoint(handlers)

    def _gather_owner_checkpoint(self, shm_handlers):
        """`shm_handlers` may be a dict keyed by rank or a list in
        backup_ranks order."""
        found_tensor, found_meta = None, {}
        for i, rank in enumerate(self.backup_ranks):
            handler = shm_handlers[rank] if isinstance(shm_handlers, dict) else shm_handlers[i]
            if handler.shared_memory:
                buffer = handler.shared_memory.buf
                meta = handler.metadata.get()
            else:
                buffer, meta = memoryview(b""), {}
            tensors, metas = self._gather_peer_ckpt(buffer, meta)
            if rank != self.rank:
                continue
            for t, m in zip(tensors, metas):
                if m:
                    found_tensor, found_meta = t, m
                    break
        return found_tensor, found_meta


class FullCkptReplicaManager(CkptReplicaManger):
    """Replicated state (DDP): every node's shm already is a replica, so
    backup is a no-op and recovery is a broadcast from any rank that still
    has the checkpoint."""

    def __init__(self, replica_count=0) -> None:
        super().__init__(replica_count)
        self.backup_ranks = self._get_backup_ranks()
        self._new_host_group()

    def _get_backup_ranks(self):
        return [node * self.local_world_size for node in range(self.node_num)]

    def backup(self, shm_handler: SharedMemoryHandler):
        pass

    def _pick_source(self, have, group):
        """all_gather one flag per rank; first rank that has the data."""
        n = dist.get_world_size(group=group)
        dev = have.device
        flags = [torch.zeros(1, dtype=torch.int8, device=dev) for _ in range(n)]
        dist.all_gather(flags, have, group=group)
        for i, f in enumerate(flags):
            if int(f) == 1:
                return i
        return None

    def gather(self, shm_handler: SharedMemoryHandler):
        """Host transport (REPL:275-332): broadcast shm bytes + meta from the
        first node rank-0 that still holds them."""
        if self.replica_count == 0 or self.rank not in self.backup_ranks:
            return None, None
        mine = _shm_bytes(shm_handler)
        have = torch.tensor([1 if mine.numel() else 0], dtype=torch.int8)
        src_index = self._pick_source(have, self._backup_group)
        if src_index is None:
            return None, None
        src_rank = self.backup_ranks[src_index]
        max_size = self._get_max_tensor_size(mine)
        if max_size == 0:
            return None, None
        payload = torch.zeros(max_size, dtype=torch.uint8)
        payload[:mine.numel()] = mine
        dist.broadcast(payload, src=src_rank, group=self._backup_group)
        metas = [shm_handler.metadata.get()]
        dist.broadcast_object_list(metas, src=src_rank, group=self._backup_group)
        return payload, metas[0]

    def recover_fused(self, shm_handler, into, group=None, healthy=True):
        """Like recover_on_device, but without any collective on the data:
        the survivor exports its staging image with CUDA IPC, every replaced
        rank maps it and runs the scatter kernel straight on the PEER pointer,
        so the bytes cross NVLink inside the kernel that verifies and unpacks
        them (transfer fused with unpack; no receive buffer).  Same-node
        ranks only.  Returns the restored step (0 if nobody had an image)."""
        if not dist.is_initialized():
            return 0
        from .. import _native
        engine = getattr(shm_handler, "_engine", None)
        staging = engine.staging if (healthy and engine is not None) else None
        meta = shm_handler.metadata.get(local=True) if healthy else {}
        has_image = staging is not None and bool(meta) and FCP_CHECKSUM_KEY in meta
        dev = torch.device("cuda", torch.cuda.current_device())
        src_index = self._pick_source(
            torch.tensor([1 if has_image else 0], dtype=torch.int8, device=dev), group)
        if src_index is None:
            return 0
        src_rank = dist.get_global_rank(group, src_index) if group else src_index
        me = dist.get_rank()
        if me == src_rank:
            torch.cuda.synchronize(dev)  # the image is complete
            handle, offset = _native.ipc_export(staging.data_ptr())
            box = [(meta, handle, offset, _meta_total_bytes(meta))]
        else:
            box = [None]
        dist.broadcast_object_list(box, src=src_rank, group=group)
        meta, handle, offset, total = box[0]
        step = meta[DLROVER_CKPT_CONFIG_KEY].step
        if me != src_rank:
            import time
            t0 = time.perf_counter()
            with torch.cuda.device(dev):
                base = _native.ipc_open(handle)  # the survivor's HBM, mapped here
            t1 = time.perf_counter()
            try:
                shm_handler.restore_from_image(
                    meta, _native.RawDeviceImage(base + offset, total), into)
            finally:
                torch.cuda.synchronize(dev)
                t2 = time.perf_counter()
                _native.ipc_close(base)
            self.last_fused_stats = {"ipc_open_s": t1 - t0, "scatter_over_nvlink_s": t2 - t1,
                                     "ipc_close_s": time.perf_counter() - t2, "bytes": total}
            logger.info("rank %s restored step %s straight from the HBM of rank %s",
                        me, step, src_rank)
        dist.barrier(group=group)  # the survivor keeps the image alive until here
        return step

    def recover_on_device(self, shm_handler, into, group=None, healthy=True):
        """Device transport: every rank of `group` (default: the main group,
        NCCL on GPU jobs) calls this with a state dict `into` of the
        checkpoint's structure.  Ranks that still hold the packed image of
        their last save in HBM staging (`healthy`) offer it; the first one
        broadcasts image + meta; every other rank scatters the received image
        into `into` on its own device, verifying the per-chunk CRC-32C.
        Returns the restored step (0 if nobody had an image)."""
        if not dist.is_initialized():
            return 0
        backend = dist.get_backend(group)
        engine = getattr(shm_handler, "_engine", None)
        staging = engine.staging if (healthy and engine is not None) else None
        meta = shm_handler.metadata.get(local=True) if healthy else {}
        has_image = staging is not None and bool(meta) and FCP_CHECKSUM_KEY in meta
        if backend == "gloo":
            dev = torch.device("cpu")
        else:
            dev = torch.device("cuda", torch.cuda.current_device())
        src_index = self._pick_source(
            torch.tensor([1 if has_image else 0], dtype=torch.int8, device=dev), group)
        if src_index is None:
            return 0
        src_rank = dist.get_global_rank(group, src_index) if group else src_index
        me = dist.get_rank()
        box = [meta if me == src_rank else None]
        dist.broadcast_object_list(box, src=src_rank, group=group)
        meta = box[0]
        total = _meta_total_bytes(meta)
        if me == src_rank:
            image = staging[:total] if dev.type == "cuda" else staging[:total].cpu()
        else:
            image = torch.empty(total, dtype=torch.uint8, device=dev)
        dist.broadcast(image, src=src_rank, group=group)
        step = meta[DLROVER_CKPT_CONFIG_KEY].step
        if me == src_rank:
            return step
        shm_handler.restore_from_image(meta, image, into)
        logger.info("rank %s restored step %s from the HBM image of rank %s", me, step, src_rank)
        return step


def _meta_total_bytes(meta):
    from ..plan import TensorMeta, traverse_state_dict
    end = 0

    def visit(m):
        nonlocal end
        if isinstance(m, TensorMeta):
            end = max(end, m.offset + m.numel * m.element_size)
        return m

    traverse_state_dict(meta, visit)
    return end
