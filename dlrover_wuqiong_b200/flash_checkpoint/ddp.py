"""DdpCheckpointer (reference: dlrover/trainer/torch/flash_checkpoint/ddp.py:25-117)."""
import os

import torch.distributed as dist

from ..common.constants import CheckpointConstant
from ..common.storage import get_checkpoint_storage
from .checkpointer import Checkpointer, StorageType
from .full_ckpt_engine import FullCheckpointEngine


class DdpCheckpointer(Checkpointer):
    """Flash checkpointer for a DDP model or any state dict with full weights.

    Args:
        checkpoint_dir: directory of the checkpoints.
        local_shard_num (int): shards per node (1 for replicated DDP; the
            number of local ranks if every rank holds its own partition).
        global_shard_num (int): shards over all nodes.
        comm_backend (str): backend of the sync group (default: main group's).
        deletion_strategy: optional CheckpointDeletionStrategy.
        save_timeout (int): seconds node 0 waits for all shards to persist.
        replica_count (int): in-memory replicas on other nodes.
        async_save (bool): return from save as soon as the tensors are
            snapshotted in HBM (extension; default off = reference semantics).
        stripe_replicas (bool): replicated state (global_shard_num == 1) on a
            GPU job: EVERY local rank calls save_checkpoint and writes 1/n of
            the node's one image over its own host link, instead of local rank
            0 writing all of it (extension; default off, or FCP_STRIPE_REPLICAS=1).
            Restores of replicated state fetch 1/n per rank and exchange the
            stripes over NVLink whenever ``load_checkpoint`` is given ``into=``
            or ``device=`` (FCP_PEER_RESTORE=0 switches that off).

    Example::

        checkpointer = DdpCheckpointer("/tmp/checkpoint/")
        for step, data in enumerate(dataloader):
            ...
            if step % 5 == 0:
                checkpointer.save_checkpoint(step, sd, path, storage_type=StorageType.MEMORY)
            elif step % 100 == 0:
                checkpointer.save_checkpoint(step, sd, path, storage_type=StorageType.DISK)
        sd = checkpointer.load_checkpoint()
    """

    def __init__(self, checkpoint_dir: str, local_shard_num=1, global_shard_num=1,
                 comm_backend="", deletion_strategy=None,
                 save_timeout=CheckpointConstant.SAVE_TIMEOUT, replica_count=0,
                 async_save=None, stripe_replicas=None):
        self.checkpoint_dir = checkpoint_dir
        self._rank = dist.get_rank() if dist.is_initialized() else 0
        self.storage = get_checkpoint_storage(deletion_strategy)
        self._engine = FullCheckpointEngine(
            checkpoint_dir=checkpoint_dir, storage=self.storage,
            local_shard_num=local_shard_num, global_shard_num=global_shard_num,
            comm_backend=comm_backend, save_timeout=save_timeout,
            replica_count=replica_count, async_save=async_save,
            stripe_replicas=stripe_replicas)

    def save_checkpoint(self, step, state_dict, path="", storage_type=StorageType.DISK):
        if path == "":
            path = os.path.join(self.checkpoint_dir, "%s/rank_%s.pt" % (step, self._rank))
        wrapped = {CheckpointConstant.MODEL_STATES_NAME: state_dict}
        paths = {CheckpointConstant.MODEL_STATES_NAME: path}
        if storage_type == StorageType.MEMORY:
            return self._engine.save_to_memory(step, wrapped, paths)
        if storage_type == StorageType.DISK:
            if not path:
                raise ValueError("path cannot be empty if storage type is disk!")
            return self._engine.save_to_storage(step, wrapped, paths)
        raise ValueError("No support storage type %s" % storage_type)

    def load_checkpoint(self, resume_path="", into=None, device=None):
        """`into`: the live state dict (same structure as the saved one) to
        restore in place on the GPU; `device`: allocate fresh tensors there.
        Default: the reference's CPU views."""
        if into is not None:
            into = {CheckpointConstant.MODEL_STATES_NAME: into}
        return self._engine.load(resume_path, into=into, device=device)

    def wait_latest_checkpoint(self):
        """Block until the last (async) save is complete in shared memory."""
        self._engine.wait_async_save()

    def close(self):
        self._engine.close()
