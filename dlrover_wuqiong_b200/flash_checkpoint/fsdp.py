"""FSDP checkpointers (reference:
dlrover/trainer/torch/flash_checkpoint/fsdp.py:36-289).

``FsdpShardCheckpointer`` saves the sharded state dict of every rank with
torch.distributed.checkpoint into the rank's shm segment;
``FsdpFullCheckpointer`` gathers full state dicts and saves them like a DDP
checkpoint.  ``FsdpCheckpointer`` (the name BASELINE.json uses) is an alias of
the sharded one.
"""
import os

import torch.distributed as dist
import torch.distributed.checkpoint as dist_cp
from torch.distributed.checkpoint.optimizer import load_sharded_optimizer_state_dict
from torch.distributed.fsdp import FullStateDictConfig
from torch.distributed.fsdp import FullyShardedDataParallel as FSDP
from torch.distributed.fsdp import (
    ShardedOptimStateDictConfig,
    ShardedStateDictConfig,
    StateDictType,
)
from torch.distributed.fsdp.api import FullOptimStateDictConfig

from ..common.constants import CheckpointConstant
from ..common.storage import get_checkpoint_storage
from .checkpointer import Checkpointer, StorageType
from .fsdp_engine import FsdpCheckpointEngine
from .full_ckpt_engine import FullCheckpointEngine


class FsdpShardCheckpointer(Checkpointer):
    """
    Example::

        checkpointer = FsdpShardCheckpointer(checkpoint_dir)
        extra_sd = {"step": step, "epoch": epoch}
        checkpointer.save_checkpoint(step, model, optimizer, extra_sd,
                                     storage_type=StorageType.MEMORY)
        ...
        extra = checkpointer.load_checkpoint(model, optimizer)
    """

    def __init__(self, checkpoint_dir: str, comm_backend="", deletion_strategy=None,
                 save_timeout=CheckpointConstant.SAVE_TIMEOUT, async_save=None):
        """`async_save` (addition; default from FCP_ASYNC_SAVE): on the GPU
        path save_checkpoint returns once the shards are snapshotted in HBM
        staging; the transfer into shm finishes behind the next steps."""
        self.checkpoint_dir = checkpoint_dir
        self.storage = get_checkpoint_storage(deletion_strategy)
        self._engine = FsdpCheckpointEngine(checkpoint_dir, self.storage, comm_backend,
                                            save_timeout, async_save=async_save)

    def wait_latest_checkpoint(self):
        """Block until the last (async) save is complete in shared memory."""
        self._engine.wait_async_save()

    def save_checkpoint(self, step, model, optimizer, extra_sd={}, path="",
                        storage_type=StorageType.DISK):
        """
        Args:
            step (int): the iteration step.
            model: an FSDP module.
            optimizer: its optimizer.
            extra_sd (dict): extra entries stored beside "model" and "optim".
            path (str): directory of this step's checkpoint
                (default ``<checkpoint_dir>/<step>``).
            storage_type: StorageType.MEMORY or StorageType.DISK.
        """
        with self._sharded_state(model):
            state_dict = {"model": model.state_dict(),
                          "optim": FSDP.optim_state_dict(model, optimizer)}
            state_dict.update(extra_sd)
            if not path:
                path = os.path.join(self.checkpoint_dir, str(step))
            paths = {CheckpointConstant.MODEL_STATES_NAME: path}
            if storage_type == StorageType.MEMORY:
                return self._engine.save_to_memory(step, state_dict, paths)
            if storage_type == StorageType.DISK:
                # (the reference passes `path` here, which only works when a
                # MEMORY save of the same step ran first; fsdp.py:114)
                return self._engine.save_to_storage(step, state_dict, paths)
            raise ValueError("No support storage type %s" % storage_type)

    @staticmethod
    def _sharded_state(model):
        """SHARDED_STATE_DICT as the reference sets it (fsdp.py:100-103), except
        that a model living on a GPU keeps its optimizer shards there: torch's
        default offloads them to the host with one pageable copy per tensor,
        and the pack kernel would have to bring them back."""
        on_gpu = any(p.is_cuda for p in model.parameters())
        if not on_gpu:
            return FSDP.state_dict_type(model, StateDictType.SHARDED_STATE_DICT)
        return FSDP.state_dict_type(
            model, StateDictType.SHARDED_STATE_DICT,
            state_dict_config=ShardedStateDictConfig(offload_to_cpu=False),
            optim_state_dict_config=ShardedOptimStateDictConfig(offload_to_cpu=False))

    def load_checkpoint(self, model, optimizer, resume_path=""):
        with FSDP.state_dict_type(model, StateDictType.SHARDED_STATE_DICT):
            # the optimizer state cannot be planned together with the model's
            state_dict = {"model": model.state_dict(), "step": 0}
            storage_reader = self._engine.load(resume_path)
            if not storage_reader:
                return {}
            dist_cp.load(state_dict, storage_reader=storage_reader,
                         no_dist=not dist.is_initialized())
            model_sd = state_dict.pop("model", None)
            if model_sd:
                model.load_state_dict(model_sd)
            optim_state = load_sharded_optimizer_state_dict(
                model_state_dict=model_sd, optimizer_key="optim",
                storage_reader=storage_reader)
            optim_sd = optim_state.pop("optim", None)
            if optim_sd:
                optimizer.load_state_dict(
                    FSDP.optim_state_dict_to_load(model, optimizer, optim_sd))
            return state_dict

    def close(self):
        self._engine.close()


FsdpCheckpointer = FsdpShardCheckpointer


class FsdpFullCheckpointer(Checkpointer):
    """Saves the FULL (gathered) model and optimizer state dicts of an FSDP
    module, one replica per node, like a DDP checkpoint.

    Example::

        checkpointer = FsdpFullCheckpointer("/tmp/checkpoint/")
        checkpointer.save_checkpoint(step, model, optimizer, {"epoch": 10}, path)
        extra = checkpointer.load_checkpoint(model, optimizer)
    """

    def __init__(self, checkpoint_dir: str, comm_backend="", deletion_strategy=None):
        self.checkpoint_dir = checkpoint_dir
        self._rank = dist.get_rank() if dist.is_initialized() else 0
        self.storage = get_checkpoint_storage(deletion_strategy)
        self._engine = FullCheckpointEngine(
            checkpoint_dir=checkpoint_dir, storage=self.storage, local_shard_num=1,
            global_shard_num=1, comm_backend=comm_backend)

    @staticmethod
    def _full_state(model):
        return FSDP.state_dict_type(
            model, StateDictType.FULL_STATE_DICT, FullStateDictConfig(rank0_only=False),
            FullOptimStateDictConfig(rank0_only=False))

    def save_checkpoint(self, step, model, optimizer, extra_sd={}, path="",
                        storage_type=StorageType.DISK):
        if path == "":
            path = os.path.join(self.checkpoint_dir, "%s/rank_%s.pt" % (step, self._rank))
        with self._full_state(model):
            msd = model.state_dict()
            osd = FSDP.optim_state_dict(model, optimizer)
        state_dict = {"model": msd, "optimizer": osd}
        state_dict.update(extra_sd)
        state_dict = {CheckpointConstant.MODEL_STATES_NAME: state_dict}
        paths = {CheckpointConstant.MODEL_STATES_NAME: path}
        if storage_type == StorageType.MEMORY:
            return self._engine.save_to_memory(step, state_dict, paths)
        if storage_type == StorageType.DISK:
            if not path:
                raise ValueError("path cannot be empty if storage type is disk!")
            if self._rank == 0:
                self.storage.safe_rmtree(os.path.dirname(path))
            return self._engine.save_to_storage(step, state_dict, paths)
        raise ValueError("No support storage type %s" % storage_type)

    def load_checkpoint(self, model, optimizer, resume_path=""):
        state_dict = self._engine.load(resume_path)
        if not state_dict:
            return {}
        model_state_dict = state_dict.pop("model", {})
        optim_state_dict = state_dict.pop("optimizer", {})
        with self._full_state(model):
            optim_state_dict = FSDP.optim_state_dict_to_load(
                model=model, optim=optimizer, optim_state_dict=optim_state_dict)
        model.load_state_dict(model_state_dict)
        optimizer.load_state_dict(optim_state_dict)
        return state_dict

    def close(self):
        self._engine.close()
