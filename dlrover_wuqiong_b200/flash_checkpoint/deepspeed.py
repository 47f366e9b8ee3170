"""DeepSpeedCheckpointer (reference:
dlrover/trainer/torch/flash_checkpoint/deepspeed.py:45-256).

DeepSpeed assembles its state dicts inside ``engine.save_checkpoint`` and
writes them with ``torch.save``; the adapter intercepts those calls to capture
``{model_states, optim_states}`` plus their paths and hands them to the flash
engine.  ``deepspeed`` itself is imported lazily, so this module is importable
(and testable with a mock engine) without it.
"""
import os
from typing import Dict

import torch.distributed as dist

from ..ckpt_saver import DeepSpeedCheckpointSaver
from ..common import env_utils
from ..common.constants import CheckpointConstant
from ..common.storage import CheckpointStorage, get_checkpoint_storage
from ._hijack import (  # noqa: F401  (torch_native_* re-exported, as the reference does)
    _read_file,
    redirected,
    torch_native_load,
    torch_native_save,
)
from .checkpointer import Checkpointer, StorageType
from .deepspeed_engine import DeepSpeedCheckpointEngine

_DS_MODEL_SD_FILE_SUFFIX = "model_states.pt"
_DS_OPTIM_SD_FILE_SUFFIX = "optim_states.pt"
_ZERO_STAGE_WEIGHTS = 3  # deepspeed ZeroStageEnum.weights


def _state_name(path, default=""):
    if path.endswith(_DS_MODEL_SD_FILE_SUFFIX):
        return CheckpointConstant.MODEL_STATES_NAME
    if path.endswith(_DS_OPTIM_SD_FILE_SUFFIX):
        return CheckpointConstant.OPTIM_STATES_NAME
    return default


class AsyncCheckpointAgent(object):
    """Stand-in for DeepSpeed's checkpoint engine / torch.save: remembers what
    DeepSpeed wanted to write where.

    Attributes:
        state_dict: state name -> object DeepSpeed passed to torch.save.
        paths: state name -> path it was meant for.
    """

    def __init__(self, storage: CheckpointStorage):
        self.state_dict: Dict[str, object] = {}
        self.paths: Dict[str, str] = {}
        self.storage = storage

    def create(self, tag):
        pass

    def save(self, state_dict, path: str):
        if not isinstance(path, str):  # file objects: not a checkpoint shard
            torch_native_save(state_dict, path)
            return
        name = _state_name(path, default=path.split("/")[-1])
        if name:
            self.state_dict[name] = state_dict
            self.paths[name] = path

    def load(self, path: str, map_location=None, **kwargs):
        name = _state_name(path)
        if name in self.state_dict:
            return self.state_dict[name]
        return self.storage.read_state_dict(
            path, lambda p: _read_file(p, map_location))

    def commit(self, tag):
        pass


class DeepSpeedCheckpointer(Checkpointer):
    """
    Example::

        engine = deepspeed.initialize(...)
        checkpointer = DeepSpeedCheckpointer(engine, save_dir)
        if step % 10 == 0:
            checkpointer.save_checkpoint(save_dir, tag, storage_type=StorageType.MEMORY)
        if step % 100 == 0:
            checkpointer.save_checkpoint(save_dir, tag, storage_type=StorageType.DISK)
    """

    def __init__(self, engine, checkpoint_dir, comm_backend="", deletion_strategy=None,
                 save_timeout=CheckpointConstant.SAVE_TIMEOUT):
        self.engine = engine
        self.checkpoint_dir = checkpoint_dir
        global_shard_num = 1
        if self.engine.zero_optimization():
            global_shard_num = dist.get_world_size(self.engine.optimizer.dp_process_group)
        zero_stage = self.engine.zero_optimization_stage()
        self.storage = get_checkpoint_storage(deletion_strategy)
        self._async_save_engine = DeepSpeedCheckpointEngine(
            checkpoint_dir, storage=self.storage, global_shard_num=global_shard_num,
            zero_stage=zero_stage, comm_backend=comm_backend, save_timeout=save_timeout)
        self._ckpt_agent = AsyncCheckpointAgent(self._async_save_engine.storage)
        self._local_rank = env_utils.get_local_rank()
        self._ds_tracer_file = os.path.join(self.checkpoint_dir,
                                            DeepSpeedCheckpointSaver.TRACER_FILE)
        self._dlrover_tracer_file = os.path.join(self.checkpoint_dir,
                                                 CheckpointConstant.TRACER_FILE_NAME)
        # below ZeRO-3 the module weights are whole on every rank: local rank 0
        # of each node saves them
        if zero_stage < _ZERO_STAGE_WEIGHTS and self._local_rank == 0:
            self.engine.save_non_zero_checkpoint = True

    def save_checkpoint(self, save_dir, tag=None, client_state={}, save_latest=True,
                        storage_type=StorageType.DISK):
        if storage_type == StorageType.MEMORY:
            return self._save_checkpoint_to_memory(save_dir, tag, client_state, save_latest)
        if storage_type == StorageType.DISK:
            return self._save_checkpoint_to_storage(save_dir, tag, client_state, save_latest)
        raise ValueError("No support storage type %s" % storage_type)

    def _capture(self, save_dir, tag, client_state, save_latest):
        with redirected(save=self._ckpt_agent.save):
            self.engine.save_checkpoint(save_dir, tag, client_state, save_latest)

    def _save_checkpoint_to_memory(self, save_dir, tag=None, client_state={}, save_latest=True):
        self._capture(save_dir, tag, client_state, save_latest)
        ok = self._async_save_engine.save_to_memory(
            tag, self._ckpt_agent.state_dict, self._ckpt_agent.paths)
        self._update_tracer_file(tag)
        return ok

    def _update_tracer_file(self, tag):
        """engine.save_checkpoint created the tag folder and moved DeepSpeed's
        "latest" file although nothing was written: undo both so "latest"
        keeps naming the last checkpoint that really is in storage."""
        if self.engine.global_rank != 0:
            return
        self.storage.safe_rmtree(os.path.join(self.checkpoint_dir, str(tag)))
        content = self.storage.read(self._dlrover_tracer_file)
        if content:
            self.storage.write(content, self._ds_tracer_file)
        else:
            self.storage.safe_remove(self._ds_tracer_file)

    def _save_checkpoint_to_storage(self, save_dir, tag=None, client_state={}, save_latest=True):
        self._capture(save_dir, tag, client_state, save_latest)
        return self._async_save_engine.save_to_storage(
            tag, self._ckpt_agent.state_dict, self._ckpt_agent.paths)

    def load_checkpoint(self, load_dir, tag=None, load_module_strict=True,
                        load_optimizer_states=True, load_lr_scheduler_states=True,
                        load_module_only=False, custom_load_fn=None):
        """Same arguments as DeepSpeedEngine.load_checkpoint; state that is
        still in shared memory is served from there."""
        # GPU jobs: tensors arrive on the current CUDA device through the
        # scatter kernel instead of as CPU views DeepSpeed would copy up one
        # by one (deepspeed.py:240-256 in the reference)
        self._ckpt_agent.state_dict = self._async_save_engine.load(
            device=self._async_save_engine.restore_device())
        with redirected(load=self._ckpt_agent.load):
            return self.engine.load_checkpoint(
                load_dir=load_dir, tag=tag, load_module_strict=load_module_strict,
                load_optimizer_states=load_optimizer_states,
                load_lr_scheduler_states=load_lr_scheduler_states,
                load_module_only=load_module_only, custom_load_fn=custom_load_fn)

    def close(self):
        self._async_save_engine.close()
