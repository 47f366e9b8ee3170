"""Megatron-LM flash checkpoint: drop-in ``save_checkpoint`` /
``load_checkpoint`` and the ``MegatronCheckpointer`` behind them (reference:
dlrover/trainer/torch/flash_checkpoint/megatron.py:54-247).

Megatron's own ``megatron.checkpointing.save_checkpoint`` builds the state
dict (args, model shards, optimizer, rng states) and calls ``torch.save(obj,
".../model_optim_rng.pt")``; while intercepted, that object goes to the flash
engine.  ``megatron`` is imported lazily so the module works (and is testable
with patched hooks) without it.
"""
import inspect
import os

import torch.distributed as dist

from ..ckpt_saver import MegatronCheckpointSaver
from ..common.constants import CheckpointConstant
from ..common.log import default_logger as logger
from ..common.singleton import Singleton
from ..common.storage import PosixDiskStorage
from ._hijack import _read_file, redirected, torch_native_load, torch_native_save
from .checkpointer import StorageType
from .megatron_engine import MegatronCheckpointEngine

try:
    from megatron import get_args
    from megatron.checkpointing import load_checkpoint as megatron_load
    from megatron.checkpointing import save_checkpoint as megatron_save
except ImportError:
    get_args = megatron_load = megatron_save = None
    logger.debug("megatron.checkpointing is not importable; patch get_args / "
                 "megatron_save / megatron_load before use")

_MODEL_SD_NAME = "model_optim_rng.pt"
_DIST_OPTIM_SD_NAME = "distrib_optim.pt"


def _get_rank():
    return dist.get_rank() if dist.is_initialized() else 0


def _state_name(path):
    if path.endswith(_MODEL_SD_NAME):
        return CheckpointConstant.MODEL_STATES_NAME
    if path.endswith(_DIST_OPTIM_SD_NAME):
        return CheckpointConstant.OPTIM_STATES_NAME
    return ""


class MegatronCheckpointer(Singleton):
    def __init__(self, checkpoint_dir, storage=None, comm_backend="",
                 save_timeout=CheckpointConstant.SAVE_TIMEOUT, replica_count=0):
        self.state_dict = {}
        self.paths = {}
        self._restored = {}  # state names of the last memory restore not yet handed out
        self.checkpoint_dir = checkpoint_dir
        self.storage = storage if storage else PosixDiskStorage()
        self.engine = MegatronCheckpointEngine(
            checkpoint_dir=checkpoint_dir, storage=self.storage, comm_backend=comm_backend,
            save_timeout=save_timeout, replica_count=replica_count)

    def save(self, state_dict, path: str):
        """torch.save stand-in."""
        if not isinstance(path, str):
            torch_native_save(state_dict, path)
            return
        name = _state_name(path)
        if not name:
            raise ValueError("MegatronCheckpointer only support the path whose suffix is "
                             "%s or %s." % (_MODEL_SD_NAME, _DIST_OPTIM_SD_NAME))
        self.state_dict[name] = state_dict
        self.paths[name] = path
        self._restored = {}

    def load(self, path: str, **kwargs):
        """torch.load stand-in: shared memory first, then the file."""
        if not isinstance(path, str):
            return torch_native_load(path)
        name = _state_name(path)
        # Megatron reads model_optim_rng.pt and then distrib_optim.pt: both
        # names come out of ONE restore of the image
        if name and name in self._restored:
            return self._restored.pop(name)
        # engine.load returns (step, state_dict); the reference treats the
        # tuple as the dict, so its memory restore never hits (megatron.py:98-108).
        # On a GPU job the tensors arrive on the current CUDA device through
        # the scatter kernel (engine.restore_device), not as CPU views that
        # Megatron would copy up one pageable transfer at a time.
        _, state_dict = self.engine.load(resume_path=path,
                                         device=self.engine.restore_device())
        if name and name in state_dict:
            self._restored = dict(state_dict)
            return self._restored.pop(name)
        return self.storage.read_state_dict(path, lambda p: _read_file(p, "cpu"))

    def update_tracer_file(self, iteration: int):
        """Megatron's save created ``iter_XXXXXXX/`` and advanced its tracker
        although the bytes only went to memory: remove the empty folder and
        point Megatron's tracker back at the last persisted step."""
        self.storage.safe_rmtree(os.path.join(self.checkpoint_dir,
                                              "iter_{:07d}".format(iteration)))
        megatron_tracer = os.path.join(self.checkpoint_dir, MegatronCheckpointSaver.TRACER_FILE)
        content = self.storage.read(os.path.join(self.checkpoint_dir,
                                                 CheckpointConstant.TRACER_FILE_NAME))
        if content:
            self.storage.write(content, megatron_tracer)
        else:
            self.storage.safe_remove(megatron_tracer)


def _call_megatron_save(iteration, model, optimizer, opt_param_scheduler, flops):
    if "num_floating_point_operations_so_far" in inspect.signature(megatron_save).parameters:
        megatron_save(iteration, model, optimizer, opt_param_scheduler, flops)
    else:
        megatron_save(iteration, model, optimizer, opt_param_scheduler)


def save_checkpoint(iteration, model, optimizer, opt_param_scheduler,
                    num_floating_point_operations_so_far=0, storage_type=StorageType.DISK,
                    storage=None, comm_backend="", save_timeout=CheckpointConstant.SAVE_TIMEOUT,
                    replica_count=0):
    """Same leading arguments as ``megatron.checkpointing.save_checkpoint``.

    Args:
        storage_type: StorageType.MEMORY (shared memory only) or DISK (shared
            memory, then persisted asynchronously by the agent).
        storage: CheckpointStorage (default PosixDiskStorage).
        comm_backend (str): backend of the sync group.
        replica_count (int): in-memory replicas on other nodes.
    """
    if storage_type not in (StorageType.MEMORY, StorageType.DISK):
        raise ValueError("No support storage type %s" % storage_type)
    saver = MegatronCheckpointer.singleton_instance(
        get_args().save, storage=storage, comm_backend=comm_backend,
        save_timeout=save_timeout, replica_count=replica_count)
    try:
        with redirected(save=saver.save):
            _call_megatron_save(iteration, model, optimizer, opt_param_scheduler,
                                num_floating_point_operations_so_far)
    finally:
        if _get_rank() == 0:
            saver.update_tracer_file(iteration)
    if storage_type == StorageType.MEMORY:
        ok = saver.engine.save_to_memory(iteration, saver.state_dict, saver.paths)
    else:
        ok = saver.engine.save_to_storage(iteration, saver.state_dict, saver.paths)
    saver.state_dict.clear()
    return ok


def load_checkpoint(model, optimizer, opt_param_scheduler, load_arg="load", strict=True,
                    storage=None, comm_backend="", save_timeout=CheckpointConstant.SAVE_TIMEOUT,
                    replica_count=0):
    """Same leading arguments as ``megatron.checkpointing.load_checkpoint``;
    returns what it returns (the iteration).  State still in shared memory is
    served from there."""
    checkpointer = MegatronCheckpointer.singleton_instance(
        get_args().save, storage=storage, comm_backend=comm_backend,
        save_timeout=save_timeout, replica_count=replica_count)
    with redirected(load=checkpointer.load):
        return megatron_load(model, optimizer, opt_param_scheduler, load_arg, strict)
