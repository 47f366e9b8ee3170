"""Engine for state dicts with full (unsharded or user-sharded) weights, used
by DdpCheckpointer (reference:
dlrover/trainer/torch/flash_checkpoint/full_ckpt_engine.py:33-215)."""
import os
from typing import Dict

import torch
import torch.distributed as dist

from ..ckpt_saver import (
    CheckpointConfig,
    CheckpointEvent,
    CheckpointEventType,
    DdpCheckpointSaver,
)
from .. import raw_ckpt
from ..common import env_utils
from ..common.constants import CheckpointConstant
from ..common.log import default_logger as logger
from .engine import CheckpointEngine, timer


class FullCheckpointEngine(CheckpointEngine):
    """
    Example::

        engine = FullCheckpointEngine(checkpoint_dir="/tmp/checkpoint/", storage=storage)
        for step, data in enumerate(dataloader):
            ...
            if step % 5 == 0:
                engine.save_to_memory(step, {"model_states": sd}, {"model_states": path})
            elif step % 100 == 0:
                engine.save_to_storage(step, {"model_states": sd}, {"model_states": path})
        state_dict = engine.load()
    """

    def __init__(self, checkpoint_dir, storage, local_shard_num=1, global_shard_num=1,
                 comm_backend="", save_timeout=CheckpointConstant.SAVE_TIMEOUT,
                 replica_count=0, async_save=None, stripe_replicas=None):
        if global_shard_num < local_shard_num:
            global_shard_num = local_shard_num
            logger.info("Set global_shard_num to %s.", local_shard_num)
        self._local_shard_num = local_shard_num
        self._global_shard_num = global_shard_num
        super().__init__(checkpoint_dir, storage, comm_backend, save_timeout,
                         replica_count=replica_count, async_save=async_save,
                         stripe_replicas=stripe_replicas)

    def get_saving_ranks(self):
        """The first `local_shard_num` local ranks of every node."""
        per_node = env_utils.get_local_world_size()
        ranks = [node * per_node + j
                 for node in range(env_utils.get_group_world_size())
                 for j in range(self._local_shard_num)]
        logger.info("The ranks to save checkpoint are %s.", ranks)
        return ranks

    def get_local_shard_num(self):
        return self._local_shard_num

    def get_global_shard_num(self):
        return self._global_shard_num

    def get_saver_class(self):
        return DdpCheckpointSaver

    @timer
    def save_to_memory(self, step, state_dict, paths: Dict[str, str]):
        conf = CheckpointConfig(step=step, paths=paths)
        return self.save_state_dict_to_memory(state_dict, conf)

    @timer
    def save_to_storage(self, step, state_dict, paths):
        success = True
        if step > self._cached_step:
            success = self.save_to_memory(step, state_dict, paths)
        if dist.is_initialized():
            dist.barrier()
        # one SAVE event per job step: global rank 0 tells its agent
        if success and self._rank == 0:
            self._event_queue.put(CheckpointEvent(type=CheckpointEventType.SAVE, step=step))
        return success

    def load(self, resume_path="", into=None, device=None):
        """Memory first, storage second.  `into` (a state dict shaped like the
        one that was saved, i.e. keyed by the single state name) / `device`
        request the in-place GPU restore."""
        _, state_dict = self.get_state_dict_from_memory(into=into, device=device)
        if state_dict:
            logger.info("Load the state dict from the CPU memory buffer.")
            names = list(state_dict.keys())
            if len(names) > 1:
                raise ValueError("The checkpoint shared memory must hold the state dict "
                                 "of one path only. Now, paths are %s" % names)
            return state_dict[names[0]]
        if into is not None:
            names = list(into.keys())
            into = into[names[0]] if len(names) == 1 else None
        restored = self._load_from_storage(resume_path, into=into)
        self._raise_if_into_wrecked(restored)
        return restored

    def _load_from_storage(self, resume_path="", into=None):
        """`into`: the live state dict to fill in place (raw image files go
        file -> pinned ring -> scatter kernel; torch.save files are loaded on
        the host and copied leaf by leaf)."""
        def read(path):
            if into is None:
                return raw_ckpt.load_any(path, map_location="cpu")
            if raw_ckpt.is_raw(path):
                return self._shm_handler.restore_from_file(path, into)
            return _copy_into(into, raw_ckpt.load_any(path, map_location="cpu"))

        if resume_path:
            return self.storage.read_state_dict(resume_path, read_func=read)
        tracker = os.path.join(self.checkpoint_dir, CheckpointConstant.TRACER_FILE_NAME)
        content = self.storage.read(tracker)
        if not content:
            return {}
        path = self._gen_restore_checkpoint_path(int(content.strip()))
        logger.info("Load the state dict from %s", path)
        return self.storage.read_state_dict(path, read_func=read)

    def _gen_restore_checkpoint_path(self, iteration):
        # unsharded: every rank reads what rank 0 wrote
        who = 0 if self._global_shard_num == 1 else self._rank
        return os.path.join(self.checkpoint_dir, "%s/rank_%s.pt" % (iteration, who))


def _copy_into(into, loaded):
    """Copy every tensor of `loaded` into the tensor at the same place of
    `into`; returns `loaded`'s structure holding `into`'s tensors."""
    if isinstance(loaded, dict):
        return {k: _copy_into(into[k], v) if isinstance(v, (dict, list)) or torch.is_tensor(v)
                else v for k, v in loaded.items()}
    if isinstance(loaded, list):
        return [_copy_into(into[i], v) if isinstance(v, (dict, list)) or torch.is_tensor(v)
                else v for i, v in enumerate(loaded)]
    if loaded.numel():
        with torch.no_grad():
            into.copy_(loaded)
    return into
