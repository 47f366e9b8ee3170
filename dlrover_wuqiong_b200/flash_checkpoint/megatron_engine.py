"""Engines behind the Megatron-LM checkpointers (reference:
dlrover/trainer/torch/flash_checkpoint/megatron_engine.py:28-275)."""
import torch.distributed as dist

from ..ckpt_saver import MegatronCheckpointSaver
from ..common.constants import CheckpointConstant
from .engine import LocalShardsEngine


def _model_parallel_coords():
    """(tp_rank, pp_rank, dp_rank, tp_world, pp_world) from Megatron's mpu, or
    the single-process defaults."""
    if not dist.is_initialized():
        return 0, 0, 0, 1, 1
    try:
        from megatron.core import mpu
    except ImportError:  # older Megatron layouts
        from megatron import mpu
    return (mpu.get_tensor_model_parallel_rank(), mpu.get_pipeline_model_parallel_rank(),
            mpu.get_data_parallel_rank(), mpu.get_tensor_model_parallel_world_size(),
            mpu.get_pipeline_model_parallel_world_size())


class MegatronCheckpointEngine(LocalShardsEngine):
    """One shard per (tensor, pipeline) model-parallel rank; data-parallel
    replicas hold the same shard and only dp rank 0 triggers persistence."""

    def __init__(self, checkpoint_dir, storage, comm_backend="",
                 save_timeout=CheckpointConstant.SAVE_TIMEOUT, replica_count=0):
        (self._tp_rank, self._pp_rank, self._dp_rank, self._tp_world_size,
         self._pp_world_size) = _model_parallel_coords()
        super().__init__(checkpoint_dir, storage, comm_backend, save_timeout,
                         replica_count=replica_count)

    def get_global_shard_num(self):
        return self._pp_world_size * self._tp_world_size

    def get_saver_class(self):
        return MegatronCheckpointSaver

    def _notifies_agent(self):
        return self._dp_rank == 0 and self._local_rank == 0

    def load(self, resume_path="", into=None, device=None):
        """(step, state_dict) from shm; (0 or step, {}) if nothing valid."""
        return self.get_state_dict_from_memory(into=into, device=device)


class MegatronDistCheckpointEngine(LocalShardsEngine):
    """Distributed-optimizer variant: every rank owns a shard."""

    def __init__(self, checkpoint_dir, storage, comm_backend="",
                 save_timeout=CheckpointConstant.SAVE_TIMEOUT):
        (self._tp_rank, self._pp_rank, self._dp_rank, self._tp_world_size,
         self._pp_world_size) = _model_parallel_coords()
        super().__init__(checkpoint_dir, storage, comm_backend, save_timeout)

    def get_saving_ranks(self):
        return None  # all ranks

    def get_global_shard_num(self):
        return dist.get_world_size() if dist.is_initialized() else 1

    def get_saver_class(self):
        return MegatronCheckpointSaver

    def load(self, resume_path="", into=None, device=None):
        return self.get_state_dict_from_memory(into=into, device=device)
