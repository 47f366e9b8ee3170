"""Engine behind DeepSpeedCheckpointer (reference:
dlrover/trainer/torch/flash_checkpoint/deepspeed_engine.py:31-159)."""
import copy

from ..ckpt_saver import DeepSpeedCheckpointSaver, SharedMemoryHandler
from ..common.constants import CheckpointConstant
from .engine import LocalShardsEngine


class DeepSpeedCheckpointEngine(LocalShardsEngine):
    """
    Args:
        checkpoint_dir (str): directory of the checkpoints.
        storage: CheckpointStorage.
        global_shard_num (int): number of ZeRO shards (1 without ZeRO).
        zero_stage (int): DeepSpeed ZeRO stage.
        comm_backend (str): backend of the sync group.
    """

    def __init__(self, checkpoint_dir, storage, global_shard_num=1, zero_stage=0,
                 comm_backend="", save_timeout=CheckpointConstant.SAVE_TIMEOUT):
        self.global_shard_num = global_shard_num
        self.zero_stage = zero_stage
        super().__init__(checkpoint_dir, storage, comm_backend, save_timeout)

    def get_global_shard_num(self):
        return self.global_shard_num

    def get_saver_class(self):
        return DeepSpeedCheckpointSaver

    def load(self, into=None, device=None):
        """State dict from shm ({} if none).  With ZeRO-1/2 the model is not
        partitioned and only local rank 0 holds "model_states"; other local
        ranks copy it out of local rank 0's segment."""
        _, state_dict = self.get_state_dict_from_memory(into=into, device=device)
        if state_dict:
            name = CheckpointConstant.MODEL_STATES_NAME
            if name not in state_dict and self.zero_stage in (1, 2):
                rank0 = SharedMemoryHandler(0, host=False)
                # views alias shm that another rank may be loading too
                state_dict[name] = copy.deepcopy(rank0.load_state_dict()[name])
        return state_dict
