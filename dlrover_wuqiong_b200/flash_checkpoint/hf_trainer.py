"""Flash checkpoint for the HuggingFace Trainer (reference:
dlrover/trainer/torch/flash_checkpoint/hf_trainer.py:59-383).

The reference re-implements ``Trainer._save_checkpoint`` of transformers 4.37
line by line with ``torch.save`` swapped out.  That private method changes
with every transformers release, so this build wraps it instead: while the
installed Trainer runs its own ``_save_checkpoint``, every ``torch.save(obj,
path)`` it issues (model weights, ``optimizer.pt``, ``scheduler.pt``,
``rng_state*.pth``) is captured; the captured objects then go to shared memory
through the flash engine and are persisted by the agent to exactly the paths
the Trainer chose.  JSON side files (config, trainer_state) are written by the
Trainer as usual.  ``transformers`` / ``deepspeed`` are imported lazily.
"""
import os
import re
import shutil
from typing import Optional

import torch.distributed as dist

from ..common.constants import CheckpointConstant
from ..common.log import default_logger as logger
from ..common.storage import PosixDiskStorage
from ._hijack import redirected, torch_native_load, torch_native_save  # noqa: F401
from .deepspeed import AsyncCheckpointAgent
from .deepspeed_engine import DeepSpeedCheckpointEngine
from .engine import CheckpointEngine
from .full_ckpt_engine import FullCheckpointEngine


class HfFlashCheckpointer(object):
    def __init__(self, checkpoint_dir, storage=None):
        self.checkpoint_dir = checkpoint_dir
        self.storage = storage if storage else PosixDiskStorage()
        self.ckpt_agent = AsyncCheckpointAgent(self.storage)
        self.async_save_engine: Optional[CheckpointEngine] = None

    def save_checkpoint_to_memory(self, step):
        return self.async_save_engine.save_to_memory(
            step, self.ckpt_agent.state_dict, self.ckpt_agent.paths)

    def save_checkpoint_to_storage(self, step):
        return self.async_save_engine.save_to_storage(
            step, self.ckpt_agent.state_dict, self.ckpt_agent.paths)


class HfDeepSpeedCheckpointer(HfFlashCheckpointer):
    def __init__(self, engine, checkpoint_dir, storage=None, comm_backend=""):
        super().__init__(checkpoint_dir, storage)
        self.engine = engine
        global_shard_num = 1
        if self.engine.zero_optimization():
            global_shard_num = dist.get_world_size(self.engine.optimizer.dp_process_group)
        self.async_save_engine = DeepSpeedCheckpointEngine(
            checkpoint_dir, storage=self.storage, global_shard_num=global_shard_num,
            zero_stage=self.engine.zero_optimization_stage(), comm_backend=comm_backend)


class HfDdpCheckpointer(HfFlashCheckpointer):
    def __init__(self, checkpoint_dir, storage=None, comm_backend=""):
        super().__init__(checkpoint_dir, storage)
        self.async_save_engine = FullCheckpointEngine(
            checkpoint_dir, storage=self.storage, comm_backend=comm_backend)


_TRAINER_CLS = None


def _build_trainer_class():
    from transformers import Trainer
    from transformers.trainer_utils import PREFIX_CHECKPOINT_DIR

    class FlashCkptTrainer(Trainer):
        """Trainer whose checkpoints go to shared memory synchronously and to
        ``output_dir/checkpoint-<step>/`` asynchronously; the last complete
        step is in ``output_dir/dlrover_latest.txt``."""

        def _flash(self, run_dir):
            if not hasattr(self, "flash_checkpointer"):
                if self.is_deepspeed_enabled:
                    self.flash_checkpointer = HfDeepSpeedCheckpointer(self.model_wrapped, run_dir)
                elif not self.is_fsdp_enabled:
                    self.flash_checkpointer = HfDdpCheckpointer(run_dir)
                else:
                    raise ValueError("Flash Checkpoint only supports DeepSpeed or DDP.")
            return self.flash_checkpointer

        def _save_checkpoint(self, model, trial, *args, **kwargs):
            run_dir = self._get_output_dir(trial=trial)
            step = self.state.global_step
            output_dir = os.path.join(run_dir, "%s-%s" % (PREFIX_CHECKPOINT_DIR, step))
            flash = self._flash(run_dir)
            agent = flash.ckpt_agent
            agent.state_dict, agent.paths = {}, {}
            # weights must pass through torch.save to be captured
            safetensors = getattr(self.args, "save_safetensors", False)
            self.args.save_safetensors = False
            try:
                with redirected(save=agent.save):
                    super()._save_checkpoint(model, trial, *args, **kwargs)
            finally:
                self.args.save_safetensors = safetensors
            success = flash.save_checkpoint_to_storage(step)
            if not success:
                logger.info("Skip saving the checkpoint of step %s because the latest "
                            "checkpoint is not finished.", step)
                shutil.rmtree(output_dir, ignore_errors=True)

        # -- the newest folders may still be waiting for the agent: only what
        # the tracker file vouches for is complete (hf_trainer.py:313-383)
        def _get_last_checkpoint_step(self):
            tracker = os.path.join(self.args.output_dir, CheckpointConstant.TRACER_FILE_NAME)
            try:
                with open(tracker) as f:
                    return int(f.read().strip() or 0)
            except (OSError, ValueError):
                return 0

        def get_last_checkpoint(self):
            """Folder of the last *complete* checkpoint, False if there is none."""
            step = self._get_last_checkpoint_step()
            if step == 0:
                return False
            return os.path.join(self.args.output_dir, "%s-%s" % (PREFIX_CHECKPOINT_DIR, step))

        def _rotate_checkpoints(self, use_mtime=False, output_dir=None):
            limit = self.args.save_total_limit
            if limit is None or limit <= 0:
                return
            last_step = self._get_last_checkpoint_step()
            complete = []
            for folder in self._sorted_checkpoints(use_mtime=use_mtime, output_dir=output_dir):
                m = re.match(r".*%s-([0-9]+)$" % PREFIX_CHECKPOINT_DIR, str(folder))
                if m and int(m.group(1)) <= last_step:
                    complete.append(folder)
            best = getattr(self.state, "best_model_checkpoint", None)
            if best is not None and limit == 1 and complete and complete[-1] != best:
                limit = 2  # never delete the checkpoint a resume would need
            for folder in complete[:max(0, len(complete) - limit)]:
                logger.info("Deleting older checkpoint [%s] due to save_total_limit = %s.",
                            folder, self.args.save_total_limit)
                shutil.rmtree(folder, ignore_errors=True)

    return FlashCkptTrainer


def __getattr__(name):
    # `from ...hf_trainer import FlashCkptTrainer` builds the class on first use,
    # so importing this module does not import transformers
    global _TRAINER_CLS
    if name == "FlashCkptTrainer":
        if _TRAINER_CLS is None:
            _TRAINER_CLS = _build_trainer_class()
        return _TRAINER_CLS
    raise AttributeError(name)
