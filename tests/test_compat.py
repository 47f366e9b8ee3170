"""CPU: the reference's import paths resolve to this package after
compat.install() (run in a child interpreter so aliases do not leak)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SNIPPET = r"""
import sys
sys.path.insert(0, %r)
from dlrover_wuqiong_b200 import compat
names = compat.install()
assert len(names) == len(compat.ALIASES), names
from dlrover.trainer.torch.flash_checkpoint.checkpointer import Checkpointer, StorageType
from dlrover.trainer.torch.flash_checkpoint.ddp import DdpCheckpointer
from dlrover.trainer.torch.flash_checkpoint.fsdp import FsdpShardCheckpointer, FsdpFullCheckpointer, FsdpCheckpointer
from dlrover.trainer.torch.flash_checkpoint.deepspeed import DeepSpeedCheckpointer
from dlrover.trainer.torch.flash_checkpoint.megatron import MegatronCheckpointer, save_checkpoint, load_checkpoint
from dlrover.trainer.torch.flash_checkpoint.megatron_dist_ckpt import MegatronDistCheckpointer
from dlrover.trainer.torch.flash_checkpoint.engine import CheckpointEngine
from dlrover.trainer.torch.flash_checkpoint.full_ckpt_engine import FullCheckpointEngine
from dlrover.trainer.torch.flash_checkpoint.replica import CkptReplicaManger
from dlrover.python.elastic_agent.torch.ckpt_saver import (AsyncCheckpointSaver, SharedMemoryHandler,
    TensorMeta, CheckpointConfig, DLROVER_CKPT_CONFIG_KEY, TempDirCheckpointSaver)
from dlrover.python.common.multi_process import SharedMemory, SharedLock, SharedQueue, SharedDict
from dlrover.python.common.storage import PosixDiskStorage, KeepLatestStepStrategy
from dlrover.python.common.constants import CheckpointConstant
import dlrover.trainer.torch.flash_checkpoint.ddp as m
assert m.__name__ == "dlrover_wuqiong_b200.flash_checkpoint.ddp"
assert FsdpCheckpointer is FsdpShardCheckpointer
assert StorageType.MEMORY != StorageType.DISK and DLROVER_CKPT_CONFIG_KEY == "_DLORVER_CKPT_CONFIG"
print("ok")
"""


def test_reference_import_paths_resolve():
    env = dict(os.environ)
    env.pop("PYTHONPATH", None)
    out = subprocess.run([sys.executable, "-c", SNIPPET % ROOT], capture_output=True, text=True,
                         env=env, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert out.stdout.strip().endswith("ok")
