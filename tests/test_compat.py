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


THIRD_PARTY = r"""
import os, pickle, sys, tempfile, time
sys.path.insert(0, %r)
os.environ["TORCHELASTIC_RUN_ID"] = "fcp_compat_%%d" %% os.getpid()
os.environ["NODE_NUM"] = "1"
os.environ["NODE_RANK"] = "0"
from dlrover_wuqiong_b200 import compat
compat.install()
# ---- from here on: only the reference's import paths, as in
# atorch/atorch/utils/fsdp_async_ckpt_util.py:1-18
import torch
import torch.distributed as dist
from dlrover.python.common import env_utils
from dlrover.python.common.singleton import singleton
from dlrover.python.common.storage import PosixDiskStorage
from dlrover.python.elastic_agent.torch.ckpt_saver import (
    AsyncCheckpointSaver, CheckpointConfig, CheckpointEvent, CheckpointEventType,
    TempDirCheckpointSaver)
from dlrover.trainer.torch.flash_checkpoint.checkpointer import StorageType
from dlrover.trainer.torch.flash_checkpoint.engine import CheckpointEngine, timer


class ThirdPartySaver(TempDirCheckpointSaver):
    tracker_file = "latest_checkpointed_iteration.txt"

    def persist_to_storage(self, local_shard_id, ckpt_config):
        state_dict = self._shm_handlers[local_shard_id].load_state_dict()
        for name, path in ckpt_config.paths.items():
            state = state_dict.get(name, None)
            if not state:
                continue
            if name in ("param_meta", "param_groups"):
                with open(path, "wb") as f:
                    pickle.dump(state, f)
            else:
                torch.save(state, path)

    def update_tracker_file(self, step):
        self.storage.write(str(step), os.path.join(self.checkpoint_dir, self.tracker_file))


@singleton
class ThirdPartyEngine(CheckpointEngine):
    def __init__(self, checkpoint_dir, storage=None, comm_backend=""):
        super().__init__(checkpoint_dir, storage or PosixDiskStorage(), comm_backend)

    def get_saving_ranks(self):
        return None

    @timer
    def save_to_memory(self, step, state_dict, paths):
        return self.save_state_dict_to_memory(state_dict, CheckpointConfig(step=step, paths=paths))

    @timer
    def save_to_storage(self, step, state_dict, paths):
        ok = True
        if step > self._cached_step:
            ok = self.save_to_memory(step, state_dict, paths)
        if dist.is_initialized():
            dist.barrier()
        if self._local_rank == 0 and ok:
            self._event_queue.put(CheckpointEvent(type=CheckpointEventType.SAVE, step=step))
        return ok

    def get_global_shard_num(self):
        return dist.get_world_size() if dist.is_initialized() else 1

    def get_local_shard_num(self):
        return env_utils.get_local_world_size()

    def load(self, resume_path=""):
        pass

    def get_saver_class(self):
        return ThirdPartySaver


AsyncCheckpointSaver.start_async_saving_ckpt()
with tempfile.TemporaryDirectory() as d:
    engine = ThirdPartyEngine(d)
    assert ThirdPartyEngine(d) is engine
    g = torch.Generator().manual_seed(1)
    state = {"params": {"flat.0": torch.randn(1000, generator=g), "flat.1": torch.randn(77, generator=g)},
             "param_meta": {"flat.0": {"shape": (10, 100)}, "n": 2},
             "optim_states": {"exp_avg.0": torch.randn(1000, generator=g)},
             "param_groups": [{"lr": 0.1}]}
    out = os.path.join(d, "iter_7")
    paths = {"params": out + "/flat_param.00000-00001", "param_meta": out + "/flat_meta.00000-00001",
             "optim_states": out + "/optim_param.00000-00001", "param_groups": out + "/optim_meta"}
    assert engine.save_to_storage(7, state, paths)
    tracker = os.path.join(d, ThirdPartySaver.tracker_file)
    t0 = time.time()
    while not (os.path.exists(tracker) and open(tracker).read() == "7"):
        assert time.time() - t0 < 60, "agent did not persist"
        time.sleep(0.1)
    assert torch.equal(torch.load(paths["params"])["flat.1"], state["params"]["flat.1"])
    assert torch.equal(torch.load(paths["optim_states"])["exp_avg.0"], state["optim_states"]["exp_avg.0"])
    assert pickle.load(open(paths["param_meta"], "rb")) == state["param_meta"]
    assert pickle.load(open(paths["param_groups"], "rb")) == state["param_groups"]
    assert type(AsyncCheckpointSaver.get_ckpt_saver()).__name__ == "ThirdPartySaver"
    engine.close()
AsyncCheckpointSaver.get_ckpt_saver().close()
print("ok")
"""


def test_third_party_engine_written_against_the_reference_runs_unchanged():
    """An engine + saver in the shape of ATorch's fsdp_async_ckpt_util
    (atorch/atorch/utils/fsdp_async_ckpt_util.py:29-133; SURVEY §8b
    "subclassing contract", §8f-3): several state names with their own paths,
    custom persistence per name, the `singleton` decorator, reference import
    paths only."""
    env = dict(os.environ)
    env.pop("PYTHONPATH", None)
    out = subprocess.run([sys.executable, "-c", THIRD_PARTY % ROOT], capture_output=True, text=True,
                         env=env, timeout=300)
    assert out.returncode == 0, (out.stdout[-1500:], out.stderr[-3000:])
    assert out.stdout.strip().endswith("ok")


def test_shim_package_resolves_the_reference_paths_in_a_fresh_process(tmp_path):
    """compat.write_shim: a process that never calls install() — a spawned
    trainer, an agent importing the saver class by module path — gets the
    aliases by importing `dlrover` from PYTHONPATH."""
    sys.path.insert(0, ROOT)
    try:
        from dlrover_wuqiong_b200 import compat
        pkg = compat.write_shim(str(tmp_path))
    finally:
        sys.path.remove(ROOT)
    assert os.path.exists(os.path.join(pkg, "__init__.py"))
    env = dict(os.environ)
    env["PYTHONPATH"] = str(tmp_path)
    code = ("import importlib\n"
            "m = importlib.import_module('dlrover.python.elastic_agent.torch.ckpt_saver')\n"
            "cls = getattr(m, 'DeepSpeedCheckpointSaver')\n"
            "from dlrover.trainer.torch.flash_checkpoint.hf_trainer import HfDdpCheckpointer\n"
            "from dlrover.python.common.grpc import find_free_port\n"
            "assert cls.__module__ == 'dlrover_wuqiong_b200.ckpt_saver' and find_free_port() > 0\n"
            "print('ok')\n")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env,
                         cwd=str(tmp_path), timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    assert out.stdout.strip().endswith("ok")
