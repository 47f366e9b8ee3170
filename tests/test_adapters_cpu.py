"""CPU: DeepSpeed and Megatron adapters, driven the way the reference's tests
drive them — a mock DeepSpeedEngine that just torch.save()s
(dlrover/trainer/tests/torch/deepspeed_ckpt_test.py:35-153) and patched
Megatron hooks (megatron_ckpt_test.py:84-186)."""
import argparse
import os
import tempfile
import time

import pytest
import torch
import torch.nn as nn
import torch.optim as optim

os.environ.setdefault("ROLE_NAME", "dlrover-trainer")
os.environ.setdefault("TORCHELASTIC_RUN_ID", "fcp_test_%d" % os.getpid())

from dlrover_wuqiong_b200.ckpt_saver import (  # noqa: E402
    AsyncCheckpointSaver,
    DeepSpeedCheckpointSaver,
    MegatronCheckpointSaver,
)
from dlrover_wuqiong_b200.common.constants import CheckpointConstant, NodeEnv  # noqa: E402
from dlrover_wuqiong_b200.common.multi_process import clear_sock_dir  # noqa: E402
from dlrover_wuqiong_b200.flash_checkpoint import megatron  # noqa: E402
from dlrover_wuqiong_b200.flash_checkpoint.checkpointer import StorageType  # noqa: E402
from dlrover_wuqiong_b200.flash_checkpoint.deepspeed import DeepSpeedCheckpointer  # noqa: E402
from dlrover_wuqiong_b200.flash_checkpoint.deepspeed_engine import (  # noqa: E402
    DeepSpeedCheckpointEngine,
)
from dlrover_wuqiong_b200.flash_checkpoint.megatron import (  # noqa: E402
    MegatronCheckpointer,
    load_checkpoint,
    save_checkpoint,
)
from dlrover_wuqiong_b200.flash_checkpoint.megatron_dist_ckpt import (  # noqa: E402
    get_dist_optimizer_checkpoint_name,
    get_parameter_state,
    load_parameter_state_from_state_dict,
)
from dlrover_wuqiong_b200.flash_checkpoint.megatron_engine import (  # noqa: E402
    MegatronCheckpointEngine,
    MegatronDistCheckpointEngine,
)
from dlrover_wuqiong_b200.common.storage import PosixDiskStorage  # noqa: E402


class SimpleNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.fc1 = nn.Linear(64, 32)
        self.fc2 = nn.Linear(32, 10)


@pytest.fixture
def saver_env():
    os.environ[NodeEnv.NODE_NUM] = "1"
    os.environ[NodeEnv.NODE_RANK] = "0"
    os.environ["LOCAL_RANK"] = "0"
    AsyncCheckpointSaver._saver_instance = None
    AsyncCheckpointSaver.start_async_saving_ckpt()
    MegatronCheckpointer._instance = None
    yield
    if AsyncCheckpointSaver._saver_instance:
        AsyncCheckpointSaver._saver_instance.close()
        AsyncCheckpointSaver._saver_instance = None
    MegatronCheckpointer._instance = None
    clear_sock_dir()


def _wait_for(pred, timeout=30):
    t0 = time.time()
    while time.time() - t0 < timeout:
        if pred():
            return True
        time.sleep(0.1)
    return False


class MockDeepSpeedEngine(object):
    def __init__(self, model, optimizer):
        self.model = model
        self.optimizer = optimizer
        self.save_non_zero_checkpoint = True
        self.global_rank = 0

    def zero_optimization(self):
        return False

    def zero_optimization_stage(self):
        return 0

    def save_checkpoint(self, save_dir, tag, client_state, save_latest):
        torch.save(self.model.state_dict(), os.path.join(save_dir, str(tag), "model_states.pt"))
        torch.save(self.optimizer.state_dict(),
                   os.path.join(save_dir, str(tag), "optim_states.pt"))

    def load_checkpoint(self, load_dir, tag, load_module_strict, load_optimizer_states,
                        load_lr_scheduler_states, load_module_only, custom_load_fn):
        self.model.load_state_dict(torch.load(os.path.join(load_dir, tag, "model_states.pt")))
        self.optimizer.load_state_dict(
            torch.load(os.path.join(load_dir, tag, "optim_states.pt")))
        return os.path.join(load_dir, tag), {}


def test_deepspeed_checkpointer(saver_env):
    model = SimpleNet()
    optimizer = optim.SGD(model.parameters(), lr=0.01, momentum=0.001)
    step = 100
    with tempfile.TemporaryDirectory() as tmpdir:
        engine = MockDeepSpeedEngine(model, optimizer)
        ckpt = DeepSpeedCheckpointer(engine, tmpdir)
        assert ckpt.save_checkpoint(tmpdir, step, storage_type=StorageType.MEMORY)
        handler = ckpt._async_save_engine._shm_handler
        assert not handler.no_checkpint_state()
        assert handler._buffer_size == 9640  # deepspeed_ckpt_test.py:119
        meta = handler.metadata.get()
        assert meta["_DLORVER_CKPT_CONFIG"].step == step
        assert meta["model_states"] is not None
        ds_tracker = os.path.join(tmpdir, "latest")
        assert not os.path.exists(ds_tracker)  # nothing persisted yet
        # served from memory: wreck the weights, then load
        want = {k: v.clone() for k, v in model.state_dict().items()}
        with torch.no_grad():
            model.fc1.weight.zero_()
        ckpt.load_checkpoint(tmpdir, str(step))
        for k, v in want.items():
            assert torch.equal(model.state_dict()[k], v)
        ckpt.save_checkpoint(tmpdir, step, storage_type=StorageType.DISK)
        assert _wait_for(lambda: os.path.exists(ds_tracker) and open(ds_tracker).read() == "100")
        assert sorted(os.listdir(os.path.join(tmpdir, "100"))) == ["model_states.pt",
                                                                  "optim_states.pt"]
        assert isinstance(AsyncCheckpointSaver.get_ckpt_saver(), DeepSpeedCheckpointSaver)
        with pytest.raises(ValueError):
            ckpt.save_checkpoint(tmpdir, str(step), storage_type=2)
        ckpt.close()


def test_deepspeed_engine_shard_policy(saver_env):
    with tempfile.TemporaryDirectory() as tmpdir:
        e = DeepSpeedCheckpointEngine(tmpdir, PosixDiskStorage(), global_shard_num=1, zero_stage=1)
        assert e.get_global_shard_num() == 1 and e.get_local_shard_num() == 1
        assert e.get_saver_class() is DeepSpeedCheckpointSaver
        assert e.load() == {}
        e.close()


def test_megatron_save_load(saver_env, monkeypatch):
    model = SimpleNet()
    optimizer = optim.SGD(model.parameters(), lr=0.01, momentum=0.001)
    suffix = "model_optim_rng.pt"
    with tempfile.TemporaryDirectory() as tmpdir:
        def get_args():
            args, _ = argparse.ArgumentParser().parse_known_args([])
            args.save = tmpdir
            return args

        def mock_save(iteration, model, optimizer, opt_scheduler):
            torch.save({"iteration": iteration, "model_states": model.state_dict(),
                        "optim_states": optimizer.state_dict()},
                       os.path.join(tmpdir, str(iteration), suffix))

        loaded = {}

        def mock_load(model, optimizer, opt_param_scheduler, load_arg="load", strict=True):
            sd = torch.load(os.path.join(tmpdir, "20", suffix))
            loaded.update(sd)
            model.load_state_dict(sd["model_states"])
            optimizer.load_state_dict(sd["optim_states"])
            return sd["iteration"]

        monkeypatch.setattr(megatron, "megatron_save", mock_save)
        monkeypatch.setattr(megatron, "megatron_load", mock_load)
        monkeypatch.setattr(megatron, "get_args", get_args)
        mgr = MegatronCheckpointer.singleton_instance(tmpdir)
        assert save_checkpoint(10, model, optimizer, None, storage_type=StorageType.MEMORY)
        handler = mgr.engine._shm_handler
        assert not handler.no_checkpint_state()
        assert handler._buffer_size == 9640  # megatron_ckpt_test.py:125
        meta = handler.metadata.get()[CheckpointConstant.MODEL_STATES_NAME]
        assert meta["iteration"] == 10 and meta["model_states"] is not None
        save_checkpoint(20, model, optimizer, None, storage_type=StorageType.DISK)
        tracker = os.path.join(tmpdir, MegatronCheckpointSaver.TRACER_FILE)
        assert _wait_for(lambda: os.path.exists(tracker) and open(tracker).read() == "20")
        assert os.listdir(os.path.join(tmpdir, "20")) == [suffix]
        # memory restore really is used (the reference's never hits): remove
        # the file and load anyway
        os.remove(os.path.join(tmpdir, "20", suffix))
        assert load_checkpoint(model, optimizer, None) == 20
        assert loaded["iteration"] == 20
        with pytest.raises(ValueError):
            save_checkpoint(20, model, optimizer, None, storage_type=2)
        mgr.engine.close()


def test_megatron_update_tracer_file(saver_env):
    iteration = 100
    with tempfile.TemporaryDirectory() as tmpdir:
        ckpt_dir = os.path.join(tmpdir, "iter_{:07d}".format(iteration))
        os.makedirs(ckpt_dir)
        mgr = MegatronCheckpointer(tmpdir)
        dl_tracker = os.path.join(tmpdir, CheckpointConstant.TRACER_FILE_NAME)
        with open(dl_tracker, "w") as f:
            f.write("50")
        mgr.update_tracer_file(iteration)
        assert not os.path.exists(ckpt_dir)
        mg_tracker = os.path.join(tmpdir, MegatronCheckpointSaver.TRACER_FILE)
        assert open(mg_tracker).read() == "50"
        os.remove(dl_tracker)
        mgr.update_tracer_file(50)
        assert not os.path.exists(mg_tracker)
        mgr.engine.close()


def test_megatron_engines_and_dist_optimizer_state(saver_env):
    assert get_dist_optimizer_checkpoint_name("/tmp", 100) == \
        "/tmp/iter_0000100/rank_00000/distrib_optim.pt"
    with tempfile.TemporaryDirectory() as tmpdir:
        e = MegatronCheckpointEngine(tmpdir, PosixDiskStorage())
        assert e.get_global_shard_num() == 1 and e.get_local_shard_num() == 1
        assert e.get_saver_class() is MegatronCheckpointSaver
        assert e.load() == (0, {})  # checkpoint_egine_test.py:222-224
        e.close()
        d = MegatronDistCheckpointEngine(tmpdir, PosixDiskStorage())
        assert d.get_saving_ranks() is None and d.get_global_shard_num() == 1

        # a stand-in for Megatron's DistributedOptimizer bookkeeping
        p0, p1 = nn.Parameter(torch.randn(6)), nn.Parameter(torch.randn(3))
        inner = optim.Adam([p0, p1], lr=0.1)
        (p0.sum() + p1.sum()).backward()
        inner.step()

        class FakeDistOpt(object):
            optimizer = inner
            gbuf_ranges = [{torch.float32: [{"param_map": {"m0": None, "m1": None}}]}]
            model_param_group_index_map = {"m0": (0, 0), "m1": (0, 1)}

        state = get_parameter_state(FakeDistOpt())
        assert state[0][0][0]["param"] is p0 and "exp_avg" in state[0][0][1]
        # save the live shards, wreck them, restore in place through the engine
        sd = {CheckpointConstant.OPTIM_STATES_NAME: state}
        paths = {CheckpointConstant.OPTIM_STATES_NAME: os.path.join(tmpdir, "o.pt")}
        assert d.save_to_memory(7, sd, paths)
        want = p0.detach().clone(), inner.state[p1]["exp_avg"].clone()
        with torch.no_grad():
            p0.zero_()
            inner.state[p1]["exp_avg"].zero_()
        step, got = d.load(into={CheckpointConstant.OPTIM_STATES_NAME:
                                 get_parameter_state(FakeDistOpt())})
        assert step == 7
        assert torch.equal(p0.detach(), want[0])
        assert torch.equal(inner.state[p1]["exp_avg"], want[1])
        # per-tensor path of the reference
        with torch.no_grad():
            p0.zero_()
        _, views = d.load()
        load_parameter_state_from_state_dict(FakeDistOpt(),
                                             views[CheckpointConstant.OPTIM_STATES_NAME])
        assert torch.equal(p0.detach(), want[0])
        d.close()


def test_hf_ddp_checkpointer_persists_every_captured_file(saver_env):
    """The HF adapter's capture + persist step (the Trainer subclass itself
    needs `accelerate`, which this image lacks): what the Trainer would
    torch.save — weights, optimizer.pt, scheduler.pt, rng_state.pth — ends up
    at exactly those paths, asynchronously."""
    from dlrover_wuqiong_b200.flash_checkpoint._hijack import redirected
    from dlrover_wuqiong_b200.flash_checkpoint.hf_trainer import HfDdpCheckpointer
    model = SimpleNet()
    optimizer = optim.AdamW(model.parameters(), lr=0.01)
    model.fc2(model.fc1(torch.randn(2, 64))).sum().backward()
    optimizer.step()
    with tempfile.TemporaryDirectory() as tmpdir:
        flash = HfDdpCheckpointer(tmpdir)
        out = os.path.join(tmpdir, "checkpoint-8")
        with redirected(save=flash.ckpt_agent.save):
            torch.save(model.state_dict(), os.path.join(out, "pytorch_model.bin"))
            torch.save(optimizer.state_dict(), os.path.join(out, "optimizer.pt"))
            torch.save({"last_epoch": 8}, os.path.join(out, "scheduler.pt"))
            torch.save({"cpu": torch.get_rng_state()}, os.path.join(out, "rng_state.pth"))
        assert sorted(flash.ckpt_agent.paths) == ["optimizer.pt", "pytorch_model.bin",
                                                  "rng_state.pth", "scheduler.pt"]
        assert flash.save_checkpoint_to_storage(8)
        tracker = os.path.join(tmpdir, CheckpointConstant.TRACER_FILE_NAME)
        assert _wait_for(lambda: os.path.exists(tracker) and open(tracker).read() == "8")
        assert sorted(os.listdir(out)) == ["optimizer.pt", "pytorch_model.bin", "rng_state.pth",
                                          "scheduler.pt"]
        back = torch.load(os.path.join(out, "pytorch_model.bin"), weights_only=False)
        assert torch.equal(back["fc1.weight"], model.fc1.weight)
        opt_back = torch.load(os.path.join(out, "optimizer.pt"), weights_only=False)
        assert torch.equal(opt_back["state"][0]["exp_avg"], optimizer.state_dict()["state"][0]["exp_avg"])
        flash.async_save_engine.close()


def test_flash_ckpt_trainer_on_a_stand_in_trainer(saver_env, monkeypatch):
    """FlashCkptTrainer wraps whatever Trainer is installed.  The real one
    cannot be built here (needs `accelerate`), so a stand-in `transformers`
    with the four members the wrapper touches plays its part: a
    `_save_checkpoint` that torch.saves weights / optimizer / scheduler / rng
    the way HF does, `_get_output_dir`, `_sorted_checkpoints`, args + state.
    hf_trainer_test.py:40-93 checks the same things on the real class."""
    import sys
    import types

    import dlrover_wuqiong_b200.flash_checkpoint.hf_trainer as hf

    class Trainer(object):
        def __init__(self, model, args):
            self.model, self.model_wrapped, self.args = model, model, args
            self.optimizer = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9)
            self.state = types.SimpleNamespace(global_step=0, best_model_checkpoint=None)
            self.is_deepspeed_enabled = self.is_fsdp_enabled = False

        def _get_output_dir(self, trial=None):
            return self.args.output_dir

        def _save_checkpoint(self, model, trial):
            out = os.path.join(self.args.output_dir, "checkpoint-%d" % self.state.global_step)
            os.makedirs(out, exist_ok=True)
            assert self.args.save_safetensors is False  # the wrapper forces torch.save
            torch.save(model.state_dict(), os.path.join(out, "pytorch_model.bin"))
            torch.save(self.optimizer.state_dict(), os.path.join(out, "optimizer.pt"))
            torch.save({"last_epoch": 3}, os.path.join(out, "scheduler.pt"))
            torch.save({"cpu": torch.get_rng_state()}, os.path.join(out, "rng_state.pth"))
            with open(os.path.join(out, "trainer_state.json"), "w") as f:
                f.write("{}")

        def _sorted_checkpoints(self, use_mtime=False, output_dir=None):
            names = [d for d in os.listdir(output_dir) if d.startswith("checkpoint-")]
            return [os.path.join(output_dir, d)
                    for d in sorted(names, key=lambda d: int(d.split("-")[1]))]

    fake = types.ModuleType("transformers")
    fake.Trainer = Trainer
    utils = types.ModuleType("transformers.trainer_utils")
    utils.PREFIX_CHECKPOINT_DIR = "checkpoint"
    fake.trainer_utils = utils
    monkeypatch.setitem(sys.modules, "transformers", fake)
    monkeypatch.setitem(sys.modules, "transformers.trainer_utils", utils)
    monkeypatch.setattr(hf, "_TRAINER_CLS", None)
    FlashCkptTrainer = hf.FlashCkptTrainer
    assert issubclass(FlashCkptTrainer, Trainer)

    with tempfile.TemporaryDirectory() as tmpdir:
        args = types.SimpleNamespace(output_dir=tmpdir, save_safetensors=True, save_total_limit=2)
        model = SimpleNet()
        trainer = FlashCkptTrainer(model, args)
        model.fc2(model.fc1(torch.randn(2, 64))).sum().backward()
        trainer.optimizer.step()
        # rotation: only folders the tracker vouches for are candidates
        for step in (10, 20, 30, 40):
            os.makedirs(os.path.join(tmpdir, "checkpoint-%d" % step))
        assert trainer._get_last_checkpoint_step() == 0 and trainer.get_last_checkpoint() is False
        with open(os.path.join(tmpdir, "dlrover_latest.txt"), "w") as f:
            f.write("30")
        assert trainer._get_last_checkpoint_step() == 30
        assert trainer.get_last_checkpoint() == os.path.join(tmpdir, "checkpoint-30")
        trainer._rotate_checkpoints(use_mtime=False, output_dir=tmpdir)
        left = sorted(d for d in os.listdir(tmpdir) if d.startswith("checkpoint-"))
        assert left == ["checkpoint-20", "checkpoint-30", "checkpoint-40"]  # 40 is not complete yet
        # a save: every torch.save of the Trainer is captured, goes to shm and is
        # persisted by the agent at the paths the Trainer chose
        trainer.state.global_step = 50
        trainer._save_checkpoint(model, None)
        assert args.save_safetensors is True  # restored
        agent = trainer.flash_checkpointer.ckpt_agent
        names = {"pytorch_model.bin", "optimizer.pt", "scheduler.pt", "rng_state.pth"}
        assert set(agent.paths) == names and names <= set(agent.state_dict)
        out = os.path.join(tmpdir, "checkpoint-50")
        tracker = os.path.join(tmpdir, "dlrover_latest.txt")
        assert _wait_for(lambda: open(tracker).read() == "50")
        assert _wait_for(lambda: all(os.path.exists(os.path.join(out, n)) for n in names))
        weights = torch.load(os.path.join(out, "pytorch_model.bin"), weights_only=False)
        assert torch.equal(weights["fc1.weight"], model.fc1.weight)
        optim = torch.load(os.path.join(out, "optimizer.pt"), weights_only=False)
        assert torch.equal(optim["state"][0]["momentum_buffer"],
                           trainer.optimizer.state_dict()["state"][0]["momentum_buffer"])
        assert os.path.exists(os.path.join(out, "trainer_state.json"))  # written by the Trainer itself
        assert trainer.get_last_checkpoint() == out
        trainer.flash_checkpointer.async_save_engine.close()
