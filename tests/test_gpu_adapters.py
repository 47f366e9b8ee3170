"""GPU: the DeepSpeed and Megatron adapters (and the distributed-optimizer
variant) on CUDA tensors, one GPU, driven like the reference's own adapter
tests — a mock DeepSpeedEngine that just torch.save()s
(dlrover/trainer/tests/torch/deepspeed_ckpt_test.py:35-72) and patched
Megatron hooks (megatron_ckpt_test.py:84-115).

For each adapter: the bytes in shm equal the oracle's image of the captured
state dict (and its per-chunk CRC-32C), the restore runs through the scatter
kernel (``last_stats["restore_path"]``) and hands the framework CUDA tensors,
and the model / optimizer end up bit-identical.
"""
import argparse
import os
import tempfile

import numpy as np
import pytest
import torch
import torch.nn as nn
import torch.optim as optim

os.environ.setdefault("ROLE_NAME", "dlrover-trainer")
os.environ.setdefault("TORCHELASTIC_RUN_ID", "fcp_test_%d" % os.getpid())

from dlrover_wuqiong_b200.ckpt_saver import FCP_CHECKSUM_KEY, AsyncCheckpointSaver  # noqa: E402
from dlrover_wuqiong_b200.common.constants import CheckpointConstant, NodeEnv  # noqa: E402
from dlrover_wuqiong_b200.common.multi_process import clear_sock_dir  # noqa: E402
from dlrover_wuqiong_b200.common.storage import PosixDiskStorage  # noqa: E402
from dlrover_wuqiong_b200.flash_checkpoint import megatron  # noqa: E402
from dlrover_wuqiong_b200.flash_checkpoint.checkpointer import StorageType  # noqa: E402
from dlrover_wuqiong_b200.flash_checkpoint.deepspeed import DeepSpeedCheckpointer  # noqa: E402
from dlrover_wuqiong_b200.flash_checkpoint.megatron import (  # noqa: E402
    MegatronCheckpointer,
    load_checkpoint,
    save_checkpoint,
)
from dlrover_wuqiong_b200.flash_checkpoint.megatron_dist_ckpt import (  # noqa: E402
    get_parameter_state,
    load_parameter_state_from_state_dict,
)
from dlrover_wuqiong_b200.flash_checkpoint.megatron_engine import (  # noqa: E402
    MegatronDistCheckpointEngine,
)
from oracle import oracle  # noqa: E402

pytestmark = pytest.mark.gpu
KERNEL_PATHS = ("registered", "bounced")


class Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.fc1 = nn.Linear(257, 129)   # odd sizes: misaligned offsets
        self.fc2 = nn.Linear(129, 31)
        self.big = nn.Linear(1024, 1500)  # a few whole regions


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


def _trained(dev):
    torch.manual_seed(11)
    model = Net().to(dev)
    opt = optim.AdamW(model.parameters(), lr=0.01)
    loss = (model.fc2(model.fc1(torch.randn(4, 257, device=dev))).sum()
            + model.big(torch.randn(2, 1024, device=dev)).sum())
    loss.backward()
    opt.step()
    return model, opt


def _assert_shm_is_oracle_image(handler, captured):
    """The segment holds exactly what the reference's copy loop would have
    written for `captured` (+ the config leaf, which has no bytes), and the
    recorded sums are the CRC-32C of those bytes."""
    _, total, expect = oracle.pack(captured)
    assert handler.shared_memory.size == total
    got = np.frombuffer(handler.shared_memory.buf, dtype=np.uint8)
    assert got.tobytes() == expect.tobytes()
    entry = handler.metadata.get()[FCP_CHECKSUM_KEY]
    assert entry["sums"] == oracle.chunk_crc32c(expect, entry["chunk_bytes"]).tolist()
    del got


def _snapshot(model, opt):
    params = {k: v.detach().clone() for k, v in model.state_dict().items()}
    state = {i: {k: (v.clone() if torch.is_tensor(v) else v) for k, v in s.items()}
             for i, s in enumerate(opt.state.values())}
    return params, state


def _wreck(model, opt):
    with torch.no_grad():
        for p in model.parameters():
            p.add_(1.0)
        for s in opt.state.values():
            s["exp_avg"].zero_()


def _assert_restored(model, opt, snap):
    params, state = snap
    for k, v in model.state_dict().items():
        assert torch.equal(v, params[k]), k
    for i, s in enumerate(opt.state.values()):
        for k, v in s.items():
            if torch.is_tensor(v):
                assert torch.equal(v, state[i][k]), (i, k)


class MockDeepSpeedEngine(object):
    """deepspeed_ckpt_test.py:35-72, with CUDA tensors."""

    def __init__(self, model, optimizer):
        self.model, self.optimizer = model, optimizer
        self.save_non_zero_checkpoint = True
        self.global_rank = 0
        self.loaded_devices = set()

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
        msd = torch.load(os.path.join(load_dir, tag, "model_states.pt"))
        self.loaded_devices |= {v.device.type for v in msd.values()}
        self.model.load_state_dict(msd)
        self.optimizer.load_state_dict(
            torch.load(os.path.join(load_dir, tag, "optim_states.pt")))
        return os.path.join(load_dir, tag), {}


def test_deepspeed_adapter_on_cuda(saver_env):
    dev = torch.device("cuda", 0)
    model, opt = _trained(dev)
    with tempfile.TemporaryDirectory() as tmpdir:
        engine = MockDeepSpeedEngine(model, opt)
        ckpt = DeepSpeedCheckpointer(engine, tmpdir)
        assert ckpt.save_checkpoint(tmpdir, 100, storage_type=StorageType.MEMORY)
        handler = ckpt._async_save_engine._shm_handler
        _assert_shm_is_oracle_image(handler, {"model_states": model.state_dict(),
                                              "optim_states": opt.state_dict()})
        snap = _snapshot(model, opt)
        _wreck(model, opt)
        ckpt.load_checkpoint(tmpdir, "100")
        assert handler.last_stats["restore_path"] in KERNEL_PATHS
        assert engine.loaded_devices == {"cuda"}  # no CPU views + per-tensor H2D
        _assert_restored(model, opt, snap)
        ckpt.close()


def test_megatron_adapter_on_cuda(saver_env, monkeypatch):
    dev = torch.device("cuda", 0)
    model, opt = _trained(dev)
    suffix = "model_optim_rng.pt"
    with tempfile.TemporaryDirectory() as tmpdir:
        def get_args():
            args, _ = argparse.ArgumentParser().parse_known_args([])
            args.save = tmpdir
            return args

        captured = {}

        def mock_save(iteration, model, optimizer, opt_scheduler):
            sd = {"iteration": iteration, "model": model.state_dict(),
                  "optimizer": optimizer.state_dict(),
                  # Megatron keeps rng states as host tensors
                  "rng_state": [{"torch_rng_state": torch.get_rng_state(),
                                 "cuda_rng_state": torch.cuda.get_rng_state(dev)}]}
            captured.update(sd)
            torch.save(sd, os.path.join(tmpdir, "iter_%07d" % iteration, suffix))

        seen = {}

        def mock_load(model, optimizer, opt_param_scheduler, load_arg="load", strict=True):
            sd = torch.load(os.path.join(tmpdir, "iter_0000010", suffix))
            seen["model_dev"] = {v.device.type for v in sd["model"].values()}
            seen["rng_dev"] = sd["rng_state"][0]["torch_rng_state"].device.type
            model.load_state_dict(sd["model"])
            optimizer.load_state_dict(sd["optimizer"])
            torch.set_rng_state(sd["rng_state"][0]["torch_rng_state"])  # needs a CPU tensor
            torch.cuda.set_rng_state(sd["rng_state"][0]["cuda_rng_state"], dev)
            return sd["iteration"]

        monkeypatch.setattr(megatron, "megatron_save", mock_save)
        monkeypatch.setattr(megatron, "megatron_load", mock_load)
        monkeypatch.setattr(megatron, "get_args", get_args)
        mgr = MegatronCheckpointer.singleton_instance(tmpdir)
        assert save_checkpoint(10, model, opt, None, storage_type=StorageType.MEMORY)
        handler = mgr.engine._shm_handler
        _assert_shm_is_oracle_image(handler, {"model_states": captured})
        snap = _snapshot(model, opt)
        _wreck(model, opt)
        assert load_checkpoint(model, opt, None) == 10
        assert handler.last_stats["restore_path"] in KERNEL_PATHS
        assert seen == {"model_dev": {"cuda"}, "rng_dev": "cpu"}
        _assert_restored(model, opt, snap)
        mgr.engine.close()


def test_megatron_dist_optimizer_shards_on_cuda(saver_env):
    dev = torch.device("cuda", 0)
    torch.manual_seed(3)
    p0 = nn.Parameter(torch.randn(70001, device=dev))
    p1 = nn.Parameter(torch.randn(333, device=dev))
    inner = optim.Adam([p0, p1], lr=0.1)
    (p0.sum() + p1.sum()).backward()
    inner.step()

    class FakeDistOpt(object):
        optimizer = inner
        gbuf_ranges = [{torch.float32: [{"param_map": {"m0": None, "m1": None}}]}]
        model_param_group_index_map = {"m0": (0, 0), "m1": (0, 1)}

    with tempfile.TemporaryDirectory() as tmpdir:
        d = MegatronDistCheckpointEngine(tmpdir, PosixDiskStorage())
        state = get_parameter_state(FakeDistOpt())
        sd = {CheckpointConstant.OPTIM_STATES_NAME: state}
        paths = {CheckpointConstant.OPTIM_STATES_NAME: os.path.join(tmpdir, "o.pt")}
        assert d.save_to_memory(7, sd, paths)
        _assert_shm_is_oracle_image(d._shm_handler, sd)
        want = (p0.detach().clone(), inner.state[p1]["exp_avg"].clone())
        # (a) the reference's per-tensor loop, fed device tensors by one scatter launch
        with torch.no_grad():
            p0.zero_()
            inner.state[p1]["exp_avg"].zero_()
        assert d.restore_device() == dev
        step, got = d.load(device=d.restore_device())
        assert step == 7 and d._shm_handler.last_stats["restore_path"] in KERNEL_PATHS
        restored = got[CheckpointConstant.OPTIM_STATES_NAME]
        assert restored[0][0][0]["param"].is_cuda
        load_parameter_state_from_state_dict(FakeDistOpt(), restored)
        assert torch.equal(p0.detach(), want[0])
        assert torch.equal(inner.state[p1]["exp_avg"], want[1])
        # (b) in place: the live shards are the scatter targets
        with torch.no_grad():
            p0.zero_()
            inner.state[p1]["exp_avg"].zero_()
        step, got = d.load(into={CheckpointConstant.OPTIM_STATES_NAME:
                                 get_parameter_state(FakeDistOpt())})
        assert step == 7
        assert torch.equal(p0.detach(), want[0])
        assert torch.equal(inner.state[p1]["exp_avg"], want[1])
        assert got[CheckpointConstant.OPTIM_STATES_NAME][0][0][0]["param"] is p0
        d.close()


def test_partial_into_allocates_the_rest_on_the_device(saver_env):
    """`into` names only the optimizer shards; the model entry is allocated on
    `device`, host-resident leaves stay on the host."""
    from dlrover_wuqiong_b200.flash_checkpoint.ddp import DdpCheckpointer
    dev = torch.device("cuda", 0)
    torch.manual_seed(5)
    sd = {"model": {"w": torch.randn(50000, device=dev), "b": torch.randn(77, device=dev)},
          "optim": {"m": torch.randn(50000, device=dev), "step": torch.tensor(3.0)},
          "rng": torch.get_rng_state(), "epoch": 4}
    with tempfile.TemporaryDirectory() as tmpdir:
        ckpt = DdpCheckpointer(tmpdir)
        assert ckpt.save_checkpoint(5, sd, storage_type=StorageType.MEMORY)
        m_want = sd["optim"]["m"].clone()
        sd["optim"]["m"].zero_()
        got = ckpt.load_checkpoint(into={"optim": {"m": sd["optim"]["m"]}}, device=dev)
        assert got["optim"]["m"] is sd["optim"]["m"] and torch.equal(sd["optim"]["m"], m_want)
        assert got["model"]["w"].is_cuda and torch.equal(got["model"]["w"], sd["model"]["w"])
        assert not got["rng"].is_cuda and torch.equal(got["rng"], sd["rng"])
        assert not got["optim"]["step"].is_cuda and got["epoch"] == 4
        with pytest.raises(ValueError):  # wrong size where `into` does name a leaf
            ckpt.load_checkpoint(into={"optim": {"m": torch.zeros(5, device=dev)}}, device=dev)
        ckpt.close()


def test_cpu_views_of_a_pinned_segment_copy_to_the_gpu(saver_env):
    """The reference's restore: load_checkpoint() returns CPU views and the
    user copies them up per tensor.  In the process that saved, the segment is
    pinned in 64 MiB pieces; a view that straddles two pieces must still be
    an ordinary (pageable) source for cuda_tensor.copy_()."""
    from dlrover_wuqiong_b200.flash_checkpoint.ddp import DdpCheckpointer
    dev = torch.device("cuda", 0)
    torch.manual_seed(8)
    sd = {"small": torch.randn(1000, device=dev),
          "big": torch.randn(48 << 20, device=dev),     # 192 MiB: spans several pieces
          "tail": torch.randn(12345, device=dev)}
    with tempfile.TemporaryDirectory() as tmpdir:
        ckpt = DdpCheckpointer(tmpdir)
        assert ckpt.save_checkpoint(1, sd, storage_type=StorageType.MEMORY)
        views = ckpt.load_checkpoint()
        assert not views["big"].is_cuda
        for k, v in sd.items():
            up = torch.empty_like(v)
            up.copy_(views[k])
            torch.cuda.synchronize()
            assert torch.equal(up, v), k
        del views
        ckpt.close()


def test_hf_flash_ckpt_trainer_on_cuda(saver_env, monkeypatch):
    """FlashCkptTrainer (stand-in Trainer, see tests/test_adapters_cpu.py) with
    the model and optimizer on the GPU: every torch.save of the Trainer is
    captured, the captured state goes through the pack kernel (shm bytes ==
    oracle image, CRC recorded) and the agent persists files that load back
    equal.  The reference's HF adapter has no restore-from-memory path
    (hf_trainer.py:123-257 only saves), so there is none to test."""
    import sys
    import time
    import types

    import dlrover_wuqiong_b200.flash_checkpoint.hf_trainer as hf
    dev = torch.device("cuda", 0)

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
            torch.save(model.state_dict(), os.path.join(out, "pytorch_model.bin"))
            torch.save(self.optimizer.state_dict(), os.path.join(out, "optimizer.pt"))
            torch.save({"cpu": torch.get_rng_state()}, os.path.join(out, "rng_state.pth"))

        def _sorted_checkpoints(self, use_mtime=False, output_dir=None):
            return []

    fake = types.ModuleType("transformers")
    fake.Trainer = Trainer
    utils = types.ModuleType("transformers.trainer_utils")
    utils.PREFIX_CHECKPOINT_DIR = "checkpoint"
    fake.trainer_utils = utils
    monkeypatch.setitem(sys.modules, "transformers", fake)
    monkeypatch.setitem(sys.modules, "transformers.trainer_utils", utils)
    monkeypatch.setattr(hf, "_TRAINER_CLS", None)
    with tempfile.TemporaryDirectory() as tmpdir:
        args = types.SimpleNamespace(output_dir=tmpdir, save_safetensors=True, save_total_limit=None)
        model = Net().to(dev)
        trainer = hf.FlashCkptTrainer(model, args)
        (model.fc2(model.fc1(torch.randn(2, 257, device=dev))).sum()
         + model.big(torch.randn(2, 1024, device=dev)).sum()).backward()
        trainer.optimizer.step()
        trainer.state.global_step = 7
        trainer._save_checkpoint(model, None)
        flash = trainer.flash_checkpointer
        handler = flash.async_save_engine._shm_handler
        captured = {k: v for k, v in flash.ckpt_agent.state_dict.items()
                    if k.endswith((".bin", ".pt", ".pth"))}  # (the engine adds its config leaf)
        assert set(captured) == {"pytorch_model.bin", "optimizer.pt", "rng_state.pth"}
        out = os.path.join(tmpdir, "checkpoint-7")
        tracker = os.path.join(tmpdir, "dlrover_latest.txt")
        t0 = time.time()
        while time.time() - t0 < 60 and not (os.path.exists(tracker)
                                             and open(tracker).read() == "7"):
            time.sleep(0.1)
        assert open(tracker).read() == "7"
        _assert_shm_is_oracle_image(handler, captured)
        assert handler.last_stats.get("save_mode") is not None  # through libfcp
        weights = torch.load(os.path.join(out, "pytorch_model.bin"), weights_only=False)
        assert torch.equal(weights["big.weight"], model.big.weight.cpu())
        optim = torch.load(os.path.join(out, "optimizer.pt"), weights_only=False)
        assert torch.equal(optim["state"][0]["momentum_buffer"],
                           trainer.optimizer.state_dict()["state"][0]["momentum_buffer"].cpu())
        flash.async_save_engine.close()
