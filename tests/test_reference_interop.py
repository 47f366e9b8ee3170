"""CPU: the stock reference agent hosting this package's saver.

The agent process of a DLRover job runs the REFERENCE's factory thread
(``AsyncCheckpointSaver.start_async_saving_ckpt``, ckpt_saver.py:409-465 there);
it builds whatever saver class the trainer names over the "factory" queue.  A
trainer that uses this package names ``dlrover_wuqiong_b200.ckpt_saver.
DdpCheckpointSaver``: the reference's thread must be able to host it, serve the
trainer's lock / meta / event queue and persist a correct ``rank_0.pt`` plus
tracker file on a DISK save.

Needs the reference's sources (/root/reference in the build container, or the
closure build() ships to baseline/_ref); skipped where neither exists.
"""
import os
import tempfile
import time

import pytest
import torch

os.environ.setdefault("ROLE_NAME", "dlrover-trainer")
os.environ.setdefault("TORCHELASTIC_RUN_ID", "fcp_test_%d" % os.getpid())

from oracle import reference_closure  # noqa: E402

pytestmark = pytest.mark.skipif(reference_closure.available_root() is None,
                                reason="the reference's sources are not available here")


def test_reference_factory_thread_hosts_this_packages_saver():
    ref = reference_closure.import_reference()
    from dlrover_wuqiong_b200.ckpt_saver import DdpCheckpointSaver
    from dlrover_wuqiong_b200.common.constants import CheckpointConstant
    from dlrover_wuqiong_b200.common.multi_process import clear_sock_dir
    from dlrover_wuqiong_b200.flash_checkpoint.ddp import DdpCheckpointer, StorageType
    os.environ["LOCAL_RANK"] = "0"
    os.environ["NODE_NUM"], os.environ["NODE_RANK"] = "1", "0"
    ref.AsyncCheckpointSaver._saver_instance = None
    ref.AsyncCheckpointSaver.start_async_saving_ckpt()  # the REFERENCE's factory thread
    with tempfile.TemporaryDirectory() as tmpdir:
        ckpt = DdpCheckpointer(tmpdir)
        try:
            torch.manual_seed(4)
            sd = {"model": {"w": torch.randn(300, 7), "b": torch.randn(7)},
                  "optimizer": {"state": {0: {"step": torch.tensor(2.0),
                                              "exp_avg": torch.randn(300, 7)}}},
                  "step": 20}
            assert ckpt.save_checkpoint(20, sd, storage_type=StorageType.DISK)
            tracker = os.path.join(tmpdir, CheckpointConstant.TRACER_FILE_NAME)
            t0 = time.time()
            while time.time() - t0 < 60 and not (os.path.exists(tracker)
                                                 and open(tracker).read() == "20"):
                time.sleep(0.1)
            assert os.path.exists(tracker) and open(tracker).read() == "20"
            hosted = ref.AsyncCheckpointSaver.get_ckpt_saver()
            assert type(hosted) is DdpCheckpointSaver  # built by the reference's thread
            back = torch.load(os.path.join(tmpdir, "20", "rank_0.pt"), weights_only=False)
            assert back["step"] == 20
            assert torch.equal(back["model"]["w"], sd["model"]["w"])
            assert torch.equal(back["optimizer"]["state"][0]["exp_avg"],
                               sd["optimizer"]["state"][0]["exp_avg"])
            # and the memory copy is served back to the trainer
            views = ckpt.load_checkpoint()
            assert torch.equal(views["model"]["b"], sd["model"]["b"])
            del views
        finally:
            ckpt.close()
            hosted = ref.AsyncCheckpointSaver.get_ckpt_saver()
            if hosted is not None:
                hosted.close()
            ref.AsyncCheckpointSaver._saver_instance = None
            clear_sock_dir()
