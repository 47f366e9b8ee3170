"""Meta IPC cost per save: full `set` of the meta tree vs the path update
(config + checksums + non-tensor leaves).  CPU only."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["TORCHELASTIC_RUN_ID"] = "ipcprobe%d" % os.getpid()
from dlrover_wuqiong_b200.ckpt_saver import CheckpointConfig, DLROVER_CKPT_CONFIG_KEY, FCP_CHECKSUM_KEY
from dlrover_wuqiong_b200.common.multi_process import SharedDict
from dlrover_wuqiong_b200.plan import TensorMeta
srv = SharedDict("m", create=True)
cli = SharedDict("m", create=False)
print("| tensors | full set ms | path update ms |\n|---|---|---|")
for n in (300, 1165, 10000):
    meta = {"model_states": {"t%d" % i: TensorMeta((1600, 1600), torch.float32, 4, 2560000, i * 10240000)
                             for i in range(n)}, "step": 1,
            DLROVER_CKPT_CONFIG_KEY: CheckpointConfig(step=1, paths={"model_states": "/x"}),
            FCP_CHECKSUM_KEY: {"algo": "crc32c", "chunk_bytes": 2 << 20, "sums": list(range(n * 5))}}
    cli.set(meta)
    t0 = time.time()
    for _ in range(5):
        cli.set(meta)
    t_set = (time.time() - t0) / 5
    upd = [((DLROVER_CKPT_CONFIG_KEY,), meta[DLROVER_CKPT_CONFIG_KEY]), (("step",), 2),
           ((FCP_CHECKSUM_KEY,), meta[FCP_CHECKSUM_KEY])]
    t0 = time.time()
    for _ in range(5):
        assert cli.update(upd)
    t_upd = (time.time() - t0) / 5
    print("| %d | %.1f | %.1f |" % (n, t_set * 1e3, t_upd * 1e3))
srv.unlink()
