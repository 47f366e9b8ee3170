"""TEST INFRASTRUCTURE.  The small sharded state used to pin the FSDP / DCP
shm writer to the reference's (oracle/gen_golden_fsdp.py writes the fixture,
tests/test_fsdp_golden.py replays it): ShardedTensor leaves (SHARD items, what
FSDP's SHARDED_STATE_DICT yields), plain tensors incl. a 0-dim AdamW ``step``
(TENSOR items) and non-tensor leaves (BYTE_IO items), with odd sizes so that
offsets are not aligned to anything."""
import torch
from torch.distributed._shard import sharded_tensor
from torch.distributed._shard.sharded_tensor import Shard, ShardMetadata


def make_inputs(seed=20240607):
    g = torch.Generator().manual_seed(seed)
    return {
        "w1": torch.randn(40, 16, generator=g),
        "w2": torch.randn(33, 7, generator=g).to(torch.bfloat16),
        "w3": torch.randn(1031, 33, generator=g),          # > 16 KiB: whole-region path
        "b": torch.randn(5, generator=g),
        "exp_avg_w1": torch.randn(40, 16, generator=g),
        "exp_avg_sq_w1": torch.rand(40, 16, generator=g),
        "exp_avg_w3": torch.randn(1031, 33, generator=g),
        "step": torch.tensor(3.0),
    }


def _sharded(t, dev):
    """One-rank ShardedTensor whose local shard is `t` (on `dev`)."""
    t = t.to(dev).contiguous()
    where = "rank:0/%s" % ("cpu" if t.device.type == "cpu" else "cuda:%d" % t.device.index)
    md = ShardMetadata(shard_offsets=[0] * t.dim(), shard_sizes=list(t.shape), placement=where)
    return sharded_tensor.init_from_local_shards([Shard(t, md)], *t.shape)


def build(inputs, dev="cpu", zeros=False):
    """The state dict handed to dist_cp.save / load.  `zeros`: same structure
    with zeroed tensors (a load target)."""
    dev = torch.device(dev)

    def v(name):
        t = inputs[name]
        return torch.zeros_like(t) if zeros else t.clone()

    return {
        "model": {"w1": _sharded(v("w1"), dev), "w2": _sharded(v("w2"), dev),
                  "w3": _sharded(v("w3"), dev), "b": v("b").to(dev)},
        "optim": {"state": {"w1": {"exp_avg": _sharded(v("exp_avg_w1"), dev),
                                   "exp_avg_sq": _sharded(v("exp_avg_sq_w1"), dev),
                                   "step": v("step").to(dev)},
                            "w3": {"exp_avg": _sharded(v("exp_avg_w3"), dev)}},
                  "param_groups": [{"lr": 0.1, "betas": (0.9, 0.95), "params": ["w1", "w3"]}]},
        "step": 7,
    }


def local_tensors(sd):
    """{dotted name: local tensor} of every tensor leaf."""
    out = {}

    def walk(v, path):
        if isinstance(v, dict):
            for k, x in v.items():
                walk(x, path + [str(k)])
        elif isinstance(v, list):
            for i, x in enumerate(v):
                walk(x, path + [str(i)])
        elif hasattr(v, "local_shards"):
            out[".".join(path)] = v.local_shards()[0].tensor
        elif torch.is_tensor(v):
            out[".".join(path)] = v

    walk(sd, [])
    return out
