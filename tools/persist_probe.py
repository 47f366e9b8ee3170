"""Agent-side persistence: torch.save of shm views (the reference's saver,
ckpt_saver.py:914-921) vs the raw image (raw_ckpt.write_raw), and the matching
loads.  CPU only.  usage: python tools/persist_probe.py [GB] [target dir]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dlrover_wuqiong_b200 import plan, raw_ckpt  # noqa: E402
from dlrover_wuqiong_b200.ckpt_saver import _host_copy_leaves, _read_state_dict_from_shm  # noqa: E402


class _Shm(object):
    def __init__(self, buf):
        self.buf = buf


def main():
    gb = float(sys.argv[1]) if len(sys.argv) > 1 else 1.6
    target = sys.argv[2] if len(sys.argv) > 2 else "/dev/shm"
    n_layers = max(1, int(gb * 1e9 / (12 * 768 * 768 * 4 * 3)))
    g = torch.Generator().manual_seed(0)
    sd = {"model_states": {"model": {}, "optimizer": {"state": {}}, "step": 7}}
    k = 0
    for layer in range(n_layers):
        for name, shape in (("attn.w", (3 * 768, 768)), ("attn.proj", (768, 768)),
                            ("mlp.fc", (4 * 768, 768)), ("mlp.proj", (768, 4 * 768)),
                            ("ln.w", (768,))):
            w = torch.randn(shape, generator=g)
            sd["model_states"]["model"]["h.%d.%s" % (layer, name)] = w
            sd["model_states"]["optimizer"]["state"][k] = {
                "step": torch.tensor(float(layer)), "exp_avg": w * 0.1, "exp_avg_sq": w * w}
            k += 1
    meta, leaves, total = plan.build_layout(sd)
    buf = bytearray(total)
    _host_copy_leaves(leaves, buf)
    print("state: %d tensors, %.2f GB -> %s" % (len(leaves), total / 1e9, target))
    views = _read_state_dict_from_shm(meta, _Shm(buf))["model_states"]
    pt, raw = os.path.join(target, "probe.pt"), os.path.join(target, "probe.raw")
    res = {}
    for rep in range(3):
        for name, fn in (("torch.save", lambda: torch.save(views, pt)),
                         ("write_raw", lambda: raw_ckpt.write_raw(
                             raw_ckpt.RawImage(meta["model_states"], buf, 7), raw))):
            t0 = time.perf_counter()
            fn()
            res.setdefault(name, []).append(time.perf_counter() - t0)
        for name, fn in (("torch.load", lambda: torch.load(pt, weights_only=False)),
                         ("torch.load(mmap)", lambda: torch.load(pt, weights_only=False, mmap=True)),
                         ("read_raw", lambda: raw_ckpt.read_raw(raw))):
            t0 = time.perf_counter()
            out = fn()
            # touch every byte so lazily mapped loads are charged too
            s = 0
            for v in out["model"].values():
                s += int(v.view(torch.int32).sum())
            for st in out["optimizer"]["state"].values():
                s += int(st["exp_avg"].view(torch.int32).sum()) + int(st["exp_avg_sq"].view(torch.int32).sum())
            res.setdefault(name + " + touch", []).append(time.perf_counter() - t0)
            del out
    for name, ts in res.items():
        print("%-24s best %.3f s  = %.2f GB/s   (runs: %s)" % (
            name, min(ts), total / 1e9 / min(ts), " ".join("%.3f" % t for t in ts)))
    os.remove(pt)
    os.remove(raw)


if __name__ == "__main__":
    main()
