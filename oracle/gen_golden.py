"""TEST INFRASTRUCTURE.  Generates tests/golden/ref_shm_cases.pt by running the
REFERENCE's own SharedMemoryHandler (imported read-only from /root/reference,
with the two-module stub of SURVEY Appendix B) on seeded state dicts.

Run in the build container only (the GPU box has no /root/reference):

    python oracle/gen_golden.py

The fixture holds, per case: the input state dict (plain tensors / scalars),
the meta tree the reference produced (TensorMeta flattened to a plain dict),
the shm size and the exact shm bytes.  Nothing in it needs the reference to be
importable when it is loaded.
"""
import os
import sys
import types

import torch
import torch.nn as nn

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests",
                   "golden", "ref_shm_cases.pt")


def import_reference():
    os.environ.setdefault("ROLE_NAME", "dlrover-trainer")
    os.environ["TORCHELASTIC_RUN_ID"] = "fcp_golden_%d" % os.getpid()
    sys.dont_write_bytecode = True
    m = types.ModuleType("dlrover.python.elastic_agent.master_client")
    m.MasterClient = type("MasterClient", (), {})
    sys.modules[m.__name__] = m
    sys.path.insert(0, REF)
    from dlrover.python.elastic_agent.torch import ckpt_saver  # noqa
    return ckpt_saver


class SimpleNet(nn.Module):  # shapes of dlrover/python/tests/test_ckpt_saver.py:57-70
    def __init__(self):
        super().__init__()
        self.fc1 = nn.Linear(64, 32)
        self.fc2 = nn.Linear(32, 10)


class ToyModel(nn.Module):  # dlrover/trainer/tests/torch/checkpoint_backup_test.py:40-47
    def __init__(self):
        super().__init__()
        self.net1 = nn.Linear(16, 16)
        self.net2 = nn.Linear(16, 8)


def make_cases():
    cases = {}
    torch.manual_seed(1234)
    m = SimpleNet()
    opt = torch.optim.SGD(m.parameters(), lr=0.01, momentum=0.001)
    cases["simplenet_sgd_9640"] = {
        "model_states": m.state_dict(), "optim_states": opt.state_dict()}

    m = ToyModel()
    opt = torch.optim.Adam(m.parameters(), lr=0.1)
    cases["toymodel_adam_1632"] = {
        "model": m.state_dict(), "optimizer": opt.state_dict()}

    # SURVEY Appendix C quirk set: odd offsets, transposed source, 0-dim,
    # bool/int8/fp64, the same tensor twice, a tuple leaf, an empty tensor
    a = torch.arange(12, dtype=torch.float32).reshape(3, 4).t()
    s = torch.tensor(3.5)
    b = torch.tensor([True, False, True])
    c = torch.randn(2, 2)
    i8 = torch.tensor([1, -2, 3], dtype=torch.int8)
    f64 = torch.tensor([2.718281828], dtype=torch.float64)
    cases["quirks"] = {
        "a_t": a, "scalar": s, "flags": b, "c": c, "c_again": c, "i8": i8,
        "f64": f64, "tuple_leaf": (torch.ones(2), 7), "empty": torch.zeros(0, 3),
        "nested": {"lst": [torch.randn(5).to(torch.bfloat16), "txt",
                           [torch.randn(3).half(), 11]],
                   "i64": torch.arange(7)},
        "u8": torch.arange(13, dtype=torch.uint8), "step": 42, "name": "x"}

    # AdamW after one step: 0-dim fp32 `step` before every exp_avg forces
    # 4-byte-shifted offsets (the common misaligned case)
    m = nn.Sequential(nn.Linear(33, 17), nn.Linear(17, 5))
    opt = torch.optim.AdamW(m.parameters(), lr=1e-3)
    m(torch.randn(4, 33)).sum().backward()
    opt.step()
    cases["mlp_adamw_stepped"] = {
        "model": m.state_dict(), "optimizer": opt.state_dict(), "step": 1}

    # strided / sliced sources
    base = torch.randn(16, 24)
    cases["noncontig"] = {
        "col_slice": base[:, 3:11], "row_step": base[::3], "t": base.t(),
        "odd_bf16": torch.randn(37).to(torch.bfloat16), "tail": torch.randn(3)}
    return cases


def plain_meta(ref, meta):
    def visit(v):
        if isinstance(v, ref.TensorMeta):
            return {"__tensor_meta__": True, "shape": tuple(v.shape),
                    "dtype": v.dtype, "element_size": v.element_size,
                    "numel": v.numel, "offset": v.offset}
        if isinstance(v, ref.CheckpointConfig):
            return {"__ckpt_config__": True, "step": v.step,
                    "writing_shm": v.writing_shm, "rank": v.rank,
                    "world_size": v.world_size, "paths": v.paths}
        return v
    return ref._traverse_state_dict(meta, visit)


def run_reference(ref, idx, state_dict, step):
    h = ref.SharedMemoryHandler(100 + idx, host=True)
    sd = dict(state_dict)
    sd[ref.DLROVER_CKPT_CONFIG_KEY] = ref.CheckpointConfig(step=step, paths={})
    h.save_state_dict(sd)
    meta = h.metadata.get()
    raw = bytes(h.shared_memory.buf[:h.shared_memory.size])
    out = {"meta": plain_meta(ref, meta), "size": h.shared_memory.size, "shm": raw}
    return h, out


def main():
    ref = import_reference()
    fixture = {"torch": torch.__version__, "cases": {}}
    for idx, (name, sd) in enumerate(make_cases().items()):
        h, out = run_reference(ref, idx, sd, step=5)
        out["state_dict"] = sd
        # second save on the same handler: offsets are reused, values and
        # non-tensor leaves refreshed (SAVER:281-290)
        if name == "mlp_adamw_stepped":
            sd2 = {"model": {k: v + 1 for k, v in sd["model"].items()},
                   "optimizer": sd["optimizer"], "step": 2}
            sd2w = dict(sd2)
            sd2w[ref.DLROVER_CKPT_CONFIG_KEY] = ref.CheckpointConfig(step=6, paths={})
            h.save_state_dict(sd2w)
            out["second"] = {
                "state_dict": sd2,
                "meta": plain_meta(ref, h.metadata.get()),
                "shm": bytes(h.shared_memory.buf[:h.shared_memory.size])}
        fixture["cases"][name] = out
        print(name, out["size"])
        h.unlink()
    assert fixture["cases"]["simplenet_sgd_9640"]["size"] == 9640
    assert fixture["cases"]["toymodel_adam_1632"]["size"] == 1632
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    torch.save(fixture, OUT)
    print("wrote", os.path.normpath(OUT), os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
