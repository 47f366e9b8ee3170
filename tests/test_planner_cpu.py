"""CPU: the product's planner / host path against the oracle on random nested
state dicts (property test), and the loud failure when the CUDA extension is
missing."""
import os

import numpy as np
import pytest
import torch
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

os.environ.setdefault("ROLE_NAME", "dlrover-trainer")
os.environ.setdefault("TORCHELASTIC_RUN_ID", "fcp_test_%d" % os.getpid())

from dlrover_wuqiong_b200 import _native, plan  # noqa: E402
from dlrover_wuqiong_b200.ckpt_saver import (  # noqa: E402
    DLROVER_CKPT_CONFIG_KEY,
    CheckpointConfig,
    SharedMemoryHandler,
)
from oracle import oracle  # noqa: E402

DTYPES = [torch.float32, torch.bfloat16, torch.float16, torch.float64, torch.int64,
          torch.int8, torch.uint8, torch.bool, torch.int16]


def _tensor(draw):
    dt = draw(st.sampled_from(DTYPES))
    shape = draw(st.lists(st.integers(0, 5), min_size=0, max_size=3))
    t = torch.randint(0, 2, shape).to(dt) if dt == torch.bool else \
        (torch.rand(shape) * 100).to(dt)
    if len(shape) == 2 and draw(st.booleans()):
        t = t.t()  # non-contiguous source
    return t


@st.composite
def state_dicts(draw, depth=0):
    n = draw(st.integers(0, 4))
    out = {}
    for i in range(n):
        kind = draw(st.integers(0, 5 if depth < 2 else 3))
        key = draw(st.sampled_from(["a", "b", "w", 0, 1, "step"])) if draw(st.booleans()) \
            else "k%d" % i
        if kind <= 1:
            out[key] = _tensor(draw)
        elif kind == 2:
            out[key] = draw(st.one_of(st.integers(-5, 5), st.text(max_size=3), st.none(),
                                      st.floats(allow_nan=False, allow_infinity=False)))
        elif kind == 3:
            out[key] = (_tensor(draw), 3)  # tuples are leaves
        elif kind == 4:
            out[key] = draw(state_dicts(depth=depth + 1))
        else:
            out[key] = [_tensor(draw), draw(state_dicts(depth=depth + 1)), "x"]
    return out


def _same_meta(o, p):
    if isinstance(o, oracle.OMeta):
        assert isinstance(p, plan.TensorMeta)
        assert (o.shape, o.dtype, o.element_size, o.numel, o.offset) == \
            (tuple(p.shape), p.dtype, p.element_size, p.numel, p.offset)
    elif isinstance(o, dict):
        assert list(o.keys()) == list(p.keys())
        for k in o:
            _same_meta(o[k], p[k])
    elif isinstance(o, list):
        assert len(o) == len(p)
        for a, b in zip(o, p):
            _same_meta(a, b)


@settings(max_examples=60, deadline=None, suppress_health_check=list(HealthCheck))
@given(state_dicts())
def test_planner_layout_equals_oracle(sd):
    o_meta, o_total = oracle.layout(sd)
    p_meta, leaves, p_total = plan.build_layout(sd)
    assert o_total == p_total
    _same_meta(o_meta, p_meta)
    flat, _, _ = oracle.flatten(sd)
    assert [m.offset for _, m in flat] == [lf.meta.offset for lf in leaves]
    # a second walk pairs the same leaves and refreshes nothing structural
    again = plan.collect_leaves(sd, p_meta)
    assert again is not None and len(again) == len(leaves)


_HANDLER = []


@settings(max_examples=25, deadline=None, suppress_health_check=list(HealthCheck))
@given(state_dicts())
def test_handler_host_path_bytes_equal_oracle(sd):
    if not _HANDLER:
        _HANDLER.append(SharedMemoryHandler(36, host=True))
    h = _HANDLER[0]
    _, total, expect = oracle.pack(sd)
    full = dict(sd)
    full[DLROVER_CKPT_CONFIG_KEY] = CheckpointConfig(step=1, paths={})
    h.save_state_dict(full)
    if total == 0:
        return
    assert h.shared_memory.size == total
    assert bytes(h.shared_memory.buf[:total]) == expect.tobytes()
    out = h.load_state_dict()
    flat_in, _, _ = oracle.flatten(sd)
    flat_out, _, _ = oracle.flatten({k: v for k, v in out.items()
                                     if k != DLROVER_CKPT_CONFIG_KEY})
    for (a, _), (b, _) in zip(flat_in, flat_out):
        assert oracle.tensor_bytes(a).tobytes() == oracle.tensor_bytes(b).tobytes()


def teardown_module(module):
    for h in _HANDLER:
        h.unlink()
        h.close()


def test_missing_extension_fails_loudly(monkeypatch):
    monkeypatch.setattr(_native, "_lib", None)
    monkeypatch.setattr(_native, "LIB_PATH", "/nonexistent/libfcp.so")
    with pytest.raises(ImportError, match="no CPU fallback"):
        _native.lib()


def test_cuda_state_without_cuda_is_an_error_not_a_fallback():
    if torch.cuda.is_available():
        pytest.skip("box has CUDA")
    from dlrover_wuqiong_b200.transfer import DeviceShardEngine
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        DeviceShardEngine("cuda:0")
