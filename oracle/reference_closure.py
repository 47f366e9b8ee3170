"""TEST / BENCH INFRASTRUCTURE, NOT PRODUCT CODE.

The reference is Python, so its own implementation of the hot path can be run
as the timed CPU baseline — but /root/reference exists only in the build
container.  ``install()`` (called from ``__graft_entry__.build()``) copies the
import closure of the reference's flash-checkpoint path (SURVEY Appendix C, 26
files, unmodified) into the git-ignored ``baseline/_ref/``, which travels to the
GPU box with the snapshot; ``import_reference()`` imports it from there (or
straight from /root/reference when that exists) with the one stub module the
closure needs: ``dlrover.python.elastic_agent.master_client`` pulls in
generated protobuf code that is not in the tree.

Only bench.py (cpu_baseline / --impl reference legs), tests/ and
oracle/gen_golden.py use this; nothing under dlrover_wuqiong_b200/ does.
"""
import os
import shutil
import sys
import types

REF_ROOT = "/root/reference"
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEST = os.path.join(REPO, "baseline", "_ref")

_FC = "dlrover/trainer/torch/flash_checkpoint/"
CLOSURE = [
    "dlrover/__init__.py",
    "dlrover/python/__init__.py",
    "dlrover/python/common/__init__.py",
    "dlrover/python/common/constants.py",
    "dlrover/python/common/env_utils.py",
    "dlrover/python/common/log.py",
    "dlrover/python/common/multi_process.py",
    "dlrover/python/common/serialize.py",
    "dlrover/python/common/singleton.py",
    "dlrover/python/common/storage.py",
    "dlrover/python/elastic_agent/__init__.py",
    "dlrover/python/elastic_agent/torch/__init__.py",
    "dlrover/python/elastic_agent/torch/ckpt_saver.py",
    "dlrover/trainer/__init__.py",
    "dlrover/trainer/torch/__init__.py",
    _FC + "__init__.py", _FC + "checkpointer.py", _FC + "ddp.py", _FC + "engine.py",
    _FC + "fsdp.py", _FC + "fsdp_engine.py", _FC + "full_ckpt_engine.py",
    _FC + "megatron.py", _FC + "megatron_dist_ckpt.py", _FC + "megatron_engine.py",
    _FC + "replica.py",
]


def install(dest=DEST, src=REF_ROOT):
    """Copy the closure, byte for byte.  No-op (returns False) where the
    reference tree is absent — the GPU box uses what the container shipped."""
    if not os.path.isdir(os.path.join(src, "dlrover")):
        return False
    for rel in CLOSURE:
        out = os.path.join(dest, rel)
        os.makedirs(os.path.dirname(out), exist_ok=True)
        shutil.copyfile(os.path.join(src, rel), out)
    return True


def available_root():
    for root in (REF_ROOT, DEST):
        if os.path.isfile(os.path.join(root, CLOSURE[12])):
            return root
    return None


def import_reference(run_id=None):
    """-> the reference's ``dlrover.python.elastic_agent.torch.ckpt_saver``
    module, or raises ImportError with the reason."""
    root = available_root()
    if root is None:
        raise ImportError("the reference closure is neither at %s nor at %s (run "
                          "__graft_entry__.build() in the build container)" % (REF_ROOT, DEST))
    os.environ.setdefault("ROLE_NAME", "dlrover-trainer")
    if run_id:
        os.environ["TORCHELASTIC_RUN_ID"] = run_id
    sys.dont_write_bytecode = True
    name = "dlrover.python.elastic_agent.master_client"
    if name not in sys.modules:
        stub = types.ModuleType(name)
        stub.MasterClient = type("MasterClient", (), {})
        sys.modules[name] = stub
    if root not in sys.path:
        sys.path.insert(0, root)
    from dlrover.python.elastic_agent.torch import ckpt_saver  # noqa: E402
    if not os.path.abspath(ckpt_saver.__file__).startswith(os.path.abspath(root)):
        raise ImportError("`dlrover` resolved to %s, not the reference" % ckpt_saver.__file__)
    return ckpt_saver
