"""Drop-in import paths: after ``install()`` the reference's module names

    from dlrover.trainer.torch.flash_checkpoint.ddp import DdpCheckpointer
    from dlrover.python.elastic_agent.torch.ckpt_saver import AsyncCheckpointSaver

resolve to this package, so training scripts and agents written against the
reference (SURVEY §8b "public Python surface to keep") run unchanged.  Nothing
is aliased when a real ``dlrover`` is importable unless ``force=True``.
"""
import importlib
import importlib.util
import sys
import types

_PKG = "dlrover_wuqiong_b200"

ALIASES = {
    "dlrover.python.common.constants": _PKG + ".common.constants",
    "dlrover.python.common.env_utils": _PKG + ".common.env_utils",
    "dlrover.python.common.grpc": _PKG + ".common.grpc",
    "dlrover.python.common.log": _PKG + ".common.log",
    "dlrover.python.common.multi_process": _PKG + ".common.multi_process",
    "dlrover.python.common.serialize": _PKG + ".common.serialize",
    "dlrover.python.common.singleton": _PKG + ".common.singleton",
    "dlrover.python.common.storage": _PKG + ".common.storage",
    "dlrover.python.elastic_agent.torch.ckpt_saver": _PKG + ".ckpt_saver",
    "dlrover.trainer.torch.flash_checkpoint.checkpointer": _PKG + ".flash_checkpoint.checkpointer",
    "dlrover.trainer.torch.flash_checkpoint.ddp": _PKG + ".flash_checkpoint.ddp",
    "dlrover.trainer.torch.flash_checkpoint.engine": _PKG + ".flash_checkpoint.engine",
    "dlrover.trainer.torch.flash_checkpoint.full_ckpt_engine":
        _PKG + ".flash_checkpoint.full_ckpt_engine",
    "dlrover.trainer.torch.flash_checkpoint.fsdp": _PKG + ".flash_checkpoint.fsdp",
    "dlrover.trainer.torch.flash_checkpoint.fsdp_engine": _PKG + ".flash_checkpoint.fsdp_engine",
    "dlrover.trainer.torch.flash_checkpoint.deepspeed": _PKG + ".flash_checkpoint.deepspeed",
    "dlrover.trainer.torch.flash_checkpoint.deepspeed_engine":
        _PKG + ".flash_checkpoint.deepspeed_engine",
    "dlrover.trainer.torch.flash_checkpoint.hf_trainer": _PKG + ".flash_checkpoint.hf_trainer",
    "dlrover.trainer.torch.flash_checkpoint.megatron": _PKG + ".flash_checkpoint.megatron",
    "dlrover.trainer.torch.flash_checkpoint.megatron_engine":
        _PKG + ".flash_checkpoint.megatron_engine",
    "dlrover.trainer.torch.flash_checkpoint.megatron_dist_ckpt":
        _PKG + ".flash_checkpoint.megatron_dist_ckpt",
    "dlrover.trainer.torch.flash_checkpoint.replica": _PKG + ".flash_checkpoint.replica",
}


def _real_dlrover_present():
    mod = sys.modules.get("dlrover")
    if mod is not None:
        return not getattr(mod, "__fcp_alias__", False)
    try:
        return importlib.util.find_spec("dlrover") is not None
    except (ImportError, ValueError):
        return False


def install(force=False):
    """Register the aliases; returns the list of names installed."""
    if _real_dlrover_present() and not force:
        return []
    done = []
    for alias, target in ALIASES.items():
        parts = alias.split(".")
        for i in range(1, len(parts)):
            name = ".".join(parts[:i])
            if name not in sys.modules:
                pkg = types.ModuleType(name)
                pkg.__path__ = []  # a namespace-like package
                pkg.__fcp_alias__ = True
                sys.modules[name] = pkg
                if i > 1:
                    setattr(sys.modules[".".join(parts[:i - 1])], parts[i - 1], pkg)
        module = importlib.import_module(target)
        sys.modules[alias] = module
        setattr(sys.modules[".".join(parts[:-1])], parts[-1], module)
        done.append(alias)
    return done


_SHIM = '''"""Stand-in for the reference's top-level package: importing it makes
``dlrover.python...`` / ``dlrover.trainer...`` resolve to dlrover_wuqiong_b200
(written by dlrover_wuqiong_b200.compat.write_shim)."""
import sys

__fcp_alias__ = True
sys.path.insert(0, %r)
from dlrover_wuqiong_b200 import compat as _compat  # noqa: E402

_compat.install(force=True)
'''


def write_shim(directory):
    """Write ``<directory>/dlrover/__init__.py``.  With `directory` on
    sys.path / PYTHONPATH every process of a job — spawned trainers, the agent
    that imports the saver class by module path — resolves the reference's
    import paths to this package without calling install() itself."""
    import os

    pkg = os.path.join(directory, "dlrover")
    os.makedirs(pkg, exist_ok=True)
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    with open(os.path.join(pkg, "__init__.py"), "w") as f:
        f.write(_SHIM % root)
    return pkg


if __name__ == "__main__":
    if len(sys.argv) == 3 and sys.argv[1] == "--write-shim":
        print(write_shim(sys.argv[2]))
    else:
        sys.exit("usage: python -m dlrover_wuqiong_b200.compat --write-shim DIR")
