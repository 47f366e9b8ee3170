"""Run the REFERENCE's own checkpoint tests against this package.

The test files are taken from a checkout of the reference (default
/root/reference; never part of this repo), copied to a scratch directory
outside the repo and run with ``dlrover.*`` resolving to this package
(compat.install).  usage: python tools/run_reference_tests.py [REFERENCE_ROOT]
[-- extra pytest args]"""
import os
import shutil
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TESTS = [
    "dlrover/python/tests/test_multi_process.py",
    "dlrover/python/tests/test_storage.py",
    "dlrover/python/tests/test_ckpt_saver.py",
    "dlrover/trainer/tests/torch/checkpoint_egine_test.py",
    "dlrover/trainer/tests/torch/ddp_checkpointer_test.py",
    "dlrover/trainer/tests/torch/checkpoint_backup_test.py",
    "dlrover/trainer/tests/torch/deepspeed_ckpt_test.py",
    "dlrover/trainer/tests/torch/megatron_ckpt_test.py",
    "dlrover/trainer/tests/torch/fsdp_ckpt_test.py",
]


def main():
    args = sys.argv[1:]
    extra = []
    if "--" in args:
        extra = args[args.index("--") + 1:]
        args = args[:args.index("--")]
    against_reference = "--against-reference" in args
    args = [a for a in args if a != "--against-reference"]
    ref = args[0] if args else "/root/reference"
    sys.path.insert(0, ROOT)
    os.environ.setdefault("ROLE_NAME", "dlrover-trainer")
    from dlrover_wuqiong_b200 import compat
    if against_reference:
        # control run: the same test files against the reference's OWN code on
        # this Python / torch (two modules stubbed), to tell
        # failures of this package from failures of the environment
        import types
        sys.dont_write_bytecode = True
        m = types.ModuleType("dlrover.python.elastic_agent.master_client")
        m.MasterClient = type("MasterClient", (), {})
        sys.modules[m.__name__] = m
        sys.path.insert(0, ref)
        import dlrover.python.common  # noqa: F401  (the real package)
        from dlrover_wuqiong_b200.common import grpc as _grpc
        sys.modules["dlrover.python.common.grpc"] = _grpc
        sys.modules["dlrover.python.common"].grpc = _grpc
    else:
        assert compat.install(force=True), "aliases not installed"
    import pytest
    with tempfile.TemporaryDirectory(prefix="fcp_reftests_") as scratch:
        files = []
        for rel in TESTS:
            src = os.path.join(ref, rel)
            if os.path.exists(src):
                dst = os.path.join(scratch, os.path.basename(rel))
                shutil.copy(src, dst)
                files.append(dst)
        # children started with mp.spawn import the test module afresh: give
        # them a `dlrover` package on sys.path that installs the aliases
        if not against_reference:
            compat.write_shim(scratch)
            sys.path.insert(0, scratch)
        rc = pytest.main(["-q", "-p", "no:cacheprovider", "--rootdir", scratch,
                          "-o", "python_files=*_test.py test_*.py", *extra, *files])
    sys.exit(int(rc))


if __name__ == "__main__":
    main()
