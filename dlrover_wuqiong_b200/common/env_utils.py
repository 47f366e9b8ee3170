"""Process / node coordinates from the environment that torchelastic and
dlrover-run export.  Mirrors dlrover/python/common/env_utils.py:19-74 (same
names, same defaults — including GROUP_RANK defaulting to 1)."""
import os

from .constants import NodeEnv


def _int_env(key, default):
    return int(os.getenv(key, default))


def get_node_rank():
    if NodeEnv.NODE_RANK in os.environ:
        return int(os.environ[NodeEnv.NODE_RANK])
    return _int_env(NodeEnv.WORKER_RANK, "0")


def get_local_world_size():
    return _int_env("LOCAL_WORLD_SIZE", 1)


def get_local_rank():
    return _int_env("LOCAL_RANK", 0)


def get_rank():
    return _int_env("RANK", 0)


def get_group_world_size():
    return _int_env("GROUP_WORLD_SIZE", 1)


def get_group_rank():
    return _int_env("GROUP_RANK", 1)


def get_torch_restart_count():
    return _int_env("TORCHELASTIC_RESTART_COUNT", 0)


def get_node_id():
    return _int_env(NodeEnv.NODE_ID, 0)


def get_node_type():
    return os.getenv(NodeEnv.NODE_TYPE, "worker")


def get_node_num():
    return _int_env(NodeEnv.NODE_NUM, 0)


def get_env(env_key):
    return os.getenv(env_key, None)
