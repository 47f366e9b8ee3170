"""The two socket helpers of dlrover/python/common/grpc.py:45-75 that
checkpoint code and its tests use; the gRPC message classes of that module
belong to the master/agent control plane and are out of scope."""
import random
import socket
from contextlib import closing


def find_free_port(port=0):
    """A TCP port that is free right now (`port` itself if it is)."""
    with closing(socket.socket(socket.AF_INET, socket.SOCK_STREAM)) as s:
        s.bind(("", port))
        s.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        return s.getsockname()[1]


def find_free_port_in_range(start=0, end=65535, random_port=True):
    """A free port in [start, end); tried in random or ascending order."""
    candidates = list(range(start, end))
    if random_port:
        random.shuffle(candidates)
    for port in candidates:
        try:
            return find_free_port(port)
        except OSError:
            continue
    raise RuntimeError("no free port in [%d, %d)" % (start, end))
