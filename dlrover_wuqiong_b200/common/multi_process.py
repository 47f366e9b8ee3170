"""Node-local IPC between the training processes and the agent that hosts the
checkpoint saver: a lock, a queue and a dict served over AF_UNIX sockets, and
a POSIX shared-memory segment that survives the death of its creator.

Same objects, names, socket paths and wire format as the reference's
dlrover/python/common/multi_process.py (SharedLock :225-307, SharedQueue
:346-438, SharedDict :453-534, SharedMemory :537-609, framing :89-109):
``/tmp/checkpoint_sock/<TORCHELASTIC_RUN_ID>/<class>_<name>.sock``, one
connection per request, 4-byte big-endian length + pickle.  Differences are
internal only: messages are received with one sized ``recv_into`` loop instead
of 1 KiB appends (the meta dict of a 10k-tensor shard is megabytes), and
servers dispatch through a method table.
"""
import ctypes
import mmap
import os
import pickle
import queue
import shutil
import socket
import threading
import time
from abc import ABCMeta, abstractmethod
from dataclasses import dataclass, field
from multiprocessing import shared_memory
from typing import Dict

import _posixshmem

from .constants import NodeEnv
from .log import default_logger as logger

SOCKET_TMP_DIR = "/tmp/checkpoint_sock/"

SUCCESS_CODE = "OK"
ERROR_CODE = "ERROR"
_APPLIED = "__applied__"


def retry_socket(func):
    """Retry while the server socket does not exist yet (the agent may still
    be starting): `retry` attempts one second apart, then one last try whose
    exception propagates."""

    def wrapper(self, *args, **kwargs):
        attempts = kwargs.get("retry", 30)
        for _ in range(attempts):
            try:
                return func(self, *args, **kwargs)
            except (FileNotFoundError, ConnectionRefusedError):
                time.sleep(1)
        return func(self, *args, **kwargs)

    return wrapper


def clear_sock_dir():
    shutil.rmtree(SOCKET_TMP_DIR, ignore_errors=True)


def _create_socket_server(path):
    os.makedirs(os.path.dirname(path), exist_ok=True)
    if os.path.exists(path):
        os.unlink(path)
    server = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    server.bind(path)
    server.listen(16)
    return server


def _create_socket_client(path):
    client = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    client.connect(path)
    return client


def _socket_send(sock, message):
    sock.sendall(len(message).to_bytes(4, "big") + message)


def _recv_exact(sock, n):
    buf = bytearray(n)
    view = memoryview(buf)
    got = 0
    while got < n:
        k = sock.recv_into(view[got:], n - got)
        if k == 0:
            raise ConnectionError("socket closed after %d of %d bytes" % (got, n))
        got += k
    return buf


def _socket_recv(sock):
    size = int.from_bytes(_recv_exact(sock, 4), "big")
    return bytes(_recv_exact(sock, size))


@dataclass
class SocketRequest(object):
    method: str = ""
    args: Dict[str, object] = field(default_factory=dict)


@dataclass
class SocketResponse(object):
    status: str = ""


@dataclass
class LockAcquireResponse(SocketResponse):
    acquired: bool = False


@dataclass
class LockedResponse(SocketResponse):
    locked: bool = False


@dataclass
class QueueGetResponse(SocketResponse):
    obj: object = None


@dataclass
class QueueSizeResponse(SocketResponse):
    size: int = 0


@dataclass
class QueueEmptyResponse(SocketResponse):
    empty: bool = False


@dataclass
class DictMessage(SocketResponse):
    meta_dict: object = None


class LocalSocketComm(metaclass=ABCMeta):
    """Base of the shared objects.  With ``create=True`` the instance owns the
    real object and serves it; otherwise every method is one request to the
    server of the same class and name."""

    def __init__(self, name="", create=False):
        self._name = name
        self._socket_file = self._create_socket_path()
        self._create = create
        self._server = None
        self._init_socket()

    def unlink(self):
        try:
            os.unlink(self._socket_file)
        except FileNotFoundError:
            pass

    def _create_socket_path(self):
        fname = "%s_%s.sock" % (type(self).__name__.lower(), self._name)
        run_id = os.getenv(NodeEnv.TORCHELASTIC_RUN_ID, "")
        root = os.path.join(SOCKET_TMP_DIR, run_id) if run_id else SOCKET_TMP_DIR
        os.makedirs(root, exist_ok=True)
        return os.path.join(root, fname)

    def _init_socket(self):
        # the socket listens from here on (clients queue in its backlog), but
        # requests are served only after the subclass constructor has set the
        # state they touch: see _serve()
        if self._create:
            self._server = _create_socket_server(self._socket_file)

    def _serve(self):
        """Last statement of every subclass constructor."""
        if self._server is not None:
            threading.Thread(target=self._sync, daemon=True,
                             name="ipc-" + os.path.basename(self._socket_file)).start()

    def _sync(self):
        """Accept loop of the server side."""
        while True:
            try:
                conn, _ = self._server.accept()
            except OSError:
                return
            try:
                try:
                    request = pickle.loads(_socket_recv(conn))
                    response = self._handle(request)
                    response.status = SUCCESS_CODE
                except Exception as e:  # noqa: BLE001 - reported to the client
                    logger.error("%s request failed: %r", type(self).__name__, e)
                    response = SocketResponse(status=ERROR_CODE)
                finally:
                    self._after_request()
                _socket_send(conn, pickle.dumps(response))
            except OSError:
                pass
            finally:
                conn.close()

    @abstractmethod
    def _handle(self, request):
        """Serve one request; return a SocketResponse subclass instance."""

    def _after_request(self):
        pass

    @retry_socket
    def _request(self, request, **_):
        client = _create_socket_client(self._socket_file)
        try:
            _socket_send(client, pickle.dumps(request))
            return pickle.loads(_socket_recv(client))
        finally:
            client.close()


class SharedLock(LocalSocketComm):
    """A lock shared by the processes of one node."""

    def __init__(self, name="", create=False):
        super().__init__(name, create)
        self._lock = threading.Lock() if self._create else None
        self._serve()

    def _handle(self, request):
        if request.method == "acquire":
            return LockAcquireResponse(acquired=self.acquire(**request.args))
        if request.method == "locked":
            return LockedResponse(locked=self.locked())
        if request.method == "release":
            self.release()
            return SocketResponse()
        raise ValueError("unknown method %s" % request.method)

    def acquire(self, blocking=True):
        if self._server:
            return self._lock.acquire(blocking=blocking)
        response = self._request(SocketRequest("acquire", {"blocking": blocking}))
        return response.acquired if response.status == SUCCESS_CODE else False

    def release(self):
        if self._server:
            if self._lock.locked():
                self._lock.release()
        else:
            self._request(SocketRequest("release", {}))

    def locked(self):
        if self._server:
            return self._lock.locked()
        # the reference returns the raw response object here (truthy even when
        # unlocked); return the flag the caller means
        response = self._request(SocketRequest("locked", {}))
        return bool(getattr(response, "locked", False))


class SharedQueue(LocalSocketComm):
    """A FIFO shared by the processes of one node."""

    def __init__(self, name="", create=False, maxsize=1):
        super().__init__(name, create)
        self._queue = queue.Queue(maxsize) if self._create else None
        self._serve()

    def _handle(self, request):
        if request.method == "put":
            self.put(**request.args)
            return SocketResponse()
        if request.method == "get":
            return QueueGetResponse(obj=self.get(**request.args))
        if request.method == "qsize":
            return QueueSizeResponse(size=self.qsize())
        if request.method == "empty":
            return QueueEmptyResponse(empty=self.empty())
        raise ValueError("unknown method %s" % request.method)

    def put(self, obj, block=True, timeout=None):
        if self._server:
            self._queue.put(obj, block=block, timeout=timeout)
        else:
            self._request(SocketRequest(
                "put", {"obj": obj, "block": block, "timeout": timeout}))

    def get(self, block=True, timeout=None):
        if self._server:
            return self._queue.get(block=block, timeout=timeout)
        response = self._request(SocketRequest("get", {"block": block, "timeout": timeout}))
        return response.obj if response.status == SUCCESS_CODE else None

    def qsize(self):
        if self._server:
            return self._queue.qsize()
        response = self._request(SocketRequest("qsize", {}))
        return response.size if response.status == SUCCESS_CODE else -1

    def empty(self):
        if self._server:
            return self._queue.empty()
        response = self._request(SocketRequest("empty", {}))
        return response.empty if response.status == SUCCESS_CODE else False


class SharedDict(LocalSocketComm):
    """A dict owned by the server (the agent) that clients replace wholesale.

    A writer first drops a token into the side queue ``shard_dict_<name>`` and
    then sends the dict; the server removes the token when the request has
    been handled, so a server-side ``get()`` issued in between waits for the
    dict that is already on its way."""

    def __init__(self, name="", create=False):
        super().__init__(name, create)
        self._dict = {}
        self._shared_queue = SharedQueue(name="shard_dict_" + name, create=self._create)
        self._serve()

    def _handle(self, request):
        if request.method == "set":
            self.set(**request.args)
            return DictMessage()
        if request.method == "get":
            return DictMessage(meta_dict=self.get(**request.args))
        if request.method == "update":
            return DictMessage(meta_dict=_APPLIED if self._apply(**request.args) else None)
        if request.method == "get_keys":
            return DictMessage(meta_dict=self.get_keys(**request.args))
        raise ValueError("unknown method %s" % request.method)

    def _apply(self, updates):
        """Assign `value` at every `path` (a tuple of keys / list indices) of
        the served dict; all-or-nothing."""
        if not self._dict:
            return False  # nothing to patch (e.g. a restarted server): needs a full set
        slots = []
        for path, value in updates:
            node = self._dict
            try:
                for key in path[:-1]:
                    try:
                        node = node[key]
                    except TypeError:  # a plain object on the way: by attribute
                        node = getattr(node, key)
                if isinstance(node, list):
                    node[path[-1]]  # must exist
                elif not isinstance(node, dict):
                    return False
                slots.append((node, path[-1], value))
            except (KeyError, IndexError, TypeError, AttributeError):
                return False
        for node, key, value in slots:
            node[key] = value
        return True

    def update(self, updates):
        """Send only what changed — `updates` = [(path, value), ...] — instead
        of the whole dict (the pickle of a 10 000-tensor meta tree costs ~60 ms
        per `set`).  Returns False if the server could not apply it (it has no
        dict of this shape yet, or does not know the method): send the whole
        dict with `set` then.  The local copy is the caller's to keep in
        step."""
        if self._server:
            return self._apply(updates)
        self._shared_queue.put(1)
        response = self._request(SocketRequest("update", {"updates": updates}))
        return (response.status == SUCCESS_CODE
                and getattr(response, "meta_dict", None) == _APPLIED)

    def _after_request(self):
        if not self._shared_queue.empty():
            self._shared_queue.get(1)

    def set(self, new_dict):
        self._dict = new_dict
        if not self._server:
            self._shared_queue.put(1)
            response = self._request(SocketRequest("set", {"new_dict": new_dict}))
            if response.status == ERROR_CODE:
                raise RuntimeError("Fail to set metadata!")

    def get(self, local=False):
        if local:
            return self._dict
        if self._server:
            while not self._shared_queue.empty():
                time.sleep(0.1)
            return self._dict
        response = self._request(SocketRequest("get", {}))
        if response.status == SUCCESS_CODE:
            self._dict = response.meta_dict
        return self._dict

    def get_keys(self, keys):
        """Only the given top-level keys (absent ones are left out).  The meta
        tree of a few thousand tensors costs tens of ms to pickle; a reader
        that wants the config alone need not pay for it.  A server that does
        not know the request is asked for the whole dict instead.  The local
        copy is not replaced."""
        if self._server:
            while not self._shared_queue.empty():
                time.sleep(0.1)
            return {k: self._dict[k] for k in keys if k in self._dict}
        response = self._request(SocketRequest("get_keys", {"keys": list(keys)}))
        if response.status == SUCCESS_CODE and isinstance(
                getattr(response, "meta_dict", None), dict):
            return response.meta_dict
        whole = self.get()
        return {k: whole[k] for k in keys if k in whole}

    def unlink(self):
        super().unlink()
        self._shared_queue.unlink()


_MADV_POPULATE_WRITE = 23  # Linux >= 5.14
_POPULATE_BLOCK = 128 << 20


class _Populator(object):
    """Makes a fresh shared mapping resident from a few threads.

    MAP_POPULATE fills the page tables of a new tmpfs mapping from one thread
    (~3 GB/s on the bench host, i.e. 6 s for a 19 GB shard before the first
    save can start); madvise(MADV_POPULATE_WRITE) on disjoint ranges from 4
    threads reaches ~9 GB/s (tools/populate_probe.py).  Blocks are handed out
    in address order, so a consumer that works front to back (pinning the
    segment piece by piece) can follow right behind via ``wait(upto)``."""

    def __init__(self, mm, size, threads):
        self._addr = ctypes.addressof(ctypes.c_char.from_buffer(mm))
        self._size = size
        self._blocks = (size + _POPULATE_BLOCK - 1) // _POPULATE_BLOCK
        self._next = 0
        self._done = [False] * self._blocks
        self._front = 0  # every block below this index is resident
        self._cv = threading.Condition()
        self._give_up = False
        self._libc = ctypes.CDLL(None, use_errno=True)
        self._libc.madvise.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
        self._threads = [threading.Thread(target=self._work, daemon=True, name="shm-populate-%d" % i)
                         for i in range(max(1, min(threads, self._blocks)))]
        for t in self._threads:
            t.start()

    def _work(self):
        while True:
            with self._cv:
                i = self._next
                if i >= self._blocks:
                    return
                self._next += 1
            lo = i * _POPULATE_BLOCK
            n = min(_POPULATE_BLOCK, self._size - lo)
            if not self._give_up and \
                    self._libc.madvise(self._addr + lo, n, _MADV_POPULATE_WRITE) != 0:
                # kernel without MADV_POPULATE_WRITE (< 5.14): leave the pages
                # to fault in on first use.  Nothing may be WRITTEN here: the
                # owner is free to fill the segment while these threads run.
                self._give_up = True
            with self._cv:
                self._done[i] = True
                while self._front < self._blocks and self._done[self._front]:
                    self._front += 1
                self._cv.notify_all()

    def wait(self, upto=None):
        need = self._blocks if upto is None else \
            min(self._blocks, (upto + _POPULATE_BLOCK - 1) // _POPULATE_BLOCK)
        with self._cv:
            while self._front < need:
                self._cv.wait()

    def join(self):
        for t in self._threads:
            t.join()


class SharedMemory(shared_memory.SharedMemory):
    """POSIX shm segment that is NOT registered with Python's resource
    tracker, so it outlives a crashed trainer and the restarted trainer (or the
    agent) can still read the last checkpoint from it.  Whoever is responsible
    must ``unlink()`` it explicitly.  Same semantics as the reference class
    (multi_process.py:537-609); on creation the mapping is made resident by
    background threads (_Populator) so the first save neither pays page faults
    in its copies nor waits for a single-threaded MAP_POPULATE."""

    _name = None
    _fd = -1
    _mmap = None
    _buf = None
    _flags = os.O_RDWR
    _mode = 0o600
    _prepend_leading_slash = True
    _populator = None

    def __init__(self, name=None, create=False, size=0):
        if not size >= 0:
            raise ValueError("'size' must be a positive integer")
        if create:
            self._flags = os.O_CREAT | os.O_EXCL | os.O_RDWR
            if size == 0:
                raise ValueError("'size' must be a positive number different from zero")
        if name is None and not self._flags & os.O_EXCL:
            raise ValueError("'name' can only be None if create=True")
        if name is None:
            while True:
                name = shared_memory._make_filename()
                try:
                    self._fd = _posixshmem.shm_open(name, self._flags, mode=self._mode)
                except FileExistsError:
                    continue
                self._name = name
                break
        else:
            name = "/" + name if self._prepend_leading_slash else name
            self._fd = _posixshmem.shm_open(name, self._flags, mode=self._mode)
            self._name = name
        try:
            if create and size:
                os.ftruncate(self._fd, size)
            size = os.fstat(self._fd).st_size
            self._mmap = mmap.mmap(self._fd, size, flags=mmap.MAP_SHARED)
        except OSError:
            self.unlink()
            raise
        self._size = size
        self._buf = memoryview(self._mmap)
        self._populator = None
        if create and size:
            threads = int(os.getenv("FCP_POPULATE_THREADS", "4"))
            if threads > 0:
                self._populator = _Populator(self._mmap, size, threads)

    def wait_populated(self, upto=None):
        """Block until the first `upto` bytes (default: all) of a segment this
        object created are resident.  Using the buffer earlier is allowed (it
        just faults pages the slow way)."""
        if self._populator is not None:
            self._populator.wait(upto)
            if upto is None or upto >= self._size:
                self._populator.join()
                self._populator = None

    def close(self):
        if self._populator is not None:
            self._populator.wait()
            self._populator.join()
            self._populator = None
        super().close()

    def unlink(self):
        if self._name:
            try:
                _posixshmem.shm_unlink(self._name)
            except FileNotFoundError:
                pass
            logger.info("Unlink the shared memory %s", self._name)
