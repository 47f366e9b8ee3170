"""Where the agent's saver persists checkpoints.  Same interface as the
reference's dlrover/python/common/storage.py:24-328 (CheckpointStorage,
PosixDiskStorage, deletion strategies, get_checkpoint_storage); kept in Python
because it is off the blocking path."""
import os
import shutil
from abc import ABCMeta, abstractmethod

from .constants import CheckpointConstant
from .log import default_logger as logger
from .serialize import ClassMeta


class CheckpointStorage(metaclass=ABCMeta):
    """Abstract storage; subclass it for object stores etc."""

    @abstractmethod
    def write(self, content, path):
        """Write str / bytes / memoryview `content` to `path`."""

    @abstractmethod
    def write_state_dict(self, state_dict, path, write_func):
        """Persist a state dict with `write_func(state_dict, path)`."""

    @abstractmethod
    def read(self, path):
        """Return the text stored at `path` ("" if absent)."""

    @abstractmethod
    def read_state_dict(self, path, read_func):
        """Return `read_func(path)` ({} if absent)."""

    @abstractmethod
    def safe_rmtree(self, dir):
        pass

    @abstractmethod
    def safe_remove(self, path):
        pass

    @abstractmethod
    def safe_makedirs(self, dir):
        pass

    @abstractmethod
    def safe_move(self, src_path, dst_path):
        pass

    @abstractmethod
    def commit(self, step, success):
        """Called once per step after all shards were (or failed to be)
        persisted."""

    @abstractmethod
    def exists(self, path):
        pass

    @abstractmethod
    def listdir(self, path):
        pass

    @abstractmethod
    def get_class_meta(self):
        """ClassMeta from which the agent can rebuild this storage."""


class PosixDiskStorage(CheckpointStorage):
    def write(self, content, path):
        binary = isinstance(content, (bytes, bytearray, memoryview))
        with open(path, "wb" if binary else "w") as f:
            f.write(content)
            f.flush()
            os.fsync(f.fileno())

    def write_state_dict(self, state_dict, path, write_func=None):
        os.makedirs(os.path.dirname(path), exist_ok=True)
        if write_func:
            write_func(state_dict, path)

    def read(self, path, mode="r"):
        if not os.path.exists(path):
            return ""
        with open(path, mode) as f:
            return f.read()

    def read_state_dict(self, path, read_func):
        if not read_func or not os.path.exists(path):
            return {}
        return read_func(path)

    def safe_rmtree(self, dir):
        if os.path.exists(dir):
            shutil.rmtree(dir, ignore_errors=True)

    def safe_remove(self, path):
        if os.path.exists(path):
            os.remove(path)

    def safe_makedirs(self, dir):
        os.makedirs(dir, exist_ok=True)

    def safe_move(self, src_path, dst_path):
        if os.path.exists(src_path) and not os.path.exists(dst_path):
            shutil.move(src_path, dst_path)

    def commit(self, step, success):
        logger.info("checkpoint step %s persisted: success=%s", step, success)

    def exists(self, path):
        return os.path.exists(path)

    def listdir(self, path):
        return os.listdir(path)

    def get_class_meta(self):
        return ClassMeta(module_path=type(self).__module__,
                         class_name=type(self).__name__)


class CheckpointDeletionStrategy(metaclass=ABCMeta):
    @abstractmethod
    def clean_up(self, step, delete_func):
        """Decide whether the checkpoint folder of `step` can go and, if so,
        call `delete_func(folder)`."""


def _try_delete(folder, delete_func):
    try:
        logger.info("Clean path %s", folder)
        delete_func(folder)
    except Exception:
        logger.warning("Fail to clean path %s!", folder)


class KeepStepIntervalStrategy(CheckpointDeletionStrategy):
    """Keep only steps that are multiples of `keep_interval`."""

    def __init__(self, keep_interval, checkpoint_dir):
        self._keep_interval = keep_interval
        self._checkpoint_dir = checkpoint_dir

    def clean_up(self, step, delete_func):
        if step % self._keep_interval != 0:
            _try_delete(os.path.join(self._checkpoint_dir, str(step)), delete_func)


class KeepLatestStepStrategy(CheckpointDeletionStrategy):
    """Keep a sliding window of the newest steps."""

    def __init__(self, max_to_keep, checkpoint_dir):
        self._max_to_keep = max(max_to_keep, 1)
        self._checkpoint_dir = checkpoint_dir
        self._steps = []

    def clean_up(self, step, delete_func):
        self._steps.append(step)
        if len(self._steps) == self._max_to_keep:
            oldest = self._steps.pop(0)
            _try_delete(os.path.join(self._checkpoint_dir, str(oldest)), delete_func)


class PosixStorageWithDeletion(PosixDiskStorage):
    """Disk storage that prunes old step folders: it remembers the step the
    tracker file held before being overwritten and hands it to the deletion
    strategy when the new step commits."""

    def __init__(self, tracker_file, deletion_strategy):
        super().__init__()
        self._tracker_file = tracker_file
        self._deletion_strategy = deletion_strategy
        self._pre_step = 0

    def write(self, content, path):
        path = str(path)
        if path.endswith(self._tracker_file):
            previous = self.read(path)
            if previous:
                self._pre_step = int(previous)
        super().write(content, path)

    def commit(self, step, success):
        super().commit(step, success)
        if success and self._pre_step != step:
            self._deletion_strategy.clean_up(self._pre_step, shutil.rmtree)

    def get_class_meta(self):
        return ClassMeta(
            module_path=type(self).__module__, class_name=type(self).__name__,
            kwargs={"tracker_file": self._tracker_file,
                    "deletion_strategy": self._deletion_strategy})


def get_checkpoint_storage(deletion_strategy=None):
    if deletion_strategy:
        return PosixStorageWithDeletion(
            tracker_file=CheckpointConstant.TRACER_FILE_NAME,
            deletion_strategy=deletion_strategy)
    return PosixDiskStorage()
