"""Public checkpointer contract (reference:
dlrover/trainer/torch/flash_checkpoint/checkpointer.py:18-65)."""
from abc import ABCMeta, abstractmethod
from enum import Enum, auto


class StorageType(Enum):
    MEMORY = auto()
    DISK = auto()


class Checkpointer(metaclass=ABCMeta):
    """Saves a training state first into node-local shared memory (fast,
    blocks training only for the device-side pack) and lets the agent persist
    it to storage asynchronously; loads from shared memory when it still
    holds a valid checkpoint, else from storage."""

    @abstractmethod
    def save_checkpoint(self, step, state_dict, path, storage_type=StorageType.DISK):
        """
        Args:
            step (int): global iteration step.
            state_dict (dict): model / optimizer / extra state to save.
            path (str): where the agent persists it (always needed: the agent
                also uses it when it saves the shm after a worker failure).
            storage_type: StorageType.MEMORY or StorageType.DISK.
        """

    @abstractmethod
    def load_checkpoint(self, resuming_path=None):
        """Return the latest state dict (from memory if possible)."""
