"""torch.save / torch.load interception used by the DeepSpeed and Megatron
adapters: the frameworks build their state dicts and "write" them with
torch.save(obj, path); while intercepted, those calls hand the objects to the
flash-checkpoint engine instead of the disk (reference:
flash_checkpoint/deepspeed.py:189-191,245-255, megatron.py:200-232)."""
import contextlib

import torch

from .. import raw_ckpt

torch_native_save = torch.save
torch_native_load = torch.load


def _read_file(path, map_location=None):
    """Read a persisted state: raw image (raw_ckpt.py) or torch.save file,
    with the un-intercepted torch.load."""
    if isinstance(path, str) and raw_ckpt.is_raw(path):
        return raw_ckpt.read_raw(path)
    return torch_native_load(path, map_location=map_location, weights_only=False)


@contextlib.contextmanager
def redirected(save=None, load=None):
    """Temporarily replace torch.save and/or torch.load."""
    prev_save, prev_load = torch.save, torch.load
    try:
        if save is not None:
            torch.save = save
        if load is not None:
            torch.load = load
        yield
    finally:
        torch.save, torch.load = prev_save, prev_load
