import torch

from . import parsing  # noqa: F401


def move_data_to_device(batch, device):
    if isinstance(batch, torch.Tensor):
        return batch.to(device)
    if isinstance(batch, (list, tuple)):
        return type(batch)(move_data_to_device(b, device) for b in batch)
    if isinstance(batch, dict):
        return {k: move_data_to_device(v, device) for k, v in batch.items()}
    return batch


def rank_zero_only(fn):
    return fn
