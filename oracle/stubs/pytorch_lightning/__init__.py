"""Minimal stand-in for ``pytorch_lightning`` (pinned 1.9.4 by the reference, absent here).

TEST INFRASTRUCTURE ONLY. It exists so that the unmodified reference package under
``/root/reference`` can be imported in the build container to pin the oracle and to generate the
golden fixtures in ``tests/golden``. It contributes no arithmetic; nothing in ``collie_b200``
imports it.
"""
import inspect

import torch


class _HParams(dict):
    """Attribute + item access, like Lightning's ``AttributeDict``."""
    def __getattr__(self, key):
        try:
            return self[key]
        except KeyError as e:
            raise AttributeError(key) from e

    def __setattr__(self, key, value):
        self[key] = value


class LightningModule(torch.nn.Module):
    def __init__(self, *args, **kwargs):
        super().__init__()
        self._hparams = _HParams()

    @property
    def hparams(self):
        return self._hparams

    def save_hyperparameters(self, *args, **kwargs):
        for a in args:
            if isinstance(a, dict):
                self._hparams.update(a)

    @property
    def device(self):
        try:
            return next(self.parameters()).device
        except StopIteration:
            return torch.device('cpu')

    def log(self, *args, **kwargs):
        return None

    def freeze(self):
        for p in self.parameters():
            p.requires_grad = False
        self.eval()

    def optimizer_step(self, epoch=None, batch_idx=None, optimizer=None, optimizer_idx=None,
                       optimizer_closure=None, **kwargs):
        if optimizer_closure is not None:
            optimizer.step(closure=optimizer_closure)
        else:
            optimizer.step()


class Trainer(object):
    """Base-class symbol only; the oracle drives ``CollieMinimalTrainer``."""
    def __init__(self, *args, **kwargs):
        """stub"""
        raise RuntimeError('pytorch_lightning.Trainer is a stub; use CollieMinimalTrainer')


from . import utilities  # noqa: E402,F401
