"""Trainers with the reference's surface, driving the fused B200 step.

Mirrors collie/model/base/trainer.py: ``CollieMinimalTrainer`` :114-547 (constructor, ``fit``,
resumable epochs, early stopping, LR schedulers, duck-typed logger, printed strings) and
``CollieTrainer`` :22-111 (Lightning kwargs accepted; Lightning itself is not a dependency, and the
step semantics are the minimal trainer's: one forward/backward per batch, then every optimizer).

One iteration of the reference loop (``zero_grad`` -> ``calculate_loss`` -> ``backward`` ->
``optimizer_step`` per optimizer, trainer.py:458-470) is ONE call of ``model.fused_step``; batches
never leave the device and the loss is accumulated on the device, read back once per epoch (and at
logging intervals).
"""
import sys
import warnings

import torch

from . import _abi
from .model import RowOptimizer


class MultiOptimizer(object):
    """layers.py:20-38."""

    def __init__(self, optimizers):
        assert isinstance(optimizers, list), f'Expected list, got {type(optimizers)}!'
        self.optimizers = optimizers

    def zero_grad(self):
        for optimizer in self.optimizers:
            optimizer.zero_grad()


class MultiLRScheduler(object):
    """layers.py:41-63."""

    def __init__(self, lr_schedulers):
        assert isinstance(lr_schedulers, list), f'Expected list, got {type(lr_schedulers)}!'
        self.lr_schedulers = [s['scheduler'] for s in lr_schedulers]

    def step(self, *args, **kwargs):
        for s in self.lr_schedulers:
            s.step(*args, **kwargs)


class CollieMinimalTrainer(object):
    def __init__(self, model, max_epochs=10, gpus=None, logger=None, early_stopping_patience=3,
                 log_every_n_steps=50, flush_logs_every_n_steps=100, enable_model_summary=True,
                 weights_summary=None, detect_anomaly=False, terminate_on_nan=None, benchmark=True,
                 deterministic=True, progress_bar_refresh_rate=None, verbosity=True):
        if gpus is None and torch.cuda.is_available():
            print('Detected GPU. Setting ``gpus`` to 1.')
            gpus = 1
        if logger is False:
            logger = None
        if early_stopping_patience is False:
            early_stopping_patience = None
        if verbosity is True:
            verbosity = 2
        elif verbosity is False:
            verbosity = 0
        if weights_summary is not None:
            warnings.warn('``weights_summary`` is deprecated and is replaced with '
                          '``enable_model_summary``.', DeprecationWarning)
        if terminate_on_nan is not None:
            warnings.warn('``terminate_on_nan`` is deprecated and is replaced with '
                          '``detect_anomaly``.', DeprecationWarning)
            if detect_anomaly is False:
                detect_anomaly = terminate_on_nan

        self.max_epochs = max_epochs
        self.gpus = gpus
        self.benchmark = benchmark
        self.deterministic = deterministic
        self.logger = logger
        self.early_stopping_patience = early_stopping_patience
        self.log_every_n_steps = log_every_n_steps
        self.flush_logs_every_n_steps = flush_logs_every_n_steps
        self.enable_model_summary = enable_model_summary
        self.weights_summary = weights_summary
        self.detect_anomaly = detect_anomaly
        self.terminate_on_nan = terminate_on_nan
        self.progress_bar_refresh_rate = progress_bar_refresh_rate
        self.verbosity = verbosity

        self.best_epoch_loss = (0, sys.maxsize)
        self.train_steps = 0
        self.val_steps = 0
        self.num_epochs_completed = 0
        self.device = 'cpu' if (self.gpus is None or self.gpus is False or self.gpus == 0) \
            else 'cuda'

    @property
    def max_epochs(self):
        return self._max_epochs

    @max_epochs.setter
    def max_epochs(self, value):
        self._max_epochs = value

    # ------------------------------------------------------------------------------------ fit
    def fit(self, model):
        """Runs the full optimization routine; resumable like the reference (trainer.py:314-334)."""
        if self.device != 'cuda':
            raise _abi.CmfError('collie_b200 trains on a CUDA device only (no CPU fallback); '
                                f'``gpus={self.gpus}`` leaves nothing to run on')
        if not getattr(self, 'first_run_pre_training_setup_complete_', False):
            self._pre_training_setup(model)
            self.first_run_pre_training_setup_complete_ = True
        self._initialize_optimizers_and_lr_schedulers(model)
        self._fit(model)

    def _fit(self, model):
        for epoch in range(self.num_epochs_completed + 1, self.max_epochs + 1):
            model.train()
            train_loss = self._train_loop_single_epoch(model, epoch)
            model.eval()
            epoch_summary = f'Epoch {epoch: >5}: train loss: {train_loss :<1.5f}'
            early_stop_loss = train_loss
            if self.logger is not None:
                self.logger.log_metrics(metrics={'train_loss_epoch': train_loss}, step=epoch)
            if self.val_dataloader is not None:
                val_loss = self._val_loop_single_epoch(model, epoch)
                epoch_summary += f', val loss: {val_loss :<1.5f}'
                early_stop_loss = val_loss
                if self.logger is not None:
                    self.logger.log_metrics(metrics={'val_loss_epoch': val_loss}, step=epoch)
            if self.logger is not None:
                self.logger.save()
            if self.verbosity >= 1:
                print(epoch_summary)
            model.hparams.num_epochs_completed += 1
            self.num_epochs_completed += 1
            if (self.early_stopping_patience is not None
                    and early_stop_loss >= self.best_epoch_loss[1]
                    and epoch >= (self.early_stopping_patience + self.best_epoch_loss[0])):
                print(f'Epoch {epoch :>5}: Early stopping activated.')
                self._finalize_training()
                return
            if early_stop_loss < self.best_epoch_loss[1]:
                self.best_epoch_loss = (epoch, early_stop_loss)
            if self.lr_scheduler is not None:
                try:
                    self.lr_scheduler.step()
                except TypeError:
                    self.lr_scheduler.step(early_stop_loss)
        self._finalize_training()

    def _pre_training_setup(self, model):
        self.train_dataloader = model.train_dataloader()
        self.val_dataloader = model.val_dataloader()
        if self.verbosity != 0 and (self.weights_summary is not None
                                    or self.enable_model_summary is True):
            print(_model_summary(model))
        if self.logger is not None:
            self.logger.log_hyperparams(model.hparams)
            self.logger.save()
        model.to(self.device)
        model._move_any_external_data_to_device()

    def _initialize_optimizers_and_lr_schedulers(self, model):
        self.lr_scheduler = None
        out = model.configure_optimizers()
        if isinstance(out, tuple):
            optimizers, lr_schedulers = out
            self.optimizer = MultiOptimizer(optimizers)
            self.lr_scheduler = MultiLRScheduler(lr_schedulers)
        elif isinstance(out, list):
            self.optimizer = MultiOptimizer(out)
        elif isinstance(out, torch.optim.Optimizer):
            self.optimizer = MultiOptimizer([out])
        else:
            raise ValueError('Unexpected output from ``model.configure_optimizers()``!')
        for o in self.optimizer.optimizers:
            if not isinstance(o, RowOptimizer):
                raise _abi.CmfError('collie_b200 trainers need RowOptimizer objects')

    def _batches(self, loader, model, epoch):
        if hasattr(loader, 'device_batches'):
            for db in loader.device_batches(model.device, epoch=epoch):
                yield db, loader.interactions
        else:  # any iterable of reference-format batches
            for batch in loader:
                yield batch, None

    def _train_loop_single_epoch(self, model, epoch):
        loader = self.train_dataloader
        if self.logger is None and hasattr(loader, 'epoch_plan'):
            # no per-step logging to do: the whole epoch is one library call (cmf_train_epoch)
            plan = loader.epoch_plan(model.device, epoch=epoch)
            if plan.users.shape[0] == 0:
                return 0.0
            losses = model.fused_step(plan, optimizers=self.optimizer.optimizers, train=True,
                                      interactions=loader.interactions, epoch=epoch,
                                      epoch_batch_size=loader.batch_size)
            self.train_steps += int(losses.shape[0])
            mean = losses.mean()
            if self.detect_anomaly and not bool(torch.isfinite(mean)):
                raise RuntimeError('non-finite training loss detected')
            return mean.item()
        total_loss = torch.zeros((), dtype=torch.float32, device=model.device)
        n = 0
        for batch_idx, (batch, data) in enumerate(self._batches(self.train_dataloader, model,
                                                                epoch)):
            loss = model.fused_step(batch, optimizers=self.optimizer.optimizers, train=True,
                                    interactions=data, epoch=epoch)
            self.train_steps += 1
            total_loss += loss
            n += 1
            self._log_step('train', self.train_steps, total_loss, batch_idx)
        if self.detect_anomaly and not bool(torch.isfinite(total_loss)):
            raise RuntimeError('non-finite training loss detected')
        return (total_loss / max(n, 1)).item()

    def _val_loop_single_epoch(self, model, epoch=0):
        loader = self.val_dataloader
        if self.logger is None and hasattr(loader, 'epoch_plan'):
            plan = loader.epoch_plan(model.device, epoch=epoch)
            if plan.users.shape[0] == 0:
                return 0.0
            losses = model.fused_step(plan, train=False, interactions=loader.interactions,
                                      epoch=epoch, epoch_batch_size=loader.batch_size)
            self.val_steps += int(losses.shape[0])
            return losses.mean().item()
        total_loss = torch.zeros((), dtype=torch.float32, device=model.device)
        n = 0
        for batch_idx, (batch, data) in enumerate(self._batches(self.val_dataloader, model,
                                                                epoch)):
            loss = model.fused_step(batch, train=False, interactions=data, epoch=epoch)
            self.val_steps += 1
            total_loss += loss
            n += 1
            self._log_step('val', self.val_steps, total_loss, batch_idx)
        return (total_loss / max(n, 1)).item()

    def _log_step(self, name, steps, total_loss, batch_idx):
        if self.logger is not None:
            if steps % self.log_every_n_steps == 0:
                batch_loss = (total_loss / (batch_idx + 1)).item()
                self.logger.log_metrics(metrics={f'{name}_loss_step': batch_loss}, step=steps)
            if steps % self.flush_logs_every_n_steps == 0:
                self.logger.save()

    def _finalize_training(self):
        if self.logger is not None:
            self.logger.save()
            self.logger.finalize(status='FINISHED')


class CollieTrainer(CollieMinimalTrainer):
    """``CollieTrainer(model, max_epochs=10, benchmark=True, deterministic=True, **kwargs)``
    (trainer.py:60-111). Lightning keyword arguments are accepted; those that only configure
    Lightning machinery (checkpointing, callbacks, sanity steps, ...) have no effect here."""

    _MINIMAL_KWARGS = ('gpus', 'logger', 'early_stopping_patience', 'log_every_n_steps',
                       'flush_logs_every_n_steps', 'enable_model_summary', 'weights_summary',
                       'detect_anomaly', 'terminate_on_nan', 'progress_bar_refresh_rate',
                       'verbosity')

    def __init__(self, model, max_epochs=10, benchmark=True, deterministic=True, **kwargs):
        minimal = {k: v for k, v in kwargs.items() if k in self._MINIMAL_KWARGS}
        self.lightning_kwargs = {k: v for k, v in kwargs.items() if k not in self._MINIMAL_KWARGS}
        if 'devices' in self.lightning_kwargs and 'gpus' not in minimal:
            minimal['gpus'] = self.lightning_kwargs['devices']
        # Lightning trains without early stopping unless a callback asks for it
        minimal.setdefault('early_stopping_patience', None)
        super().__init__(model, max_epochs=max_epochs, benchmark=benchmark,
                         deterministic=deterministic, **minimal)
        self.model = model

    def fit(self, model=None):
        super().fit(self.model if model is None else model)


def _model_summary(model):
    rows = ['  | Name            | Type            | Params', '-' * 47]
    for i, (name, module) in enumerate(model.named_children()):
        n = sum(p.numel() for p in module.parameters())
        rows.append(f'{i} | {name:<15} | {type(module).__name__:<15} | {n}')
    rows.append(f'{sum(p.numel() for p in model.parameters())} total params')
    return '\n'.join(rows)
