"""Trainers with the reference's surface, driving the fused B200 step.

Mirrors collie/model/base/trainer.py: ``CollieMinimalTrainer`` :114-547 (constructor, ``fit``,
resumable epochs, early stopping, LR schedulers, duck-typed logger, printed strings) and
``CollieTrainer`` :22-111 (Lightning kwargs accepted; Lightning itself is not a dependency, and the
step semantics are the minimal trainer's: one forward/backward per batch, then every optimizer).

How an epoch runs is decided once per ``fit`` by picking an engine:

* ``_FusedEngine``    one GPU. One iteration of the reference loop (``zero_grad`` ->
  ``calculate_loss`` -> ``backward`` -> ``optimizer_step`` per optimizer, trainer.py:458-470) is ONE
  ``cmf_train_step``; without a logger the whole epoch is ONE library call (``cmf_train_epoch``)
  over the loader's device-side epoch plan. Losses stay on the device until the epoch ends.
* ``_ShardedEngine``  several GPUs, one process each (launched with ``torch.distributed.run``;
  the reference's ``Trainer(gpus=N)``, trainer.py:60-79): users partitioned over the ranks, item
  table row-sharded (``collie_b200.distributed``); the trained tables are gathered back into the
  caller's model when ``fit`` returns, so ``save_model`` / evaluation work as on one GPU.
* ``_AutogradEngine`` a callable loss or a foreign optimizer: the reference's own four-call loop
  over the stand-alone CUDA kernels (differentiable ``forward``, ``RowOptimizer.step``).
"""
import sys
import warnings

import numpy as np
import torch
from torch.optim.lr_scheduler import ReduceLROnPlateau

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


def _world():
    """(rank, world size) of the default process group, (0, 1) outside ``torch.distributed``."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


class _Patience(object):
    """Early stopping as the reference counts it (trainer.py:379-390): the best loss and the epoch
    it was seen at; stop once an epoch fails to improve on it ``patience`` epochs later."""

    def __init__(self, patience, best):
        self.patience, self.best = patience, best

    def should_stop(self, epoch, loss):
        return (self.patience is not None and loss >= self.best[1]
                and epoch >= self.patience + self.best[0])

    def note(self, epoch, loss):
        if loss < self.best[1]:
            self.best = (epoch, loss)


# ------------------------------------------------------------------------------------- engines
class _FusedEngine(object):
    """Epochs on one GPU through ``model.fused_step``."""

    def __init__(self, trainer, model):
        self.t, self.model = trainer, model
        self.optimizers = trainer.optimizer.optimizers

    def _plan_route(self, loader):
        # no per-step logging to do: the whole epoch is one library call (cmf_train_epoch)
        return self.t.logger is None and hasattr(loader, 'epoch_plan')

    def _batches(self, loader, epoch):
        if hasattr(loader, 'device_batches'):
            for db in loader.device_batches(self.model.device, epoch=epoch):
                yield db, loader.interactions
        else:  # any iterable of reference-format batches
            for batch in loader:
                yield batch, None

    def _pass(self, loader, epoch, train):
        """Mean of the per-batch losses of one pass (trainer.py:485, :504), as a float."""
        t, model = self.t, self.model
        name = 'train' if train else 'val'
        kw = dict(optimizers=self.optimizers) if train else {}
        if self._plan_route(loader):
            plan = loader.epoch_plan(model.device, epoch=epoch)
            if plan.users.shape[0] == 0:
                return 0.0
            losses = model.fused_step(plan, train=train, interactions=loader.interactions,
                                      epoch=epoch, epoch_batch_size=loader.batch_size, **kw)
            t._count(name, int(losses.shape[0]))
            mean = losses.mean()
        else:
            total = torch.zeros((), dtype=torch.float32, device=model.device)
            n = 0
            for batch_idx, (batch, data) in enumerate(self._batches(loader, epoch)):
                total += model.fused_step(batch, train=train, interactions=data, epoch=epoch, **kw)
                n += 1
                t._count(name, 1)
                t._log_step(name, total, batch_idx)
            mean = total / max(n, 1)
        if train and t.detect_anomaly and not bool(torch.isfinite(mean)):
            raise RuntimeError('non-finite training loss detected')
        return mean.item()

    def train_epoch(self, epoch):
        return self._pass(self.t.train_dataloader, epoch, True)

    def val_epoch(self, epoch):
        return self._pass(self.t.val_dataloader, epoch, False)

    def finish(self):
        pass


class _AutogradEngine(_FusedEngine):
    """The reference's own loop, call for call (trainer.py:458-470), for what cannot be fused: a
    user-supplied loss callable or optimizer. Forward and backward are the stand-alone CUDA
    kernels behind ``model.forward``; the optimizer is whatever ``configure_optimizers`` built."""

    def _pass(self, loader, epoch, train):
        t, model = self.t, self.model
        name = 'train' if train else 'val'
        model._unfused = True           # calculate_loss builds the autograd graph
        total = torch.zeros((), dtype=torch.float32, device=model.device)
        n = 0
        for batch_idx, batch in enumerate(self._reference_batches(loader, epoch)):
            if train:
                t.optimizer.zero_grad()
                loss = model.calculate_loss(batch)
                loss.backward()
                for i, optimizer in enumerate(self.optimizers):
                    model.optimizer_step(epoch=epoch, batch_idx=batch_idx, optimizer=optimizer,
                                         optimizer_idx=i)
            else:
                with torch.no_grad():
                    loss = model.calculate_loss(batch)
            total += loss.detach()
            n += 1
            t._count(name, 1)
            t._log_step(name, total, batch_idx)
        if train and t.detect_anomaly and not bool(torch.isfinite(total)):
            raise RuntimeError('non-finite training loss detected')
        return (total / max(n, 1)).item()

    def _reference_batches(self, loader, epoch):
        """Batches in the reference's formats, drawn exactly like the fused route draws them (the
        same epoch plan, negatives keyed by the trainer's epoch), so that the two routes see the
        same data."""
        if not hasattr(loader, 'device_batches'):
            yield from loader
            return
        data = loader.interactions
        for db in loader.device_batches(self.model.device, epoch=epoch):
            if db.ratings is not None:
                yield db.users, db.items, db.ratings
            else:
                yield (db.users, db.items), data.sample_negatives(db.users, db.gidx, epoch=epoch)

    def finish(self):
        self.model._unfused = False


class _ShardedEngine(object):
    """Epochs over ``world`` GPUs, one process each. Every rank calls ``fit`` with the same model
    and data; the engine cuts the data by user owner, builds the rank's local model (its users'
    rows, the padded item table), trains through ``ShardedMatrixFactorization`` -- one optimizer
    step per unique row per GLOBAL batch of ``batch_size`` interactions -- and copies the trained
    tables back into the caller's model in ``finish``."""

    def __init__(self, trainer, model, rank, world):
        import torch.distributed as dist
        from . import distributed as D
        self.t, self.full, self.rank, self.world, self.D, self.dist = (trainer, model, rank, world,
                                                                      D, dist)
        dev = torch.device('cuda', torch.cuda.current_device())
        self.dev = dev
        loader = trainer.train_dataloader
        if not hasattr(loader, 'interactions'):
            raise _abi.CmfError('multi-GPU training needs the package\'s own data loaders')
        self.local_train = self._local_loader(loader)
        self.local_val = self._local_loader(trainer.val_dataloader) \
            if trainer.val_dataloader is not None else None
        self.local = self._local_model(model)
        self.sharded = D.ShardedMatrixFactorization(self.local, model.hparams.num_items,
                                                    max_batch=self.local_train.batch_size + 1)
        self.optimizers = self.local.configure_optimizers()
        if isinstance(self.optimizers, tuple):
            self.optimizers, schedulers = self.optimizers
            trainer.lr_scheduler = MultiLRScheduler(schedulers)
        if not isinstance(self.optimizers, (list, tuple)):
            self.optimizers = [self.optimizers]
        trainer.optimizer = MultiOptimizer(list(self.optimizers))

    def _local_loader(self, loader):
        from .interactions import ExplicitInteractions, Interactions
        D, data = self.D, loader.interactions
        per_rank, padded = D.shard_rows(data.num_items, self.world)
        n_local = D.num_local_users(data.num_users, self.rank, self.world)
        explicit = isinstance(data, ExplicitInteractions)
        parts = D.partition_interactions(data.mat.row, data.mat.col, self.rank, self.world,
                                         extra=data.mat.data if explicit else None)
        if explicit:
            local = ExplicitInteractions(users=parts[0], items=parts[1], ratings=parts[3],
                                         num_users=n_local, num_items=padded,
                                         allow_missing_ids=True,
                                         remove_duplicate_user_item_pairs=False)
        else:
            local = Interactions(
                users=parts[0], items=parts[1], num_users=n_local, num_items=padded,
                num_negative_samples=data.num_negative_samples, allow_missing_ids=True,
                remove_duplicate_user_item_pairs=False, check_num_negative_samples_is_valid=False,
                max_number_of_samples_to_consider=data.max_number_of_samples_to_consider,
                seed=data.seed)
        local.global_positions = parts[2]
        # the global batch of ``batch_size`` interactions, split evenly
        per_gpu = max(1, -(-int(loader.batch_size) // self.world))
        return type(loader)(interactions=local, batch_size=per_gpu, shuffle=loader.shuffle)

    def _local_model(self, model):
        h = model.hparams
        per_rank, padded = self.D.shard_rows(h.num_items, self.world)
        meta = h.get('metadata_for_loss')
        if meta:    # labels of the padding rows are never looked at
            meta = {k: torch.cat([torch.as_tensor(v).reshape(-1),
                                  torch.zeros(padded - h.num_items,
                                              dtype=torch.as_tensor(v).dtype)])
                    for k, v in meta.items()}
        local = type(model)(
            train=self.local_train, val=self.local_val, embedding_dim=h.embedding_dim,
            dropout_p=h.dropout_p, sparse=h.sparse, lr=h.lr, bias_lr=h.bias_lr,
            lr_scheduler_func=model.lr_scheduler_func, weight_decay=h.weight_decay,
            optimizer=model.optimizer, bias_optimizer=model.bias_optimizer, loss=model.loss,
            metadata_for_loss=meta, metadata_for_loss_weights=h.get('metadata_for_loss_weights'),
            y_range=h.y_range).to(self.dev)
        mine = torch.arange(self.rank, h.num_users, self.world, device=self.dev)
        src = model.to(self.dev)
        with torch.no_grad():
            local.user_embeddings.weight.copy_(src.user_embeddings.weight[mine])
            local.user_biases.weight.copy_(src.user_biases.weight[mine])
            local.item_embeddings.weight.zero_()
            local.item_biases.weight.zero_()
            local.item_embeddings.weight[:h.num_items].copy_(src.item_embeddings.weight)
            local.item_biases.weight[:h.num_items].copy_(src.item_biases.weight)
        return local

    def _pass(self, loader, epoch, train):
        dist, t = self.dist, self.t
        name = 'train' if train else 'val'
        steps = self.D.common_steps_per_epoch(loader)
        total = torch.zeros((), dtype=torch.float32, device=self.dev)
        it = loader.device_batches(self.dev, epoch=epoch)
        for j in range(steps):
            db = next(it)
            n_here = torch.tensor([int(db.users.shape[0])], dtype=torch.int64, device=self.dev)
            dist.all_reduce(n_here)          # (the last batches of the ranks differ in length)
            if train:
                total += self.sharded.train_step(db, self.optimizers,
                                                 interactions=loader.interactions, epoch=epoch,
                                                 global_batch=int(n_here.item()),
                                                 reduce_loss=False)
            else:
                total += self.local.fused_step(db, train=False, interactions=loader.interactions,
                                               epoch=epoch, global_batch=int(n_here.item()),
                                               num_items_logical=self.full.hparams.num_items)
            t._count(name, 1)
        dist.all_reduce(total)               # the ranks' shares of every step's global mean
        mean = total / max(steps, 1)
        if train and t.detect_anomaly and not bool(torch.isfinite(mean)):
            raise RuntimeError('non-finite training loss detected')
        return mean.item()

    def train_epoch(self, epoch):
        return self._pass(self.local_train, epoch, True)

    def val_epoch(self, epoch):
        return self._pass(self.local_val, epoch, False)

    def finish(self):
        """Every rank's copy of the caller's model receives the trained tables (all-gather of the
        user rows, the owners' item rows are everywhere already)."""
        dist, full, local = self.dist, self.full, self.local
        h = full.hparams
        n_max = -(-h.num_users // self.world)
        with torch.no_grad():
            for name in ('user_embeddings', 'user_biases'):
                src = getattr(local, name).weight.data
                pad = torch.zeros((n_max,) + tuple(src.shape[1:]), dtype=src.dtype, device=self.dev)
                pad[:src.shape[0]] = src
                parts = [torch.empty_like(pad) for _ in range(self.world)]
                dist.all_gather(parts, pad)
                dst = getattr(full, name).weight.data
                for q in range(self.world):
                    n_q = self.D.num_local_users(h.num_users, q, self.world)
                    dst[q::self.world] = parts[q][:n_q]
            I, bi = self.sharded.gather_item_table()
            full.item_embeddings.weight.data.copy_(I)
            full.item_biases.weight.data.copy_(bi.view(-1, 1))
        torch.cuda.synchronize(self.dev)


# ------------------------------------------------------------------------------------ trainers
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
        """Runs the full optimization routine; resumable like the reference (trainer.py:314-334):
        epochs already completed are not repeated, optimizers and schedulers are built anew."""
        if self.device != 'cuda':
            raise _abi.CmfError('collie_b200 trains on a CUDA device only (no CPU fallback); '
                                f'``gpus={self.gpus}`` leaves nothing to run on')
        self._prepare_once(model)
        engine = self._engine = self._pick_engine(model)
        patience = _Patience(self.early_stopping_patience, self.best_epoch_loss)
        try:
            for epoch in range(self.num_epochs_completed + 1, self.max_epochs + 1):
                model.train()
                report = {'train_loss_epoch': self._train_loop_single_epoch(model, epoch)}
                model.eval()
                if self.val_dataloader is not None:
                    report['val_loss_epoch'] = self._val_loop_single_epoch(model, epoch)
                self._report(epoch, report)
                model.hparams.num_epochs_completed += 1
                self.num_epochs_completed += 1
                # validation loss when there is one, else training loss (trainer.py:352-372)
                watched = report.get('val_loss_epoch', report['train_loss_epoch'])
                if patience.should_stop(epoch, watched):
                    print(f'Epoch {epoch :>5}: Early stopping activated.')
                    break
                patience.note(epoch, watched)
                self._step_schedulers(watched)
        finally:
            self.best_epoch_loss = patience.best
            engine.finish()
            self._engine = None
        self._finalize_training()

    def _prepare_once(self, model):
        """First ``fit`` of this trainer only (trainer.py:404-425)."""
        if getattr(self, 'first_run_pre_training_setup_complete_', False):
            return
        self.first_run_pre_training_setup_complete_ = True
        self.train_dataloader, self.val_dataloader = model.train_dataloader(), \
            model.val_dataloader()
        wants_summary = self.weights_summary is not None or self.enable_model_summary is True
        if self.verbosity != 0 and wants_summary:
            print(_model_summary(model))
        if self.logger is not None:
            self.logger.log_hyperparams(model.hparams)
            self.logger.save()
        model.to(self.device)
        model._move_any_external_data_to_device()

    def _pick_engine(self, model):
        """Fresh optimizers and schedulers (trainer.py:427-443) and the engine that fits them."""
        self.lr_scheduler = None
        rank, world = _world()
        many = isinstance(self.gpus, int) and not isinstance(self.gpus, bool) and self.gpus > 1
        if many and world == 1:
            raise _abi.CmfError(
                f'``gpus={self.gpus}`` trains with one process per GPU: launch the script with '
                f'``python -m torch.distributed.run --nproc-per-node {self.gpus} ...`` '
                '(torch.distributed initialised with the nccl backend)')
        if world > 1 and (many or self.gpus in (-1, 'auto')):
            if many and self.gpus != world:
                raise _abi.CmfError(f'``gpus={self.gpus}`` but {world} processes were launched')
            return _ShardedEngine(self, model, rank, world)
        built = model.configure_optimizers()
        if isinstance(built, tuple):
            optimizers, schedulers = built
            self.lr_scheduler = MultiLRScheduler(schedulers)
        elif isinstance(built, torch.optim.Optimizer):
            optimizers = [built]
        elif isinstance(built, list):
            optimizers = built
        else:
            raise ValueError('Unexpected output from ``model.configure_optimizers()``!')
        self.optimizer = MultiOptimizer(list(optimizers))
        fused = (getattr(model, '_fusable', True) and model._loss_name is not None
                 and all(isinstance(o, RowOptimizer) for o in optimizers))
        return _FusedEngine(self, model) if fused else _AutogradEngine(self, model)

    def _report(self, epoch, report):
        line = f'Epoch {epoch: >5}: train loss: {report["train_loss_epoch"] :<1.5f}'
        if 'val_loss_epoch' in report:
            line += f', val loss: {report["val_loss_epoch"] :<1.5f}'
        if self.logger is not None:
            for key, value in report.items():
                self.logger.log_metrics(metrics={key: value}, step=epoch)
            self.logger.save()
        if self.verbosity >= 1:
            print(line)

    def _step_schedulers(self, watched):
        """Plateau schedulers follow the watched loss, the others the epoch count
        (trainer.py:393-399)."""
        if self.lr_scheduler is None:
            return
        for scheduler in self.lr_scheduler.lr_schedulers:
            if isinstance(scheduler, ReduceLROnPlateau):
                scheduler.step(watched)
            else:
                scheduler.step()

    def _count(self, name, n):
        if name == 'train':
            self.train_steps += n
        else:
            self.val_steps += n

    def _log_step(self, name, total_loss, batch_idx):
        if self.logger is None:
            return
        steps = self.train_steps if name == 'train' else self.val_steps
        if steps % self.log_every_n_steps == 0:
            batch_loss = (total_loss / (batch_idx + 1)).item()
            self.logger.log_metrics(metrics={f'{name}_loss_step': batch_loss}, step=steps)
        if steps % self.flush_logs_every_n_steps == 0:
            self.logger.save()

    # the two hooks of the reference's loop (trainer.py:445-504); ``fit`` goes through them
    def _train_loop_single_epoch(self, model, epoch):
        return self._current_engine(model).train_epoch(epoch)

    def _val_loop_single_epoch(self, model, epoch=0):
        return self._current_engine(model).val_epoch(epoch)

    def _current_engine(self, model):
        if getattr(self, '_engine', None) is None:      # driven from outside ``fit``
            self._prepare_once(model)
            self._engine = self._pick_engine(model)
        return self._engine

    def _finalize_training(self):
        if self.logger is not None:
            self.logger.save()
            self.logger.finalize(status='FINISHED')


class CollieTrainer(CollieMinimalTrainer):
    """``CollieTrainer(model, max_epochs=10, benchmark=True, deterministic=True, **kwargs)``
    (trainer.py:60-111). Lightning keyword arguments are accepted; those that only configure
    Lightning machinery (checkpointing, callbacks, sanity steps, ...) have no effect here.
    ``gpus=N`` (or ``devices=N``) with N > 1 trains over N GPUs when the script runs as N processes
    (``torch.distributed.run``), see ``_ShardedEngine``."""

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
