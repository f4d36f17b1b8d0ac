"""``MatrixFactorizationModel`` with the reference's constructor, attributes and checkpoint format,
running on the fused sm_100a kernels.

Mirrors:
  collie/model/matrix_factorization.py   MatrixFactorizationModel :12-167
  collie/model/base/layers.py            ScaledEmbedding :6-10, ZeroEmbedding :13-17
  collie/model/base/base_pipeline.py     __init__ :131-243, _configure_loss :277-340,
                                         configure_optimizers :342-479, calculate_loss :582-654,
                                         get_item_predictions :656-718, save_model :880-900

Lightning is not a dependency: the small pieces of ``LightningModule`` the reference API relies on
(``hparams``, ``device``, ``freeze``, ``optimizer_step``) are provided here.
"""
import textwrap
import warnings
from collections.abc import Iterable
from functools import partial

import numpy as np
import torch
from torch.optim.lr_scheduler import ReduceLROnPlateau

from . import _abi
from .interactions import (ApproximateNegativeSamplingInteractionsDataLoader, DeviceBatch,
                           ExplicitInteractions, Interactions, InteractionsDataLoader)


class HParams(dict):
    """Attribute + item access, like Lightning's ``AttributeDict`` (what lands in checkpoints)."""

    def __getattr__(self, key):
        try:
            return self[key]
        except KeyError as e:
            raise AttributeError(key) from e

    def __setattr__(self, key, value):
        self[key] = value


class ScaledEmbedding(torch.nn.Embedding):
    """N(0, 1 / (2.5 * embedding_dim)) initialisation (layers.py:6-10)."""

    def reset_parameters(self):
        self.weight.data.normal_(0, 1.0 / (self.embedding_dim * 2.5))


class ZeroEmbedding(torch.nn.Embedding):
    """All-zero initialisation (layers.py:13-17)."""

    def reset_parameters(self):
        self.weight.data.zero_()


class RowOptimizer(torch.optim.Optimizer):
    """``torch.optim.Optimizer``-shaped holder of a row-update rule, its hyper-parameters and its
    state tensors. LR schedulers act on ``param_groups`` as usual; the arithmetic happens inside the
    fused step (``cmf_train_step``) or, for gradients produced by autograd, in ``step()`` through
    ``cmf_optimizer_rows``. Rules and their torch counterparts: ``sparse_adam`` ->
    ``torch.optim.SparseAdam``, ``adam`` -> ``Adam``, ``sgd`` -> ``SGD``, ``adagrad`` -> ``Adagrad``
    (the four optimizers of base_pipeline.py:418-433)."""

    def __init__(self, params, rule, lr=1e-3, weight_decay=0.0, betas=(0.9, 0.999), eps=1e-8):
        if rule not in _abi.OPT_IDS:
            raise ValueError('{} is not a valid optimizer!'.format(rule))
        defaults = dict(lr=lr, weight_decay=weight_decay, betas=betas, eps=eps)
        super().__init__(params, defaults)
        self.rule = rule
        if rule == 'sparse_adam':
            # step() moves only the rows a step gathered: the autograd functions of this package
            # note them on parameters that carry this list
            for group in self.param_groups:
                for p in group['params']:
                    if not hasattr(p, '_cmf_rows'):
                        p._cmf_rows = []

    def state_for(self, p):
        st = self.state[p]
        if 'step' not in st:
            st['step'] = 0
            if self.rule != 'sgd':
                st['exp_avg'] = torch.zeros_like(p, memory_format=torch.preserve_format)
            if self.rule in ('adam', 'sparse_adam'):
                st['exp_avg_sq'] = torch.zeros_like(p, memory_format=torch.preserve_format)
        return st

    def group_of(self, p):
        for g in self.param_groups:
            if any(q is p for q in g['params']):
                return g
        raise KeyError('parameter not managed by this optimizer')

    def abi_opt(self, p, advance=True):
        """``cmf_opt_t`` for parameter ``p``; advances its step counter like torch does."""
        st = self.state_for(p)
        if advance:
            st['step'] += 1
        g = self.group_of(p)
        return _abi.make_opt(self.rule, g['lr'], max(st['step'], 1), g.get('betas', (0.9, 0.999)),
                             g.get('eps', 1e-8), g.get('weight_decay', 0.0)), st

    @torch.no_grad()
    def step(self, closure=None):
        """Apply the rule to dense ``.grad`` tensors produced by autograd (every row)."""
        loss = closure() if closure is not None else None
        for group in self.param_groups:
            for p in group['params']:
                if p.grad is None:
                    continue
                if p.grad.is_sparse:
                    raise _abi.CmfError('RowOptimizer.step() takes dense gradients; sparse-row '
                                        'updates happen inside the fused training step')
                _abi.require_cuda(p.data, 'parameter')
                opt, st = self.abi_opt(p)
                rows, dim = (p.shape[0], p.shape[1]) if p.dim() == 2 else (p.numel(), 1)
                grad = p.grad.contiguous()
                # SparseAdam moves the rows of the (coalesced) sparse gradient only: the autograd
                # functions of this package note the rows they gathered on the parameter
                touched, n_touched = None, 0
                noted = getattr(p, '_cmf_rows', None)
                if noted:
                    if self.rule == 'sparse_adam':
                        touched = torch.unique(torch.cat([r.reshape(-1).long() for r in noted])) \
                            .int().contiguous()
                        n_touched = int(touched.numel())
                    p._cmf_rows = []
                with _abi.on_device(p.device):
                    rc = _abi.lib().cmf_optimizer_rows(
                        _abi.ptr(p.data), _abi.ptr(st.get('exp_avg')),
                        _abi.ptr(st.get('exp_avg_sq')), _abi.ptr(grad), rows, dim,
                        _abi.ptr(touched), n_touched, opt, 0, _abi.stream_ptr(p.device))
                _abi.check(rc, 'cmf_optimizer_rows')
        return loss


class _ScoresFunction(torch.autograd.Function):
    """``forward`` of the model as one kernel, differentiable through ``cmf_backward_scores``."""

    @staticmethod
    def forward(ctx, U, I, bu, bi, users, items, y_range):
        out = torch.empty(users.shape[0], dtype=torch.float32, device=U.device)
        yr = (_abi.C.c_float * 2)(*y_range) if y_range is not None else None
        with _abi.on_device(U.device):
            rc = _abi.lib().cmf_forward_scores(
                _abi.ptr(U), _abi.ptr(I), _abi.ptr(bu), _abi.ptr(bi), U.shape[1], _abi.ptr(users),
                _abi.ptr(items), users.element_size(), users.shape[0], yr, _abi.ptr(out),
                _abi.stream_ptr(U.device))
        _abi.check(rc, 'cmf_forward_scores')
        ctx.save_for_backward(U, I, users, items, out)
        ctx.y_range = y_range
        ctx.shapes = (bu.shape, bi.shape)
        ctx.params = (U, I, bu, bi)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        U, I, users, items, out = ctx.saved_tensors
        gU, gI = torch.zeros_like(U), torch.zeros_like(I)
        gbu = torch.zeros(ctx.shapes[0], dtype=torch.float32, device=U.device)
        gbi = torch.zeros(ctx.shapes[1], dtype=torch.float32, device=U.device)
        yr = (_abi.C.c_float * 2)(*ctx.y_range) if ctx.y_range is not None else None
        g = grad_out.contiguous().float()
        with _abi.on_device(U.device):
            rc = _abi.lib().cmf_backward_scores(
                _abi.ptr(U), _abi.ptr(I), U.shape[1], _abi.ptr(users), _abi.ptr(items),
                users.element_size(), users.shape[0], _abi.ptr(g), _abi.ptr(out), yr, _abi.ptr(gU),
                _abi.ptr(gI), _abi.ptr(gbu), _abi.ptr(gbi), _abi.stream_ptr(U.device))
        _abi.check(rc, 'cmf_backward_scores')
        # the rows this call gathered (what torch's sparse embedding gradient would contain):
        # RowOptimizer('sparse_adam').step() moves those rows only
        for p, ids in zip(ctx.params, (users, items, users, items)):
            rows = getattr(p, '_cmf_rows', None)
            if rows is not None:
                rows.append(ids)
        return gU, gI, gbu, gbi, None, None, None


class _GatherConcat(torch.autograd.Function):
    """``X[r] = [table_0[ids_0[r]] | table_1[ids_1[r]] | ...]`` through ``cmf_gather_rows`` (one
    pass per source table straight into its column block); backward adds the column blocks of
    ``dX`` into dense gradient tables of the sources that require one
    (``cmf_scatter_add_rows``)."""

    @staticmethod
    def forward(ctx, ids_list, *tables):
        n = int(ids_list[0].shape[0])
        widths = [int(t.shape[1]) for t in tables]
        out = torch.empty((n, sum(widths)), dtype=torch.float32, device=tables[0].device)
        lib = _abi.lib()
        col = 0
        with _abi.on_device(out.device):
            for ids, table, w in zip(ids_list, tables, widths):
                _abi.require_cuda(table, 'table')
                _abi.check(lib.cmf_gather_rows(
                    _abi.ptr(table), table.shape[0], w, _abi.ptr(ids), ids.element_size(), n,
                    _abi.ptr(out), out.shape[1], col, _abi.stream_ptr(out.device)),
                    'cmf_gather_rows')
                col += w
        ctx.ids_list, ctx.widths = ids_list, widths
        ctx.shapes = [tuple(t.shape) for t in tables]
        ctx.needs = [t.requires_grad for t in tables]
        return out

    @staticmethod
    def backward(ctx, grad_out):
        grad_out = grad_out.contiguous()
        lib = _abi.lib()
        grads, col = [], 0
        n = int(grad_out.shape[0])
        with _abi.on_device(grad_out.device):
            for ids, shape, w, need in zip(ctx.ids_list, ctx.shapes, ctx.widths, ctx.needs):
                g = None
                if need:
                    g = torch.zeros(shape, dtype=torch.float32, device=grad_out.device)
                    _abi.check(lib.cmf_scatter_add_rows(
                        _abi.ptr(g), shape[0], w, _abi.ptr(ids), ids.element_size(), n,
                        _abi.ptr(grad_out), grad_out.shape[1], col,
                        _abi.stream_ptr(grad_out.device)), 'cmf_scatter_add_rows')
                grads.append(g)
                col += w
        return (None,) + tuple(grads)


def _as_ids(t, device):
    if not torch.is_tensor(t):
        t = torch.as_tensor(np.asarray(t))
    if t.dtype not in (torch.int32, torch.int64):
        t = t.long()
    return t.to(device).contiguous()


class MatrixFactorizationModel(torch.nn.Module):
    """Matrix factorization with user/item biases, trained by the fused B200 step.

    Same constructor as the reference (matrix_factorization.py:73-94).
    """

    def __init__(self, train=None, val=None, embedding_dim=30, dropout_p=0.0, sparse=False,
                 lr=1e-3, bias_lr=1e-2,
                 lr_scheduler_func=partial(ReduceLROnPlateau, patience=1),
                 weight_decay=0.0, optimizer='adam', bias_optimizer='sgd', loss='hinge',
                 metadata_for_loss=None, metadata_for_loss_weights=None, y_range=None,
                 load_model_path=None, map_location=None):
        super().__init__()
        if isinstance(train, (Interactions, ExplicitInteractions)):
            train = InteractionsDataLoader(interactions=train, shuffle=True)
        if isinstance(val, (Interactions, ExplicitInteractions)):
            val = InteractionsDataLoader(interactions=val, shuffle=False)
        self.train_loader = train
        self.val_loader = val
        self.lr_scheduler_func = lr_scheduler_func
        self.loss = loss
        self.optimizer = optimizer
        self.bias_optimizer = bias_optimizer
        self._hparams = HParams()
        self._workspace = None
        self._meta_dev = None

        if load_model_path is not None:
            self._load_model_init_helper(load_model_path, map_location)
            return

        if self.train_loader is None:
            raise TypeError('``train`` must be provided to all newly-instantiated models!')
        if self.val_loader is not None:
            assert self.train_loader.num_users == self.val_loader.num_users, (
                'Both training and val ``num_users`` must equal: '
                f'{self.train_loader.num_users} != {self.val_loader.num_users}.')
            assert self.train_loader.num_items == self.val_loader.num_items, (
                'Both training and val ``num_items`` must equal: '
                f'{self.train_loader.num_items} != {self.val_loader.num_items}.')
            if (hasattr(self.train_loader, 'num_negative_samples')
                    and _has_negs(self.train_loader)):
                msg = ('Training and val ``num_negative_samples`` property must both equal ``1``'
                       ' or both be greater than ``1``.')
                if self.train_loader.num_negative_samples == 1:
                    assert self.val_loader.num_negative_samples == 1, msg
                elif self.train_loader.num_negative_samples > 1:
                    assert self.val_loader.num_negative_samples > 1, msg
                else:
                    raise ValueError('``self.train_loader.num_negative_samples`` must be greater '
                                     f'than ``0``, not {self.train_loader.num_negative_samples}.')

        self.hparams.update(dict(
            embedding_dim=embedding_dim, dropout_p=dropout_p, sparse=sparse, lr=lr, bias_lr=bias_lr,
            weight_decay=weight_decay,
            optimizer=optimizer if isinstance(optimizer, str) else None,
            bias_optimizer=bias_optimizer if isinstance(bias_optimizer, str) else None,
            loss=loss if isinstance(loss, str) else None,
            metadata_for_loss=metadata_for_loss,
            metadata_for_loss_weights=metadata_for_loss_weights, y_range=y_range,
            load_model_path=None, map_location=None))
        self.hparams.num_users = self.train_loader.num_users
        self.hparams.num_items = self.train_loader.num_items
        self.hparams.num_epochs_completed = 0

        self._configure_loss()

        if self.hparams.sparse and self.hparams.weight_decay != 0:
            warnings.warn(textwrap.dedent(f'''
                ``weight_decay`` value must be 0 when ``sparse`` is flagged, not
                {self.hparams.weight_decay}. Setting to 0.
                ''').replace('\n', ' ').strip())
            self.hparams.weight_decay = 0.0

        self._setup_model()

    # ------------------------------------------------------------------ Lightning-ish surface
    @property
    def hparams(self):
        return self._hparams

    @property
    def device(self):
        return self.user_embeddings.weight.device

    def freeze(self):
        for p in self.parameters():
            p.requires_grad = False
        self.eval()

    def optimizer_step(self, epoch=None, batch_idx=None, optimizer=None, optimizer_idx=None,
                       optimizer_closure=None, **kwargs):
        """Lightning hook signature kept (trainer.py:465-470)."""
        if optimizer_closure is not None:
            optimizer.step(closure=optimizer_closure)
        else:
            optimizer.step()

    def train_dataloader(self):
        return self.train_loader

    def val_dataloader(self):
        return self.val_loader

    def _move_any_external_data_to_device(self):
        """Metadata tensors used by the losses follow the model (base_pipeline.py:259-275)."""
        self._meta_dev = None

    # ------------------------------------------------------------------ set-up
    def _load_model_init_helper(self, load_model_path, map_location):
        loaded = _load_checkpoint(load_model_path, map_location)
        for key, value in loaded['hparams'].items():
            self.hparams[key] = value
        self.hparams['load_model_path'] = load_model_path
        self.hparams['map_location'] = map_location
        self.loss_function = None
        self._loss_name = None
        self._setup_model()
        self.load_state_dict(loaded['state_dict'])
        self.eval()

    def _configure_loss(self):
        """String -> loss, with the reference's errors and warnings (base_pipeline.py:277-340)."""
        self.loss_function = None
        self._loss_name = None
        if callable(self.loss):
            # (the reference leaves ``_is_implicit`` unset for callables, base_pipeline.py:281)
            self.loss_function = self.loss
            return
        self.hparams._is_implicit = False
        if self.loss in ('mse', 'mae'):
            self._loss_name = self.loss
            self.loss_function = self.loss
            return
        self.hparams._is_implicit = True
        if not _has_negs(self.train_loader):
            raise ValueError('``num_negative_samples`` attribute not found in ``train_loader`` - '
                             'are you using explicit data with an implicit loss function?')
        n = self.train_loader.num_negative_samples
        if self.loss == 'warp':
            if n > 1:
                self._loss_name = 'warp'
            else:
                raise ValueError('Cannot use WARP loss with a single negative sample!')
        elif 'bpr' in self.loss:
            if n > 1:
                self._loss_name = 'adaptive_bpr'
            else:
                if 'adaptive' in self.loss:
                    warnings.warn('Adaptive BPR loss specified, but ``num_negative_samples`` == 1.'
                                  ' Using standard BPR loss instead.')
                self._loss_name = 'bpr'
        elif 'hinge' in self.loss or self.loss == 'adaptive':
            if n > 1:
                self._loss_name = 'adaptive_hinge'
            else:
                if 'adaptive' in self.loss:
                    warnings.warn('Adaptive hinge loss specified, but ``num_negative_samples`` =='
                                  ' 1. Using standard hinge loss instead.')
                self._loss_name = 'hinge'
        else:
            raise ValueError('{} is not a valid loss function.'.format(self.loss))
        self.loss_function = self._loss_name

    def _setup_model(self):
        h = self.hparams
        self.user_biases = ZeroEmbedding(h.num_users, 1, sparse=h.sparse)
        self.item_biases = ZeroEmbedding(h.num_items, 1, sparse=h.sparse)
        self.user_embeddings = ScaledEmbedding(h.num_users, h.embedding_dim, sparse=h.sparse)
        self.item_embeddings = ScaledEmbedding(h.num_items, h.embedding_dim, sparse=h.sparse)
        self.dropout = torch.nn.Dropout(p=h.dropout_p)

    # ------------------------------------------------------------------ optimizers
    def configure_optimizers(self):
        """Same return shapes as the reference (base_pipeline.py:342-404)."""
        if self.bias_optimizer is not None:
            if self.bias_optimizer == 'infer':
                self.bias_optimizer = self.optimizer
            if self.hparams.bias_lr == 'infer':
                self.hparams.bias_lr = self.hparams.lr
            bias_optimizer = self._get_optimizer(self.bias_optimizer, 'bias')
            optimizer = self._get_optimizer(self.optimizer, 'all_but_bias')
        else:
            optimizer = self._get_optimizer(self.optimizer, 'all')
        if self.lr_scheduler_func is not None:
            monitor = 'val_loss_epoch' if self.val_loader is not None else 'train_loss_epoch'
            opts = [optimizer]
            scheds = [{'scheduler': self.lr_scheduler_func(optimizer), 'monitor': monitor}]
            if self.bias_optimizer is not None:
                opts.append(bias_optimizer)
                scheds.append({'scheduler': self.lr_scheduler_func(bias_optimizer),
                               'monitor': monitor})
            return (opts, scheds)
        if self.bias_optimizer is not None:
            return [optimizer, bias_optimizer]
        return optimizer

    def _get_optimizer(self, optimizer, optimizer_type):
        named = list(self.named_parameters())
        if optimizer_type == 'bias':
            params, lr = [p for n, p in named if 'bias' in n], self.hparams.bias_lr
        elif optimizer_type == 'all_but_bias':
            params, lr = [p for n, p in named if 'bias' not in n], self.hparams.lr
        else:
            params, lr = [p for _, p in named], self.hparams.lr
        if callable(optimizer):
            # a user-supplied optimizer class / factory (base_pipeline.py:409-416): the trainers
            # then run the reference's unfused loop (differentiable forward, optimizer.step())
            try:
                return optimizer(params, lr=lr, weight_decay=self.hparams.weight_decay)
            except TypeError:
                return optimizer(params, lr=lr)
        if optimizer not in _abi.OPT_IDS:
            raise ValueError('{} is not a valid optimizer!'.format(optimizer))
        wd = 0.0 if optimizer == 'sparse_adam' else self.hparams.weight_decay
        return RowOptimizer([{'params': params, 'lr': lr}], optimizer, lr=lr, weight_decay=wd)

    # ------------------------------------------------------------------ forward
    def forward(self, users, items):
        """``preds = <U[u], I[i]> + bu[u] + bi[i]`` (+ ``y_range`` sigmoid), one fused kernel
        (matrix_factorization.py:120-159)."""
        U = _abi.require_cuda(self.user_embeddings.weight, 'user_embeddings.weight')
        users = _as_ids(users, U.device)
        items = _as_ids(items, U.device)
        if users.dtype != items.dtype:
            users, items = users.long(), items.long()
        if users.shape != items.shape or users.dim() != 1:
            raise ValueError('``users`` and ``items`` must be 1-d tensors of equal length')
        yr = tuple(float(v) for v in self.hparams.y_range) if self.hparams.y_range is not None \
            else None
        if self.training and self.hparams.dropout_p > 0:
            # the stand-alone forward of a model in training mode (matrix_factorization.py:143-157:
            # dropout on each embedding, then the product): rows through the gather kernel, the
            # masks from torch's generator. The fused training step draws its own masks in-kernel.
            D = U.shape[1]
            rows = _GatherConcat.apply([users, items], U, self.item_embeddings.weight)
            bias = _GatherConcat.apply([users, items], self.user_biases.weight,
                                       self.item_biases.weight)
            preds = (self.dropout(rows[:, :D]) * self.dropout(rows[:, D:])).sum(dim=1) \
                + bias[:, 0] + bias[:, 1]
            if yr is not None:
                preds = torch.sigmoid(preds) * (yr[1] - yr[0]) + yr[0]
            return preds
        return _ScoresFunction.apply(U, self.item_embeddings.weight, self.user_biases.weight,
                                     self.item_biases.weight, users, items, yr)

    def _get_item_embeddings(self):
        return self.item_embeddings.weight.data

    def _get_user_embeddings(self):
        return self.user_embeddings.weight.data

    # ------------------------------------------------------------------ fused step
    def _workspace_for(self, B, N):
        h = self.hparams
        need = _abi.lib().cmf_train_workspace_bytes(B, N, h.num_users, h.num_items,
                                                    h.embedding_dim)
        ws = self._workspace
        if ws is None or ws.numel() < need or ws.device != self.device:
            ws = torch.empty(int(need), dtype=torch.uint8, device=self.device)
            _abi.check(_abi.lib().cmf_workspace_init(_abi.ptr(ws), ws.numel(),
                                                     _abi.stream_ptr(self.device)),
                       'cmf_workspace_init')
            self._workspace = ws
            self._workspace_B = B
        elif B > self._workspace_B:
            # the per-step part of the layout grows with B; a bigger batch needs a fresh layout
            self._workspace = None
            return self._workspace_for(B, N)
        return ws

    def _metadata_device(self):
        if self._meta_dev is None:
            meta = self.hparams.get('metadata_for_loss') or {}
            weights = self.hparams.get('metadata_for_loss_weights') or {}
            if len(meta) > _abi.CMF_MAX_METADATA:
                raise ValueError(f'at most {_abi.CMF_MAX_METADATA} metadata keys are supported')
            total = sum(weights.values()) if len(meta) else 0
            if total > 1:
                raise ValueError(f'sum of metadata weights was {total}, must be <=1')
            self._meta_dev = [(torch.as_tensor(v).reshape(-1).to(self.device, torch.int32)
                               .contiguous(), float(weights[k])) for k, v in meta.items()]
        return self._meta_dev

    def _normalise_batch(self, batch):
        """Reference batch formats (base_pipeline.py:603, :635) or a ``DeviceBatch``."""
        dev = self.device
        if isinstance(batch, DeviceBatch):
            return batch.users, batch.items, None, batch.gidx, batch.ratings
        if len(batch) == 2 and isinstance(batch[0], Iterable) and len(batch[0]) == 2:
            if self.hparams.get('_is_implicit') is False:
                raise ValueError('Explicit loss with implicit data is invalid!')
            (users, pos), negs = batch
            negs = torch.as_tensor(negs).to(dev).long()
            if negs.dim() == 1:
                negs = negs.view(-1, 1)
            return (torch.as_tensor(users).to(dev, torch.int32).contiguous(),
                    torch.as_tensor(pos).to(dev, torch.int32).contiguous(), negs.contiguous(),
                    None, None)
        if len(batch) == 3:
            if self.hparams.get('_is_implicit') is True:
                raise ValueError('Implicit loss with explicit data is invalid!')
            users, items, ratings = batch
            return (torch.as_tensor(users).to(dev, torch.int32).contiguous(),
                    torch.as_tensor(items).to(dev, torch.int32).contiguous(), None, None,
                    torch.as_tensor(ratings).to(dev).float().contiguous())
        raise ValueError(f'Unexpected format for batch: {batch}. See docs for expected format.')

    def fused_step(self, batch, optimizers=None, train=True, interactions=None, epoch=0,
                   outputs=None, users_sorted=False, global_batch=0, skip_item_update=False,
                   num_items_logical=0, epoch_batch_size=None, item_grads_event=None):
        """One step through ``cmf_train_step``; returns the loss as a 0-d device tensor.

        ``optimizers``: what ``configure_optimizers`` returned (one optimizer, or [main, bias]).
        ``interactions``: the dataset behind a ``DeviceBatch`` (seed / exact-sampling CSR).
        ``outputs``: optional dict; keys among ``negs``, ``sel``, ``pos_scores``, ``neg_scores`` are
        filled with device tensors for inspection.
        ``epoch_batch_size``: ``batch`` is a whole planned epoch (``loader.epoch_plan``); all its
        steps of that batch size run in ONE library call (``cmf_train_epoch``) and the per-step
        losses come back as a 1-d device tensor.
        ``global_batch`` / ``skip_item_update``: multi-GPU hooks (``collie_b200.distributed``): the
        loss is a mean over the global batch, and the item-row gradients stay in the workspace
        for the caller to reduce across ranks; ``item_grads_event`` (a ``torch.cuda.Event`` that
        has been recorded once) is recorded again as soon as those gradients are complete.
        """
        if self._loss_name is None:
            raise _abi.CmfError('a custom loss callable cannot run inside the fused step; use one '
                                "of 'bpr', 'hinge', 'adaptive', 'adaptive_bpr', 'warp', 'mse', "
                                "'mae'")
        h = self.hparams
        users, items, negs, gidx, ratings = self._normalise_batch(batch)
        U = _abi.require_cuda(self.user_embeddings.weight.data, 'user_embeddings.weight')
        I = _abi.require_cuda(self.item_embeddings.weight.data, 'item_embeddings.weight')
        bu = _abi.require_cuda(self.user_biases.weight.data, 'user_biases.weight')
        bi = _abi.require_cuda(self.item_biases.weight.data, 'item_biases.weight')
        implicit = self._loss_name not in ('mse', 'mae')
        n_total = int(users.shape[0])
        B = n_total if epoch_batch_size is None else int(epoch_batch_size)
        if epoch_batch_size is not None:
            if B < 1:
                raise ValueError('epoch_batch_size must be positive')
            if outputs is not None:
                raise ValueError('inspection outputs are per batch; not available for a whole epoch')
            if skip_item_update:
                raise ValueError('a whole epoch cannot leave its item updates to the caller')
        n_steps = (n_total + B - 1) // B if epoch_batch_size is not None else 1
        a = _abi.cmf_step_t()
        a.U, a.I, a.bu, a.bi = U.data_ptr(), I.data_ptr(), bu.data_ptr(), bi.data_ptr()
        a.num_users, a.num_items, a.D = h.num_users, h.num_items, h.embedding_dim
        a.users, a.items = users.data_ptr(), items.data_ptr()
        a.B = B
        a.users_sorted = 1 if (users_sorted or getattr(batch, 'sorted', False)) else 0
        keep = [users, items]
        if implicit:
            if ratings is not None and negs is None and gidx is None:
                raise ValueError('Implicit loss with explicit data is invalid!')
            if negs is not None:
                if negs.shape[0] != n_total:
                    raise ValueError('negative samples must have one row per interaction')
                a.N = int(negs.shape[1])
                a.negs_in = negs.data_ptr()
                keep.append(negs)
            else:
                data = interactions if interactions is not None else self.train_loader.interactions
                a.N = int(data.num_negative_samples)
                a.seed = int(data.seed)
                a.epoch = int(epoch)
                a.max_tries = int(data.max_number_of_samples_to_consider)
                if gidx is None:
                    raise ValueError('a batch without negatives needs dataset positions (gidx)')
                a.gidx = gidx.data_ptr()
                keep.append(gidx)
                if data.exact_sampling:
                    st = data.device_state(self.device)
                    a.csr_indptr, a.csr_items = st['indptr'].data_ptr(), st['csr_items'].data_ptr()
            meta = self._metadata_device()
            a.n_meta = len(meta)
            for k, (labels, w) in enumerate(meta):
                a.meta[k] = labels.data_ptr()
                a.meta_w[k] = w
        else:
            if ratings is None:
                raise ValueError('Explicit loss with implicit data is invalid!')
            a.N = 0
            a.ratings = ratings.data_ptr()
            keep.append(ratings)
            if h.y_range is not None:
                a.has_y_range, a.y_lo, a.y_hi = 1, float(h.y_range[0]), float(h.y_range[1])
        a.loss = _abi.LOSS_IDS[self._loss_name]
        a.train = 1 if train else 0
        if train:
            if optimizers is None:
                raise ValueError('a training step needs the optimizers from configure_optimizers()')
            if isinstance(optimizers, tuple):
                optimizers = optimizers[0]
            if not isinstance(optimizers, (list, tuple)):
                optimizers = [optimizers]
            main = optimizers[0]
            bias = optimizers[1] if len(optimizers) > 1 else optimizers[0]
            for o in (main, bias):
                if not isinstance(o, RowOptimizer):
                    raise _abi.CmfError('the fused step needs the RowOptimizer objects returned by '
                                        'configure_optimizers()')
            pU, pI = self.user_embeddings.weight, self.item_embeddings.weight
            pbu, pbi = self.user_biases.weight, self.item_biases.weight
            a.emb_opt, sU = main.abi_opt(pU)
            _, sI = main.abi_opt(pI)
            a.bias_opt, sbu = bias.abi_opt(pbu)
            _, sbi = bias.abi_opt(pbi)
            for st_ in (sU, sI, sbu, sbi):       # one optimizer step per batch of the epoch
                st_['step'] += n_steps - 1

            def sp(st, key):
                t = st.get(key)
                return t.data_ptr() if t is not None else None
            a.U_m, a.U_v = sp(sU, 'exp_avg'), sp(sU, 'exp_avg_sq')
            a.I_m, a.I_v = sp(sI, 'exp_avg'), sp(sI, 'exp_avg_sq')
            a.bu_m, a.bu_v = sp(sbu, 'exp_avg'), sp(sbu, 'exp_avg_sq')
            a.bi_m, a.bi_v = sp(sbi, 'exp_avg'), sp(sbi, 'exp_avg_sq')
        loss_out = torch.empty(max(n_steps, 1), dtype=torch.float32, device=self.device)
        a.loss_out = loss_out.data_ptr()
        if outputs is not None:
            N = max(int(a.N), 1)
            if 'negs' in outputs and implicit:
                outputs['negs'] = torch.empty((B, N), dtype=torch.int64, device=self.device)
                a.negs_out = outputs['negs'].data_ptr()
            if 'sel' in outputs and implicit:
                outputs['sel'] = torch.empty(B, dtype=torch.int32, device=self.device)
                a.sel_out = outputs['sel'].data_ptr()
            if 'pos_scores' in outputs:
                outputs['pos_scores'] = torch.empty(B, dtype=torch.float32, device=self.device)
                a.pos_scores_out = outputs['pos_scores'].data_ptr()
            if 'neg_scores' in outputs and implicit:
                outputs['neg_scores'] = torch.empty((B, N), dtype=torch.float32,
                                                    device=self.device)
                a.neg_scores_out = outputs['neg_scores'].data_ptr()
        ws = self._workspace_for(B, int(a.N))
        a.workspace, a.workspace_bytes = ws.data_ptr(), ws.numel()
        a.B_global = int(global_batch)
        a.skip_item_update = 1 if skip_item_update else 0
        a.num_items_logical = int(num_items_logical)
        # torch.use_deterministic_algorithms(True) (or model.deterministic = True) asks for the
        # bit-repeatable form of the step (cmf_step_t.deterministic)
        det = getattr(self, 'deterministic', None)
        if det is None:
            det = torch.are_deterministic_algorithms_enabled()
        a.deterministic = 1 if (det and train) else 0
        if self.training and h.dropout_p > 0:
            # nn.Dropout semantics: active whenever the module is in training mode
            a.dropout_p = float(h.dropout_p)
            data = interactions if interactions is not None else self.train_loader.interactions
            a.seed = int(getattr(data, 'seed', 0) or 0)
            a.epoch = int(epoch)
            if not train:       # (no optimizer step to key the masks with)
                a.emb_opt.step = 1
        if item_grads_event is not None:
            a.item_grads_ready_event = int(item_grads_event.cuda_event)
        with _abi.on_device(self.device):
            if epoch_batch_size is not None:
                rc = _abi.lib().cmf_train_epoch(_abi.C.byref(a), n_total, loss_out.data_ptr(),
                                                _abi.stream_ptr(self.device))
                _abi.check(rc, 'cmf_train_epoch')
                return loss_out[:n_steps]
            rc = _abi.lib().cmf_train_step(_abi.C.byref(a), _abi.stream_ptr(self.device))
            _abi.check(rc, 'cmf_train_step')
        return loss_out[0]

    def calculate_loss(self, batch, differentiable=None):
        """Loss of one batch in the reference's formats (base_pipeline.py:582-654).

        Named losses: the fused kernel's value, no autograd graph -- training goes through the
        trainers (``fused_step``), which compute the same loss and apply its gradient in the same
        kernels. ``differentiable=True`` (the default for a callable ``loss``, and what the
        trainers' unfused loop asks for) builds the reference's own graph instead: ``forward`` for
        the positives, ``forward`` for ``users.repeat(N)`` and the flattened negatives
        (base_pipeline.py:616-623), the loss callable with the reference's keyword arguments
        (:626-634); ``backward()`` then runs ``cmf_backward_scores``."""
        if differentiable is None:
            differentiable = self._loss_name is None or getattr(self, '_unfused', False)
        if not differentiable:
            return self.fused_step(batch, train=False)
        h = self.hparams
        users, items, negs, gidx, ratings = self._normalise_batch(batch)
        if negs is None and ratings is None:      # a DeviceBatch of implicit data: draw them
            data = self.train_loader.interactions
            negs = data.sample_negatives(users, gidx)
        if ratings is not None:
            if h.get('_is_implicit') is True:
                raise ValueError('Implicit loss with explicit data is invalid!')
            preds = self(users, items)
            fn = self.loss_function
            if fn == 'mse':
                return torch.nn.functional.mse_loss(preds, ratings.float())
            if fn == 'mae':
                return torch.nn.functional.l1_loss(preds, ratings.float())
            return fn(preds, ratings.float())
        if h.get('_is_implicit') is False:
            raise ValueError('Explicit loss with implicit data is invalid!')
        n_neg = int(negs.shape[1])
        neg_t = negs.t().contiguous()                       # [N, B], like the reference
        pos_preds = self(users.long(), items.long())
        neg_preds = self(users.long().repeat(n_neg), neg_t.reshape(-1)).view(n_neg, len(users))
        fn = self.loss_function
        if isinstance(fn, str):
            from . import loss as L
            fn = L.NAMED[fn]
            if self._loss_name in ('bpr', 'hinge'):
                neg_preds, neg_t = neg_preds[0], neg_t[0]
        elif n_neg == 1:
            neg_preds, neg_t = neg_preds.view(-1), neg_t.view(-1)
        meta = h.get('metadata_for_loss')
        if meta:
            meta = {k: torch.as_tensor(v).to(self.device) for k, v in meta.items()}
        return fn(pos_preds, neg_preds, num_items=h.num_items, positive_items=items.long(),
                  negative_items=neg_t, metadata=meta,
                  metadata_weights=h.get('metadata_for_loss_weights'))

    # ------------------------------------------------------------------ predictions / checkpoints
    def get_item_predictions(self, user_id=0, unseen_items_only=False, sort_values=True):
        """base_pipeline.py:656-718."""
        import pandas as pd
        h = self.hparams
        if user_id >= h.num_users:
            raise ValueError(
                f'``user_id`` {user_id} is not in the model. Expected ID between ``0`` and '
                f'``self.hparams.num_users - 1`` (``{h.num_users - 1}``), not ``{user_id}``')
        users = torch.full((h.num_items,), int(user_id), dtype=torch.long, device=self.device)
        items = torch.arange(h.num_items, dtype=torch.long, device=self.device)
        with torch.no_grad():
            preds = pd.Series(self(users, items).detach().cpu().numpy())
        if sort_values:
            preds = preds.sort_values(ascending=False)
        if unseen_items_only:
            seen = [self.train_loader.mat.tocsr()[user_id, :].nonzero()[1]]
            if self.val_loader is not None:
                seen.append(self.val_loader.mat.tocsr()[user_id, :].nonzero()[1])
            preds = preds.drop(np.concatenate(seen))
        return preds

    def get_user_predictions(self, item_id=0, unseen_users_only=False, sort_values=True):
        """base_pipeline.py:720-782."""
        import pandas as pd
        h = self.hparams
        if item_id >= h.num_items:
            raise ValueError(
                f'``item_id`` {item_id} is not in the model. Expected ID between ``0`` and '
                f'``self.hparams.num_items - 1`` (``{h.num_items - 1}``), not ``{item_id}``')
        users = torch.arange(h.num_users, dtype=torch.long, device=self.device)
        items = torch.full((h.num_users,), int(item_id), dtype=torch.long, device=self.device)
        with torch.no_grad():
            preds = pd.Series(self(users, items).detach().cpu().numpy())
        if sort_values:
            preds = preds.sort_values(ascending=False)
        if unseen_users_only:
            seen = [self.train_loader.mat.tocsc()[:, item_id].nonzero()[0]]
            if self.val_loader is not None:
                seen.append(self.val_loader.mat.tocsc()[:, item_id].nonzero()[0])
            preds = preds.drop(np.concatenate(seen))
        return preds

    def item_item_similarity(self, item_id):
        """base_pipeline.py:783-815: cosine similarity of ``item_id`` to every item, sorted
        descending (the item itself comes first)."""
        h = self.hparams
        if item_id >= h.num_items:
            raise ValueError(
                f'``item_id`` {item_id} is not in the model. Expected ID between ``0`` and '
                f'``self.hparams.num_items - 1`` (``{h.num_items - 1}``), not ``{item_id}``')
        return self._calculate_embedding_similarity(self._get_item_embeddings(), item_id)

    def user_user_similarity(self, user_id):
        """base_pipeline.py:817-847."""
        h = self.hparams
        if user_id >= h.num_users:
            raise ValueError(
                f'``user_id`` {user_id} is not in the model. Expected ID between ``0`` and '
                f'``self.hparams.num_users - 1`` (``{h.num_users - 1}``), not ``{user_id}``')
        return self._calculate_embedding_similarity(self._get_user_embeddings(), user_id)

    def _calculate_embedding_similarity(self, embeddings, id):
        """base_pipeline.py:849-859 through ``cmf_row_similarity``."""
        import pandas as pd
        E = _abi.require_cuda(embeddings, 'embeddings')
        out = torch.empty(E.shape[0], dtype=torch.float32, device=E.device)
        with torch.cuda.device(E.device):
            _abi.check(_abi.lib().cmf_row_similarity(_abi.ptr(E), E.shape[0], E.shape[1], int(id),
                                                     _abi.ptr(out), _abi.stream_ptr(E.device)),
                       'cmf_row_similarity')
        return pd.Series(out.cpu().numpy()).sort_values(ascending=False)

    def save_model(self, filename='model.pth'):
        """``{'state_dict', 'hparams'}`` exactly as the reference writes it
        (base_pipeline.py:880-900); optimizer state and datasets are not saved."""
        torch.save({'state_dict': self.state_dict(), 'hparams': dict(self.hparams)}, str(filename))


def _load_checkpoint(path, map_location=None):
    """``{'state_dict', 'hparams'}`` as ``save_model`` of either library wrote it
    (collie/model/base/base_pipeline.py:880-900). The reference pickles ``hparams`` as PyTorch
    Lightning's ``AttributeDict``; Lightning is not a dependency here, so classes of its packages
    that cannot be imported are read as plain attribute dictionaries."""
    import pickle
    import types

    class _Unpickler(pickle.Unpickler):
        def find_class(self, module, name):
            try:
                return super().find_class(module, name)
            except (ImportError, AttributeError):
                if module.split('.')[0] in ('pytorch_lightning', 'lightning', 'lightning_fabric',
                                            'lightning_utilities'):
                    return HParams
                raise

    shim = types.ModuleType('collie_b200_checkpoint_pickle')
    shim.Unpickler = _Unpickler
    shim.load = lambda f, **kw: _Unpickler(f, **kw).load()
    shim.__dict__.update({k: getattr(pickle, k) for k in ('UnpicklingError', 'PickleError')})
    return torch.load(str(path), map_location=map_location, weights_only=False,
                      pickle_module=shim)


def _has_negs(loader):
    try:
        loader.num_negative_samples
        return True
    except AttributeError:
        return False
