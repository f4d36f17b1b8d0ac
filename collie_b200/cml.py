"""``CollaborativeMetricLearningModel`` (the second embedding model of the path's
family): two embedding tables, no biases, the prediction of a (user, item) pair is the pairwise
distance of their rows.

Reference: collie/model/collaborative_metric_learning.py (constructor :63-84, tables :86-98,
forward :100-124). The forward is ONE kernel (both gathers + the reduction,
``cmf_pairwise_distance``), its backward one more (``cmf_pairwise_distance_backward``: dense
gradient tables filled with RED, the rows that were gathered remembered for SparseAdam); training
runs through the trainers' unfused loop (``calculate_loss -> backward -> optimizer step``) with
the losses of ``collie_b200.loss`` -- the fused step scores dot products, not distances -- and
evaluation through ``forward`` over all pairs. ``state_dict`` keys, hyper-parameters and the
checkpoint format are the reference's."""
from functools import partial

import torch
from torch.optim.lr_scheduler import ReduceLROnPlateau

from . import _abi
from .model import MatrixFactorizationModel, RowOptimizer, ScaledEmbedding, _as_ids


class _PairwiseDistance(torch.autograd.Function):
    """``|| U[users] - I[items] + 1e-6 ||_2`` per pair (torch.nn.functional.pairwise_distance)."""

    @staticmethod
    def forward(ctx, U, I, users, items):
        out = torch.empty(users.shape[0], dtype=torch.float32, device=U.device)
        with _abi.on_device(U.device):
            rc = _abi.lib().cmf_pairwise_distance(
                _abi.ptr(U), _abi.ptr(I), U.shape[1], _abi.ptr(users), _abi.ptr(items),
                users.element_size(), users.shape[0], _abi.ptr(out), _abi.stream_ptr(U.device))
        _abi.check(rc, 'cmf_pairwise_distance')
        ctx.save_for_backward(U, I, users, items, out)
        ctx.params = (U, I)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        U, I, users, items, out = ctx.saved_tensors
        need_u, need_i = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        gU = torch.zeros_like(U) if need_u else None
        gI = torch.zeros_like(I) if need_i else None
        g = grad_out.contiguous().float()
        with _abi.on_device(U.device):
            rc = _abi.lib().cmf_pairwise_distance_backward(
                _abi.ptr(U), _abi.ptr(I), U.shape[1], _abi.ptr(users), _abi.ptr(items),
                users.element_size(), users.shape[0], _abi.ptr(g), _abi.ptr(out), _abi.ptr(gU),
                _abi.ptr(gI), _abi.stream_ptr(U.device))
        _abi.check(rc, 'cmf_pairwise_distance_backward')
        # the rows this step gathered: what torch's sparse embedding gradient would contain
        for p, ids, need in ((ctx.params[0], users, need_u), (ctx.params[1], items, need_i)):
            rows = getattr(p, '_cmf_rows', None)     # (set by RowOptimizer('sparse_adam'))
            if need and rows is not None:
                rows.append(ids)
        return gU, gI, None, None


class CollaborativeMetricLearningModel(MatrixFactorizationModel):
    """Same constructor as the reference (collaborative_metric_learning.py:63-84)."""

    _fusable = False        # the trainers take the unfused loop; metrics score through forward()

    def __init__(self, train=None, val=None, embedding_dim=30, sparse=False, lr=1e-3,
                 lr_scheduler_func=partial(ReduceLROnPlateau, patience=1), weight_decay=0.0,
                 optimizer='adam', loss='hinge', metadata_for_loss=None,
                 metadata_for_loss_weights=None, y_range=None, load_model_path=None,
                 map_location=None):
        super().__init__(train=train, val=val, embedding_dim=embedding_dim, dropout_p=0.0,
                         sparse=sparse, lr=lr, bias_lr=None, lr_scheduler_func=lr_scheduler_func,
                         weight_decay=weight_decay, optimizer=optimizer, bias_optimizer=None,
                         loss=loss, metadata_for_loss=metadata_for_loss,
                         metadata_for_loss_weights=metadata_for_loss_weights, y_range=y_range,
                         load_model_path=load_model_path, map_location=map_location)

    def _setup_model(self):
        h = self.hparams
        for k in ('dropout_p', 'bias_lr', 'bias_optimizer'):
            h.pop(k, None)                         # (not hyper-parameters of this model)
        self.user_embeddings = ScaledEmbedding(h.num_users, h.embedding_dim, sparse=h.sparse)
        self.item_embeddings = ScaledEmbedding(h.num_items, h.embedding_dim, sparse=h.sparse)

    def configure_optimizers(self):
        """One optimizer over both tables (the reference passes no bias optimizer:
        base_pipeline.py:372-373)."""
        params = [p for p in self.parameters() if p.requires_grad]
        lr = self.hparams.lr
        if callable(self.optimizer):
            try:
                optimizer = self.optimizer(params, lr=lr, weight_decay=self.hparams.weight_decay)
            except TypeError:
                optimizer = self.optimizer(params, lr=lr)
        elif self.optimizer in _abi.OPT_IDS:
            optimizer = RowOptimizer([{'params': params, 'lr': lr}], self.optimizer, lr=lr,
                                     weight_decay=self.hparams.weight_decay)
        else:
            raise ValueError('{} is not a valid optimizer!'.format(self.optimizer))
        if self.lr_scheduler_func is not None:
            monitor = 'val_loss_epoch' if self.val_loader is not None else 'train_loss_epoch'
            return ([optimizer], [{'scheduler': self.lr_scheduler_func(optimizer),
                                   'monitor': monitor}])
        return optimizer

    def forward(self, users, items):
        U = _abi.require_cuda(self.user_embeddings.weight, 'user_embeddings.weight')
        I = _abi.require_cuda(self.item_embeddings.weight, 'item_embeddings.weight')
        users, items = _as_ids(users, U.device), _as_ids(items, U.device)
        if users.shape != items.shape or users.dim() != 1:
            raise ValueError('``users`` and ``items`` must be 1-d tensors of equal length')
        return _PairwiseDistance.apply(U, I, users, items)

    def fused_step(self, *args, **kwargs):
        raise _abi.CmfError('CollaborativeMetricLearningModel trains through the trainers\' '
                            'unfused loop (calculate_loss -> backward -> optimizer step); the '
                            'fused step scores dot products')

    def calculate_loss(self, batch, differentiable=True):
        return super().calculate_loss(batch, differentiable=True)

    def _get_item_embeddings(self):
        return self.item_embeddings.weight.data

    def _get_user_embeddings(self):
        return self.user_embeddings.weight.data
