"""``HybridPretrainedModel``: frozen (or fine-tuned) matrix-factorization embeddings + item / user
metadata -> dense layers -> score (BASELINE config C4).

Mirrors collie/model/hybrid_pretrained_matrix_factorization.py: constructor :101-158, layer set-up
and xavier initialisation :227-325, metadata layers :327-361, prediction :363-399, forward
:401-464, freeze / unfreeze :476-484, directory checkpoints :486-534, ``load_from_hybrid_model``
:536-556. Module and ``state_dict`` names are the reference's (``embeddings.0/1``, ``biases.0/1``,
``combined_layer_i``, ``item_metadata_layer_i``, ``user_metadata_layer_i``), so checkpoints
interchange.

What runs where: the input of the first dense layer,
``[user_meta[u] | U[u] | I[i] | item_meta[i]]``, is written by the library's row-gather kernel
straight into its column blocks (one pass per source table, no ``torch.cat`` of four gathers) and
its gradient returns to the embedding tables through the vector-RED scatter kernel
(``cmf_gather_rows`` / ``cmf_scatter_add_rows``). The dense layers themselves are plain GEMMs with
a bias and ``leaky_relu``: they go to cuBLAS through ``torch.nn.functional.linear`` (fp32, TF32
off), with autograd for their backward; the losses are the reference's formulas on the score
vectors (``collie_b200.loss``). Training goes through the trainers' unfused loop
(``zero_grad -> calculate_loss -> backward -> optimizer_step``, trainer.py:458-470).
"""
import os
from collections import OrderedDict
from functools import partial
from pathlib import Path

import numpy as np
import torch
from torch.optim.lr_scheduler import ReduceLROnPlateau

from . import _abi
from .model import (HParams, MatrixFactorizationModel, RowOptimizer, ScaledEmbedding,
                    ZeroEmbedding, _GatherConcat, _as_ids, _load_checkpoint)


def _as_float_table(x):
    """DataFrame / ndarray / tensor -> float32 tensor [rows, cols] (reference :138-152)."""
    if x is None:
        return None
    if hasattr(x, 'to_numpy'):
        x = x.to_numpy()
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(x)
    return x.float()


class HybridPretrainedModel(MatrixFactorizationModel):
    """Same constructor as the reference (hybrid_pretrained_matrix_factorization.py:101-126)."""

    _fusable = False        # the trainers take the unfused loop (dense layers through autograd)

    def __init__(self, train=None, val=None, item_metadata=None, user_metadata=None,
                 trained_model=None, item_metadata_layers_dims=None, user_metadata_layers_dims=None,
                 combined_layers_dims=[128, 64, 32], freeze_embeddings=True, dropout_p=0.0,
                 lr=1e-3, lr_scheduler_func=partial(ReduceLROnPlateau, patience=1),
                 weight_decay=0.0, optimizer='adam', loss='hinge', metadata_for_loss=None,
                 metadata_for_loss_weights=None, load_model_path=None, map_location=None):
        self._hybrid_init = None
        if load_model_path is None:
            if trained_model is None:
                raise ValueError('Must provide ``trained_model`` for ``HybridPretrainedModel``.')
            if item_metadata is None and user_metadata is None:
                raise ValueError('Must provide item metadata and/or user_metadata for '
                                 '``HybridPretrainedModel``.')
            item_metadata = _as_float_table(item_metadata)
            user_metadata = _as_float_table(user_metadata)
            self._hybrid_init = dict(
                trained_model=trained_model, item_metadata=item_metadata,
                user_metadata=user_metadata,
                extra=dict(item_metadata_layers_dims=item_metadata_layers_dims,
                           user_metadata_layers_dims=user_metadata_layers_dims,
                           combined_layers_dims=list(combined_layers_dims),
                           freeze_embeddings=freeze_embeddings,
                           item_metadata_num_cols=(None if item_metadata is None
                                                   else int(item_metadata.shape[1])),
                           user_metadata_num_cols=(None if user_metadata is None
                                                   else int(user_metadata.shape[1]))))
        self.item_metadata = item_metadata if load_model_path is None else None
        self.user_metadata = user_metadata if load_model_path is None else None
        super().__init__(train=train, val=val, embedding_dim=None, dropout_p=dropout_p,
                         sparse=False, lr=lr, bias_lr=None, lr_scheduler_func=lr_scheduler_func,
                         weight_decay=weight_decay, optimizer=optimizer, bias_optimizer=None,
                         loss=loss, metadata_for_loss=metadata_for_loss,
                         metadata_for_loss_weights=metadata_for_loss_weights, y_range=None,
                         load_model_path=load_model_path, map_location=map_location)

    # ------------------------------------------------------------------ set-up
    def _setup_model(self):
        h = self.hparams
        if self._hybrid_init is not None and 'combined_layers_dims' not in h:
            h.update(self._hybrid_init['extra'])
            for k in ('embedding_dim', 'sparse', 'bias_lr', 'bias_optimizer', 'y_range'):
                h.pop(k, None)                     # (not hyper-parameters of this model)
        trained = self._hybrid_init['trained_model'] if self._hybrid_init else None
        if h.get('load_model_path') is None and trained is not None:
            import copy
            # the trained model's weights are copied, never modified (reference :253-262)
            self.embeddings = torch.nn.Sequential(copy.deepcopy(trained.user_embeddings),
                                                  copy.deepcopy(trained.item_embeddings))
            self.biases = torch.nn.Sequential(copy.deepcopy(trained.user_biases),
                                              copy.deepcopy(trained.item_biases))
            h.user_num_embeddings = self.embeddings[0].num_embeddings
            h.user_embeddings_dim = self.embeddings[0].embedding_dim
            h.item_num_embeddings = self.embeddings[1].num_embeddings
            h.item_embeddings_dim = self.embeddings[1].embedding_dim
        else:
            self.embeddings = torch.nn.Sequential(
                ScaledEmbedding(h.user_num_embeddings, h.user_embeddings_dim),
                ScaledEmbedding(h.item_num_embeddings, h.item_embeddings_dim))
            self.biases = torch.nn.Sequential(ZeroEmbedding(h.user_num_embeddings, 1),
                                              ZeroEmbedding(h.item_num_embeddings, 1))
        if h.freeze_embeddings:
            self.freeze_embeddings()
        else:
            self.unfreeze_embeddings()
        self.dropout = torch.nn.Dropout(p=h.dropout_p)

        def dense_stack(prefix, dims):
            layers = []
            for i in range(1, len(dims)):
                layer = torch.nn.Linear(dims[i - 1], dims[i])
                torch.nn.init.xavier_normal_(layer.weight)
                self.add_module(f'{prefix}_{i - 1}', layer)
                layers.append(layer)
            return layers

        item_out, user_out = h.item_metadata_num_cols, h.user_metadata_num_cols
        self.item_metadata_layers = self.user_metadata_layers = None
        if h.item_metadata_layers_dims is not None:
            self.item_metadata_layers = dense_stack(
                'item_metadata_layer', [h.item_metadata_num_cols] + list(h.item_metadata_layers_dims))
            item_out = h.item_metadata_layers_dims[-1]
        if h.user_metadata_layers_dims is not None:
            self.user_metadata_layers = dense_stack(
                'user_metadata_layer', [h.user_metadata_num_cols] + list(h.user_metadata_layers_dims))
            user_out = h.user_metadata_layers_dims[-1]
        width = h.user_embeddings_dim + h.item_embeddings_dim + (item_out or 0) + (user_out or 0)
        self.combined_layers = dense_stack('combined_layer',
                                           [width] + list(h.combined_layers_dims) + [1])

    def freeze_embeddings(self):
        """Remove gradient requirement from the embeddings (reference :476-479)."""
        self.embeddings[0].weight.requires_grad = False
        self.embeddings[1].weight.requires_grad = False

    def unfreeze_embeddings(self):
        self.embeddings[0].weight.requires_grad = True
        self.embeddings[1].weight.requires_grad = True

    # the attribute names the base class (losses, similarity helpers, metrics) looks for
    @property
    def user_embeddings(self):
        return self.embeddings[0]

    @property
    def item_embeddings(self):
        return self.embeddings[1]

    @property
    def user_biases(self):
        return self.biases[0]

    @property
    def item_biases(self):
        return self.biases[1]

    def _move_any_external_data_to_device(self):
        super()._move_any_external_data_to_device()
        dev = self.device
        if self.item_metadata is not None:
            self.item_metadata = self.item_metadata.to(dev).contiguous()
        if self.user_metadata is not None:
            self.user_metadata = self.user_metadata.to(dev).contiguous()

    # ------------------------------------------------------------------ optimizers
    def configure_optimizers(self):
        """One optimizer over every trainable parameter (the reference passes no bias optimizer:
        base_pipeline.py:372-373)."""
        params = [p for p in self.parameters() if p.requires_grad]
        lr = self.hparams.lr
        if callable(self.optimizer):
            try:
                optimizer = self.optimizer(params, lr=lr, weight_decay=self.hparams.weight_decay)
            except TypeError:
                optimizer = self.optimizer(params, lr=lr)
        elif self.optimizer in _abi.OPT_IDS:
            rule = 'adam' if self.optimizer == 'sparse_adam' else self.optimizer
            optimizer = RowOptimizer([{'params': params, 'lr': lr}], rule, lr=lr,
                                     weight_decay=self.hparams.weight_decay)
        else:
            raise ValueError('{} is not a valid optimizer!'.format(self.optimizer))
        if self.lr_scheduler_func is not None:
            monitor = 'val_loss_epoch' if self.val_loader is not None else 'train_loss_epoch'
            return ([optimizer], [{'scheduler': self.lr_scheduler_func(optimizer),
                                   'monitor': monitor}])
        return optimizer

    # ------------------------------------------------------------------ forward
    def _metadata_output(self, kind, table, ids):
        """Metadata rows of ``ids`` through the metadata-only layers, if any (reference :327-361)."""
        layers = getattr(self, f'{kind}_metadata_layers')
        out = _GatherConcat.apply([ids], table)
        for layer in layers:
            out = self.dropout(torch.nn.functional.leaky_relu(layer(out)))
        return out

    def forward(self, users, items):
        U = _abi.require_cuda(self.embeddings[0].weight, 'embeddings.0.weight')
        dev = U.device
        users, items = _as_ids(users, dev), _as_ids(items, dev)
        if users.shape != items.shape or users.dim() != 1:
            raise ValueError('``users`` and ``items`` must be 1-d tensors of equal length')
        if self.item_metadata is None and self.user_metadata is None:
            raise _abi.CmfError('the model has no metadata (a loaded model needs its '
                                '``item_metadata.pkl`` / ``user_metadata.pkl``)')
        if (self.item_metadata is not None and not self.item_metadata.is_cuda) or \
                (self.user_metadata is not None and not self.user_metadata.is_cuda):
            self._move_any_external_data_to_device()
        I = self.embeddings[1].weight
        with torch.autocast('cuda', enabled=False):
            if self.item_metadata_layers is None and self.user_metadata_layers is None:
                # the whole input of the first dense layer in one buffer, one pass per source
                ids, tables = [], []
                if self.user_metadata is not None:
                    ids.append(users)
                    tables.append(self.user_metadata)
                ids += [users, items]
                tables += [U, I]
                if self.item_metadata is not None:
                    ids.append(items)
                    tables.append(self.item_metadata)
                x = _GatherConcat.apply(ids, *tables)
            else:
                parts = []
                if self.user_metadata is not None:
                    parts.append(self._metadata_output('user', self.user_metadata, users)
                                 if self.user_metadata_layers is not None
                                 else _GatherConcat.apply([users], self.user_metadata))
                parts.append(_GatherConcat.apply([users, items], U, I))
                if self.item_metadata is not None:
                    parts.append(self._metadata_output('item', self.item_metadata, items)
                                 if self.item_metadata_layers is not None
                                 else _GatherConcat.apply([items], self.item_metadata))
                x = torch.cat(parts, dim=1)
            # dense layers: plain GEMMs -> cuBLAS (fp32; TF32 stays off unless the user turns it on)
            for layer in self.combined_layers[:-1]:
                x = self.dropout(torch.nn.functional.leaky_relu(layer(x)))
            bias = _GatherConcat.apply([users, items], self.biases[0].weight,
                                       self.biases[1].weight)
            preds = self.combined_layers[-1](x) + bias[:, :1] + bias[:, 1:]
        return preds.squeeze()

    def fused_step(self, *args, **kwargs):
        raise _abi.CmfError('HybridPretrainedModel trains through the trainers\' unfused loop '
                            '(calculate_loss -> backward -> optimizer step); it has no fused step')

    def calculate_loss(self, batch, differentiable=True):
        return super().calculate_loss(batch, differentiable=True)

    def _get_item_embeddings(self):
        return self.embeddings[1].weight.data

    def _get_user_embeddings(self):
        return self.embeddings[0].weight.data

    # ------------------------------------------------------------------ checkpoints
    def _load_model_init_helper(self, load_model_path, map_location):
        """A DIRECTORY with ``model.pth`` (+ ``item_metadata.pkl`` / ``user_metadata.pkl``),
        reference :203-225."""
        import joblib
        import warnings
        loaded = _load_checkpoint(os.path.join(str(load_model_path), 'model.pth'), map_location)
        for key, value in loaded['hparams'].items():
            self.hparams[key] = value
        self.hparams['load_model_path'] = load_model_path
        self.hparams['map_location'] = map_location
        self.loss_function = None
        self._loss_name = None
        self._setup_model()
        self.load_state_dict(loaded['state_dict'])
        for kind in ('item', 'user'):
            try:
                setattr(self, f'{kind}_metadata', _as_float_table(
                    joblib.load(os.path.join(str(load_model_path), f'{kind}_metadata.pkl'))))
            except FileNotFoundError:
                if self.hparams.get(f'{kind}_metadata_layers_dims') is not None:
                    warnings.warn(f'``{kind}_metadata.pkl`` not found')
        self.eval()

    def save_model(self, path='model', overwrite=False):
        """Directory checkpoint like the reference's (:486-534): ``model.pth`` with
        ``{'state_dict', 'hparams'}`` and the metadata tables as joblib pickles."""
        import joblib
        path = str(path)
        if os.path.exists(path) and os.listdir(path) and overwrite is False:
            raise ValueError(f'Data exists in ``path`` at {path} and ``overwrite`` is False.')
        Path(path).mkdir(parents=True, exist_ok=True)
        if self.item_metadata is not None:
            joblib.dump(self.item_metadata.cpu(), os.path.join(path, 'item_metadata.pkl'))
        if self.user_metadata is not None:
            joblib.dump(self.user_metadata.cpu(), os.path.join(path, 'user_metadata.pkl'))
        state = OrderedDict((k, v) for k, v in self.state_dict().items()
                            if '_trained_model' not in k)
        hparams = {k: v for k, v in self.hparams.items()}
        torch.save({'state_dict': state, 'hparams': hparams}, os.path.join(path, 'model.pth'))

    def load_from_hybrid_model(self, hybrid_model):
        """Copy hyper-parameters and weights of another instance (reference :536-556)."""
        for key, value in hybrid_model.hparams.items():
            self.hparams[key] = value
        self._hybrid_init = None
        self._setup_model()
        self.load_state_dict(hybrid_model.state_dict())
        self.eval()
