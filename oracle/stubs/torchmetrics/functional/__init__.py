import numpy as np
import torch
from sklearn.metrics import roc_auc_score


def auroc(preds, target, task=None, **kwargs):
    if 'pos_label' in kwargs:
        # old-API call; the reference then falls through to ``task='binary'`` (metrics.py:276-279)
        raise TypeError('pos_label is not accepted')
    p = preds.detach().cpu().numpy()
    t = target.detach().cpu().numpy()
    if t.min() == t.max():
        return torch.tensor(0.0)
    return torch.tensor(float(roc_auc_score(t, p)), dtype=torch.float32)


def mean_squared_error(preds, target):
    return ((preds.double() - target.double()) ** 2).mean()


def mean_absolute_error(preds, target):
    return (preds.double() - target.double()).abs().mean()
