"""Loss callables with the reference's signatures (collie/loss/*.py), on torch tensors.

The trainers do not call these for the named losses -- ``'bpr'``, ``'hinge'``, ``'adaptive'``,
``'warp'`` ... are computed inside the fused training kernel, gradient included. They exist for
the unfused route (``MatrixFactorizationModel.calculate_loss`` with autograd: a callable loss, a
foreign optimizer) and for user code that imports them like from the reference. Every function
takes positive scores ``[B]`` and negative scores ``[B]`` (or ``[N, B]`` for the adaptive / WARP
forms) and returns ``(sum l + sum l^2) / B``.

Mirrors: bpr_loss bpr.py:8-89, adaptive_bpr_loss :92-167, hinge_loss hinge.py:8-85,
adaptive_hinge_loss :88-171, warp_loss warp.py:8-133, ideal_difference_from_metadata
metadata_utils.py:6-71.
"""
import torch


def ideal_difference_from_metadata(positive_items, negative_items, metadata, metadata_weights):
    """``1 - sum_k w_k [meta_k[pos] == meta_k[neg]]`` with the shape of ``negative_items``."""
    weights = metadata_weights or {}
    total = sum(weights.values())
    if total > 1:
        raise ValueError(f'sum of metadata weights was {total}, must be <=1')
    credit = torch.zeros(negative_items.shape, dtype=torch.float32,
                         device=negative_items.device)
    for key, labels in (metadata or {}).items():
        labels = torch.as_tensor(labels).reshape(-1).to(negative_items.device)
        same = labels[positive_items.long()] == labels[negative_items.long()]
        credit = credit + same.float() * float(weights[key])
    return 1.0 - credit


def _ideal(positive_items, negative_items, metadata, metadata_weights):
    if metadata is None or len(metadata) == 0 or positive_items is None or negative_items is None:
        return 1.0
    return ideal_difference_from_metadata(positive_items, negative_items, metadata,
                                          metadata_weights)


def _reduce(per_pair, batch):
    return (per_pair.sum() + per_pair.pow(2).sum()) / batch


def bpr_loss(positive_scores, negative_scores, num_items=None, positive_items=None,
             negative_items=None, metadata=None, metadata_weights=None):
    ideal = _ideal(positive_items, negative_items, metadata, metadata_weights)
    return _reduce(ideal - torch.sigmoid(positive_scores - negative_scores), len(positive_scores))


def hinge_loss(positive_scores, negative_scores, num_items=None, positive_items=None,
               negative_items=None, metadata=None, metadata_weights=None):
    ideal = _ideal(positive_items, negative_items, metadata, metadata_weights)
    return _reduce(torch.clamp(ideal - (positive_scores - negative_scores), min=0.0),
                   len(positive_scores))


def _hardest(many_negative_scores, negative_items):
    """The best-scored negative of every interaction (first one on ties) and its item id."""
    top, which = torch.max(many_negative_scores.view(-1, many_negative_scores.shape[-1]), dim=0)
    chosen = None
    if negative_items is not None:
        chosen = negative_items.view(-1, negative_items.shape[-1]).gather(
            0, which.view(1, -1)).view(-1)
    return top, chosen


def adaptive_bpr_loss(positive_scores, many_negative_scores, num_items=None, positive_items=None,
                      negative_items=None, metadata=None, metadata_weights=None):
    top, chosen = _hardest(many_negative_scores, negative_items)
    return bpr_loss(positive_scores, top, num_items, positive_items, chosen, metadata,
                    metadata_weights)


def adaptive_hinge_loss(positive_scores, many_negative_scores, num_items=None,
                        positive_items=None, negative_items=None, metadata=None,
                        metadata_weights=None):
    top, chosen = _hardest(many_negative_scores, negative_items)
    return hinge_loss(positive_scores, top, num_items, positive_items, chosen, metadata,
                      metadata_weights)


def warp_loss(positive_scores, many_negative_scores, num_items, positive_items=None,
              negative_items=None, metadata=None, metadata_weights=None):
    """WARP: the first negative (in sampling order) that violates the margin, weighted by
    ``log(num_items / tries)``; interactions without a violation contribute nothing."""
    n_neg, batch = many_negative_scores.shape
    ideal = 1.0
    if metadata is not None and len(metadata) > 0 and negative_items is not None:
        ideal = ideal_difference_from_metadata(
            positive_items.view(1, -1).expand(n_neg, batch), negative_items, metadata,
            metadata_weights)
    margin = ideal - positive_scores.view(1, -1) + many_negative_scores        # [N, B]
    violated = margin > 0
    any_violation = violated.any(dim=0)
    first = torch.argmax(violated.int(), dim=0)            # first True (0 when there is none)
    tries = (first + 1).float()
    weight = torch.log(float(num_items) / tries)
    picked = margin.gather(0, first.view(1, -1)).view(-1)
    per_pair = torch.where(any_violation, weight * picked, torch.zeros_like(picked))
    return _reduce(per_pair, batch)


NAMED = {'bpr': bpr_loss, 'hinge': hinge_loss, 'adaptive_bpr': adaptive_bpr_loss,
         'adaptive_hinge': adaptive_hinge_loss, 'warp': warp_loss}
