"""CPU oracle for the Collie matrix-factorization hot path.

THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE. Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it. Nothing under
``collie_b200/`` imports, calls or links anything under ``oracle/``.

What it is: a restatement, in plain torch-CPU fp32 / numpy, of what the reference computes on the
path ``MatrixFactorizationModel`` training step + ``evaluate_in_batches``. Paths are relative to
``/root/reference`` (ShopRunner/collie v1.3.1):

* forward                   collie/model/matrix_factorization.py:143-157
* batch -> loss             collie/model/base/base_pipeline.py:603-652
* losses                    collie/loss/bpr.py:75-89,153-167  hinge.py:71-85,157-171
                            warp.py:73-133,136-154  metadata_utils.py:60-71
* explicit losses           torch.nn.MSELoss / L1Loss selected at base_pipeline.py:287-292
* optimizers                torch.optim.{Adam,SparseAdam,SGD,Adagrad} built at base_pipeline.py:418-433
* negative sampling         collie/interactions/datasets.py:390-445 (contract; the stream is ours)
* metrics                   collie/metrics.py:167-198 (mapk) :224-240 (mrr) :266-282 (auc)
                            :367-386 (evaluate_in_batches)

The arithmetic of the reference lives in a third-party dependency that is not under
``/root/reference``: ``torch`` (pinned ``torch==1.13.1``, requirements.txt:103; this image has
2.11.0) and ``torchmetrics==0.11.4`` ``functional.auroc`` (requirements.txt:108; absent; the
reference's own test pins ``auc`` to ``sklearn.metrics.roc_auc_score``,
tests/test_metrics.py:140-154). The optimizer formulas below restate the published torch
algorithms and are checked against ``torch.optim`` in ``tests/test_oracle.py``.

Parity pinning: every function here is checked (``tests/test_oracle.py``) against the reference's
own known-answer tests (tests/test_losses.py:80-216, tests/test_metrics.py:101-154,230-256, values
re-stated in ``tests/golden/reference_known_answers.json``) and against outputs of the unmodified
reference run in the build container (``tests/golden/*.npz``, produced by
``oracle/make_golden.py``). Negative-sampling *streams* and ``torch.topk`` tie order are "parity
unpinned" in the reference itself (no test pins them); the canonical rules used here are: the
counter-based Philox stream specified in ``philox_negatives`` and "score descending, item id
ascending".
"""
from typing import Dict, Optional, Sequence, Tuple

import numpy as np
import torch

# ------------------------------------------------------------------------------------------------
# forward
# ------------------------------------------------------------------------------------------------


def mf_forward(U: torch.Tensor, I: torch.Tensor, bu: torch.Tensor, bi: torch.Tensor,
               users: torch.Tensor, items: torch.Tensor,
               y_range: Optional[Tuple[float, float]] = None) -> torch.Tensor:
    """score = <U[u], I[i]> + bu[u] + bi[i]; optional sigmoid range squashing.

    Reference: collie/model/matrix_factorization.py:143-157 (dropout_p = 0).
    ``bu`` / ``bi`` are ``[rows, 1]`` like the reference's ``ZeroEmbedding`` weights.
    """
    users = users.long()
    items = items.long()
    s = (U[users] * I[items]).sum(dim=1) + bu[users].squeeze(1) + bi[items].squeeze(1)
    if y_range is not None:
        s = torch.sigmoid(s) * (y_range[1] - y_range[0]) + y_range[0]
    return s


# ------------------------------------------------------------------------------------------------
# losses (implicit). pos: [B], neg: [N, B] (or [B] for the one-negative forms)
# ------------------------------------------------------------------------------------------------


def embedding_similarity(E: torch.Tensor, id: int) -> torch.Tensor:
    """BasePipeline._calculate_embedding_similarity (collie/model/base/base_pipeline.py:849-859):
    rows divided by their L2 norm, seed row times the transposed table (before the host sort)."""
    En = E / E.norm(dim=1)[:, None]
    return torch.matmul(En[[id], :], En.transpose(1, 0)).squeeze(0)


def ideal_difference(pos_items: torch.Tensor, neg_items: torch.Tensor,
                     metadata: Optional[Dict[str, torch.Tensor]],
                     weights: Optional[Dict[str, float]]):
    """1 - sum_k w_k [meta_k[pos] == meta_k[neg]]   (collie/loss/metadata_utils.py:60-71)."""
    if metadata is None or len(metadata) == 0:
        return 1.0
    total = sum(weights.values())
    if total > 1:
        raise ValueError(f'sum of metadata weights was {total}, must be <=1')
    frac = torch.zeros(pos_items.shape, dtype=torch.float32)
    for key, labels in metadata.items():
        labels = labels.reshape(-1)
        frac = frac + (labels[pos_items.long()] == labels[neg_items.long()]).int() * weights[key]
    return 1.0 - frac


def _mean_l_plus_l2(l: torch.Tensor, batch: int) -> torch.Tensor:
    # every implicit loss in the reference ends with (sum l + sum l^2) / len(positive_scores)
    return (l.sum() + (l * l).sum()) / batch


def bpr(pos, neg, pos_items=None, neg_items=None, metadata=None, weights=None):
    """collie/loss/bpr.py:75-89."""
    ideal = ideal_difference(pos_items, neg_items, metadata, weights)
    return _mean_l_plus_l2(ideal - torch.sigmoid(pos - neg), len(pos))


def hinge(pos, neg, pos_items=None, neg_items=None, metadata=None, weights=None):
    """collie/loss/hinge.py:71-85."""
    ideal = ideal_difference(pos_items, neg_items, metadata, weights)
    return _mean_l_plus_l2(torch.clamp(ideal - (pos - neg), min=0), len(pos))


def _hardest_negative(neg, pos_items, neg_items):
    """max over the N axis, first index on ties (bpr.py:153-158, hinge.py:157-162)."""
    top, which = torch.max(neg, 0)
    chosen_items = None
    if neg_items is not None and pos_items is not None:
        chosen_items = neg_items[which, torch.arange(neg.shape[1])]
    return top, which, chosen_items


def adaptive_bpr(pos, neg, pos_items=None, neg_items=None, metadata=None, weights=None):
    """collie/loss/bpr.py:153-167."""
    top, _, chosen = _hardest_negative(neg, pos_items, neg_items)
    return bpr(pos, top, pos_items, chosen, metadata, weights)


def adaptive_hinge(pos, neg, pos_items=None, neg_items=None, metadata=None, weights=None):
    """collie/loss/hinge.py:157-171."""
    top, _, chosen = _hardest_negative(neg, pos_items, neg_items)
    return hinge(pos, top, pos_items, chosen, metadata, weights)


def warp_first_violation(h: torch.Tensor) -> torch.Tensor:
    """Index of the first column with h > 0 (strict), or N when there is none.

    Reference: collie/loss/warp.py:136-154 with the sentinel column of ones of :101-105.
    """
    n = h.shape[1]
    bad = h > 0
    idx = torch.arange(n).expand_as(h)
    first = torch.where(bad, idx, torch.full_like(idx, n)).min(dim=1).values
    return first


def warp(pos, neg, num_items, pos_items=None, neg_items=None, metadata=None, weights=None):
    """collie/loss/warp.py:73-133.  neg: [N, B]."""
    n, b = neg.shape
    if metadata is not None and len(metadata) > 0:
        ideal = ideal_difference(pos_items.repeat([n, 1]), neg_items, metadata, weights).t()
    else:
        ideal = 1.0
    h = ideal - pos.view(b, 1) + neg.t()                       # [B, N]
    t = warp_first_violation(h)                                # [B] in 0..N
    tries = (t + 1).float()
    w = torch.log(num_items / tries)
    h_ext = torch.cat([h, torch.ones(b, 1)], dim=1)
    picked = h_ext.gather(1, t.view(b, 1)).squeeze(1)
    l = w * picked * (tries <= n).float()
    return _mean_l_plus_l2(l, b)


IMPLICIT_LOSSES = {
    'bpr': bpr, 'hinge': hinge, 'adaptive_bpr': adaptive_bpr, 'adaptive_hinge': adaptive_hinge,
    'warp': warp,
}


def resolve_loss_name(loss: str, num_negative_samples: int) -> str:
    """String -> concrete loss, as collie/model/base/base_pipeline.py:277-340 does."""
    if loss in ('mse', 'mae'):
        return loss
    if 'warp' in loss:
        if num_negative_samples <= 1:
            raise ValueError('Cannot use WARP loss with a single negative sample!')
        return 'warp'
    if 'bpr' in loss:
        return 'adaptive_bpr' if num_negative_samples > 1 else 'bpr'
    if 'hinge' in loss or 'adaptive' in loss:
        return 'adaptive_hinge' if num_negative_samples > 1 else 'hinge'
    raise ValueError(f'{loss} is not a valid loss')


def implicit_batch_loss(tables, users, pos_items, neg_items_bn, loss_name, num_items,
                        metadata=None, weights=None):
    """``calculate_loss`` for ``((users, pos), negs[B, N])``  (base_pipeline.py:603-634)."""
    U, I, bu, bi = tables
    users = users.long()
    pos_items = pos_items.long()
    neg_nb = neg_items_bn.t().long()                            # [N, B]
    pos = mf_forward(U, I, bu, bi, users, pos_items)
    neg = mf_forward(U, I, bu, bi, users.repeat(neg_nb.shape[0]), neg_nb.flatten())
    neg = neg.view(neg_nb.shape[0], len(users))
    if loss_name in ('bpr', 'hinge'):
        # single-negative losses receive the [1, B] matrix and broadcast (as the reference does)
        fn = IMPLICIT_LOSSES[loss_name]
        return fn(pos, neg, pos_items, neg_nb, metadata, weights)
    if loss_name == 'warp':
        return warp(pos, neg, num_items, pos_items, neg_nb, metadata, weights)
    return IMPLICIT_LOSSES[loss_name](pos, neg, pos_items, neg_nb, metadata, weights)


def explicit_batch_loss(tables, users, items, ratings, loss_name, y_range=None):
    """``calculate_loss`` for ``(users, items, ratings)``  (base_pipeline.py:635-650)."""
    U, I, bu, bi = tables
    p = mf_forward(U, I, bu, bi, users, items, y_range)
    r = ratings.float()
    if loss_name == 'mse':
        return ((p - r) ** 2).mean()
    if loss_name == 'mae':
        return (p - r).abs().mean()
    raise ValueError(loss_name)


# ------------------------------------------------------------------------------------------------
# optimizers: restated published torch algorithms (defaults beta=(.9,.999), eps=1e-8)
# ------------------------------------------------------------------------------------------------


def sparse_adam_rows(p, m, v, rows, g_rows, step, lr, b1=0.9, b2=0.999, eps=1e-8):
    """torch.optim.SparseAdam on the coalesced gradient: only ``rows`` move (even where g == 0)."""
    m[rows] += (1 - b1) * (g_rows - m[rows])
    v[rows] += (1 - b2) * (g_rows * g_rows - v[rows])
    step_size = lr * np.sqrt(1 - b2 ** step) / (1 - b1 ** step)
    p[rows] -= step_size * m[rows] / (v[rows].sqrt() + eps)


def dense_adam(p, m, v, g, step, lr, b1=0.9, b2=0.999, eps=1e-8, weight_decay=0.0):
    """torch.optim.Adam (single-tensor path): every row moves, eps outside the bias-corrected sqrt."""
    if weight_decay != 0:
        g = g + weight_decay * p
    m.mul_(b1).add_(g, alpha=1 - b1)
    v.mul_(b2).addcmul_(g, g, value=1 - b2)
    denom = v.sqrt() / np.sqrt(1 - b2 ** step) + eps
    p.addcdiv_(m, denom, value=-lr / (1 - b1 ** step))


def sgd(p, g, lr, weight_decay=0.0):
    """torch.optim.SGD without momentum."""
    if weight_decay != 0:
        g = g + weight_decay * p
    p.add_(g, alpha=-lr)


# ------------------------------------------------------------------------------------------------
# negative sampling: the reference's contract + this project's counter-based stream
# ------------------------------------------------------------------------------------------------

_M0, _M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_W0, _W1 = 0x9E3779B9, 0xBB67AE85


def philox4x32_10(counter: Sequence[int], key: Sequence[int]) -> Tuple[int, int, int, int]:
    """Philox4x32-10 (Salmon et al., SC'11). Scalar python-int implementation."""
    c0, c1, c2, c3 = [int(x) & 0xFFFFFFFF for x in counter]
    k0, k1 = [int(x) & 0xFFFFFFFF for x in key]
    for _ in range(10):
        p0 = 0xD2511F53 * c0
        p1 = 0xCD9E8D57 * c2
        hi0, lo0 = p0 >> 32, p0 & 0xFFFFFFFF
        hi1, lo1 = p1 >> 32, p1 & 0xFFFFFFFF
        c0, c1, c2, c3 = (hi1 ^ c1 ^ k0) & 0xFFFFFFFF, lo1, (hi0 ^ c3 ^ k1) & 0xFFFFFFFF, lo0
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return c0, c1, c2, c3


def _draw(seed: int, epoch: int, gidx: int, d: int, num_items: int) -> int:
    """d-th uniform item id of interaction ``gidx``: word d%4 of Philox block d//4.

    key = (seed_lo, seed_hi); counter = (gidx_lo, gidx_hi, epoch, d // 4);
    id = (word * num_items) >> 32.
    """
    words = philox4x32_10((gidx & 0xFFFFFFFF, gidx >> 32, epoch, d // 4),
                          (seed & 0xFFFFFFFF, seed >> 32))
    return (words[d % 4] * num_items) >> 32


def philox_negatives(seed: int, epoch: int, global_idx: np.ndarray, users: np.ndarray, n_neg: int,
                     num_items: int, indptr: Optional[np.ndarray] = None,
                     indices: Optional[np.ndarray] = None, max_tries: int = 200) -> np.ndarray:
    """Negatives ``[B, N]`` for interactions with dataset positions ``global_idx``.

    Approximate mode (``indptr is None``): N iid uniform ids  (datasets.py:433-443).
    Exact mode: the reference's rejection loop (datasets.py:399-426) restated over the draw stream:
    a candidate is rejected if the user has interacted with it or it was already taken; rejections
    share a per-row budget of ``max_tries``; once the budget is spent the remaining slots are filled
    with unconditioned draws and the last (rejected) candidate is kept as the row's final id.
    """
    out = np.zeros((len(global_idx), n_neg), dtype=np.int64)
    for r, (g, u) in enumerate(zip(global_idx.tolist(), np.asarray(users).tolist())):
        d = 0
        if indptr is None:
            for n in range(n_neg):
                out[r, n] = _draw(seed, epoch, g, n, num_items)
            continue
        seen = set(indices[indptr[u]:indptr[u + 1]].tolist())
        taken = []
        checked = 0
        while len(taken) < n_neg:
            c = _draw(seed, epoch, g, d, num_items)
            d += 1
            while c in seen or c in taken:
                if checked >= max_tries:
                    left = n_neg - len(taken) - 1
                    for _ in range(left):
                        taken.append(_draw(seed, epoch, g, d, num_items))
                        d += 1
                    break
                c = _draw(seed, epoch, g, d, num_items)
                d += 1
                checked += 1
            taken.append(c)
        out[r] = taken[:n_neg]
    return out


# ------------------------------------------------------------------------------------------------
# metrics (canonical order: score descending, item id ascending)
# ------------------------------------------------------------------------------------------------


def canonical_order(scores: np.ndarray) -> np.ndarray:
    """Item ids per row sorted by (score desc, id asc) -- stable sort of the negated scores."""
    return np.argsort(-scores, axis=1, kind='stable')


def topk_ids(scores: np.ndarray, k: int) -> np.ndarray:
    if k > scores.shape[1]:
        raise ValueError(f'Ensure ``k`` ({k}) is less than the number of items ({scores.shape[1]})')
    return canonical_order(scores)[:, :k]


def mapk_per_user(targets_csr, user_ids, scores: np.ndarray, k: int) -> np.ndarray:
    """AP@k per user  (collie/metrics.py:167-198; NaN -> 0 when the user has no targets)."""
    ids = topk_ids(scores, k)
    out = np.zeros(len(user_ids), dtype=np.float64)
    for r, u in enumerate(user_ids):
        row = targets_csr[u]
        pos = set(row.indices[row.data > 0].tolist())
        nnz = targets_csr[u].getnnz()
        hits = 0
        acc = 0.0
        for j, it in enumerate(ids[r]):
            if int(it) in pos:
                hits += 1
                acc += hits / (j + 1)
        denom = min(k, nnz)
        out[r] = acc / denom if denom > 0 else 0.0
    return out


def mrr_per_user(targets_csr, user_ids, scores: np.ndarray) -> np.ndarray:
    """1 / rank of the best-ranked target, 0 without targets  (collie/metrics.py:224-240)."""
    order = canonical_order(scores)
    out = np.zeros(len(user_ids), dtype=np.float64)
    for r, u in enumerate(user_ids):
        row = targets_csr[u]
        pos = set(row.indices[row.data > 0].tolist())
        for j, it in enumerate(order[r]):
            if int(it) in pos:
                out[r] = 1.0 / (j + 1)
                break
    return out


def auc_per_user(targets_csr, user_ids, scores: np.ndarray) -> np.ndarray:
    """Mann-Whitney AUROC of sigmoid(score) in fp32 vs. binary targets  (collie/metrics.py:266-282).

    Degenerate rows (no positives or no negatives) give 0, as torchmetrics 0.11 does.
    """
    out = np.zeros(len(user_ids), dtype=np.float64)
    sig = torch.sigmoid(torch.from_numpy(np.ascontiguousarray(scores, dtype=np.float32))).numpy()
    for r, u in enumerate(user_ids):
        y = np.zeros(scores.shape[1], dtype=bool)
        row = targets_csr[u]
        y[row.indices[row.data > 0]] = True
        p, n = sig[r][y], sig[r][~y]
        if len(p) == 0 or len(n) == 0:
            continue
        n_sorted = np.sort(n)
        below = np.searchsorted(n_sorted, p, side='left')
        ties = np.searchsorted(n_sorted, p, side='right') - below
        out[r] = (below.sum() + 0.5 * ties.sum()) / (len(p) * len(n))
    return out


def evaluate_in_batches(metric_names, test_coo, tables, k=10, batch_size=20):
    """collie/metrics.py:367-395: unique test users, slices of ``batch_size``, user-weighted mean."""
    U, I, bu, bi = [t.detach().float() for t in tables]
    users = np.unique(test_coo.row)
    targets = test_coo.tocsr()
    n_items = test_coo.shape[1]
    batch_size = min(batch_size, len(users))
    acc = [0.0] * len(metric_names)
    for s in range(0, len(users), batch_size):
        ub = users[s:s + batch_size]
        uu = torch.from_numpy(np.repeat(ub, n_items)).long()
        ii = torch.arange(n_items).repeat(len(ub))
        scores = mf_forward(U, I, bu, bi, uu, ii).view(len(ub), n_items).numpy()
        for m, name in enumerate(metric_names):
            if name == 'mapk':
                val = mapk_per_user(targets, ub, scores, k).mean()
            elif name == 'mrr':
                val = mrr_per_user(targets, ub, scores).mean()
            elif name == 'auc':
                val = auc_per_user(targets, ub, scores).mean()
            else:
                raise ValueError(name)
            acc[m] += val * len(ub)
    out = [a / len(users) for a in acc]
    return out[0] if len(out) == 1 else out


# ------------------------------------------------------------------------------------------------
# one training step, the way CollieMinimalTrainer drives it (trainer.py:458-470)
# ------------------------------------------------------------------------------------------------


class OracleMF(torch.nn.Module):
    """The four parameter tables of collie/model/matrix_factorization.py:106-117."""

    def __init__(self, num_users, num_items, dim, sparse=False, seed=0):
        super().__init__()
        g = torch.Generator().manual_seed(seed)
        self.user_biases = torch.nn.Embedding(num_users, 1, sparse=sparse)
        self.item_biases = torch.nn.Embedding(num_items, 1, sparse=sparse)
        self.user_embeddings = torch.nn.Embedding(num_users, dim, sparse=sparse)
        self.item_embeddings = torch.nn.Embedding(num_items, dim, sparse=sparse)
        with torch.no_grad():
            # ScaledEmbedding / ZeroEmbedding init (collie/model/base/layers.py:6-17)
            self.user_biases.weight.zero_()
            self.item_biases.weight.zero_()
            self.user_embeddings.weight.normal_(0, 1.0 / (dim * 2.5), generator=g)
            self.item_embeddings.weight.normal_(0, 1.0 / (dim * 2.5), generator=g)

    def tables(self):
        return (self.user_embeddings.weight, self.item_embeddings.weight,
                self.user_biases.weight, self.item_biases.weight)

    def forward(self, users, items, y_range=None):
        u, i = users.long(), items.long()
        s = ((self.user_embeddings(u) * self.item_embeddings(i)).sum(dim=1)
             + self.user_biases(u).squeeze(1) + self.item_biases(i).squeeze(1))
        if y_range is not None:
            s = torch.sigmoid(s) * (y_range[1] - y_range[0]) + y_range[0]
        return s

    def implicit_loss(self, users, pos, negs_bn, loss_name, metadata=None, weights=None):
        neg_nb = negs_bn.t().long()
        p = self(users, pos)
        n = self(users.long().repeat(neg_nb.shape[0]), neg_nb.flatten()).view(neg_nb.shape[0], -1)
        num_items = self.item_embeddings.num_embeddings
        if loss_name == 'warp':
            return warp(p, n, num_items, pos.long(), neg_nb, metadata, weights)
        return IMPLICIT_LOSSES[loss_name](p, n, pos.long(), neg_nb, metadata, weights)

    def explicit_loss(self, users, items, ratings, loss_name, y_range=None):
        p = self(users, items, y_range)
        r = ratings.float()
        return ((p - r) ** 2).mean() if loss_name == 'mse' else (p - r).abs().mean()


def make_optimizers(model: OracleMF, optimizer='adam', bias_optimizer='sgd', lr=1e-3, bias_lr=1e-2,
                    weight_decay=0.0):
    """Parameter grouping of base_pipeline.py:342-479: names containing 'bias' -> bias optimizer."""
    def build(name, params, rate):
        if name == 'sgd':
            return torch.optim.SGD(params, lr=rate, weight_decay=weight_decay)
        if name == 'adam':
            return torch.optim.Adam(params, lr=rate, weight_decay=weight_decay)
        if name == 'adagrad':
            return torch.optim.Adagrad(params, lr=rate, weight_decay=weight_decay)
        if name == 'sparse_adam':
            return torch.optim.SparseAdam(params, lr=rate)
        raise ValueError(f'{name} is not a valid optimizer!')

    named = list(model.named_parameters())
    if bias_optimizer is None:
        return [build(optimizer, [p for _, p in named], lr)]
    if bias_optimizer == 'infer':
        bias_optimizer = optimizer
    if bias_lr == 'infer':
        bias_lr = lr
    main = build(optimizer, [p for n, p in named if 'bias' not in n], lr)
    bias = build(bias_optimizer, [p for n, p in named if 'bias' in n], bias_lr)
    return [main, bias]


def train_step(model: OracleMF, optimizers, loss_closure) -> float:
    """zero_grad -> loss -> backward -> step each optimizer (trainer.py:459-470)."""
    for opt in optimizers:
        opt.zero_grad()
    loss = loss_closure()
    loss.backward()
    for opt in optimizers:
        opt.step()
    return float(loss.detach())


# ------------------------------------------------------------------------------------------------
# HybridPretrainedModel
# ------------------------------------------------------------------------------------------------


def hybrid_forward(state: Dict[str, torch.Tensor], item_metadata: torch.Tensor,
                   users: torch.Tensor, items: torch.Tensor,
                   user_metadata: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Scores of ``HybridPretrainedModel`` in eval mode from its ``state_dict`` (reference names):
    ``[user_meta_out | U[u] | I[i] | item_meta_out]`` -> ``combined_layer_*`` with ``leaky_relu``
    between them -> + both biases (collie/model/hybrid_pretrained_matrix_factorization.py
    :327-361 metadata layers, :363-399 prediction, :401-464 forward). Pinned by
    tests/golden/hybrid_*.npz, which the unmodified reference produced."""
    users, items = users.long(), items.long()

    def stack(prefix, x, last_linear):
        i = 0
        while f'{prefix}_{i}.weight' in state:
            x = torch.nn.functional.linear(x, state[f'{prefix}_{i}.weight'],
                                           state[f'{prefix}_{i}.bias'])
            if not (last_linear and f'{prefix}_{i + 1}.weight' not in state):
                x = torch.nn.functional.leaky_relu(x)
            i += 1
        return x
    parts = []
    if user_metadata is not None:
        parts.append(stack('user_metadata_layer', user_metadata[users], False))
    parts += [state['embeddings.0.weight'][users], state['embeddings.1.weight'][items]]
    if item_metadata is not None:
        parts.append(stack('item_metadata_layer', item_metadata[items], False))
    out = stack('combined_layer', torch.cat(parts, dim=1), True)
    return (out + state['biases.0.weight'][users] + state['biases.1.weight'][items]).squeeze()


# ------------------------------------------------------------------------------------------------
# dropout (collie/model/matrix_factorization.py:118, :147)
# ------------------------------------------------------------------------------------------------


def dropout_mask(seed: int, step: int, epoch: int, gidx: int, slot: int, side: int, dim: int,
                 p: float) -> np.ndarray:
    """The mask ``nn.Dropout(p)`` would apply to one embedding row of one scored pair, drawn from
    this project's counter-based stream (the reference draws from torch's global generator, which
    no test pins): interaction at dataset position ``gidx``, ``slot`` 0 = the positive / 1 + n =
    negative n, ``side`` 0 = user row / 1 = item row. Element d takes word d % 4 of the Philox
    block with counter (gidx_lo, gidx_hi, epoch, slot << 16 | side << 15 | d // 4) and key
    (seed_lo ^ 0x85EBCA6B, seed_hi ^ step); it is kept, scaled by 1 / (1 - p), when the word is
    >= floor(p * 2^32)."""
    thr = max(1, int(p * 4294967296.0))
    scale = np.float32(1.0) / (np.float32(1.0) - np.float32(p))
    out = np.zeros(dim, dtype=np.float32)
    key = ((seed & 0xFFFFFFFF) ^ 0x85EBCA6B, ((seed >> 32) & 0xFFFFFFFF) ^ (step & 0xFFFFFFFF))
    for blk in range((dim + 3) // 4):
        words = philox4x32_10((gidx & 0xFFFFFFFF, gidx >> 32, epoch,
                               (slot << 16) | (side << 15) | blk), key)
        for j in range(4):
            d = blk * 4 + j
            if d < dim and words[j] >= thr:
                out[d] = scale
    return out


def mf_forward_dropout(U, I, bu, bi, users, items, mask_u, mask_i, y_range=None):
    """``forward`` in training mode: masks on both gathered embeddings, then the product
    (matrix_factorization.py:143-157)."""
    users, items = users.long(), items.long()
    s = ((U[users] * mask_u) * (I[items] * mask_i)).sum(dim=1) + bu[users].squeeze(1) \
        + bi[items].squeeze(1)
    if y_range is not None:
        s = torch.sigmoid(s) * (y_range[1] - y_range[0]) + y_range[0]
    return s
