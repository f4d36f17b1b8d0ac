"""Evaluation with the reference's API (``collie/metrics.py``), computed by ``cmf_eval_tile``.

Mirrors: ``get_preds`` :77-107, ``mapk`` :144-198, ``mrr`` :201-240, ``auc`` :243-282,
``evaluate_in_batches`` :285-395 (implicit data; the explicit variant is outside this path).

``evaluate_in_batches`` takes the fused route when every metric is one of ``mapk`` / ``mrr`` /
``auc`` and the model is this package's ``MatrixFactorizationModel``: the score tile is produced and
ranked on the device, and only three accumulators come back to the host, once. Any other metric
callable receives a materialised ``preds`` tile exactly like in the reference. Canonical ranking
order is score descending, item id ascending (``torch.topk``'s tie order is unpinned there).
"""
import numpy as np
import torch
from scipy.sparse import csr_matrix

from . import _abi
from .interactions import Interactions

# users per fused tile; the result does not depend on it (per-user values are summed)
_TILE_USERS = 512
_MAX_TILE_BYTES = 1 << 30
_TC_TILE_USERS = 32768   # users per tensor-core call (128-user CTA tiles inside)
# what the last tensor-core evaluation reported (see cmf_eval_t.tc_status): informative
LAST_TC_STATUS = {}


def _targets_device(targets: csr_matrix, user_ids, device):
    """CSR rows of ``user_ids`` restricted to entries > 0, item ids sorted, on ``device``."""
    sub = csr_matrix(targets[np.asarray(user_ids)]).copy()
    sub.sum_duplicates()
    sub.data = np.where(sub.data > 0, sub.data, 0)
    sub.eliminate_zeros()
    sub.sort_indices()
    return (torch.from_numpy(sub.indptr.astype(np.int64)).to(device),
            torch.from_numpy(sub.indices.astype(np.int32)).to(device),
            torch.from_numpy(np.diff(sub.indptr).astype(np.int64)).to(device))


def _targets_all_users_device(targets: csr_matrix, test_users, device):
    """Same as ``_targets_device(targets, test_users, ...)`` when ``test_users`` are exactly the rows
    of ``targets`` that hold entries: dropping the empty rows of a CSR leaves ``indices`` as it is,
    so no row gather is needed on the host."""
    sub = targets
    if not sub.has_canonical_format:
        sub = sub.copy()
        sub.sum_duplicates()
    if sub.nnz and (sub.data <= 0).any():
        return _targets_device(targets, test_users, device)
    counts = np.diff(sub.indptr)[np.asarray(test_users)]
    if int(counts.sum()) != sub.nnz:
        return _targets_device(targets, test_users, device)
    indptr = np.zeros(len(counts) + 1, dtype=np.int64)
    np.cumsum(counts, out=indptr[1:])
    return (torch.from_numpy(indptr).to(device),
            torch.from_numpy(sub.indices.astype(np.int32, copy=False)).to(device),
            torch.from_numpy(counts.astype(np.int64)).to(device))


def _eval_tile(device, n_users, n_items, k, indptr, tgt_items, want, tables=None, user_ids=None,
               scores=None, precision=0, workspace=None, reuse_item_prep=False, item_offset=0,
               ext_best=None, into=None, tc_status=None, tc_overflow=None, tc_terms=0):
    """One ``cmf_eval_tile`` call; returns dict of per-user device tensors. ``item_offset`` /
    ``n_items`` select an item shard of the table (global ids come back); ``ext_best`` = (score,
    id) of every user's best target over ALL shards, against which the MRR counter counts;
    ``into`` = (dict of whole-evaluation output tensors, first row): the tile's per-user outputs are
    written there instead of into fresh tensors. ``tc_status`` (precision 1): int32[2] device
    tensor the call adds to -- [0] users ranked by the exact path inside the call, [1] users it
    could not take over (their outputs are not written: repeat the tile with precision 0)."""
    a = _abi.cmf_eval_t()
    keep = []
    if scores is not None:
        scores = _abi.require_cuda(scores.float().contiguous(), 'preds')
        a.scores_in = scores.data_ptr()
        keep.append(scores)
    else:
        U, I, bu, bi = tables
        a.U, a.I, a.bu, a.bi = U.data_ptr(), I.data_ptr(), bu.data_ptr(), bi.data_ptr()
        a.D = U.shape[1]
        a.user_ids = user_ids.data_ptr()
    a.n_users_tile, a.item_offset, a.num_items_shard, a.k = n_users, int(item_offset), n_items, k
    if ext_best is not None:
        a.ext_best_score, a.ext_best_id = ext_best[0].data_ptr(), ext_best[1].data_ptr()
        keep.extend(ext_best)
    a.tgt_indptr, a.tgt_items = indptr.data_ptr(), tgt_items.data_ptr()
    a.precision = precision
    a.reuse_item_prep = 1 if reuse_item_prep else 0
    a.tc_terms = tc_terms
    if tc_status is not None:
        a.tc_status = tc_status.data_ptr()
        keep.append(tc_status)
        if tc_overflow is not None:          # (rows tensor, row number of the tile's first user)
            a.tc_overflow_rows = tc_overflow[0].data_ptr()
            a.tc_overflow_cap = int(tc_overflow[0].numel())
            a.tc_row_base = int(tc_overflow[1])
            keep.append(tc_overflow[0])
    out = {}

    def vec(name, dtype):
        if into is not None and name in into[0]:
            return into[0][name][into[1]:into[1] + n_users]
        return torch.empty(n_users, dtype=dtype, device=device)
    if 'mapk' in want:
        out['ap'] = vec('ap', torch.float64)
        a.ap = out['ap'].data_ptr()
    if 'topk' in want:
        out['topk_ids'] = torch.empty((n_users, k), dtype=torch.int32, device=device)
        out['topk_scores'] = torch.empty((n_users, k), dtype=torch.float32, device=device)
        a.topk_ids, a.topk_scores = out['topk_ids'].data_ptr(), out['topk_scores'].data_ptr()
    if 'mrr' in want:
        a.want_mrr = 1
        out['n_before_best'] = vec('n_before_best', torch.int32)
        a.n_before_best = out['n_before_best'].data_ptr()
    if 'auc' in want:
        a.want_auc = 1
        out['auc_pairs'] = vec('auc_pairs', torch.float64)
        a.auc_pairs = out['auc_pairs'].data_ptr()
    need = _abi.lib().cmf_eval_workspace_bytes(n_users, n_items, int(a.D), k, precision)
    if workspace is None or workspace.numel() < need:
        workspace = torch.empty(int(need), dtype=torch.uint8, device=device)
    a.workspace, a.workspace_bytes = workspace.data_ptr(), workspace.numel()
    with _abi.on_device(device):
        rc = _abi.lib().cmf_eval_tile(_abi.C.byref(a), _abi.stream_ptr(device))
    _abi.check(rc, 'cmf_eval_tile')
    out['_workspace'] = workspace
    return out


def _check_k(k, n_items):
    if k > n_items:
        raise ValueError(f'Ensure ``k`` ({k}) is less than the number of items ({n_items})')


def _per_user(name, res, nnz, n_items):
    """Per-user metric values (float64, on device) from the kernel's ingredients."""
    if name == 'mapk':
        return res['ap']
    if name == 'mrr':
        nb = res['n_before_best']
        return torch.where(nb >= 0, 1.0 / (nb.double() + 1.0), torch.zeros_like(nb,
                                                                                dtype=torch.float64))
    if name == 'auc':
        pos = nnz.double()
        neg = float(n_items) - pos
        denom = pos * neg
        return torch.where(denom > 0, res['auc_pairs'] / denom.clamp(min=1.0),
                           torch.zeros_like(denom))
    raise ValueError(name)


def _metric_on_preds(name, targets, user_ids, preds, k):
    preds = torch.as_tensor(preds)
    if not preds.is_cuda:
        raise _abi.CmfError('metrics run on the GPU: ``preds`` must be a CUDA tensor '
                            '(collie_b200 has no CPU path)')
    n_users, n_items = preds.shape
    if name == 'mapk':
        _check_k(k, n_items)
    indptr, items, nnz = _targets_device(targets, user_ids, preds.device)
    res = _eval_tile(preds.device, n_users, n_items, k if name == 'mapk' else 1, indptr, items,
                     {name}, scores=preds)
    return _per_user(name, res, nnz, n_items).mean().item()


def mapk(targets: csr_matrix, user_ids, preds, k: int = 10) -> float:
    """Mean average precision at ``k`` (collie/metrics.py:144-198)."""
    return _metric_on_preds('mapk', targets, user_ids, preds, k)


def mrr(targets: csr_matrix, user_ids, preds, k=None) -> float:
    """Mean reciprocal rank of the best-ranked target (collie/metrics.py:201-240)."""
    return _metric_on_preds('mrr', targets, user_ids, preds, k)


def auc(targets: csr_matrix, user_ids, preds, k=None) -> float:
    """Mean per-user AUROC of ``sigmoid(preds)`` (collie/metrics.py:243-282)."""
    return _metric_on_preds('auc', targets, user_ids, preds, k)


_FUSED = {mapk: 'mapk', mrr: 'mrr', auc: 'auc'}


def _model_tables(model):
    return tuple(_abi.require_cuda(t.data, n) for t, n in (
        (model.user_embeddings.weight, 'user_embeddings.weight'),
        (model.item_embeddings.weight, 'item_embeddings.weight'),
        (model.user_biases.weight, 'user_biases.weight'),
        (model.item_biases.weight, 'item_biases.weight')))


def get_preds(model, user_ids, n_items, device=None):
    """Scores of every item for ``user_ids``: tensor ``n_users x n_items``
    (collie/metrics.py:77-107), one tiled kernel instead of ``n_users * n_items`` gathers."""
    if not getattr(model, '_fusable', True):
        # a model whose score is not <u, i> + biases (dense layers on top): every (user, item)
        # pair through its forward, like the reference (collie/metrics.py:60-72, :104-107)
        dev = model.device
        users = torch.as_tensor(np.asarray(user_ids)).to(dev).long()
        items = torch.arange(int(n_items), device=dev)
        with torch.no_grad():
            return model(users.repeat_interleave(int(n_items)),
                         items.repeat(users.shape[0])).view(-1, int(n_items)).float()
    U, I, bu, bi = _model_tables(model)
    users = torch.as_tensor(np.asarray(user_ids)).to(U.device, torch.int32).contiguous()
    out = torch.empty((users.shape[0], int(n_items)), dtype=torch.float32, device=U.device)
    with _abi.on_device(U.device):
        rc = _abi.lib().cmf_score_tile(_abi.ptr(U), _abi.ptr(I), _abi.ptr(bu), _abi.ptr(bi),
                                       U.shape[1], _abi.ptr(users), users.shape[0], int(n_items),
                                       _abi.ptr(out), _abi.stream_ptr(U.device))
    _abi.check(rc, 'cmf_score_tile')
    return out


def evaluate_in_batches(metric_list, test_interactions, model, k=10, batch_size=20, logger=None,
                        verbose=True, precision=0):
    """Evaluate a model with potentially several metrics (collie/metrics.py:285-395).

    Returns a float for one metric, else a list, exactly like the reference. ``precision`` selects
    the scoring arithmetic of the fused route (0: exact fp32; 1: tcgen05 tensor cores).
    """
    if not isinstance(test_interactions, Interactions):
        raise ValueError('``test_interactions`` must be of type ``Interactions``, not '
                         f'{type(test_interactions)}. Try using ``explicit_evaluate_in_batches`` '
                         'instead.')
    if not torch.cuda.is_available():
        raise _abi.CmfError('evaluate_in_batches runs on a CUDA device; none is available and '
                            'collie_b200 has no CPU fallback')
    if not next(model.parameters()).is_cuda:
        model.to('cuda')
    model._move_any_external_data_to_device()
    device = model.device

    # the test set in the form the kernels take it (users with targets, their targets as CSR on
    # the device) is derived once per Interactions object and device: a validation set is evaluated
    # every epoch, and the host-side derivation costs far more than the evaluation itself
    cache = test_interactions.__dict__.setdefault('_cmf_eval_cache', {})
    key = (str(device), id(test_interactions.mat), test_interactions.mat.nnz)
    if cache.get('key') != key:
        cache.clear()
        cache['key'] = key
        cache['test_users'] = np.unique(test_interactions.mat.row)
        cache['targets'] = test_interactions.mat.tocsr()
    test_users, targets = cache['test_users'], cache['targets']
    n_items = test_interactions.num_items
    n_test = len(test_users)
    names = [_FUSED.get(m) for m in metric_list]

    if all(n is not None for n in names) and getattr(model, '_fusable', True):
        if 'mapk' in names:
            _check_k(k, n_items)
        tables = _model_tables(model)
        if 'dev' not in cache:
            cache['dev'] = _targets_all_users_device(targets, test_users, device) + (
                torch.from_numpy(test_users.astype(np.int32)).to(device),)
        indptr, tgt_items, nnz, users_dev = cache['dev']
        D = tables[0].shape[1]
        # tensor-core route for what it supports (top-k / MAP / MRR); AUC stays on the exact path
        use_tc = precision == 1 and D <= 128 and k <= 16
        if precision == 1 and not use_tc:
            raise _abi.CmfError('precision=1 needs embedding_dim <= 128 and k <= 16')
        tc_names = [n for n in names if n != 'auc'] if use_tc else []
        ex_names = [n for n in names if n not in tc_names]
        sums = torch.zeros(len(names), dtype=torch.float64, device=device)
        # per-user ingredients of the whole evaluation, filled tile by tile, reduced once
        _ING = {'mapk': ('ap', torch.float64), 'mrr': ('n_before_best', torch.int32),
                'auc': ('auc_pairs', torch.float64)}
        whole = {_ING[n_][0]: torch.empty(n_test, dtype=_ING[n_][1], device=device)
                 for n_ in names}
        if ex_names:
            tile = max(1, min(_TILE_USERS, _MAX_TILE_BYTES // max(1, 12 * n_items)))
            ws = None
            for s in range(0, n_test, tile):
                e = min(s + tile, n_test)
                ws = _eval_tile(device, e - s, n_items, k, indptr[s:e + 1], tgt_items,
                                set(ex_names), tables=tables, user_ids=users_dev[s:e],
                                precision=0, workspace=ws, into=(whole, s))['_workspace']
        status = None
        if tc_names:
            tile = _TC_TILE_USERS
            ws = None
            status = torch.zeros(4, dtype=torch.int32, device=device)
            overflow = torch.empty(n_test, dtype=torch.int32, device=device)
            terms, fresh_prep, s = 0, True, 0
            while s < n_test:
                e = min(s + tile, n_test)
                ws = _eval_tile(device, e - s, n_items, k, indptr[s:e + 1], tgt_items,
                                set(tc_names), tables=tables, user_ids=users_dev[s:e],
                                precision=1, workspace=ws, reuse_item_prep=not fresh_prep,
                                into=(whole, s), tc_status=status,
                                tc_overflow=(overflow, s), tc_terms=terms)['_workspace']
                fresh_prep = False
                if s == 0 and e < n_test:
                    # the first tile of several tells how the rest will go (one read-back)
                    handed, lost = (int(x) for x in status[:2].tolist())
                    if terms == 0 and 'mrr' not in tc_names and handed > (e - s) // 50:
                        # one bf16 term left more than 2 % of the tile to the exact path (top
                        # scores closer together than its bound can tell): three terms cost less
                        # than that many exact re-rankings; a tile that lost a quarter is redone
                        terms, fresh_prep = 3, True
                        if lost > (e - s) // 4:
                            status.zero_()
                            continue
                    elif lost > (e - s) // 4:
                        # a quarter of the users left to the exact path even so (e.g. an
                        # untrained model: all scores within the bound of each other): do not
                        # spend the remaining tiles on the tensor cores
                        LAST_TC_STATUS.clear()
                        LAST_TC_STATUS.update(users=n_test, handed_to_exact_path=handed,
                                              not_taken_over=lost, gave_up_after=e)
                        return evaluate_in_batches(metric_list, test_interactions, model, k=k,
                                                   batch_size=batch_size, logger=logger,
                                                   verbose=verbose, precision=0)
                s = e
        for j, name in enumerate(names):
            sums[j] = _per_user(name, whole, nnz, n_items).sum()
        if status is not None:
            # one read-back for the whole evaluation: the metric sums and the tensor-core status
            n_over = int(status[2].item()) if tc_names else 0
            if n_over:
                # users beyond what the calls could hand to the exact path themselves: ONLY they
                # go through it now (their rows of the per-user arrays are overwritten)
                rows = torch.sort(overflow[:n_over].long())[0]
                sub_nnz = nnz[rows]
                sub_indptr = torch.zeros(n_over + 1, dtype=torch.int64, device=device)
                torch.cumsum(sub_nnz, 0, out=sub_indptr[1:])
                take = torch.repeat_interleave(indptr[rows] - sub_indptr[:-1], sub_nnz) \
                    + torch.arange(int(sub_indptr[-1]), device=device)
                sub_items = tgt_items[take]
                sub_users = users_dev[rows].contiguous()
                tile_x = max(1, min(_TILE_USERS, _MAX_TILE_BYTES // max(1, 12 * n_items)))
                ws_x = None
                for s in range(0, n_over, tile_x):
                    e = min(s + tile_x, n_over)
                    res = _eval_tile(device, e - s, n_items, k, sub_indptr[s:e + 1], sub_items,
                                     set(tc_names), tables=tables, user_ids=sub_users[s:e],
                                     precision=0, workspace=ws_x)
                    ws_x = res['_workspace']
                    for n_ in tc_names:
                        key_ = _ING[n_][0]
                        whole[key_][rows[s:e]] = res[key_]
                for j, name in enumerate(names):
                    sums[j] = _per_user(name, whole, nnz, n_items).sum()
            host = torch.cat([sums, status.double()]).tolist()
            LAST_TC_STATUS.clear()
            LAST_TC_STATUS.update(users=n_test, handed_to_exact_path=int(host[-4]),
                                  not_taken_over=int(host[-3]), re_ranked_afterwards=n_over)
            all_scores = [v / n_test for v in host[:len(names)]]
        else:
            all_scores = (sums / n_test).tolist()
    else:
        if n_test < batch_size:
            batch_size = n_test
        acc = [0.0] * len(metric_list)
        for i in range(int(np.ceil(n_test / batch_size))):
            user_range = test_users[i * batch_size:(i + 1) * batch_size]
            preds = get_preds(model, user_range, n_items, device)
            for j, metric in enumerate(metric_list):
                acc[j] += metric(targets=targets, user_ids=user_range, preds=preds, k=k) \
                    * len(user_range)
        all_scores = [a / n_test for a in acc]

    if logger is not None:
        _log_metrics(model, logger, metric_list, all_scores, verbose)
    return all_scores[0] if len(all_scores) == 1 else all_scores


class _ErrorMetric(object):
    """Stand-ins for ``torchmetrics.MeanSquaredError`` / ``MeanAbsoluteError`` (the package is not
    in this image): ``metric(preds, target)`` accumulates, ``compute()`` returns the mean,
    ``reset()`` clears. The sums stay on the device of the inputs (float64)."""
    def __init__(self):
        self.reset()

    def reset(self):
        self._sum, self._n = None, 0

    def _err(self, d):
        raise NotImplementedError

    def __call__(self, preds, target):
        d = preds.detach().double().flatten() - target.detach().double().flatten().to(preds.device)
        s = self._err(d).sum()
        self._sum = s if self._sum is None else self._sum + s
        self._n += d.numel()

    update = __call__

    def compute(self):
        if self._n == 0:
            return torch.tensor(float('nan'))
        return (self._sum / self._n).float().cpu()


class MeanSquaredError(_ErrorMetric):
    def _err(self, d):
        return d * d


class MeanAbsoluteError(_ErrorMetric):
    def _err(self, d):
        return d.abs()


def explicit_evaluate_in_batches(metric_list, test_interactions, model, logger=None, verbose=True,
                                 **kwargs):
    """Evaluate an explicit model with ``torchmetrics``-style metric objects
    (collie/metrics.py:398-502): batches of ``(users, items, ratings)`` from an
    ``InteractionsDataLoader`` over ``test_interactions`` (``**kwargs`` go to it), ``preds =
    model(users, items)`` through the CUDA forward, every metric called as ``metric(preds, ratings)``
    and ``compute()``d at the end; metrics are ``reset()`` on the way out. The stand-ins of this
    module take the predictions on the device; any other metric object gets CPU tensors like in the
    reference."""
    from .interactions import ExplicitInteractions, InteractionsDataLoader
    if not isinstance(test_interactions, ExplicitInteractions):
        raise ValueError('``test_interactions`` must be of type ``ExplicitInteractions``, not '
                         f'{type(test_interactions)}. Try using ``evaluate_in_batches`` instead.')
    if not torch.cuda.is_available():
        raise _abi.CmfError('explicit_evaluate_in_batches runs on a CUDA device; none is available '
                            'and collie_b200 has no CPU fallback')
    try:
        if not next(model.parameters()).is_cuda:
            model.to('cuda')
        model._move_any_external_data_to_device()
        kwargs.setdefault('batch_size', 1 << 16)
        loader = InteractionsDataLoader(interactions=test_interactions, **kwargs)
        with torch.no_grad():
            for users, items, ratings in loader:
                preds = model(users, items)
                for metric in metric_list:
                    if isinstance(metric, _ErrorMetric):
                        metric(preds, ratings)
                    else:
                        metric(preds.cpu(), ratings.cpu())
        all_scores = [metric.compute() for metric in metric_list]
        if logger is not None:
            _log_metrics(model, logger, metric_list, all_scores, verbose)
        return all_scores[0] if len(all_scores) == 1 else all_scores
    finally:
        for metric in metric_list:
            metric.reset()


def _log_metrics(model, logger, metric_list, all_scores, verbose):
    """collie/metrics.py:524-543."""
    try:
        step = model.hparams.get('num_epochs_completed')
    except Exception:
        step = None
    metrics_dict = dict(zip([getattr(m, '__name__', type(m).__name__) for m in metric_list],
                            all_scores))
    if verbose:
        print(f'Logging metrics {metrics_dict} to ``logger``...')
    logger.log_metrics(metrics=metrics_dict, step=step)
