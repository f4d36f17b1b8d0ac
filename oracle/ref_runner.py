"""Runs the UNMODIFIED reference (ShopRunner/collie v1.3.1) on the host CPU cores: the timed CPU
baseline of ``bench.py`` and a cross-check of the restatement in ``collie_oracle.py``.

TEST / MEASUREMENT INFRASTRUCTURE ONLY -- nothing under ``collie_b200/`` imports this.

The reference is pure Python. ``__graft_entry__.build()`` copies its package directory, untouched,
from ``/root/reference/collie`` into the git-ignored ``oracle/_ref/collie`` (``oracle/build_ref.sh``
is the same recipe as a script), so that it travels to the GPU box with the repository snapshot.
``oracle/stubs`` stands in for ``pytorch_lightning`` / ``torchmetrics`` / ``fire``, which are absent
from the image and contribute no arithmetic (``auroc`` is sklearn's ``roc_auc_score``, the yardstick of
the reference's own test, tests/test_metrics.py:140-154).

Call sites exercised, as BASELINE.md section 3 describes:
  collie/model/base/trainer.py:314-402      CollieMinimalTrainer.fit (the plain-PyTorch step loop)
  collie/interactions/dataloaders.py:176-294 ApproximateNegativeSamplingInteractionsDataLoader
  collie/interactions/dataloaders.py:70-173  InteractionsDataLoader (exact negative sampling)
  collie/metrics.py:285-395                  evaluate_in_batches([mapk, mrr, auc])
"""
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, '_ref')
STUBS = os.path.join(HERE, 'stubs')


def available():
    return os.path.isfile(os.path.join(REF_DIR, 'collie', '__init__.py'))


def import_reference():
    """``import collie`` from ``oracle/_ref`` (through the stubs). Returns the module."""
    if not available():
        raise RuntimeError('oracle/_ref/collie is missing: run __graft_entry__.build() where '
                           '/root/reference exists')
    for p in (REF_DIR, STUBS):
        if p not in sys.path:
            sys.path.insert(0, p)
    import importlib.util     # oracle/stubs/sitecustomize.py: scipy dok_matrix compat patch
    spec = importlib.util.spec_from_file_location('_collie_ref_compat',
                                                  os.path.join(STUBS, 'sitecustomize.py'))
    spec.loader.exec_module(importlib.util.module_from_spec(spec))
    import collie
    if not os.path.abspath(collie.__file__).startswith(REF_DIR):
        raise RuntimeError(f'collie was imported from {collie.__file__}, not from {REF_DIR}')
    return collie


def _implicit_data(collie, users, items, num_users, num_items, n_neg, exact):
    import numpy as np
    return collie.interactions.Interactions(
        users=np.asarray(users), items=np.asarray(items), ratings=np.ones(len(users)),
        num_negative_samples=n_neg, num_users=num_users, num_items=num_items,
        allow_missing_ids=True, remove_duplicate_user_item_pairs=False,
        check_num_negative_samples_is_valid=False,
        max_number_of_samples_to_consider=200 if exact else 0)


def train_seconds(users, items, num_users, num_items, dim, n_neg, batch, steps, warmup, lr,
                  bias_lr, loss='bpr', optimizer='sparse_adam', sparse=True, loader='approximate',
                  threads=None, seed=0, gpus=0):
    """Seconds the reference's ``CollieMinimalTrainer.fit`` takes for ``steps`` batches of
    ``batch`` interactions (one epoch over ``steps * batch`` interactions), after a warm-up fit of
    ``warmup`` batches on the same model. ``users`` / ``items`` must hold at least
    ``(steps + warmup) * batch`` entries. ``gpus=0`` keeps the reference on the host cores (its
    trainer moves the model to a GPU by itself when one is visible, trainer.py:234-237);
    ``gpus=1`` is the reference's stock-torch GPU path, timed as an informative extra."""
    import numpy as np
    import torch
    collie = import_reference()
    from collie.model import CollieMinimalTrainer, MatrixFactorizationModel
    if threads:
        torch.set_num_threads(int(threads))
    torch.manual_seed(seed)
    np.random.seed(seed)
    users, items = np.asarray(users), np.asarray(items)
    need = (steps + warmup) * batch
    if len(users) < need:
        raise ValueError(f'{need} interactions needed, {len(users)} given')
    if loader == 'approximate':
        Loader = collie.interactions.ApproximateNegativeSamplingInteractionsDataLoader
    elif loader == 'exact':
        Loader = collie.interactions.InteractionsDataLoader
    else:
        raise ValueError(loader)

    def loader_of(a, b):
        data = _implicit_data(collie, users[a:b], items[a:b], num_users, num_items, n_neg,
                              exact=(loader == 'exact'))
        return Loader(data, batch_size=batch, shuffle=True)

    first = loader_of(0, max(warmup, 1) * batch)
    model = MatrixFactorizationModel(train=first, embedding_dim=dim, sparse=sparse, lr=lr,
                                     bias_lr=bias_lr, optimizer=optimizer, bias_optimizer='sgd',
                                     loss=loss, lr_scheduler_func=None)

    def fit(train_loader):
        model.train_loader = train_loader
        trainer = CollieMinimalTrainer(model, max_epochs=1, gpus=gpus, logger=None, verbosity=0,
                                       enable_model_summary=False, early_stopping_patience=None)
        t0 = time.perf_counter()
        trainer.fit(model)
        if gpus:
            torch.cuda.synchronize()
        return time.perf_counter() - t0

    if warmup > 0:
        fit(first)
    return fit(loader_of(warmup * batch, (warmup + steps) * batch))


def eval_seconds(tables, test_users, test_items, num_users, num_items, k=10, batch_size=20,
                 threads=None, metrics=('mapk', 'mrr', 'auc')):
    """Seconds (and the metric values) of the reference's
    ``evaluate_in_batches([mapk, mrr, auc], test, model, k)`` on the CPU. ``tables`` =
    (U, I, bu, bi) torch CPU tensors loaded into a reference ``MatrixFactorizationModel``."""
    import numpy as np
    import torch
    collie = import_reference()
    from collie import metrics as M
    from collie.model import MatrixFactorizationModel
    if threads:
        torch.set_num_threads(int(threads))
    test = _implicit_data(collie, test_users, test_items, num_users, num_items, 1, exact=False)
    U, I, bu, bi = tables
    model = MatrixFactorizationModel(train=test, embedding_dim=int(U.shape[1]), sparse=False,
                                     lr_scheduler_func=None)
    with torch.no_grad():
        model.user_embeddings.weight.copy_(U)
        model.item_embeddings.weight.copy_(I)
        model.user_biases.weight.copy_(bu.view(-1, 1))
        model.item_biases.weight.copy_(bi.view(-1, 1))
    model.eval()
    fns = [getattr(M, m) for m in metrics]
    t0 = time.perf_counter()
    out = M.evaluate_in_batches(fns, test, model, k=k, batch_size=batch_size, verbose=False)
    dt = time.perf_counter() - t0
    n_users_eval = int(len(np.unique(np.asarray(test_users))))
    return dt, (out if isinstance(out, list) else [out]), n_users_eval
