"""CPU-only checks: the C-ABI library loads and exports every symbol the header declares, and the
host-side mirror keeps the reference's argument checking and error behaviour. No compute calls."""
import os
import re

import numpy as np
import pytest
import torch

import collie_b200 as cb
from collie_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, 'include', 'collie_b200.h')).read()
    declared = set(re.findall(r'CMF_API\s+[\w\s\*]+?\b(cmf_\w+)\s*\(', header))
    assert declared, 'no declarations found'
    assert declared == set(_abi.SIGNATURES), declared ^ set(_abi.SIGNATURES)
    lib = _abi.lib()
    for name in declared:
        assert getattr(lib, name) is not None
    assert lib.cmf_version() >= 100
    assert lib.cmf_launch_count() == 0 or lib.cmf_launch_count() > 0


def test_struct_sizes_match_header_layout():
    # natural alignment, 8-byte pointers: catches a field added on one side only
    assert _abi.C.sizeof(_abi.cmf_opt_t) == 56
    assert _abi.C.sizeof(_abi.cmf_step_t) % 8 == 0
    assert _abi.C.sizeof(_abi.cmf_eval_t) % 8 == 0


def test_argument_errors_without_gpu():
    lib = _abi.lib()
    assert lib.cmf_train_workspace_bytes(0, 5, 10, 10, 8) == 0
    assert lib.cmf_train_workspace_bytes(1024, 5, 100, 50, 16) > 0
    assert lib.cmf_eval_workspace_bytes(8, 50, 16, 10, 0) > 0
    assert lib.cmf_set_tuning(99, 1) < 0
    assert b'unknown key' in lib.cmf_last_error()
    rc = lib.cmf_forward_scores(None, None, None, None, 8, None, None, 4, 0, None, None, None)
    assert rc == -1 and b'null' in lib.cmf_last_error()
    # the entry points added in round 2 refuse bad arguments before they touch a device
    rc = lib.cmf_merge_topk(None, None, 2, 10, 5, None, None, None, None, None, None)
    assert rc != 0 and b'null' in lib.cmf_last_error()
    one = (_abi.C.c_int32 * 1)()
    onef = (_abi.C.c_float * 1)()
    rc = lib.cmf_merge_topk(one, onef, 2, 10, 64, None, None, None, None, None, None)
    assert rc != 0 and b'k <= 32' in lib.cmf_last_error()
    rc = lib.cmf_pairwise_distance(None, None, 8, None, None, 4, 0, None, None)
    assert rc != 0 and b'null' in lib.cmf_last_error()
    rc = lib.cmf_pairwise_distance(onef, onef, 8, one, one, 2, 1, onef, None)
    assert rc != 0 and b'int32 or int64' in lib.cmf_last_error()
    assert lib.cmf_exchange_pad_bytes() == 128
    info = (_abi.C.c_double * 4)()
    assert lib.cmf_eval_tc_info(128, info) == 0
    assert info[0] == 144 and info[1] == 400 and 0 < info[3] < info[2] < 1e-2


def _pairs(n_users=20, n_items=30, n=200, seed=0):
    rng = np.random.default_rng(seed)
    flat = rng.choice(n_users * n_items, size=n, replace=False)
    return flat // n_items, flat % n_items


def test_interactions_validation_like_reference():
    users, items = _pairs()
    with pytest.raises(ValueError, match='Lengths'):
        cb.Interactions(users=users, items=items[:-1])
    with pytest.raises(ValueError, match='every integer'):
        cb.Interactions(users=users * 2, items=items)
    with pytest.raises(AssertionError):
        cb.Interactions()
    with pytest.raises(AssertionError, match='must be less than'):
        cb.Interactions(users=users, items=items, num_negative_samples=30, allow_missing_ids=True)
    with pytest.raises(ValueError, match='Ratings must be provided'):
        cb.ExplicitInteractions(users=users, items=items)
    with pytest.warns(UserWarning, match='contain ``0``s'):
        d = cb.Interactions(users=users, items=items, ratings=np.r_[0, np.ones(len(users) - 1)],
                            allow_missing_ids=True)
    assert d.num_interactions == len(users) - 1
    d = cb.Interactions(users=np.r_[users, users[:5]], items=np.r_[items, items[:5]],
                        allow_missing_ids=True)
    assert d.num_interactions == len(users)          # duplicates removed
    assert len(d) == len(users) and d.toarray().shape == (d.num_users, d.num_items)
    assert 'Interactions object with' in repr(d)
    e = cb.ExplicitInteractions(users=users, items=items, ratings=np.full(len(users), 3),
                                allow_missing_ids=True)
    with pytest.raises(AttributeError):
        e.num_negative_samples
    assert e[3] == (e.mat.row[3], e.mat.col[3], 3)


def test_loaders_and_model_configuration():
    users, items = _pairs()
    data = cb.Interactions(users=users, items=items, allow_missing_ids=True,
                           num_negative_samples=4)
    loader = cb.ApproximateNegativeSamplingInteractionsDataLoader(data, batch_size=64)
    assert data.max_number_of_samples_to_consider == 0 and len(loader) == 4
    assert loader.num_users == data.num_users and loader.num_negative_samples == 4
    with pytest.raises(ValueError, match='does not support explicit'):
        cb.ApproximateNegativeSamplingInteractionsDataLoader(
            cb.ExplicitInteractions(users=users, items=items, ratings=np.ones(len(users)),
                                    allow_missing_ids=True))
    with pytest.raises(TypeError, match='``train`` must be provided'):
        cb.MatrixFactorizationModel()
    model = cb.MatrixFactorizationModel(train=loader, embedding_dim=12, loss='bpr')
    assert model._loss_name == 'adaptive_bpr' and model.hparams._is_implicit is True
    assert model.hparams.num_users == data.num_users and model.hparams.embedding_dim == 12
    assert set(model.state_dict()) == {'user_biases.weight', 'item_biases.weight',
                                       'user_embeddings.weight', 'item_embeddings.weight'}
    assert float(model.user_biases.weight.abs().sum()) == 0.0
    std = model.item_embeddings.weight.std().item()
    assert 0.5 / (12 * 2.5) < std < 2.0 / (12 * 2.5)
    out = model.configure_optimizers()
    assert isinstance(out, tuple) and len(out[0]) == 2 and out[1][0]['monitor'] == \
        'train_loss_epoch'
    assert out[0][0].rule == 'adam' and out[0][1].rule == 'sgd'
    assert out[0][1].param_groups[0]['lr'] == 1e-2
    one = cb.MatrixFactorizationModel(train=loader, bias_optimizer=None, lr_scheduler_func=None)
    assert isinstance(one.configure_optimizers(), torch.optim.Optimizer)
    single = cb.Interactions(users=users, items=items, allow_missing_ids=True,
                             num_negative_samples=1)
    with pytest.warns(UserWarning, match='Adaptive hinge loss specified'):
        m1 = cb.MatrixFactorizationModel(train=single, loss='adaptive')
    assert m1._loss_name == 'hinge'
    with pytest.warns(UserWarning, match='weight_decay'):
        m2 = cb.MatrixFactorizationModel(train=data, sparse=True, weight_decay=0.1)
    assert m2.hparams.weight_decay == 0.0
    val = cb.Interactions(users=users, items=items, allow_missing_ids=True, num_items=31)
    with pytest.raises(AssertionError, match='num_items'):
        cb.MatrixFactorizationModel(train=data, val=val)
    ex = cb.ExplicitInteractions(users=users, items=items, ratings=np.ones(len(users)),
                                 allow_missing_ids=True)
    with pytest.raises(ValueError, match='explicit data with an implicit loss'):
        cb.MatrixFactorizationModel(train=ex, loss='hinge')


def test_no_cpu_fallback_anywhere():
    users, items = _pairs()
    data = cb.Interactions(users=users, items=items, allow_missing_ids=True,
                           max_number_of_samples_to_consider=0)
    model = cb.MatrixFactorizationModel(train=data, lr_scheduler_func=None)
    with pytest.raises(_abi.CmfError):
        model(torch.zeros(2, dtype=torch.long), torch.zeros(2, dtype=torch.long))
    if not torch.cuda.is_available():
        with pytest.raises(_abi.CmfError):
            cb.evaluate_in_batches([cb.mapk], data, model)
        with pytest.raises(_abi.CmfError):
            cb.CollieMinimalTrainer(model, verbosity=0).fit(model)
        with pytest.raises(_abi.CmfError):
            next(iter(model.train_loader))
    # nothing in the product imports the oracle
    for root, _, files in os.walk(os.path.join(ROOT, 'collie_b200')):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                text = open(os.path.join(root, f)).read()
                assert not re.search(r'^\s*(from|import)\s+oracle|#include.*oracle|'
                                     r'import_module\(.oracle', text, flags=re.M), f


def test_duplicate_pairs_keep_first_position_and_last_value():
    """collie/interactions/datasets.py:137-146: the reference pushes the entries through a dict, so
    a duplicated (user, item) pair stays where it FIRST appeared and takes its LAST value (checked
    against the reference itself when this was written: rows [0 1 2 0], cols [5 3 1 2], data
    [3 5 4 6] for the input below)."""
    import collie_b200 as cb
    users = np.array([0, 1, 0, 2, 1, 0])
    items = np.array([5, 3, 5, 1, 3, 2])
    ratings = np.array([1., 2., 3., 4., 5., 6.])
    d = cb.ExplicitInteractions(users=users, items=items, ratings=ratings, num_users=3, num_items=6,
                                allow_missing_ids=True)
    assert d.mat.row.tolist() == [0, 1, 2, 0]
    assert d.mat.col.tolist() == [5, 3, 1, 2]
    assert d.mat.data.tolist() == [3., 5., 4., 6.]
