"""GPU parity: the CUDA path (through the C ABI) against the oracle and against golden vectors
produced by the unmodified reference. Tolerances: bit-exact for ids / indices, 1e-5 relative for
losses, gradients and metrics (north star), 1e-5 relative (+1e-6 absolute) for parameters after k
optimizer steps."""
import glob
import os

import numpy as np
import pytest
import scipy.sparse as sp
import torch

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
IMPLICIT = sorted(glob.glob(os.path.join(GOLDEN, 'implicit_*.npz')))
EXPLICIT = sorted(glob.glob(os.path.join(GOLDEN, 'explicit_*.npz')))
METRICS = sorted(glob.glob(os.path.join(GOLDEN, 'metrics_*.npz')))

GOLDEN_OPT = {
    'implicit_bpr_n1_sparse_adam': 'sparse_adam', 'implicit_hinge_n1_adam': 'adam',
    'implicit_adaptive_bpr_n5_sparse_adam': 'sparse_adam',
    'implicit_adaptive_hinge_n10_adam': 'adam', 'implicit_warp_n7_sparse_adam': 'sparse_adam',
    'implicit_warp_n4_sgd': 'sgd', 'implicit_adaptive_hinge_meta_n5': 'sparse_adam',
    'implicit_warp_meta_n6': 'adam', 'implicit_adaptive_bpr_d10_n5': 'adam',
    'implicit_adaptive_bpr_d30_n3': 'sparse_adam', 'implicit_adaptive_bpr_d128_n5': 'sparse_adam',
    'implicit_hinge_adagrad': 'adagrad', 'explicit_mse_adam': 'adam',
    'explicit_mse_sparse_adam_d64': 'sparse_adam', 'explicit_mae_sgd': 'sgd',
    'explicit_mse_yrange': 'adam',
}
LOSS_ARG = {'bpr_loss': 'bpr', 'hinge_loss': 'hinge', 'adaptive_bpr_loss': 'bpr',
            'adaptive_hinge_loss': 'adaptive', 'warp_loss': 'warp'}


def _ids(paths):
    return [os.path.basename(p)[:-4] for p in paths]


def _implicit_data(n_users, n_items, n_neg, seed=0, exact=False, n=400):
    import collie_b200 as cb
    rng = np.random.default_rng(seed)
    flat = rng.choice(n_users * n_items, size=n, replace=False)
    return cb.Interactions(users=flat // n_items, items=flat % n_items, num_users=n_users,
                           num_items=n_items, num_negative_samples=n_neg, allow_missing_ids=True,
                           check_num_negative_samples_is_valid=False,
                           max_number_of_samples_to_consider=200 if exact else 0, seed=seed + 1)


def _model_from_golden(g, opt_name, loss, explicit=False, meta=None):
    import collie_b200 as cb
    n_users, n_items, dim = int(g['n_users']), int(g['n_items']), int(g['dim'])
    if explicit:
        rng = np.random.default_rng(0)
        data = cb.ExplicitInteractions(users=rng.integers(0, n_users, 50),
                                       items=rng.integers(0, n_items, 50),
                                       ratings=rng.integers(1, 6, 50), num_users=n_users,
                                       num_items=n_items, allow_missing_ids=True)
        yr = None if np.isnan(g['y_range']).any() else tuple(g['y_range'].tolist())
        kw = dict(y_range=yr)
    else:
        data = _implicit_data(n_users, n_items, int(g['n_neg']))
        kw = {}
        if meta:
            kw = dict(metadata_for_loss={k: torch.from_numpy(v) for k, v in meta.items()},
                      metadata_for_loss_weights={'a': 0.2, 'b': 0.3})
    model = cb.MatrixFactorizationModel(
        train=data, embedding_dim=dim, sparse=(opt_name == 'sparse_adam'), lr=float(g['lr']),
        bias_lr=float(g['bias_lr']), optimizer=opt_name, bias_optimizer='sgd', loss=loss,
        lr_scheduler_func=None, **kw).cuda()
    with torch.no_grad():
        model.user_embeddings.weight.copy_(torch.from_numpy(g['init_U']))
        model.item_embeddings.weight.copy_(torch.from_numpy(g['init_I']))
        model.user_biases.weight.copy_(torch.from_numpy(g['init_bu']))
        model.item_biases.weight.copy_(torch.from_numpy(g['init_bi']))
    return model


def _tables(model):
    return {'U': model.user_embeddings.weight, 'I': model.item_embeddings.weight,
            'bu': model.user_biases.weight, 'bi': model.item_biases.weight}


# ------------------------------------------------------------------------------------ forward
@pytest.mark.parametrize('dim', [1, 7, 10, 30, 32, 64, 96, 128, 160, 256, 384])
@pytest.mark.parametrize('idx_dtype', [torch.int64, torch.int32])
def test_forward_scores_match_oracle(dim, idx_dtype):
    import collie_b200 as cb
    from oracle import collie_oracle as O
    rng = np.random.default_rng(dim)
    n_users, n_items, m = 61, 77, 1001
    model = cb.MatrixFactorizationModel(train=_implicit_data(n_users, n_items, 3),
                                        embedding_dim=dim, lr_scheduler_func=None).cuda()
    with torch.no_grad():
        for p in model.parameters():
            p.copy_(torch.from_numpy(rng.normal(0, 0.5, tuple(p.shape)).astype(np.float32)))
    users = torch.from_numpy(rng.integers(0, n_users, m)).to(idx_dtype)
    items = torch.from_numpy(rng.integers(0, n_items, m)).to(idx_dtype)
    got = model(users.cuda(), items.cuda()).detach().cpu()
    tabs = [t.detach().cpu() for t in _tables(model).values()]
    exp = O.mf_forward(*tabs, users, items)
    np.testing.assert_allclose(got.numpy(), exp.numpy(), rtol=1e-5, atol=1e-6)
    # empty input
    assert model(users[:0].cuda(), items[:0].cuda()).shape == (0,)


def test_forward_y_range_and_autograd_match_oracle():
    import collie_b200 as cb
    from oracle import collie_oracle as O
    rng = np.random.default_rng(3)
    n_users, n_items, dim, m = 23, 31, 24, 300
    data = cb.ExplicitInteractions(users=rng.integers(0, n_users, 60),
                                   items=rng.integers(0, n_items, 60),
                                   ratings=rng.integers(1, 6, 60), num_users=n_users,
                                   num_items=n_items, allow_missing_ids=True)
    model = cb.MatrixFactorizationModel(train=data, embedding_dim=dim, loss='mse', y_range=(1, 5),
                                        lr_scheduler_func=None).cuda()
    ref = O.OracleMF(n_users, n_items, dim)
    with torch.no_grad():
        for p in model.parameters():
            p.copy_(torch.from_numpy(rng.normal(0, 0.5, tuple(p.shape)).astype(np.float32)))
        for name in ('user_embeddings', 'item_embeddings', 'user_biases', 'item_biases'):
            getattr(ref, name).weight.copy_(getattr(model, name).weight.cpu())
    users = torch.from_numpy(rng.integers(0, n_users, m))
    items = torch.from_numpy(rng.integers(0, n_items, m))
    w = torch.from_numpy(rng.normal(size=m).astype(np.float32))
    model.train()
    out = model(users.cuda(), items.cuda())
    (out * w.cuda()).sum().backward()
    exp = ref(users, items, (1.0, 5.0))
    (exp * w).sum().backward()
    np.testing.assert_allclose(out.detach().cpu().numpy(), exp.detach().numpy(), rtol=1e-5)
    for name in ('user_embeddings', 'item_embeddings', 'user_biases', 'item_biases'):
        np.testing.assert_allclose(getattr(model, name).weight.grad.cpu().numpy(),
                                   getattr(ref, name).weight.grad.numpy(), rtol=1e-5, atol=1e-6,
                                   err_msg=name)


# ------------------------------------------------------------------------------------ train steps
@pytest.mark.parametrize('chunk_len', [0, 1, 5])
@pytest.mark.parametrize('path', IMPLICIT, ids=_ids(IMPLICIT))
def test_golden_implicit_steps(path, chunk_len):
    """Losses and final parameters after 3 steps equal what the unmodified reference produced."""
    from collie_b200 import _abi
    g = np.load(path)
    name = os.path.basename(path)[:-4]
    meta = {k[5:]: g[k] for k in g.files if k.startswith('meta_')}
    model = _model_from_golden(g, GOLDEN_OPT[name], LOSS_ARG[str(g['loss_name'])], meta=meta)
    opts = model.configure_optimizers()
    prev = _abi.lib().cmf_set_tuning(0, chunk_len)
    try:
        for s in range(int(g['steps'])):
            batch = ((torch.from_numpy(g[f'step{s}_users'].astype(np.int32)).cuda(),
                      torch.from_numpy(g[f'step{s}_pos'].astype(np.int32)).cuda()),
                     torch.from_numpy(g[f'step{s}_negs']).cuda())
            val = model.calculate_loss(batch).item()      # loss-only path
            loss = model.fused_step(batch, optimizers=opts).item()
            np.testing.assert_allclose(loss, float(g[f'step{s}_loss']), rtol=1e-5)
            np.testing.assert_allclose(val, loss, rtol=1e-6)
    finally:
        _abi.lib().cmf_set_tuning(0, prev)
    for k, t in _tables(model).items():
        np.testing.assert_allclose(t.detach().cpu().numpy(), g['final_' + k], rtol=1e-5, atol=1e-6,
                                   err_msg=k)


@pytest.mark.parametrize('path', EXPLICIT, ids=_ids(EXPLICIT))
def test_golden_explicit_steps(path):
    g = np.load(path)
    name = os.path.basename(path)[:-4]
    model = _model_from_golden(g, GOLDEN_OPT[name], 'mae' if 'mae' in name else 'mse',
                               explicit=True)
    opts = model.configure_optimizers()
    for s in range(int(g['steps'])):
        batch = (torch.from_numpy(g[f'step{s}_users'].astype(np.int32)).cuda(),
                 torch.from_numpy(g[f'step{s}_items'].astype(np.int32)).cuda(),
                 torch.from_numpy(g[f'step{s}_ratings']).cuda())
        loss = model.fused_step(batch, optimizers=opts).item()
        np.testing.assert_allclose(loss, float(g[f'step{s}_loss']), rtol=1e-5)
    for k, t in _tables(model).items():
        np.testing.assert_allclose(t.detach().cpu().numpy(), g['final_' + k], rtol=1e-5, atol=1e-6,
                                   err_msg=k)


@pytest.mark.parametrize('path', IMPLICIT, ids=_ids(IMPLICIT))
def test_golden_implicit_gradients(path):
    """Gradients of all four tables: with plain SGD at lr=1 the fused step moves every parameter by
    exactly -grad, which is compared with autograd's gradient from the reference run."""
    g = np.load(path)
    meta = {k[5:]: g[k] for k in g.files if k.startswith('meta_')}
    model = _model_from_golden(g, 'sgd', LOSS_ARG[str(g['loss_name'])], meta=meta)
    model.hparams.lr = 1.0
    model.hparams.bias_lr = 1.0
    opts = model.configure_optimizers()
    before = {k: t.detach().clone() for k, t in _tables(model).items()}
    batch = ((torch.from_numpy(g['step0_users'].astype(np.int32)).cuda(),
              torch.from_numpy(g['step0_pos'].astype(np.int32)).cuda()),
             torch.from_numpy(g['step0_negs']).cuda())
    model.fused_step(batch, optimizers=opts)
    for k, t in _tables(model).items():
        grad = (before[k] - t.detach()).cpu().numpy()
        ref = g['step0_grad_' + k]
        scale = np.abs(ref).max() + 1e-12
        # the gradient is recovered as before - after of an SGD step with lr = 1: the updated
        # parameter is rounded to fp32 at ITS magnitude (|p| < 4 here, half an ulp = 2.4e-7), so
        # every recovered element carries up to that absolute error whatever its own size
        np.testing.assert_allclose(grad, ref, rtol=1e-5, atol=5e-7 * max(1.0, scale), err_msg=k)


def test_selected_negative_indices_bit_exact():
    """argmax slot (adaptive) / first-violation slot (warp) equal the oracle's, bit for bit, on
    exactly representable weights (multiples of 2^-3) so that ties are genuine."""
    import collie_b200 as cb
    from oracle import collie_oracle as O
    rng = np.random.default_rng(9)
    n_users, n_items, dim, n_neg, b = 40, 60, 16, 8, 512
    for loss, oname in (('bpr', 'adaptive_bpr'), ('adaptive', 'adaptive_hinge'), ('warp', 'warp')):
        model = cb.MatrixFactorizationModel(train=_implicit_data(n_users, n_items, n_neg),
                                            embedding_dim=dim, loss=loss,
                                            lr_scheduler_func=None).cuda()
        with torch.no_grad():
            for p in model.parameters():
                p.copy_(torch.from_numpy((rng.integers(-4, 5, tuple(p.shape)) / 8.0)
                                         .astype(np.float32)))
        users = torch.from_numpy(rng.integers(0, n_users, b).astype(np.int32))
        pos = torch.from_numpy(rng.integers(0, n_items, b).astype(np.int32))
        negs = torch.from_numpy(rng.integers(0, n_items, (b, n_neg)))
        out = {'sel': None, 'pos_scores': None, 'neg_scores': None}
        loss_val = model.fused_step(((users.cuda(), pos.cuda()), negs.cuda()), train=False,
                                    outputs=out).item()
        tabs = [t.detach().cpu() for t in _tables(model).values()]
        ps = O.mf_forward(*tabs, users, pos)
        ns = O.mf_forward(*tabs, users.long().repeat(n_neg), negs.t().flatten()).view(n_neg, b)
        assert torch.equal(out['pos_scores'].cpu(), ps)
        assert torch.equal(out['neg_scores'].cpu(), ns.t())
        if oname == 'warp':
            exp_sel = O.warp_first_violation(1.0 - ps.view(b, 1) + ns.t())
            exp_loss = O.warp(ps, ns, n_items)
        else:
            exp_sel = torch.max(ns, 0)[1]
            exp_loss = O.IMPLICIT_LOSSES[oname](ps, ns)
        assert torch.equal(out['sel'].cpu().long(), exp_sel.long())
        np.testing.assert_allclose(loss_val, exp_loss.item(), rtol=1e-5)


# ------------------------------------------------------------------------------------ sampler
@pytest.mark.parametrize('exact', [False, True])
def test_sampled_negatives_bit_exact(exact):
    """Device sampler == the oracle's statement of the stream, for the stand-alone kernel and for
    the negatives the fused step draws in-kernel."""
    import collie_b200 as cb
    from oracle import collie_oracle as O
    n_users, n_items, n_neg = 30, 45, 6
    data = _implicit_data(n_users, n_items, n_neg, seed=4, exact=exact, n=500)
    st = data.device_state('cuda')
    gidx = torch.arange(100, 356, device='cuda')
    users = st['users'][gidx]
    got = data.sample_negatives(users, gidx, epoch=3).cpu().numpy()
    csr = data.mat.tocsr()
    csr.sort_indices()
    exp = O.philox_negatives(data.seed, 3, gidx.cpu().numpy(), users.cpu().numpy(), n_neg, n_items,
                             csr.indptr if exact else None, csr.indices if exact else None,
                             max_tries=200)
    assert np.array_equal(got, exp)
    if exact:
        dense = data.mat.toarray() > 0
        u = users.cpu().numpy()
        for r in range(len(u)):
            assert not dense[u[r], got[r]].any()
            assert len(set(got[r].tolist())) == n_neg
    model = cb.MatrixFactorizationModel(train=data, embedding_dim=8, loss='bpr',
                                        lr_scheduler_func=None).cuda()
    out = {'negs': None}
    batch = cb.interactions.DeviceBatch(users, st['items'][gidx], gidx, None)
    model.fused_step(batch, train=False, interactions=data, epoch=3, outputs=out)
    assert np.array_equal(out['negs'].cpu().numpy(), exp)


def test_exact_sampler_budget_exhaustion_matches_oracle():
    import collie_b200 as cb
    from oracle import collie_oracle as O
    # user 0 has interacted with every item but two: the rejection budget runs out
    n_items = 12
    users = np.zeros(10, dtype=np.int64)
    items = np.arange(10)
    data = cb.Interactions(users=users, items=items, num_users=1, num_items=n_items,
                           num_negative_samples=4, allow_missing_ids=True,
                           check_num_negative_samples_is_valid=False,
                           max_number_of_samples_to_consider=5, seed=11)
    gidx = torch.arange(10, device='cuda')
    u = torch.zeros(10, dtype=torch.int32, device='cuda')
    got = data.sample_negatives(u, gidx, epoch=0).cpu().numpy()
    csr = data.mat.tocsr()
    exp = O.philox_negatives(11, 0, np.arange(10), users, 4, n_items, csr.indptr, csr.indices,
                             max_tries=5)
    assert np.array_equal(got, exp)


# ------------------------------------------------------------------------------------ metrics
@pytest.mark.parametrize('path', METRICS, ids=_ids(METRICS))
def test_golden_metrics(path):
    """evaluate_in_batches equals the reference's numbers (fused route and materialised route)."""
    import collie_b200 as cb
    g = np.load(path)
    n_users, n_items, dim = int(g['n_users']), int(g['n_items']), int(g['dim'])
    test = cb.Interactions(users=g['test_users'], items=g['test_items'], num_users=n_users,
                           num_items=n_items, allow_missing_ids=True,
                           check_num_negative_samples_is_valid=False,
                           max_number_of_samples_to_consider=0)
    model = cb.MatrixFactorizationModel(train=test, embedding_dim=dim,
                                        lr_scheduler_func=None).cuda()
    with torch.no_grad():
        for k, t in _tables(model).items():
            t.copy_(torch.from_numpy(g[k]))
    has_ties = 'exact_grid' in path
    scores = cb.get_preds(model, np.arange(n_users), n_items)
    np.testing.assert_allclose(scores.cpu().numpy(), g['scores'], rtol=1e-5, atol=1e-5)

    def custom(targets, user_ids, preds, k):   # forces the materialised route
        return cb.mapk(targets, user_ids, preds, k)
    for k in (1, 4, 10):
        exp = g[f'k{k}_bs20']
        got = cb.evaluate_in_batches([cb.mapk, cb.mrr, cb.auc], test, model, k=k, verbose=False)
        np.testing.assert_allclose(got[2], exp[2], rtol=1e-5)
        if not has_ties:
            np.testing.assert_allclose(got[:2], exp[:2], rtol=1e-5)
            got2 = cb.evaluate_in_batches([custom, cb.mrr], test, model, k=k, batch_size=7,
                                          verbose=False)
            np.testing.assert_allclose(got2, exp[:2], rtol=1e-5)
        assert isinstance(cb.evaluate_in_batches([cb.mapk], test, model, k=k, verbose=False), float)


def test_reference_metric_known_answers_on_gpu():
    """tests/test_metrics.py:101-154 of the reference, through the CUDA metric callables."""
    import json
    import collie_b200 as cb
    ka = json.load(open(os.path.join(GOLDEN, 'reference_known_answers.json')))
    users, items = np.array(ka['metrics_users']), np.array(ka['metrics_items'])
    targets = sp.coo_matrix((np.ones(len(users)), (users, items)), shape=(3, 4)).tocsr()
    scores = torch.tensor(ka['metrics_scores'], dtype=torch.float32).cuda()
    assert cb.mapk(targets, np.array([1, 2]), scores[[1, 2]], k=4) == ka['mapk_k4_users_1_2']
    np.testing.assert_almost_equal(cb.mapk(targets, np.arange(3), scores, k=1), ka['mapk_k1'])
    np.testing.assert_almost_equal(cb.mrr(targets, np.arange(3), scores), ka['mrr'])
    from sklearn.metrics import roc_auc_score
    exp = np.mean([roc_auc_score(targets[i].toarray()[0], scores[i].cpu().numpy())
                   for i in range(3)])
    np.testing.assert_almost_equal(cb.auc(targets, np.arange(3), scores), exp)
    with pytest.raises(ValueError):
        cb.mapk(targets, np.arange(3), scores, k=5)


def test_topk_ids_bit_exact_with_ties():
    """Top-k item ids equal the oracle's canonical order (score desc, id asc) bit for bit on the
    exact-arithmetic fixture, which is full of genuine ties."""
    import collie_b200 as cb
    from collie_b200 import metrics as M
    from oracle import collie_oracle as O
    g = np.load(os.path.join(GOLDEN, 'metrics_exact_grid.npz'))
    n_users, n_items = int(g['n_users']), int(g['n_items'])
    tabs = tuple(torch.from_numpy(g[k]).cuda() for k in ('U', 'I', 'bu', 'bi'))
    users = torch.arange(n_users, dtype=torch.int32, device='cuda')
    coo = sp.coo_matrix((np.ones(len(g['test_users'])), (g['test_users'], g['test_items'])),
                        shape=(n_users, n_items))
    indptr, items, _ = M._targets_device(coo.tocsr(), np.arange(n_users), 'cuda')
    for k in (1, 10, 37, n_items):
        res = M._eval_tile('cuda', n_users, n_items, k, indptr, items, {'topk', 'mapk', 'mrr'},
                           tables=tabs, user_ids=users)
        exp_ids = O.topk_ids(g['scores'], k)
        assert np.array_equal(res['topk_ids'].cpu().numpy(), exp_ids)
        exp_scores = np.take_along_axis(g['scores'], exp_ids, axis=1)
        assert np.array_equal(res['topk_scores'].cpu().numpy(), exp_scores)
        np.testing.assert_allclose(res['ap'].cpu().numpy(),
                                   O.mapk_per_user(coo.tocsr(), np.arange(n_users), g['scores'], k),
                                   rtol=1e-12)
        mrr_exp = O.mrr_per_user(coo.tocsr(), np.arange(n_users), g['scores'])
        nb = res['n_before_best'].cpu().numpy()
        got = np.where(nb >= 0, 1.0 / (nb + 1.0), 0.0)
        np.testing.assert_allclose(got, mrr_exp, rtol=1e-12)


# ------------------------------------------------------------------------------------ end to end
def test_trainer_end_to_end_and_checkpoint(tmp_path, capsys):
    import collie_b200 as cb
    from oracle import collie_oracle as O
    rng = np.random.default_rng(5)
    n_users, n_items = 120, 90
    # planted structure: user u likes items with the same parity
    users = rng.integers(0, n_users, 6000)
    items = (rng.integers(0, n_items // 2, 6000) * 2 + users % 2) % n_items
    key = np.unique(users * n_items + items)
    users, items = key // n_items, key % n_items
    cut = int(0.9 * len(users))
    perm = rng.permutation(len(users))
    tr, te = perm[:cut], perm[cut:]
    kw = dict(num_users=n_users, num_items=n_items, allow_missing_ids=True, seed=7,
              num_negative_samples=5)
    train = cb.Interactions(users=users[tr], items=items[tr], **kw)
    val = cb.Interactions(users=users[te], items=items[te], **kw)
    model = cb.MatrixFactorizationModel(train=cb.InteractionsDataLoader(train, batch_size=256,
                                                                        shuffle=True),
                                        val=cb.InteractionsDataLoader(val, batch_size=256),
                                        embedding_dim=16, lr=5e-2, sparse=True,
                                        optimizer='sparse_adam', loss='adaptive')
    before = None
    trainer = cb.CollieMinimalTrainer(model, max_epochs=8, verbosity=1,
                                      early_stopping_patience=None)
    out = capsys.readouterr().out
    assert 'Detected GPU. Setting ``gpus`` to 1.' in out
    trainer.fit(model)
    out = capsys.readouterr().out
    assert '| Name            | Type            | Params' in out
    assert 'Epoch     1: train loss:' in out and 'val loss:' in out
    assert model.hparams.num_epochs_completed == 8 and trainer.num_epochs_completed == 8
    trainer.fit(model)                       # same max_epochs: no-op
    assert trainer.num_epochs_completed == 8
    trainer.max_epochs += 1
    trainer.fit(model)
    assert trainer.num_epochs_completed == 9
    got = cb.evaluate_in_batches([cb.mapk, cb.mrr, cb.auc], val, model, k=10, verbose=False)
    tabs = tuple(t.detach().cpu() for t in _tables(model).values())
    exp = O.evaluate_in_batches(['mapk', 'mrr', 'auc'], val.mat, tabs, k=10)
    np.testing.assert_allclose(got, exp, rtol=1e-5)
    assert got[2] > 0.55, got                # better than chance on held-out pairs
    # checkpoint round trip with the reference's keys
    path = tmp_path / 'model.pth'
    model.save_model(path)
    blob = torch.load(path, weights_only=False)
    assert set(blob['state_dict']) == {'user_biases.weight', 'item_biases.weight',
                                       'user_embeddings.weight', 'item_embeddings.weight'}
    assert tuple(blob['state_dict']['user_biases.weight'].shape) == (n_users, 1)
    loaded = cb.MatrixFactorizationModel(load_model_path=path).cuda()
    assert loaded.hparams.num_epochs_completed == 9
    u = torch.arange(50, device='cuda')
    assert torch.equal(loaded(u, u), model(u, u))
    assert before is None


def test_explicit_training_reduces_mse():
    import collie_b200 as cb
    rng = np.random.default_rng(6)
    n_users, n_items, n = 200, 150, 8000
    users, items = rng.integers(0, n_users, n), rng.integers(0, n_items, n)
    ratings = (1 + (users % 3) + (items % 3)).astype(np.float32)
    data = cb.ExplicitInteractions(users=users, items=items, ratings=ratings, num_users=n_users,
                                   num_items=n_items, allow_missing_ids=True)
    model = cb.MatrixFactorizationModel(train=cb.InteractionsDataLoader(data, batch_size=512,
                                                                        shuffle=True),
                                        embedding_dim=64, lr=1e-2, loss='mse', y_range=(1, 5),
                                        lr_scheduler_func=None)
    trainer = cb.CollieTrainer(model, max_epochs=1, verbosity=0, enable_model_summary=False)
    trainer.fit(model)
    first = model.calculate_loss(next(iter(model.train_loader))).item()
    trainer.max_epochs = 15
    trainer.fit(model)
    last = model.calculate_loss(next(iter(model.train_loader))).item()
    assert last < 0.5 * first, (first, last)
    # explicit_evaluate_in_batches (collie/metrics.py:398-502) against the same numbers on the host
    with torch.no_grad():
        preds = model(torch.from_numpy(data.mat.row).cuda(),
                      torch.from_numpy(data.mat.col).cuda()).cpu().numpy().astype(np.float64)
    truth = data.mat.data.astype(np.float64)

    class HostMSE(object):       # a foreign metric object: gets CPU tensors like in the reference
        def __init__(self):
            self.reset()

        def reset(self):
            self.s, self.n = 0.0, 0

        def __call__(self, p, t):
            assert not p.is_cuda and not t.is_cuda
            self.s += float(((p.double() - t.double()) ** 2).sum())
            self.n += p.numel()

        def compute(self):
            return self.s / self.n
    host = HostMSE()
    mse, mae, mse_host = cb.explicit_evaluate_in_batches(
        [cb.MeanSquaredError(), cb.MeanAbsoluteError(), host], data, model, verbose=False,
        batch_size=1000)
    np.testing.assert_allclose(float(mse), np.mean((preds - truth) ** 2), rtol=1e-5)
    np.testing.assert_allclose(float(mae), np.mean(np.abs(preds - truth)), rtol=1e-5)
    np.testing.assert_allclose(mse_host, float(mse), rtol=1e-6)
    assert host.n == 0                                        # metrics are reset on the way out
    one = cb.explicit_evaluate_in_batches([cb.MeanSquaredError()], data, model, verbose=False)
    np.testing.assert_allclose(float(one), float(mse), rtol=1e-6)
    implicit = cb.Interactions(users=users, items=items, num_users=n_users, num_items=n_items,
                               allow_missing_ids=True, check_num_negative_samples_is_valid=False)
    with pytest.raises(ValueError, match='must be of type ``ExplicitInteractions``'):
        cb.explicit_evaluate_in_batches([cb.MeanSquaredError()], implicit, model)


def test_errors_are_loud():
    import collie_b200 as cb
    from collie_b200 import _abi
    data = _implicit_data(10, 12, 1, n=60)
    with pytest.raises(ValueError, match='WARP'):
        cb.MatrixFactorizationModel(train=data, loss='warp', lr_scheduler_func=None)
    with pytest.raises(ValueError, match='not a valid loss'):
        cb.MatrixFactorizationModel(train=data, loss='nope', lr_scheduler_func=None)
    model = cb.MatrixFactorizationModel(train=data, lr_scheduler_func=None, optimizer='nope')
    with pytest.raises(ValueError, match='not a valid optimizer'):
        model.configure_optimizers()
    model = cb.MatrixFactorizationModel(train=data, lr_scheduler_func=None)
    with pytest.raises(_abi.CmfError):          # CPU tensors: no fallback
        model(torch.zeros(3, dtype=torch.long), torch.zeros(3, dtype=torch.long))
    model.cuda()
    with pytest.raises(ValueError, match='Implicit loss with explicit data'):
        model.calculate_loss((torch.zeros(3), torch.zeros(3), torch.zeros(3)))
    with pytest.raises(ValueError):
        model.calculate_loss((torch.zeros(3),))
    # workspace too small is reported, not ignored
    a = _abi.cmf_step_t()
    assert _abi.lib().cmf_train_step(_abi.C.byref(a), None) == -1
    assert b'null parameter table' in _abi.lib().cmf_last_error()


# ------------------------------------------------------------------------------------ tensor cores
def test_tensor_core_eval_bit_exact_on_exact_fixture():
    """precision=1 (tcgen05 GEMM over bf16-split operands): on the exact-arithmetic fixture every
    value is representable in bf16, so scores and top-k ids equal the oracle's bit for bit."""
    from collie_b200 import metrics as M
    from oracle import collie_oracle as O
    g = np.load(os.path.join(GOLDEN, 'metrics_exact_grid.npz'))
    n_users, n_items = int(g['n_users']), int(g['n_items'])
    tabs = tuple(torch.from_numpy(g[k]).cuda() for k in ('U', 'I', 'bu', 'bi'))
    users = torch.arange(n_users, dtype=torch.int32, device='cuda')
    coo = sp.coo_matrix((np.ones(len(g['test_users'])), (g['test_users'], g['test_items'])),
                        shape=(n_users, n_items))
    indptr, items, _ = M._targets_device(coo.tocsr(), np.arange(n_users), 'cuda')
    for k in (1, 10, 16):
        res = M._eval_tile('cuda', n_users, n_items, k, indptr, items, {'topk', 'mapk', 'mrr'},
                           tables=tabs, user_ids=users, precision=1)
        exp_ids = O.topk_ids(g['scores'], k)
        assert np.array_equal(res['topk_ids'].cpu().numpy(), exp_ids)
        assert np.array_equal(res['topk_scores'].cpu().numpy(),
                              np.take_along_axis(g['scores'], exp_ids, axis=1))
        np.testing.assert_allclose(res['ap'].cpu().numpy(),
                                   O.mapk_per_user(coo.tocsr(), np.arange(n_users), g['scores'], k),
                                   rtol=1e-12)
        nb = res['n_before_best'].cpu().numpy().astype(np.float64)
        np.testing.assert_allclose(np.where(nb >= 0, 1.0 / (nb + 1.0), 0.0),
                                   O.mrr_per_user(coo.tocsr(), np.arange(n_users), g['scores']),
                                   rtol=1e-12)


@pytest.mark.parametrize('shape', [(300, 1000, 128), (77, 5000, 64), (1000, 333, 32), (130, 26744, 128),
                                   (50, 700, 5), (200, 300, 33), (64, 257, 100), (3, 40, 1)])
def test_tensor_core_eval_matches_exact_path(shape):
    """Random fp32 weights: the tensor-core path returns the exact path's ids, scores, AP@k and
    MRR counter bit for bit (candidates selected on the tensor cores, re-scored in fp32)."""
    import collie_b200 as cb
    from collie_b200 import metrics as M
    n_users, n_items, dim = shape
    rng = np.random.default_rng(n_items)
    tabs = tuple(torch.from_numpy(rng.normal(0, 1.0 / np.sqrt(dim), s).astype(np.float32)).cuda()
                 for s in ((n_users, dim), (n_items, dim), (n_users, 1), (n_items, 1)))
    tu = rng.integers(0, n_users, 6 * n_users)
    ti = rng.integers(0, n_items, 6 * n_users)
    coo = sp.coo_matrix((np.ones(len(tu)), (tu, ti)), shape=(n_users, n_items))
    coo.sum_duplicates()
    tile = np.arange(n_users)
    indptr, items, nnz = M._targets_device(coo.tocsr(), tile, 'cuda')
    users = torch.from_numpy(tile.astype(np.int32)).cuda()
    k = min(10, n_items)
    ex = M._eval_tile('cuda', n_users, n_items, k, indptr, items, {'topk', 'mapk', 'mrr'},
                      tables=tabs, user_ids=users, precision=0)
    # the tensor cores select, fp32 decides: ids, scores, AP and the MRR counter are the exact
    # path's, bit for bit -- with the MRR counter (three bf16 terms) and without (one term)
    for want, terms in (({'topk', 'mapk', 'mrr'}, 0), ({'topk', 'mapk'}, 0), ({'topk', 'mapk'}, 3)):
        status = torch.zeros(2, dtype=torch.int32, device='cuda')
        tc = M._eval_tile('cuda', n_users, n_items, k, indptr, items, want, tables=tabs,
                          user_ids=users, precision=1, tc_status=status, tc_terms=terms)
        assert int(status[1]) == 0
        # ... and by the tensor-core route itself: the exact path may take over the odd user whose
        # scores are too close to call, not the fixture
        assert int(status[0]) <= max(2, n_users // 20), status.tolist()
        assert torch.equal(tc['topk_ids'], ex['topk_ids'])
        assert torch.equal(tc['topk_scores'], ex['topk_scores'])
        assert torch.equal(tc['ap'], ex['ap'])
        if 'mrr' in want:
            assert torch.equal(tc['n_before_best'], ex['n_before_best'])
    # through the public API
    test = cb.Interactions(mat=coo, num_users=n_users, num_items=n_items, allow_missing_ids=True,
                           check_num_negative_samples_is_valid=False,
                           max_number_of_samples_to_consider=0)
    model = cb.MatrixFactorizationModel(train=test, embedding_dim=dim, lr_scheduler_func=None).cuda()
    with torch.no_grad():
        for t, src in zip(_tables(model).values(), tabs):
            t.copy_(src)
    got = cb.evaluate_in_batches([cb.mapk, cb.mrr, cb.auc], test, model, k=k, verbose=False,
                                 precision=1)
    exp = cb.evaluate_in_batches([cb.mapk, cb.mrr, cb.auc], test, model, k=k, verbose=False)
    np.testing.assert_allclose(got, exp, rtol=1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize('shape', [(1024, 26744, 128), (300, 3000, 64)])
def test_tensor_core_eval_with_uneven_item_norms(shape):
    """Item rows whose norms spread over two orders of magnitude (what a trained model has:
    popular items grow), biases of mixed size: the error bound of a pair follows THAT pair's
    norms, so the tensor-core route still settles (nearly) every user itself, and what it returns
    is the exact path's result bit for bit."""
    from collie_b200 import metrics as M
    n_users, n_items, dim = shape
    rng = np.random.default_rng(7 * n_items)
    iscale = np.exp(rng.normal(0.0, 1.0, (n_items, 1))).astype(np.float32)     # lognormal norms
    uscale = np.exp(rng.normal(0.0, 0.5, (n_users, 1))).astype(np.float32)
    tabs = (rng.normal(0, 0.3 / np.sqrt(dim), (n_users, dim)).astype(np.float32) * uscale,
            rng.normal(0, 0.3 / np.sqrt(dim), (n_items, dim)).astype(np.float32) * iscale,
            rng.normal(0, 0.05, (n_users, 1)).astype(np.float32),
            (rng.normal(0, 0.05, (n_items, 1)) * iscale).astype(np.float32))
    tabs = tuple(torch.from_numpy(np.ascontiguousarray(t)).cuda() for t in tabs)
    tu = rng.integers(0, n_users, 6 * n_users)
    ti = rng.integers(0, n_items, 6 * n_users)
    coo = sp.coo_matrix((np.ones(len(tu)), (tu, ti)), shape=(n_users, n_items))
    coo.sum_duplicates()
    tile = np.arange(n_users)
    indptr, items, nnz = M._targets_device(coo.tocsr(), tile, 'cuda')
    users = torch.from_numpy(tile.astype(np.int32)).cuda()
    k = 10
    ex = M._eval_tile('cuda', n_users, n_items, k, indptr, items, {'topk', 'mapk', 'mrr'},
                      tables=tabs, user_ids=users, precision=0)
    for want in ({'topk', 'mapk', 'mrr'}, {'topk', 'mapk'}):
        status = torch.zeros(2, dtype=torch.int32, device='cuda')
        tc = M._eval_tile('cuda', n_users, n_items, k, indptr, items, want, tables=tabs,
                          user_ids=users, precision=1, tc_status=status)
        assert int(status[1]) == 0, status.tolist()
        assert int(status[0]) <= n_users // 20, status.tolist()
        assert torch.equal(tc['topk_ids'], ex['topk_ids'])
        assert torch.equal(tc['topk_scores'], ex['topk_scores'])
        assert torch.equal(tc['ap'], ex['ap'])
        if 'mrr' in want:
            assert torch.equal(tc['n_before_best'], ex['n_before_best'])


@pytest.mark.gpu
@pytest.mark.parametrize('n_shards,n_users,k', [(2, 300, 10), (8, 1000, 10), (5, 77, 1), (8, 64, 16),
                                                (12, 50, 12)])
def test_merge_of_sharded_topk_lists(n_shards, n_users, k):
    """cmf_merge_topk (item-sharded evaluation): global top-k in canonical order and AP@k from the
    shards' lists, against the torch statement of the same merge; scores drawn from a few values
    so that ties across shards decide by item id, short lists padded with -1."""
    from collie_b200 import _abi
    from collie_b200.distributed import ap_from_topk_ids, merge_topk
    rng = np.random.default_rng(n_shards * 1000 + k)
    block = 40
    n_items = n_shards * block
    ids = np.full((n_shards, n_users, k), -1, dtype=np.int32)
    sc = np.full((n_shards, n_users, k), -np.inf, dtype=np.float32)
    for s_ in range(n_shards):
        for u in range(n_users):
            m = int(rng.integers(0, k + 1)) if u % 7 == 0 else k
            pick = rng.choice(block, size=m, replace=False) + s_ * block
            val = rng.integers(-3, 4, size=m).astype(np.float32) * 0.5
            order = np.lexsort((pick, -val))
            ids[s_, u, :m], sc[s_, u, :m] = pick[order], val[order]
    counts = rng.integers(0, 6, n_users)
    tgt = [np.sort(rng.choice(n_items, size=c, replace=False)).astype(np.int32) for c in counts]
    indptr = torch.from_numpy(np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)).cuda()
    tgt_items = torch.from_numpy(np.concatenate(tgt + [np.zeros(0, np.int32)])).cuda()
    d_ids, d_sc = torch.from_numpy(ids).cuda(), torch.from_numpy(sc).cuda()
    out_ids = torch.empty((n_users, k), dtype=torch.int32, device='cuda')
    out_sc = torch.empty((n_users, k), dtype=torch.float32, device='cuda')
    ap = torch.empty(n_users, dtype=torch.float64, device='cuda')
    _abi.check(_abi.lib().cmf_merge_topk(d_ids.data_ptr(), d_sc.data_ptr(), n_shards, n_users, k,
                                         indptr.data_ptr(), tgt_items.data_ptr(),
                                         out_ids.data_ptr(), out_sc.data_ptr(), ap.data_ptr(),
                                         torch.cuda.current_stream().cuda_stream))
    exp_s, exp_i = merge_topk(d_sc.permute(1, 0, 2).reshape(n_users, -1),
                              d_ids.permute(1, 0, 2).reshape(n_users, -1), k)
    assert torch.equal(out_ids, exp_i)
    assert torch.equal(out_sc, torch.where(exp_i < 0, torch.full_like(exp_s, -float('inf')), exp_s))
    row_of = torch.repeat_interleave(torch.arange(n_users, device='cuda'),
                                     torch.from_numpy(counts).cuda())
    keys = row_of.long() * n_items + tgt_items.long()
    exp_ap = ap_from_topk_ids(exp_i, torch.arange(n_users, device='cuda'), keys,
                              torch.from_numpy(counts).cuda(), n_items)
    np.testing.assert_allclose(ap.cpu().numpy(), exp_ap.cpu().numpy(), rtol=1e-12, atol=0)


@pytest.mark.gpu
@pytest.mark.parametrize('n_shards,precision', [(2, 1), (3, 1), (2, 0)])
def test_item_shards_merge_to_the_unsharded_result(n_shards, precision):
    """What evaluate_item_sharded does, on one GPU: every shard of item rows is scored by its own
    cmf_eval_tile call (item_offset, the users' best target over ALL shards given from outside),
    the shards' top-k lists are merged by cmf_merge_topk and the MRR counters summed: ids, AP@k and
    the counter equal the unsharded exact evaluation bit for bit."""
    from collie_b200 import _abi
    from collie_b200 import metrics as M
    n_users, n_items, dim, k = 700, 5003, 128, 10
    rng = np.random.default_rng(n_shards)
    tabs = tuple(torch.from_numpy(rng.normal(0, 0.3 / np.sqrt(dim), s).astype(np.float32)).cuda()
                 for s in ((n_users, dim), (n_items, dim), (n_users, 1), (n_items, 1)))
    tu = rng.integers(0, n_users, 6 * n_users)
    ti = rng.integers(0, n_items, 6 * n_users)
    coo = sp.coo_matrix((np.ones(len(tu)), (tu, ti)), shape=(n_users, n_items))
    coo.sum_duplicates()
    tile = np.arange(n_users)
    indptr, items, nnz = M._targets_device(coo.tocsr(), tile, 'cuda')
    users = torch.from_numpy(tile.astype(np.int32)).cuda()
    ex = M._eval_tile('cuda', n_users, n_items, k, indptr, items, {'topk', 'mapk', 'mrr'},
                      tables=tabs, user_ids=users, precision=0)
    # the best target of every user over all shards: exact scores of the targets
    U, I, bu, bi = tabs
    row_of = torch.repeat_interleave(torch.arange(n_users, device='cuda'), nnz.long())
    S = torch.full((n_users, n_items), -float('inf'), device='cuda')
    full = M._eval_tile('cuda', n_users, n_items, n_items if n_items <= 16 else 16, indptr, items,
                        {'topk'}, tables=tabs, user_ids=users, precision=0)
    del S, full
    from collie_b200 import _abi as A
    sc_all = torch.empty((n_users, n_items), dtype=torch.float32, device='cuda')
    A.check(A.lib().cmf_score_tile(U.data_ptr(), I.data_ptr(), bu.data_ptr(), bi.data_ptr(), dim,
                                   users.data_ptr(), n_users, n_items, sc_all.data_ptr(),
                                   torch.cuda.current_stream().cuda_stream))
    tsc = sc_all[row_of, items.long()]
    best_s = torch.full((n_users,), -float('inf'), device='cuda')
    best_s.scatter_reduce_(0, row_of, tsc, reduce='amax', include_self=True)
    big = torch.iinfo(torch.int32).max
    best_id = torch.full((n_users,), big, dtype=torch.int32, device='cuda')
    at = tsc == best_s[row_of]
    best_id.scatter_reduce_(0, row_of[at], items[at].int(), reduce='amin', include_self=True)
    best_id = torch.where(best_id == big, torch.full_like(best_id, -1), best_id)
    per = (n_items + n_shards - 1) // n_shards
    all_ids = torch.full((n_shards, n_users, k), -1, dtype=torch.int32, device='cuda')
    all_s = torch.full((n_shards, n_users, k), -float('inf'), dtype=torch.float32, device='cuda')
    before = torch.zeros(n_users, dtype=torch.int64, device='cuda')
    status = torch.zeros(2, dtype=torch.int32, device='cuda')
    for sh in range(n_shards):
        lo, hi = sh * per, min((sh + 1) * per, n_items)
        res = M._eval_tile('cuda', n_users, hi - lo, k, indptr, items, {'topk', 'mrr'},
                           tables=tabs, user_ids=users, precision=precision, item_offset=lo,
                           ext_best=(best_s, best_id),
                           tc_status=status if precision == 1 else None)
        all_ids[sh], all_s[sh] = res['topk_ids'], res['topk_scores']
        before += res['n_before_best'].clamp(min=0).long()
    torch.cuda.synchronize()
    assert int(status[1]) == 0
    out_ids = torch.empty((n_users, k), dtype=torch.int32, device='cuda')
    out_sc = torch.empty((n_users, k), dtype=torch.float32, device='cuda')
    ap = torch.empty(n_users, dtype=torch.float64, device='cuda')
    _abi.check(_abi.lib().cmf_merge_topk(all_ids.data_ptr(), all_s.data_ptr(), n_shards, n_users,
                                         k, indptr.data_ptr(), items.data_ptr(),
                                         out_ids.data_ptr(), out_sc.data_ptr(), ap.data_ptr(),
                                         torch.cuda.current_stream().cuda_stream))
    assert torch.equal(out_ids, ex['topk_ids'])
    assert torch.equal(out_sc, ex['topk_scores'])
    assert torch.equal(ap, ex['ap'])
    has = ex['n_before_best'] >= 0
    assert torch.equal(before[has], ex['n_before_best'][has].long())


@pytest.mark.gpu
def test_tensor_core_eval_over_tiles_of_unequal_size():
    """An evaluation of 40,000 users runs as a 32,768-user tile and a smaller one that reuses the
    item side of the workspace (bf16 copy, per-item bounds, norm order): what the second call
    finds again must not depend on how many users the tile has."""
    import collie_b200 as cb
    n_users, n_items, dim = 40_000, 2_003, 32
    rng = np.random.default_rng(5)
    tu = np.repeat(np.arange(n_users), 2)
    ti = rng.integers(0, n_items, len(tu))
    test = cb.Interactions(users=tu, items=ti, num_users=n_users, num_items=n_items,
                           allow_missing_ids=True, check_num_negative_samples_is_valid=False,
                           max_number_of_samples_to_consider=0)
    model = cb.MatrixFactorizationModel(train=test, embedding_dim=dim, lr_scheduler_func=None).cuda()
    with torch.no_grad():
        for t in _tables(model).values():
            t.copy_(torch.from_numpy(rng.normal(0, 0.3 / np.sqrt(dim), tuple(t.shape))
                                     .astype(np.float32)))
    for metrics in ([cb.mapk], [cb.mapk, cb.mrr]):
        got = cb.evaluate_in_batches(metrics, test, model, k=10, verbose=False, precision=1)
        exp = cb.evaluate_in_batches(metrics, test, model, k=10, verbose=False)
        np.testing.assert_allclose(np.atleast_1d(got), np.atleast_1d(exp), rtol=1e-12)
        from collie_b200 import metrics as M
        assert M.LAST_TC_STATUS.get('not_taken_over', 0) == 0, M.LAST_TC_STATUS


@pytest.mark.gpu
def test_device_ingest_equals_the_host_rule(monkeypatch):
    """cmf_ingest_pairs (large inputs): a duplicated (user, item) pair keeps its FIRST position and
    its LAST value, like the reference's dict-backed construction (datasets.py:137-146), and the
    positives CSR left on the device equals scipy's."""
    import collie_b200 as cb
    from collie_b200 import interactions as X
    rng = np.random.default_rng(8)
    n_users, n_items, n = 3000, 700, 400_000                   # ~17 % duplicates
    users, items = rng.integers(0, n_users, n), rng.integers(0, n_items, n)
    ratings = rng.integers(1, 6, n).astype(np.float64)
    dev = cb.ExplicitInteractions(users=users, items=items, ratings=ratings, num_users=n_users,
                                  num_items=n_items, allow_missing_ids=True)
    assert dev._ingest_csr is not None
    monkeypatch.setattr(X, '_DEVICE_INGEST_MIN', 1 << 60)      # the host rule
    host = cb.ExplicitInteractions(users=users, items=items, ratings=ratings, num_users=n_users,
                                   num_items=n_items, allow_missing_ids=True)
    assert host._ingest_csr is None and host.num_interactions < n
    np.testing.assert_array_equal(dev.mat.row, host.mat.row)
    np.testing.assert_array_equal(dev.mat.col, host.mat.col)
    np.testing.assert_array_equal(dev.mat.data, host.mat.data)
    monkeypatch.undo()
    imp = cb.Interactions(users=users, items=items, num_users=n_users, num_items=n_items,
                          allow_missing_ids=True, check_num_negative_samples_is_valid=False)
    st = imp.device_state(torch.device('cuda', torch.cuda.current_device()))
    csr = imp.mat.tocsr()
    csr.sort_indices()
    np.testing.assert_array_equal(st['indptr'].cpu().numpy(), csr.indptr)
    np.testing.assert_array_equal(st['csr_items'].cpu().numpy(), csr.indices)
    # no duplicates at all: everything is kept, in order
    flat = rng.choice(n_users * n_items, size=300_000, replace=False)
    uniq = cb.Interactions(users=flat // n_items, items=flat % n_items, num_users=n_users,
                           num_items=n_items, allow_missing_ids=True,
                           check_num_negative_samples_is_valid=False)
    np.testing.assert_array_equal(uniq.mat.row, flat // n_items)
    np.testing.assert_array_equal(uniq.mat.col, flat % n_items)


# ------------------------------------------------------------------------------------ epoch plan
@pytest.mark.gpu
@pytest.mark.parametrize('n,batch,n_users,with_gpos', [(10_000, 256, 300, False),
                                                       (100_003, 4096, 70_000, True),
                                                       (777, 1000, 5, False)])
def test_epoch_plan_properties(n, batch, n_users, with_gpos):
    """cmf_epoch_plan (loader): a permutation of the dataset, cut into batches ordered by user,
    reproducible per (seed, epoch) and different across epochs (samplers.py:51-60 shuffles on the
    host; the order inside a batch is free there)."""
    import collie_b200 as cb
    rng = np.random.default_rng(n)
    users = rng.integers(0, n_users, n).astype(np.int32)
    items = rng.integers(0, 1000, n).astype(np.int32)
    data = cb.Interactions(users=users, items=items, num_users=n_users, num_items=1000,
                           allow_missing_ids=True, remove_duplicate_user_item_pairs=False,
                           check_num_negative_samples_is_valid=False,
                           max_number_of_samples_to_consider=0, seed=11)
    d_users, d_items = data.mat.row.astype(np.int32), data.mat.col.astype(np.int32)
    gpos = None
    if with_gpos:
        gpos = rng.permutation(10 * n)[:n].astype(np.int64)
        data.global_positions = gpos
    loader = cb.InteractionsDataLoader(data, batch_size=batch, shuffle=True, drop_last=False)

    def epoch(e, sort=True):
        bs = list(loader.device_batches('cuda', epoch=e, sort_within_batch=sort))
        return (torch.cat([b.users for b in bs]).cpu().numpy(),
                torch.cat([b.items for b in bs]).cpu().numpy(),
                torch.cat([b.gidx for b in bs]).cpu().numpy(), bs)

    u, i, g, bs = epoch(3)
    assert len(bs) == (n + batch - 1) // batch and all(b.sorted for b in bs)
    pos = g if gpos is None else np.argsort(gpos)[np.searchsorted(np.sort(gpos), g)]
    assert np.array_equal(np.sort(pos), np.arange(n))            # a permutation of the dataset
    assert np.array_equal(u, d_users[pos]) and np.array_equal(i, d_items[pos])
    for s in range(0, n, batch):                                 # every batch ordered by user
        assert np.all(np.diff(u[s:s + batch]) >= 0)
    u2, i2, g2, _ = epoch(3)
    assert np.array_equal(g, g2)                                 # same (seed, epoch), same plan
    _, _, g3, _ = epoch(4)
    assert n <= batch or not np.array_equal(g, g3)             # (one batch: nothing to shuffle)
    # equal users keep their dataset order inside a batch
    for s in range(0, n, batch):
        uu, pp = u[s:s + batch], pos[s:s + batch]
        same = uu[1:] == uu[:-1]
        assert np.all(pp[1:][same] > pp[:-1][same])
    # every batch is a spread-out sample of the dataset, and different batches hold different
    # users' interactions in similar proportion
    if n >= 10_000:
        first = pos[:batch]
        assert abs(first.mean() / n - 0.5) < 0.05
        assert abs(np.mean(d_users[pos[:batch]]) - d_users.mean()) < 0.1 * n_users
    # without the in-batch sort (reference-format iteration): a permutation in its own order
    _, _, g4, bs4 = epoch(3, sort=False)
    assert not any(b.sorted for b in bs4)
    p4 = g4 if gpos is None else np.argsort(gpos)[np.searchsorted(np.sort(gpos), g4)]
    assert np.array_equal(np.sort(p4), np.arange(n))
    assert np.mean(p4 == np.arange(n)) < 0.01 + 2.0 / n
    if n >= 10_000:
        assert abs(p4[:batch].mean() / n - 0.5) < 0.05
    # unshuffled: dataset order, batch by batch
    plain = cb.InteractionsDataLoader(data, batch_size=batch, shuffle=False)
    gp = torch.cat([b.gidx for b in plain.device_batches('cuda', epoch=0,
                                                         sort_within_batch=False)]).cpu().numpy()
    assert np.array_equal(gp, np.arange(n) if gpos is None else gpos)


# ------------------------------------------------------------------------------------ similarity
@pytest.mark.gpu
@pytest.mark.parametrize('name', ['similarity_random', 'similarity_d128'])
def test_golden_similarity(name):
    """item_item_similarity / user_user_similarity equal the reference's series
    (base_pipeline.py:783-859): same values per id, sorted descending, the seed first."""
    import collie_b200 as cb
    from oracle import collie_oracle as O
    g = np.load(os.path.join(GOLDEN, name + '.npz'))
    n_users, n_items, dim = int(g['n_users']), int(g['n_items']), int(g['dim'])
    data = cb.Interactions(users=np.arange(n_users) % n_users, items=np.arange(n_users) % n_items,
                           num_users=n_users, num_items=n_items, allow_missing_ids=True,
                           check_num_negative_samples_is_valid=False,
                           max_number_of_samples_to_consider=0)
    model = cb.MatrixFactorizationModel(train=data, embedding_dim=dim,
                                        lr_scheduler_func=None).cuda()
    with torch.no_grad():
        for k, t in _tables(model).items():
            t.copy_(torch.from_numpy(g[k]))
    for i in g['item_ids']:
        got = model.item_item_similarity(int(i))
        assert got.index[0] == i and np.all(np.diff(got.values) <= 0)
        np.testing.assert_allclose(got.sort_index().values, g[f'item_sim_{i}'], rtol=1e-5,
                                   atol=1e-6)
        np.testing.assert_allclose(got.sort_index().values,
                                   O.embedding_similarity(torch.from_numpy(g['I']), int(i)).numpy(),
                                   rtol=1e-5, atol=1e-6)
    for u in g['user_ids']:
        got = model.user_user_similarity(int(u))
        assert got.index[0] == u and len(got) == n_users
        np.testing.assert_allclose(got.sort_index().values, g[f'user_sim_{u}'], rtol=1e-5,
                                   atol=1e-6)
    with pytest.raises(ValueError, match='is not in the model'):
        model.item_item_similarity(n_items)
    with pytest.raises(ValueError, match='is not in the model'):
        model.user_user_similarity(n_users)


@pytest.mark.gpu
def test_batch_sorter_orders_by_user_and_keeps_rows_together():
    """BatchSorter (host-fed pipelines): a stable sort of one batch by user on the current stream;
    a step on the sorted batch equals the step on the unsorted one."""
    import collie_b200 as cb
    from collie_b200.interactions import BatchSorter, DeviceBatch
    rng = np.random.default_rng(3)
    n_users, n_items, n = 500, 300, 4000
    users = rng.integers(0, n_users, n).astype(np.int32)
    items = rng.integers(0, n_items, n).astype(np.int32)
    gidx = rng.permutation(10 * n)[:n].astype(np.int64)
    u, i, g = (torch.from_numpy(x).cuda() for x in (users, items, gidx))
    sorter = BatchSorter(n + 7, n_users, 'cuda')
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        db = sorter(u, i, g)
    torch.cuda.current_stream().wait_stream(side)
    assert db.sorted and db.ratings is None
    order = np.argsort(users, kind='stable')
    assert np.array_equal(db.users.cpu().numpy(), users[order])
    assert np.array_equal(db.items.cpu().numpy(), items[order])
    assert np.array_equal(db.gidx.cpu().numpy(), gidx[order])
    with pytest.raises(ValueError, match='exceeds this sorter'):
        sorter(torch.cat([u, u]), torch.cat([i, i]), torch.cat([g, g]))
    # the same training step either way (negatives are keyed by gidx, not by batch position)
    data = cb.Interactions(users=users, items=items, num_users=n_users, num_items=n_items,
                           num_negative_samples=3, allow_missing_ids=True, seed=9,
                           remove_duplicate_user_item_pairs=False,
                           check_num_negative_samples_is_valid=False,
                           max_number_of_samples_to_consider=0)
    losses, tables = [], []
    for batch in (DeviceBatch(u, i, g, None), db):
        torch.manual_seed(0)
        model = cb.MatrixFactorizationModel(train=data, embedding_dim=32, sparse=True, lr=1e-2,
                                            optimizer='sparse_adam', loss='bpr',
                                            lr_scheduler_func=None).cuda()
        opts = model.configure_optimizers()
        losses.append(model.fused_step(batch, optimizers=opts, interactions=data, epoch=0).item())
        tables.append(model.item_embeddings.weight.detach().cpu().numpy().copy())
    np.testing.assert_allclose(losses[0], losses[1], rtol=1e-6)
    np.testing.assert_allclose(tables[0], tables[1], rtol=1e-4, atol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize('explicit', [False, True])
def test_whole_epoch_call_equals_the_step_loop(explicit):
    """cmf_train_epoch (one call per epoch, trainer.py:451-478's loop moved into the library) =
    the same steps issued one by one: per-step losses, tables, optimizer step counters."""
    import collie_b200 as cb
    rng = np.random.default_rng(12)
    n_users, n_items, n, bsz = 300, 200, 5000, 512            # 9 full batches + a ragged one
    users, items = rng.integers(0, n_users, n), rng.integers(0, n_items, n)
    if explicit:
        data = cb.ExplicitInteractions(users=users, items=items,
                                       ratings=rng.integers(1, 6, n).astype(np.float32),
                                       num_users=n_users, num_items=n_items,
                                       allow_missing_ids=True,
                                       remove_duplicate_user_item_pairs=False)
        kw = dict(loss='mse', optimizer='adam', sparse=False)
    else:
        data = cb.Interactions(users=users, items=items, num_users=n_users, num_items=n_items,
                               num_negative_samples=4, allow_missing_ids=True, seed=2,
                               remove_duplicate_user_item_pairs=False,
                               check_num_negative_samples_is_valid=False,
                               max_number_of_samples_to_consider=0)
        kw = dict(loss='bpr', optimizer='sparse_adam', sparse=True)
    loader = cb.InteractionsDataLoader(data, batch_size=bsz, shuffle=True, drop_last=False)

    def fresh():
        torch.manual_seed(4)
        m = cb.MatrixFactorizationModel(train=loader, embedding_dim=24, lr=1e-2,
                                        lr_scheduler_func=None, **kw).cuda()
        return m, m.configure_optimizers()
    a, oa = fresh()
    b, ob = fresh()
    for epoch in (1, 2):        # the trainers number epochs from 1
        step_losses = [a.fused_step(db, optimizers=oa, interactions=data, epoch=epoch)
                       for db in loader.device_batches('cuda', epoch=epoch)]
        plan = loader.epoch_plan('cuda', epoch=epoch)
        epoch_losses = b.fused_step(plan, optimizers=ob, interactions=data, epoch=epoch,
                                    epoch_batch_size=bsz)
        assert epoch_losses.shape == (10,)
        np.testing.assert_allclose(epoch_losses.cpu().numpy(),
                                   torch.stack(step_losses).cpu().numpy(), rtol=2e-5)
    for ta, tb in zip(_tables(a).values(), _tables(b).values()):
        np.testing.assert_allclose(tb.detach().cpu().numpy(), ta.detach().cpu().numpy(),
                                   rtol=1e-4, atol=1e-6, err_msg='epoch call vs step loop')
    opt_a = oa[0] if isinstance(oa, (list, tuple)) else oa
    opt_b = ob[0] if isinstance(ob, (list, tuple)) else ob
    pa, pb = a.user_embeddings.weight, b.user_embeddings.weight
    assert opt_a.state_for(pa)['step'] == opt_b.state_for(pb)['step'] == 20
    # the trainer takes the one-call route when nothing has to be logged per step
    c, _ = fresh()
    t = cb.CollieTrainer(c, max_epochs=2, verbosity=0, enable_model_summary=False)
    t.fit(c)
    assert t.train_steps == 20
    for ta, tc in zip(_tables(a).values(), _tables(c).values()):
        np.testing.assert_allclose(tc.detach().cpu().numpy(), ta.detach().cpu().numpy(),
                                   rtol=1e-4, atol=1e-6)


# ------------------------------------------------------------------------------------ size limits
@pytest.mark.gpu
@pytest.mark.parametrize('dim,n_neg,loss,oname', [(256, 3, 'bpr', 'adaptive_bpr'),
                                                  (1024, 2, 'adaptive', 'adaptive_hinge'),
                                                  (64, 64, 'warp', 'warp'),
                                                  (12, 64, 'bpr', 'adaptive_bpr')])
def test_steps_at_the_size_limits_match_oracle(dim, n_neg, loss, oname):
    """The widest rows (embedding_dim 1024), rows wider than one 128-float slice, and the largest
    negative count (64) go through other kernel instantiations than the golden cases: two
    sparse-Adam steps against the oracle."""
    import collie_b200 as cb
    from oracle import collie_oracle as O
    rng = np.random.default_rng(dim + n_neg)
    n_users, n_items, batch = 90, 150, 700
    data = _implicit_data(n_users, n_items, n_neg)
    torch.manual_seed(dim)
    model = cb.MatrixFactorizationModel(train=data, embedding_dim=dim, sparse=True, lr=1e-2,
                                        optimizer='sparse_adam', loss=loss,
                                        lr_scheduler_func=None).cuda()
    ref = O.OracleMF(n_users, n_items, dim, sparse=True)
    names = ('user_embeddings', 'item_embeddings', 'user_biases', 'item_biases')
    with torch.no_grad():
        for name in names:
            w = torch.from_numpy(rng.normal(0, 0.4 / np.sqrt(dim) if 'emb' in name else 0.1,
                                            tuple(getattr(model, name).weight.shape))
                                 .astype(np.float32))
            getattr(model, name).weight.copy_(w)
            getattr(ref, name).weight.copy_(w)
    opts = model.configure_optimizers()
    ref_opts = O.make_optimizers(ref, 'sparse_adam', 'sgd', 1e-2, 1e-2)
    for step in range(2):
        bu = torch.from_numpy(rng.integers(0, n_users, batch).astype(np.int32))
        bp = torch.from_numpy(rng.integers(0, n_items, batch).astype(np.int32))
        bn = torch.from_numpy(rng.integers(0, n_items, (batch, n_neg)))
        got = model.fused_step(((bu.cuda(), bp.cuda()), bn.cuda()), optimizers=opts).item()
        exp = O.train_step(ref, ref_opts, lambda: ref.implicit_loss(bu, bp, bn, oname))
        np.testing.assert_allclose(got, exp, rtol=2e-5)
    for name in names:
        np.testing.assert_allclose(getattr(model, name).weight.detach().cpu().numpy(),
                                   getattr(ref, name).weight.detach().numpy(), rtol=2e-4,
                                   atol=5e-5, err_msg=name)   # Adam divides by sqrt(v): where a
    # row's contributions nearly cancel, the summation order shows up as a fraction of lr = 1e-2


@pytest.mark.gpu
@pytest.mark.parametrize('dim', [10, 128, 200])
def test_score_tile_kernels_agree_bit_for_bit(dim):
    """The 128 x 128 contraction (full tiles) and the 64 x 64 one (small tiles) accumulate every
    score in the same order: identical bits; both equal the oracle's scores to fp32 tolerance."""
    import collie_b200 as cb
    from oracle import collie_oracle as O
    rng = np.random.default_rng(dim)
    n_users, n_items = 230, 333
    model = cb.MatrixFactorizationModel(train=_implicit_data(n_users, n_items, 3),
                                        embedding_dim=dim, lr_scheduler_func=None).cuda()
    with torch.no_grad():
        for p in model.parameters():
            p.copy_(torch.from_numpy(rng.normal(0, 0.5, tuple(p.shape)).astype(np.float32)))
    users = rng.permutation(n_users)
    whole = cb.get_preds(model, users, n_items)                       # 230 users: 128-wide tiles
    parts = torch.cat([cb.get_preds(model, users[s:s + 50], n_items)  # 50 users: 64-wide tiles
                       for s in range(0, n_users, 50)])
    assert torch.equal(whole, parts)
    tabs = [t.detach().cpu() for t in _tables(model).values()]
    uu = torch.from_numpy(users).repeat_interleave(n_items)
    ii = torch.arange(n_items).repeat(n_users)
    exp = O.mf_forward(*tabs, uu, ii).view(n_users, n_items)
    np.testing.assert_allclose(whole.cpu().numpy(), exp.numpy(), rtol=1e-5, atol=1e-5)


# ------------------------------------------------------------------------------------ memory safety
def _guarded(nbytes, pad=4096):
    """``nbytes`` of device scratch with canary bytes on both sides: (whole buffer, interior view)."""
    whole = torch.full((nbytes + 2 * pad,), 0xA5, dtype=torch.uint8, device='cuda')
    return whole, whole[pad:pad + nbytes], pad


def _intact(whole, pad):
    return bool((whole[:pad] == 0xA5).all()) and bool((whole[-pad:] == 0xA5).all())


@pytest.mark.gpu
def test_kernels_stay_inside_their_workspaces():
    """Every scratch area is handed over at exactly the size the library asks for, with canary
    bytes around it: training step (split kernels and the generic kernel), epoch plan (both
    forms), exact and tensor-core evaluation tiles. (Stands in for a memory checker.)"""
    import collie_b200 as cb
    from collie_b200 import _abi, metrics as M
    lib = _abi.lib()
    rng = np.random.default_rng(77)
    n_users, n_items, n = 333, 517, 6000
    users, items = rng.integers(0, n_users, n), rng.integers(0, n_items, n)
    for dim, n_neg in ((128, 5), (10, 3), (260, 2)):
        data = cb.Interactions(users=users, items=items, num_users=n_users, num_items=n_items,
                               num_negative_samples=n_neg, allow_missing_ids=True, seed=1,
                               remove_duplicate_user_item_pairs=False,
                               check_num_negative_samples_is_valid=False,
                               max_number_of_samples_to_consider=0)
        loader = cb.InteractionsDataLoader(data, batch_size=1000, shuffle=True)
        model = cb.MatrixFactorizationModel(train=loader, embedding_dim=dim, sparse=True,
                                            optimizer='sparse_adam', loss='bpr',
                                            lr_scheduler_func=None).cuda()
        opts = model.configure_optimizers()
        # ---- epoch plan workspace (shuffled + ordered by user: the presorted form)
        need = int(lib.cmf_epoch_plan_workspace_bytes(n))
        pw, pv, pad = _guarded(need)
        loader._plan_ws = {'cuda:0': pv, str(torch.device('cuda', 0)): pv}
        plan = loader.epoch_plan(torch.device('cuda', 0), epoch=1)
        torch.cuda.synchronize()
        assert _intact(pw, pad) and plan.users.shape[0] == n
        # ---- training workspace, whole epoch through cmf_train_epoch (ragged last batch)
        need = int(lib.cmf_train_workspace_bytes(1000, n_neg, n_users, n_items, dim))
        tw, tv, pad = _guarded(need)
        _abi.check(lib.cmf_workspace_init(_abi.ptr(tv), need, _abi.stream_ptr('cuda')), 'init')
        model._workspace, model._workspace_B = tv, 1000
        losses = model.fused_step(plan, optimizers=opts, interactions=data, epoch=1,
                                  epoch_batch_size=1000)
        # an unsorted batch too: the in-step sort uses the per-step part of the workspace
        st = data.device_state('cuda')
        from collie_b200.interactions import DeviceBatch
        idx = torch.arange(700, device='cuda')
        model.fused_step(DeviceBatch(st['users'][:700], st['items'][:700], idx, None),
                         optimizers=opts, interactions=data, epoch=2)
        torch.cuda.synchronize()
        assert model._workspace is tv and _intact(tw, pad)
        assert bool(torch.isfinite(losses).all()) and losses.shape == (6,)
        # ---- the general plan form (no shuffle)
        plain = cb.InteractionsDataLoader(data, batch_size=1000, shuffle=False)
        pw2, pv2, pad = _guarded(int(lib.cmf_epoch_plan_workspace_bytes(n)))
        plain._plan_ws = {'cuda:0': pv2}
        plain.epoch_plan(torch.device('cuda', 0), epoch=0)
        torch.cuda.synchronize()
        assert _intact(pw2, pad)
        # ---- evaluation tiles
        tabs = M._model_tables(model)
        import scipy.sparse as sp
        coo = sp.coo_matrix((np.ones(n), (users, items)), shape=(n_users, n_items))
        coo.sum_duplicates()
        tile = np.arange(200)
        indptr, tgt, nnz = M._targets_device(coo.tocsr(), tile, 'cuda')
        uid = torch.arange(200, dtype=torch.int32, device='cuda')
        for precision in ((0, 1) if dim <= 128 else (0,)):
            need = int(lib.cmf_eval_workspace_bytes(200, n_items, dim, 7, precision))
            ew, ev, pad = _guarded(need)
            want = {'topk', 'mapk', 'mrr'} | ({'auc'} if precision == 0 else set())
            res = M._eval_tile('cuda', 200, n_items, 7, indptr, tgt, want, tables=tabs,
                               user_ids=uid, precision=precision, workspace=ev)
            torch.cuda.synchronize()
            assert res['_workspace'] is ev and _intact(ew, pad)
            assert int(res['topk_ids'].min()) >= 0 and int(res['topk_ids'].max()) < n_items


@pytest.mark.gpu
def test_batch_sorter_carries_ratings():
    """Explicit batches: the ratings travel with their interactions through the sorter."""
    from collie_b200.interactions import BatchSorter
    rng = np.random.default_rng(8)
    n = 3000
    users = rng.integers(0, 50, n).astype(np.int32)
    items = rng.integers(0, 90, n).astype(np.int32)
    ratings = rng.integers(1, 6, n).astype(np.float32)
    gidx = np.arange(n, dtype=np.int64) + 10
    sorter = BatchSorter(n, 50, 'cuda', with_ratings=True)
    db = sorter(*(torch.from_numpy(x).cuda() for x in (users, items, gidx)),
                ratings=torch.from_numpy(ratings).cuda())
    order = np.argsort(users, kind='stable')
    assert db.sorted and np.array_equal(db.users.cpu().numpy(), users[order])
    assert np.array_equal(db.items.cpu().numpy(), items[order])
    assert np.array_equal(db.ratings.cpu().numpy(), ratings[order])
    assert np.array_equal(db.gidx.cpu().numpy(), gidx[order])
    with pytest.raises(ValueError, match='ratings'):
        sorter(*(torch.from_numpy(x).cuda() for x in (users, items, gidx)))
