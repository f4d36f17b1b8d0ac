"""Pin the oracle: reference known answers, golden vectors made by the unmodified reference
(oracle/make_golden.py), and the third-party torch.optim algorithms. CPU only."""
import glob
import json
import os

import numpy as np
import pytest
import scipy.sparse as sp
import torch

from oracle import collie_oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
KA = json.load(open(os.path.join(GOLDEN, 'reference_known_answers.json')))


def _scores():
    s = torch.tensor(KA['loss_scores'])
    pos_items = torch.tensor(KA['positive_items'])
    neg_items = torch.tensor(KA['negative_items'])
    many = torch.tensor(KA['many_negative_items'])
    return s, pos_items, neg_items, many


def test_reference_loss_known_answers():
    s, pi, ni, many = _scores()
    ans = KA['losses']
    meta_a = {'a': torch.tensor(KA['metadata_a'])}
    meta_ab = {'a': torch.tensor(KA['metadata_a']), 'b': torch.tensor(KA['metadata_b'])}
    got = {
        'bpr': O.bpr(s[pi], s[ni]),
        'hinge': O.hinge(s[pi], s[ni]),
        'adaptive_bpr': O.adaptive_bpr(s[pi], s[many]),
        'adaptive_hinge': O.adaptive_hinge(s[pi], s[many]),
        'warp_num_items_4': O.warp(s[pi], s[many], 4),
        'bpr_meta_a02': O.bpr(s[pi], s[ni], pi, ni, meta_a, {'a': 0.2}),
        'hinge_meta_a02': O.hinge(s[pi], s[ni], pi, ni, meta_a, {'a': 0.2}),
        'adaptive_bpr_meta_a02_b03': O.adaptive_bpr(s[pi], s[many], pi, many, meta_ab,
                                                    {'a': 0.2, 'b': 0.3}),
        'adaptive_hinge_meta_a02_b03': O.adaptive_hinge(s[pi], s[many], pi, many, meta_ab,
                                                        {'a': 0.2, 'b': 0.3}),
        'warp_meta_a02_b03_num_items_4': O.warp(s[pi], s[many], 4, pi, many, meta_ab,
                                                {'a': 0.2, 'b': 0.3}),
    }
    for name, expected in ans.items():
        np.testing.assert_almost_equal(got[name].item(), expected, decimal=3, err_msg=name)


def test_reference_ideal_difference_known_answers():
    _, pi, ni, many = _scores()
    a = O.ideal_difference(pi, ni, {'a': torch.tensor(KA['metadata_a'])}, {'a': 0.2})
    np.testing.assert_allclose(a.numpy(), KA['ideal_difference_a02'], rtol=1e-6)
    ab = O.ideal_difference(pi.repeat(4, 1), many,
                            {'a': torch.tensor(KA['metadata_a']),
                             'b': torch.tensor(KA['metadata_b'])}, {'a': 0.2, 'b': 0.3})
    np.testing.assert_allclose(ab.numpy(), KA['ideal_difference_a02_b03_many'], rtol=1e-6)
    with pytest.raises(ValueError):
        O.ideal_difference(pi, ni, {'a': torch.tensor(KA['metadata_a'])}, {'a': 1.2})


def _metric_fixture():
    users, items = np.array(KA['metrics_users']), np.array(KA['metrics_items'])
    coo = sp.coo_matrix((np.ones(len(users)), (users, items)), shape=(3, 4))
    return coo, np.array(KA['metrics_scores'], dtype=np.float32)


def test_reference_metric_known_answers():
    coo, scores = _metric_fixture()
    csr = coo.tocsr()
    assert O.mapk_per_user(csr, [1, 2], scores[[1, 2]], 4).mean() == KA['mapk_k4_users_1_2']
    np.testing.assert_almost_equal(O.mapk_per_user(csr, [0, 1, 2], scores, 1).mean(), KA['mapk_k1'])
    np.testing.assert_almost_equal(O.mrr_per_user(csr, [0, 1, 2], scores).mean(), KA['mrr'])
    with pytest.raises(ValueError):
        O.mapk_per_user(csr, [0, 1, 2], scores, 5)
    from sklearn.metrics import roc_auc_score
    exp = np.mean([roc_auc_score(csr[i].toarray()[0], scores[i]) for i in range(3)])
    np.testing.assert_almost_equal(O.auc_per_user(csr, [0, 1, 2], scores).mean(), exp)


class _ScoreTables:
    """tables whose forward reproduces a fixed score matrix: U = one-hot, I = scores^T."""
    def __init__(self, scores):
        n_u, n_i = scores.shape
        self.t = (torch.eye(n_u), torch.from_numpy(scores.T.copy()), torch.zeros(n_u, 1),
                  torch.zeros(n_i, 1))


@pytest.mark.parametrize('batch_size', [20, 2, 1])
def test_reference_evaluate_in_batches_known_answers(batch_size):
    coo, scores = _metric_fixture()
    got = O.evaluate_in_batches(['mapk', 'mrr', 'auc'], coo, _ScoreTables(scores).t, k=4,
                                batch_size=batch_size)
    exp = KA['evaluate_in_batches_k4']
    np.testing.assert_almost_equal(got[0], exp['mapk'])
    np.testing.assert_almost_equal(got[1], exp['mrr'])
    np.testing.assert_almost_equal(got[2], exp['auc'])


# ---------------------------------------------------------------- golden vectors from the reference

IMPLICIT = sorted(glob.glob(os.path.join(GOLDEN, 'implicit_*.npz')))
EXPLICIT = sorted(glob.glob(os.path.join(GOLDEN, 'explicit_*.npz')))
METRICS = sorted(glob.glob(os.path.join(GOLDEN, 'metrics_*.npz')))


def _optim_names(path):
    base = os.path.basename(path)
    for name in ('sparse_adam', 'adagrad', 'adam', 'sgd'):
        if name in base:
            return name
    return 'adam' if 'd10' in base or 'meta_n6' in base else 'sparse_adam'


def _load_model(g, sparse):
    m = O.OracleMF(int(g['n_users']), int(g['n_items']), int(g['dim']), sparse=sparse)
    with torch.no_grad():
        m.user_embeddings.weight.copy_(torch.from_numpy(g['init_U']))
        m.item_embeddings.weight.copy_(torch.from_numpy(g['init_I']))
        m.user_biases.weight.copy_(torch.from_numpy(g['init_bu']))
        m.item_biases.weight.copy_(torch.from_numpy(g['init_bi']))
    return m


def _meta(g):
    meta = {k[5:]: torch.from_numpy(g[k]) for k in g.files if k.startswith('meta_')}
    return (meta, {'a': 0.2, 'b': 0.3}) if meta else (None, None)


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


@pytest.mark.parametrize('path', IMPLICIT, ids=[os.path.basename(p)[:-4] for p in IMPLICIT])
def test_golden_implicit_steps(path):
    g = np.load(path)
    opt_name = GOLDEN_OPT[os.path.basename(path)[:-4]]
    model = _load_model(g, sparse=(opt_name == 'sparse_adam'))
    opts = O.make_optimizers(model, opt_name, 'sgd', float(g['lr']), float(g['bias_lr']))
    meta, w = _meta(g)
    loss_name = str(g['loss_name']).replace('_loss', '')
    for s in range(int(g['steps'])):
        users = torch.from_numpy(g[f'step{s}_users'])
        pos = torch.from_numpy(g[f'step{s}_pos'])
        negs = torch.from_numpy(g[f'step{s}_negs'])
        if s == 0:
            np.testing.assert_allclose(model(users, pos).detach().numpy(), g['step0_pos_scores'],
                                       rtol=1e-6, atol=1e-6)
        for o in opts:
            o.zero_grad()
        loss = model.implicit_loss(users, pos, negs, loss_name, meta, w)
        loss.backward()
        np.testing.assert_allclose(loss.item(), float(g[f'step{s}_loss']), rtol=1e-6)
        if s == 0:
            for name, p in zip(('U', 'I', 'bu', 'bi'), model.tables()):
                gr = p.grad.to_dense() if p.grad.is_sparse else p.grad
                np.testing.assert_allclose(gr.numpy(), g['step0_grad_' + name], rtol=1e-5,
                                           atol=1e-7, err_msg=name)
        for o in opts:
            o.step()
    for name, p in zip(('U', 'I', 'bu', 'bi'), model.tables()):
        np.testing.assert_allclose(p.detach().numpy(), g['final_' + name], rtol=1e-5, atol=1e-6,
                                   err_msg=name)


@pytest.mark.parametrize('path', EXPLICIT, ids=[os.path.basename(p)[:-4] for p in EXPLICIT])
def test_golden_explicit_steps(path):
    g = np.load(path)
    opt_name = GOLDEN_OPT[os.path.basename(path)[:-4]]
    model = _load_model(g, sparse=(opt_name == 'sparse_adam'))
    opts = O.make_optimizers(model, opt_name, 'sgd', float(g['lr']), float(g['bias_lr']))
    yr = None if np.isnan(g['y_range']).any() else tuple(g['y_range'].tolist())
    loss_name = 'mae' if 'mae' in path else 'mse'
    for s in range(int(g['steps'])):
        for o in opts:
            o.zero_grad()
        loss = model.explicit_loss(torch.from_numpy(g[f'step{s}_users']),
                                   torch.from_numpy(g[f'step{s}_items']),
                                   torch.from_numpy(g[f'step{s}_ratings']), loss_name, yr)
        loss.backward()
        np.testing.assert_allclose(loss.item(), float(g[f'step{s}_loss']), rtol=1e-6)
        for o in opts:
            o.step()
    for name, p in zip(('U', 'I', 'bu', 'bi'), model.tables()):
        np.testing.assert_allclose(p.detach().numpy(), g['final_' + name], rtol=1e-5, atol=1e-6,
                                   err_msg=name)


@pytest.mark.parametrize('path', METRICS, ids=[os.path.basename(p)[:-4] for p in METRICS])
def test_golden_metrics(path):
    g = np.load(path)
    coo = sp.coo_matrix((np.ones(len(g['test_users'])), (g['test_users'], g['test_items'])),
                        shape=(int(g['n_users']), int(g['n_items'])))
    tables = tuple(torch.from_numpy(g[k]) for k in ('U', 'I', 'bu', 'bi'))
    has_ties = 'exact_grid' in path
    for k in (1, 4, 10):
        for bs in (20, 7):
            got = O.evaluate_in_batches(['mapk', 'mrr', 'auc'], coo, tables, k=k, batch_size=bs)
            exp = g[f'k{k}_bs{bs}']
            # AUC does not depend on tie order; MAP/MRR do (torch.topk's tie order is unpinned in
            # the reference), so on the tie-heavy fixture only AUC is compared exactly
            np.testing.assert_allclose(got[2], exp[2], rtol=1e-6)
            if not has_ties:
                np.testing.assert_allclose(got[:2], exp[:2], rtol=1e-6)


@pytest.mark.parametrize('name', ['similarity_random', 'similarity_d128'])
def test_golden_similarity(name):
    """item_item_similarity / user_user_similarity of the reference (base_pipeline.py:783-859)."""
    g = np.load(os.path.join(GOLDEN, name + '.npz'))
    U, I = torch.from_numpy(g['U']), torch.from_numpy(g['I'])
    for i in g['item_ids']:
        np.testing.assert_allclose(O.embedding_similarity(I, int(i)).numpy(), g[f'item_sim_{i}'],
                                   rtol=1e-6, atol=1e-7)
    for u in g['user_ids']:
        np.testing.assert_allclose(O.embedding_similarity(U, int(u)).numpy(), g[f'user_sim_{u}'],
                                   rtol=1e-6, atol=1e-7)


# ---------------------------------------------------------------- third-party optimizer algorithms

def test_optimizer_formulas_match_torch_optim():
    rng = np.random.default_rng(5)
    p0 = torch.from_numpy(rng.normal(size=(9, 6)).astype(np.float32))
    # SparseAdam: rows 1,4,4,7 gathered; row 7 has an all-zero gradient but must still move
    pt = torch.nn.Parameter(p0.clone())
    opt = torch.optim.SparseAdam([pt], lr=0.01)
    p, m, v = p0.clone(), torch.zeros_like(p0), torch.zeros_like(p0)
    for step in range(1, 5):
        rows = torch.tensor([1, 4, 4, 7])
        vals = torch.from_numpy(rng.normal(size=(4, 6)).astype(np.float32))
        vals[3] = 0
        pt.grad = torch.sparse_coo_tensor(rows.view(1, -1), vals, p0.shape)
        opt.step()
        uniq = torch.tensor([1, 4, 7])
        g = torch.stack([vals[0], vals[1] + vals[2], vals[3]])
        O.sparse_adam_rows(p, m, v, uniq, g, step, 0.01)
    np.testing.assert_allclose(p.numpy(), pt.detach().numpy(), rtol=2e-6, atol=1e-7)
    # dense Adam with weight decay, SGD
    pt = torch.nn.Parameter(p0.clone())
    opt = torch.optim.Adam([pt], lr=0.02, weight_decay=0.1)
    p, m, v = p0.clone(), torch.zeros_like(p0), torch.zeros_like(p0)
    for step in range(1, 5):
        g = torch.from_numpy(rng.normal(size=(9, 6)).astype(np.float32))
        pt.grad = g.clone()
        opt.step()
        O.dense_adam(p, m, v, g, step, 0.02, weight_decay=0.1)
    np.testing.assert_allclose(p.numpy(), pt.detach().numpy(), rtol=2e-6, atol=1e-7)


# ---------------------------------------------------------------- sampler contract

def test_philox_known_answer():
    # Random123 known-answer vectors for philox4x32-10
    assert O.philox4x32_10((0, 0, 0, 0), (0, 0)) == (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)
    assert O.philox4x32_10((0xffffffff,) * 4, (0xffffffff, 0xffffffff)) == (
        0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)
    assert O.philox4x32_10((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344),
                           (0xa4093822, 0x299f31d0)) == (
        0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)


def test_sampler_contract():
    """Properties the reference tests pin (tests/test_interactions.py:356-456): shapes, range,
    exact negatives never hit the user's positives and are distinct, approximate = plain draws."""
    rng = np.random.default_rng(0)
    n_users, n_items, n_neg = 11, 40, 6
    dense = rng.random((n_users, n_items)) < 0.3
    csr = sp.csr_matrix(dense.astype(np.float32))
    users = rng.integers(0, n_users, size=50)
    gidx = np.arange(50) + 1000
    approx = O.philox_negatives(7, 1, gidx, users, n_neg, n_items)
    assert approx.shape == (50, n_neg) and approx.min() >= 0 and approx.max() < n_items
    exact = O.philox_negatives(7, 1, gidx, users, n_neg, n_items, csr.indptr, csr.indices)
    for r, u in enumerate(users):
        assert not dense[u, exact[r]].any()
        assert len(set(exact[r].tolist())) == n_neg
    # same counters -> same stream; another epoch -> another stream
    assert (O.philox_negatives(7, 1, gidx, users, n_neg, n_items) == approx).all()
    assert (O.philox_negatives(7, 2, gidx, users, n_neg, n_items) != approx).any()
    # exhausted budget: unconditioned fill, still N ids in range
    full = sp.csr_matrix(np.ones((1, 5), dtype=np.float32))
    out = O.philox_negatives(1, 0, np.array([0]), np.array([0]), 3, 5, full.indptr, full.indices,
                             max_tries=4)
    assert out.shape == (1, 3) and out.max() < 5
