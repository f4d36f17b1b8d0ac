"""``CollaborativeMetricLearningModel`` against golden files produced by the unmodified reference
(``oracle/make_golden.py cml``): predictions (pairwise distances), the losses of three iterations
of the reference's own loop with injected negatives, both tables afterwards.

Reference: collie/model/collaborative_metric_learning.py:86-124."""
import os

import numpy as np
import pytest
import torch

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden')
CASES = {'cml_hinge_adam': dict(optimizer='adam', sparse=False, loss='hinge'),
         'cml_adaptive_sparse_adam': dict(optimizer='sparse_adam', sparse=True,
                                          loss='adaptive_hinge'),
         'cml_bpr_sgd': dict(optimizer='sgd', sparse=False, loss='bpr')}


def test_cml_goldens_are_present_and_shaped():
    for name in CASES:
        g = np.load(os.path.join(GOLD, name + '.npz'))
        assert g['init_user_embeddings.weight'].shape == (int(g['n_users']), int(g['dim']))
        assert g['preds'].shape == (80,)
        assert sorted(k for k in g.files if k.startswith('final_')) == \
            ['final_item_embeddings.weight', 'final_user_embeddings.weight']


@pytest.mark.gpu
@pytest.mark.parametrize('name', sorted(CASES))
def test_cml_predictions_losses_and_steps_match_the_reference(name):
    import collie_b200 as cb
    g = np.load(os.path.join(GOLD, name + '.npz'))
    n_users, n_items, dim, n_neg = (int(g[k]) for k in ('n_users', 'n_items', 'dim', 'n_neg'))
    rng = np.random.default_rng(0)
    train = cb.Interactions(users=rng.integers(0, n_users, 100), items=rng.integers(0, n_items, 100),
                            num_users=n_users, num_items=n_items, num_negative_samples=n_neg,
                            allow_missing_ids=True, remove_duplicate_user_item_pairs=False,
                            check_num_negative_samples_is_valid=False,
                            max_number_of_samples_to_consider=0)
    model = cb.CollaborativeMetricLearningModel(train=train, embedding_dim=dim, lr=float(g['lr']),
                                                lr_scheduler_func=None, **CASES[name]).cuda()
    assert sorted(model.state_dict()) == ['item_embeddings.weight', 'user_embeddings.weight']
    assert 'bias_lr' not in model.hparams and 'dropout_p' not in model.hparams
    with torch.no_grad():
        for k, v in model.state_dict().items():
            v.copy_(torch.from_numpy(g['init_' + k]))
    model.eval()
    with torch.no_grad():
        preds = model(torch.from_numpy(g['pred_users']), torch.from_numpy(g['pred_items']))
    np.testing.assert_allclose(preds.cpu().numpy(), g['preds'], rtol=1e-5, atol=1e-6)
    model.train()
    opt = model.configure_optimizers()
    for s in range(3):
        batch = ((torch.from_numpy(g[f'step{s}_users']).cuda(),
                  torch.from_numpy(g[f'step{s}_pos']).cuda()),
                 torch.from_numpy(g[f'step{s}_negs']).cuda())
        opt.zero_grad()
        loss = model.calculate_loss(batch)
        loss.backward()
        opt.step()
        np.testing.assert_allclose(float(loss.detach()), float(g[f'step{s}_loss']), rtol=1e-5)
    for k, v in model.state_dict().items():
        got, exp = v.detach().cpu().numpy(), g['final_' + k]
        # (Adam's first steps follow the relative error of a gradient element: DESIGN.md 4)
        close = np.isclose(got, exp, rtol=1e-5, atol=1e-6)
        assert close.mean() >= 0.98, (k, close.mean())
        np.testing.assert_allclose(got, exp, rtol=1e-4, atol=1e-5, err_msg=k)


@pytest.mark.gpu
def test_cml_trains_evaluates_and_round_trips(tmp_path):
    """Through the trainer's unfused loop, evaluate_in_batches (forward over all pairs) and a
    checkpoint in the reference's format."""
    import collie_b200 as cb
    rng = np.random.default_rng(4)
    n_users, n_items = 60, 80
    data = cb.Interactions(users=rng.integers(0, n_users, 1500), items=rng.integers(0, n_items, 1500),
                           num_users=n_users, num_items=n_items, num_negative_samples=3,
                           allow_missing_ids=True, check_num_negative_samples_is_valid=False,
                           max_number_of_samples_to_consider=0)
    model = cb.CollaborativeMetricLearningModel(train=data, embedding_dim=16, lr=1e-2,
                                                lr_scheduler_func=None).cuda()
    before = model.user_embeddings.weight.detach().clone()
    cb.CollieTrainer(model, max_epochs=2, verbosity=0, enable_model_summary=False).fit(model)
    assert not torch.equal(before, model.user_embeddings.weight.detach())
    m1 = cb.evaluate_in_batches([cb.mapk, cb.mrr, cb.auc], data, model, k=5, verbose=False)
    path = str(tmp_path / 'cml.pth')
    model.save_model(path)
    again = cb.CollaborativeMetricLearningModel(load_model_path=path).cuda()
    m2 = cb.evaluate_in_batches([cb.mapk, cb.mrr, cb.auc], data, again, k=5, verbose=False)
    np.testing.assert_allclose(m1, m2, rtol=1e-12)
    with pytest.raises(Exception, match='unfused'):
        model.fused_step(None)
