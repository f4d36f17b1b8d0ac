"""``HybridPretrainedModel`` (BASELINE config C4) against golden files produced by the
unmodified reference (``oracle/make_golden.py hybrid``): predictions, the losses of three
iterations of the reference's loop and every parameter after them -- frozen embeddings, fine-tuned
embeddings, and metadata-only layers. Reference:
collie/model/hybrid_pretrained_matrix_factorization.py:227-464."""
import os

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, 'tests', 'golden')
CASES = ['hybrid_frozen', 'hybrid_unfrozen', 'hybrid_meta_layers']


def _load(name):
    return np.load(os.path.join(GOLDEN, name + '.npz'), allow_pickle=False)


@pytest.mark.parametrize('name', CASES)
def test_oracle_hybrid_forward_matches_the_reference(name):
    from oracle import collie_oracle as O
    g = _load(name)
    state = {k[5:]: torch.from_numpy(g[k]) for k in g.files if k.startswith('init_')}
    got = O.hybrid_forward(state, torch.from_numpy(g['item_metadata']),
                           torch.from_numpy(g['pred_users']), torch.from_numpy(g['pred_items']))
    np.testing.assert_allclose(got.numpy(), g['preds'], rtol=1e-6, atol=1e-6)


def _build(cb, g, device):
    n_users, n_items, dim = int(g['n_users']), int(g['n_items']), int(g['dim'])
    data = cb.Interactions(users=np.arange(n_users) % n_users, items=np.arange(n_users) % n_items,
                           num_users=n_users, num_items=n_items,
                           num_negative_samples=int(g['n_neg']), allow_missing_ids=True,
                           check_num_negative_samples_is_valid=False)
    mf = cb.MatrixFactorizationModel(train=data, embedding_dim=dim, lr_scheduler_func=None)
    layers = g['meta_layers'].tolist() or None
    model = cb.HybridPretrainedModel(train=data, item_metadata=g['item_metadata'],
                                     trained_model=mf, item_metadata_layers_dims=layers,
                                     combined_layers_dims=[16, 8],
                                     freeze_embeddings=bool(int(g['freeze'])), lr=float(g['lr']),
                                     optimizer='adam', loss='adaptive_hinge',
                                     lr_scheduler_func=None)
    init = {k[5:]: torch.from_numpy(g[k]) for k in g.files if k.startswith('init_')}
    assert set(init) == set(model.state_dict()), (sorted(init), sorted(model.state_dict()))
    model.load_state_dict(init)
    return model.to(device)


@pytest.mark.gpu
@pytest.mark.parametrize('name', CASES)
def test_hybrid_predictions_losses_and_steps_match_the_reference(name):
    import collie_b200 as cb
    g = _load(name)
    model = _build(cb, g, 'cuda')
    model.eval()
    with torch.no_grad():
        preds = model(torch.from_numpy(g['pred_users']).cuda(), torch.from_numpy(g['pred_items']).cuda())
    np.testing.assert_allclose(preds.cpu().numpy(), g['preds'], rtol=1e-5, atol=1e-6)
    model.train()
    opt = model.configure_optimizers()
    assert isinstance(opt, cb.RowOptimizer)          # the library's dense-gradient Adam kernel
    s = 0
    while f'step{s}_users' in g.files:
        batch = ((torch.from_numpy(g[f'step{s}_users']).cuda(),
                  torch.from_numpy(g[f'step{s}_pos']).cuda()),
                 torch.from_numpy(g[f'step{s}_negs']).cuda())
        opt.zero_grad()
        loss = model.calculate_loss(batch)
        loss.backward()
        opt.step()
        np.testing.assert_allclose(loss.item(), float(g[f'step{s}_loss']), rtol=1e-5,
                                   err_msg=f'loss of step {s}')
        s += 1
    assert s == 3
    frozen = bool(int(g['freeze']))
    last = len(model.combined_layers) - 1
    for k, v in model.state_dict().items():
        exp = g['final_' + k]
        if frozen and k.startswith('embeddings'):
            assert np.array_equal(v.cpu().numpy(), g['init_' + k]), k      # never touched
        if k in ('biases.0.weight', f'combined_layer_{last}.bias'):
            # The true gradient of a user bias -- and of the last layer's bias, which enters every
            # score alike -- is exactly zero: every interaction sends it +g (the positive) and -g
            # (the chosen negative). What reaches Adam is the rounding residue of that sum
            # (~1e-9, its sign a matter of summation order), which Adam's g / |g| turns into a
            # step of up to +-lr: noise in the reference and here alike. Bounded, not compared.
            assert np.abs(v.cpu().numpy() - g['init_' + k]).max() <= s * float(g['lr']) * 1.001
            continue
        np.testing.assert_allclose(v.cpu().numpy(), exp, rtol=1e-4, atol=1e-5, err_msg=k)
        tight = np.abs(v.cpu().numpy() - exp) <= 1e-6 + 1e-5 * np.abs(exp)
        assert tight.mean() >= 0.98, (k, float(tight.mean()))


@pytest.mark.gpu
def test_hybrid_trains_through_the_trainer_and_round_trips(tmp_path, capsys):
    """The reference's fixture in small (tests/fixtures/model_fixtures.py:371-428): pretrain MF,
    train the hybrid on top with frozen embeddings, evaluate, save to a directory, load back."""
    import collie_b200 as cb
    rng = np.random.default_rng(3)
    n_users, n_items = 60, 90
    flat = rng.choice(n_users * n_items, size=1500, replace=False)
    data = cb.Interactions(users=flat // n_items, items=flat % n_items, num_users=n_users,
                           num_items=n_items, num_negative_samples=4, allow_missing_ids=True,
                           check_num_negative_samples_is_valid=False, seed=1)
    mf = cb.MatrixFactorizationModel(train=data, embedding_dim=8, lr=1e-1, lr_scheduler_func=None)
    cb.CollieTrainer(mf, max_epochs=1, verbosity=0, enable_model_summary=False,
                     logger=False).fit(mf)
    meta = rng.normal(size=(n_items, 5)).astype(np.float32)
    model = cb.HybridPretrainedModel(
        train=cb.InteractionsDataLoader(data, batch_size=256, shuffle=True), item_metadata=meta,
        trained_model=mf, combined_layers_dims=[16, 8], freeze_embeddings=True, lr=1e-2,
        optimizer=torch.optim.Adam, loss='warp', lr_scheduler_func=None)
    before = model.embeddings[1].weight.detach().clone()
    trainer = cb.CollieTrainer(model, max_epochs=3, verbosity=1, enable_model_summary=False,
                               logger=False)
    trainer.fit(model)
    losses = [float(ln.split('train loss:')[1]) for ln in capsys.readouterr().out.splitlines()
              if 'train loss' in ln]
    assert len(losses) == 3 and losses[-1] < losses[0]
    assert torch.equal(model.embeddings[1].weight.detach().cpu(), before.cpu())
    m = cb.evaluate_in_batches([cb.mapk, cb.mrr, cb.auc], data, model, k=5, verbose=False)
    assert all(np.isfinite(v) and 0 <= v <= 1 for v in m)
    path = str(tmp_path / 'hybrid')
    model.save_model(path)
    again = cb.HybridPretrainedModel(load_model_path=path).cuda()
    users = torch.from_numpy(rng.integers(0, n_users, 30)).cuda()
    items = torch.from_numpy(rng.integers(0, n_items, 30)).cuda()
    model.eval()
    with torch.no_grad():
        assert torch.equal(again(users, items), model(users, items))
    # item-item similarity keeps working on the embeddings (reference :466-474)
    assert len(model.item_item_similarity(3)) == n_items
