"""The bit-repeatable form of the training step (cmf_step_t.deterministic; requested from Python by
torch.use_deterministic_algorithms(True) or model.deterministic = True): item-row gradients summed
in batch order from a sorted contribution list, every user summed by one warp.

Reference: the CPU path of the reference is deterministic (torch index_add on one thread,
collie/model/base/base_pipeline.py:418-433 via torch.optim); its CUDA path is not. Here the fast
step is not either (vector RED in arrival order), the deterministic one is."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _tables(model):
    return {'U': model.user_embeddings.weight, 'I': model.item_embeddings.weight,
            'bu': model.user_biases.weight, 'bi': model.item_biases.weight}


def _data(cb, rng, n_users, n_items, n, n_neg, explicit=False):
    # a skewed item side (hot rows with hundreds of contributions per batch) and a few heavy users
    items = np.minimum((rng.pareto(1.2, n) * 3).astype(np.int64), n_items - 1)
    users = np.minimum((rng.pareto(1.5, n) * 40).astype(np.int64), n_users - 1)
    if explicit:
        return cb.ExplicitInteractions(users=users, items=items,
                                       ratings=rng.integers(1, 6, n).astype(np.float32),
                                       num_users=n_users, num_items=n_items,
                                       allow_missing_ids=True,
                                       remove_duplicate_user_item_pairs=False)
    return cb.Interactions(users=users, items=items, num_users=n_users, num_items=n_items,
                           num_negative_samples=n_neg, allow_missing_ids=True, seed=5,
                           remove_duplicate_user_item_pairs=False,
                           check_num_negative_samples_is_valid=False,
                           max_number_of_samples_to_consider=0)


def _train(cb, loader, data, kw, dim, deterministic, chunk_len=0, epochs=2):
    from collie_b200 import _abi
    torch.manual_seed(9)
    m = cb.MatrixFactorizationModel(train=loader, embedding_dim=dim, lr=1e-2,
                                    lr_scheduler_func=None, **kw).cuda()
    m.deterministic = deterministic
    opts = m.configure_optimizers()
    prev = _abi.lib().cmf_set_tuning(0, chunk_len)
    try:
        losses = []
        for epoch in range(1, epochs + 1):
            for db in loader.device_batches('cuda', epoch=epoch):
                losses.append(m.fused_step(db, optimizers=opts, interactions=data, epoch=epoch))
        torch.cuda.synchronize()
    finally:
        _abi.lib().cmf_set_tuning(0, prev)
    return torch.stack(losses), {k: v.detach().clone() for k, v in _tables(m).items()}


@pytest.mark.parametrize('case', [
    dict(dim=128, n_neg=5, kw=dict(loss='bpr', optimizer='sparse_adam', sparse=True)),
    dict(dim=32, n_neg=10, kw=dict(loss='warp', optimizer='sparse_adam', sparse=True)),
    dict(dim=64, n_neg=3, kw=dict(loss='hinge', optimizer='adam', sparse=False)),
    dict(dim=260, n_neg=4, kw=dict(loss='adaptive', optimizer='sgd', sparse=True)),
    dict(dim=64, n_neg=0, explicit=True, kw=dict(loss='mse', optimizer='sparse_adam', sparse=True)),
], ids=lambda c: f"{c['kw']['loss']}_{c['kw']['optimizer']}_d{c['dim']}")
def test_step_is_bit_repeatable(case):
    """Twelve steps, three times: the deterministic runs agree bit for bit with each other --
    also across chunk lengths of the backward kernel, which change which warp sums which user
    but not the order of any sum -- and with the fast step to rounding."""
    import collie_b200 as cb
    rng = np.random.default_rng(31)
    n_users, n_items, n, bsz = 4000, 900, 24576, 4096
    data = _data(cb, rng, n_users, n_items, n, case['n_neg'], explicit=case.get('explicit', False))
    loader = cb.InteractionsDataLoader(data, batch_size=bsz, shuffle=True, drop_last=False)
    la, a = _train(cb, loader, data, case['kw'], case['dim'], True)
    lb, b = _train(cb, loader, data, case['kw'], case['dim'], True)
    lc, c = _train(cb, loader, data, case['kw'], case['dim'], True, chunk_len=4)
    assert torch.equal(la, lb) and torch.equal(la, lc)
    for k in a:
        assert torch.equal(a[k], b[k]), k
        assert torch.equal(a[k], c[k]), k
    lf, f = _train(cb, loader, data, case['kw'], case['dim'], False)
    np.testing.assert_allclose(lf.cpu().numpy(), la.cpu().numpy(), rtol=1e-5)
    for k in a:
        x, y = f[k].cpu().numpy(), a[k].cpu().numpy()
        if k == 'bu' and not case.get('explicit', False):
            continue        # zero true gradient: rounding residue under Adam
        # (Adam follows the relative error of a gradient element, DESIGN.md 4: elements whose
        # terms nearly cancel differ between two summation orders by up to ~1e-3 * lr per step)
        close = np.isclose(x, y, rtol=1e-5, atol=1e-6)
        assert close.mean() >= 0.98, (k, close.mean())
        np.testing.assert_allclose(x, y, rtol=1e-3, atol=12 * 1e-2 * 2e-3, err_msg=k)


def test_deterministic_step_refuses_shapes_it_does_not_cover():
    import collie_b200 as cb
    rng = np.random.default_rng(3)
    data = _data(cb, rng, 300, 200, 2048, 3)
    loader = cb.InteractionsDataLoader(data, batch_size=512, shuffle=True)
    m = cb.MatrixFactorizationModel(train=loader, embedding_dim=10, lr_scheduler_func=None).cuda()
    m.deterministic = True
    opts = m.configure_optimizers()
    db = next(iter(loader.device_batches('cuda', epoch=1)))
    with pytest.raises(Exception, match='deterministic'):
        m.fused_step(db, optimizers=opts, interactions=data, epoch=1)
