"""``dropout_p > 0`` (collie/model/matrix_factorization.py:118, :147: independent masks on the user
and the item embedding of every scored pair, training mode only) inside the fused training step.
The masks come from the counter-based stream (``oracle.dropout_mask`` states it on the CPU), so the
oracle can be given the very same masks: loss and every table after one step must agree."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _masks(O, seed, step, epoch, n_rows, slot, dim, p):
    mu = np.stack([O.dropout_mask(seed, step, epoch, b, slot, 0, dim, p) for b in range(n_rows)])
    mi = np.stack([O.dropout_mask(seed, step, epoch, b, slot, 1, dim, p) for b in range(n_rows)])
    return torch.from_numpy(mu), torch.from_numpy(mi)


@pytest.mark.parametrize('dim', [8, 6, 5])
@pytest.mark.parametrize('loss', ['adaptive_hinge', 'warp', 'mse'])
def test_fused_step_with_dropout_matches_oracle_with_the_same_masks(loss, dim):
    import collie_b200 as cb
    from oracle import collie_oracle as O
    rng = np.random.default_rng(dim)
    n_users, n_items, n_neg, B, p, lr = 30, 40, 3, 24, 0.35, 0.5
    explicit = loss == 'mse'
    flat = rng.choice(n_users * n_items, size=300, replace=False)
    if explicit:
        data = cb.ExplicitInteractions(users=flat // n_items, items=flat % n_items,
                                       ratings=rng.integers(1, 6, 300).astype(np.float32),
                                       num_users=n_users, num_items=n_items, allow_missing_ids=True)
        data.seed = 77
    else:
        data = cb.Interactions(users=flat // n_items, items=flat % n_items, num_users=n_users,
                               num_items=n_items, num_negative_samples=n_neg,
                               allow_missing_ids=True, check_num_negative_samples_is_valid=False,
                               max_number_of_samples_to_consider=0, seed=77)
    model = cb.MatrixFactorizationModel(train=data, embedding_dim=dim, dropout_p=p, lr=lr,
                                        bias_lr=lr, optimizer='sgd', bias_optimizer='sgd',
                                        loss=loss, lr_scheduler_func=None).cuda()
    ref = O.OracleMF(n_users, n_items, dim)
    names = ('user_embeddings', 'item_embeddings', 'user_biases', 'item_biases')
    with torch.no_grad():
        for name in names:
            w = getattr(model, name).weight
            w.copy_(torch.from_numpy(rng.normal(0, 0.5, tuple(w.shape)).astype(np.float32)))
            getattr(ref, name).weight.copy_(w.cpu())
    bu = torch.from_numpy(rng.integers(0, n_users, B).astype(np.int32))
    bp = torch.from_numpy(rng.integers(0, n_items, B).astype(np.int32))
    opts = model.configure_optimizers()
    model.train()
    epoch = 4
    U, I, bU, bI = ref.tables()
    if explicit:
        br = torch.from_numpy(rng.integers(1, 6, B).astype(np.float32))
        got = model.fused_step((bu.cuda(), bp.cuda(), br.cuda()), optimizers=opts, epoch=epoch)
        mu, mi = _masks(O, 77, 1, epoch, B, 0, dim, p)
        preds = O.mf_forward_dropout(U, I, bU, bI, bu, bp, mu, mi)
        exp = ((preds - br) ** 2).mean()
    else:
        bn = torch.from_numpy(rng.integers(0, n_items, (B, n_neg)))
        got = model.fused_step(((bu.cuda(), bp.cuda()), bn.cuda()), optimizers=opts, epoch=epoch)
        mu, mi = _masks(O, 77, 1, epoch, B, 0, dim, p)
        pos = O.mf_forward_dropout(U, I, bU, bI, bu, bp, mu, mi)
        negs = []
        for n in range(n_neg):
            mu_n, mi_n = _masks(O, 77, 1, epoch, B, 1 + n, dim, p)
            negs.append(O.mf_forward_dropout(U, I, bU, bI, bu, bn[:, n], mu_n, mi_n))
        neg = torch.stack(negs)                                   # [N, B]
        if model._loss_name == 'warp':
            exp = O.warp(pos, neg, n_items, bp.long(), bn.t().long(), None, None)
        else:
            exp = O.IMPLICIT_LOSSES[model._loss_name](pos, neg, bp.long(), bn.t().long(), None,
                                                      None)
    exp.backward()
    np.testing.assert_allclose(got.item(), exp.item(), rtol=1e-5)
    with torch.no_grad():
        for t in ref.tables():
            t -= lr * t.grad
    for name in names:
        np.testing.assert_allclose(getattr(model, name).weight.detach().cpu().numpy(),
                                   getattr(ref, name).weight.detach().numpy(), rtol=1e-5,
                                   atol=1e-6, err_msg=name)
    # evaluation mode: no dropout (the fused value equals the plain loss)
    model.eval()
    batch = (bu.cuda(), bp.cuda(), br.cuda()) if explicit else ((bu.cuda(), bp.cuda()), bn.cuda())
    a = model.calculate_loss(batch).item()
    b = model.calculate_loss(batch).item()
    assert a == b


def test_trainer_runs_with_dropout_and_masks_change_between_steps():
    """The reference's own MovieLens example trains with dropout (collie/movielens/run.py)."""
    import collie_b200 as cb
    rng = np.random.default_rng(0)
    flat = rng.choice(50 * 70, size=1500, replace=False)
    data = cb.Interactions(users=flat // 70, items=flat % 70, num_users=50, num_items=70,
                           num_negative_samples=5, allow_missing_ids=True,
                           check_num_negative_samples_is_valid=False, seed=5)
    model = cb.MatrixFactorizationModel(
        train=cb.InteractionsDataLoader(data, batch_size=256, shuffle=True), embedding_dim=10,
        dropout_p=0.05, lr=5e-2, loss='adaptive', optimizer='adam', lr_scheduler_func=None).cuda()
    model.train()
    db = next(model.train_loader.device_batches(model.device, epoch=0))
    l1 = model.fused_step(db, train=False, interactions=data, epoch=0).item()
    l2 = model.fused_step(db, train=False, interactions=data, epoch=1).item()
    assert l1 != l2                       # another epoch, other negatives and other masks
    t = cb.CollieMinimalTrainer(model, max_epochs=3, verbosity=0, enable_model_summary=False,
                                logger=None)
    t.fit(model)
    assert model.hparams.num_epochs_completed == 3
    m = cb.evaluate_in_batches([cb.mapk], data, model, k=10, verbose=False)
    assert 0 < m <= 1
