"""The unfused route (callable losses / optimizers are part of the reference's surface,
collie/model/base/base_pipeline.py:281, :409): ``calculate_loss`` with an autograd graph through
the stand-alone CUDA kernels, the reference's four-call loop in the trainers, and agreement with
the fused step."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _data(cb, n_users=40, n_items=60, n=700, n_neg=4, seed=0):
    rng = np.random.default_rng(seed)
    flat = rng.choice(n_users * n_items, size=n, replace=False)
    return cb.Interactions(users=flat // n_items, items=flat % n_items, num_users=n_users,
                           num_items=n_items, num_negative_samples=n_neg, allow_missing_ids=True,
                           check_num_negative_samples_is_valid=False,
                           max_number_of_samples_to_consider=0, seed=3)


@pytest.mark.parametrize('loss', ['bpr', 'adaptive_hinge', 'warp'])
def test_differentiable_loss_and_its_gradients_match_the_oracle(loss):
    import collie_b200 as cb
    from oracle import collie_oracle as O
    data = _data(cb, n_neg=1 if loss == 'bpr' else 4)
    model = cb.MatrixFactorizationModel(train=data, embedding_dim=16, loss=loss,
                                        lr_scheduler_func=None).cuda()
    rng = np.random.default_rng(1)
    names = ('user_embeddings', 'item_embeddings', 'user_biases', 'item_biases')
    ref = O.OracleMF(40, 60, 16)
    with torch.no_grad():
        for name in names:
            p = getattr(model, name).weight
            p.copy_(torch.from_numpy(rng.normal(0, 0.5, tuple(p.shape)).astype(np.float32)))
            getattr(ref, name).weight.copy_(p.cpu())
    B, N = 96, data.num_negative_samples
    bu = torch.from_numpy(rng.integers(0, 40, B).astype(np.int32))
    bp = torch.from_numpy(rng.integers(0, 60, B).astype(np.int32))
    bn = torch.from_numpy(rng.integers(0, 60, (B, N)))
    model.train()
    got = model.calculate_loss(((bu.cuda(), bp.cuda()), bn.cuda()), differentiable=True)
    assert got.requires_grad
    got.backward()
    exp = ref.implicit_loss(bu, bp, bn, model._loss_name)
    exp.backward()
    np.testing.assert_allclose(got.item(), exp.item(), rtol=1e-5)
    # ... and equals the fused kernel's value
    np.testing.assert_allclose(model.calculate_loss(((bu.cuda(), bp.cuda()), bn.cuda())).item(),
                               exp.item(), rtol=1e-5)
    for name in names:
        np.testing.assert_allclose(getattr(model, name).weight.grad.cpu().numpy(),
                                   getattr(ref, name).weight.grad.numpy(), rtol=1e-5, atol=1e-6,
                                   err_msg=name)


def test_callable_loss_and_callable_optimizer_train_through_the_trainer(capsys):
    """A user loss (here: the package's own ``hinge_loss`` wrapped in a lambda, so nothing can
    recognise it) and ``torch.optim.Adam`` as optimizer callable: the trainer takes the
    reference's loop, the loss goes down, and one epoch equals the same epoch of a model that
    names the loss and the optimizer (fused route) -- both start from the same weights and see
    the same batches and negatives."""
    import collie_b200 as cb
    from collie_b200.loss import adaptive_hinge_loss
    data = _data(cb, n=1200)
    kw = dict(embedding_dim=8, lr=2e-2, bias_lr=2e-2, lr_scheduler_func=None)
    custom = cb.MatrixFactorizationModel(
        train=cb.InteractionsDataLoader(data, batch_size=256, shuffle=False),
        loss=lambda *a, **k: adaptive_hinge_loss(*a, **k), optimizer=torch.optim.Adam,
        bias_optimizer=torch.optim.SGD, **kw).cuda()
    named = cb.MatrixFactorizationModel(
        train=cb.InteractionsDataLoader(data, batch_size=256, shuffle=False),
        loss='adaptive_hinge', optimizer='adam', bias_optimizer='sgd', **kw).cuda()
    named.load_state_dict(custom.state_dict())
    assert custom._loss_name is None and callable(custom.loss_function)
    for m in (custom, named):
        t = cb.CollieMinimalTrainer(m, max_epochs=1, verbosity=0, enable_model_summary=False,
                                    logger=None)
        t.fit(m)
    for (k, a), b in zip(custom.state_dict().items(), named.state_dict().values()):
        np.testing.assert_allclose(a.cpu().numpy(), b.cpu().numpy(), rtol=1e-4, atol=1e-5,
                                   err_msg=k)
    t2 = cb.CollieMinimalTrainer(custom, max_epochs=4, verbosity=1, enable_model_summary=False,
                                 logger=None)
    t2.fit(custom)
    lines = [ln for ln in capsys.readouterr().out.splitlines() if 'train loss' in ln]
    losses = [float(ln.split('train loss:')[1]) for ln in lines]
    assert len(losses) == 4 and losses[-1] < losses[0]
