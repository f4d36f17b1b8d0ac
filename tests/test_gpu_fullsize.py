"""The training step and the evaluation at BASELINE.json's full size (config C3: 138,493 users x
26,744 items, D = 128, 5 negatives, batches of 262,144), checked through properties that do not
need an oracle run of that size: invariants of the loss gradients, linearity in the batch,
reproducibility of the sampler, locality of the sparse update, agreement of the two evaluation
paths. Everything goes through the C ABI."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

N_USERS, N_ITEMS, DIM, N_NEG, BATCH = 138_493, 26_744, 128, 5, 262_144


@pytest.fixture(scope='module')
def world():
    import collie_b200 as cb
    rng = np.random.default_rng(20)
    n = 2 * BATCH
    users = rng.integers(0, N_USERS, n).astype(np.int32)
    items = rng.integers(0, N_ITEMS, n).astype(np.int32)
    data = cb.Interactions(users=users, items=items, num_users=N_USERS, num_items=N_ITEMS,
                           num_negative_samples=N_NEG, allow_missing_ids=True, seed=4,
                           remove_duplicate_user_item_pairs=False,
                           check_num_negative_samples_is_valid=False,
                           max_number_of_samples_to_consider=0)
    st = data.device_state('cuda')
    return cb, data, st


def _model(cb, data, optimizer, bias_optimizer='sgd', lr=1e-2, loss='bpr'):
    torch.manual_seed(1)
    return cb.MatrixFactorizationModel(train=data, embedding_dim=DIM, sparse=optimizer != 'adam',
                                       lr=lr, bias_lr=lr, optimizer=optimizer,
                                       bias_optimizer=bias_optimizer, loss=loss,
                                       lr_scheduler_func=None).cuda()


def _batch(st, lo, hi):
    from collie_b200.interactions import DeviceBatch
    idx = torch.arange(lo, hi, device='cuda')
    return DeviceBatch(st['users'][lo:hi], st['items'][lo:hi], idx, None)


def _tables(m):
    return [m.user_embeddings.weight.data, m.item_embeddings.weight.data,
            m.user_biases.weight.data, m.item_biases.weight.data]


@pytest.mark.parametrize('loss', ['bpr', 'warp'])
def test_bias_gradients_cancel_at_full_size(world, loss):
    """Every implicit loss sends +g to the positive and -g to one negative: user biases receive
    exactly zero and the item-bias gradients sum to zero (SGD, so the update IS the gradient)."""
    cb, data, st = world
    m = _model(cb, data, 'sgd', loss=loss, lr=0.5)
    before = [t.clone() for t in _tables(m)]
    loss_v = m.fused_step(_batch(st, 0, BATCH), optimizers=m.configure_optimizers(),
                          interactions=data, epoch=0)
    assert torch.isfinite(loss_v) and loss_v.item() > 0
    after = _tables(m)
    assert torch.equal(after[2], before[2])                       # user biases: bit for bit
    moved = (after[3] - before[3]).double()
    assert moved.abs().max().item() > 0
    assert abs(moved.sum().item()) < 1e-6 * moved.abs().sum().item()
    assert not torch.equal(after[0], before[0]) and not torch.equal(after[1], before[1])


def test_loss_and_gradients_are_linear_in_the_batch(world):
    """mean-loss(batch) = weighted mean over its halves, and the SGD update of the whole batch =
    the sum of the halves' updates computed for the same global batch size (same start)."""
    cb, data, st = world
    whole, h1, h2 = _batch(st, 0, BATCH), _batch(st, 0, BATCH // 2), _batch(st, BATCH // 2, BATCH)
    lr = 1e3            # the updates must stand clear of the parameters' own fp32 resolution
    ref = _model(cb, data, 'sgd', lr=lr)
    start = [t.clone() for t in _tables(ref)]
    l_whole = ref.fused_step(whole, train=False, interactions=data, epoch=3).item()
    l_1 = ref.fused_step(h1, train=False, interactions=data, epoch=3).item()
    l_2 = ref.fused_step(h2, train=False, interactions=data, epoch=3).item()
    np.testing.assert_allclose(l_whole, 0.5 * (l_1 + l_2), rtol=1e-5)

    def delta(batches):
        m = _model(cb, data, 'sgd', lr=lr)
        total = [torch.zeros_like(t, dtype=torch.float64) for t in start]
        for b in batches:
            with torch.no_grad():
                for t, s in zip(_tables(m), start):
                    t.copy_(s)
            m.fused_step(b, optimizers=m.configure_optimizers(), interactions=data, epoch=3,
                         global_batch=BATCH)
            for acc, t, s in zip(total, _tables(m), start):
                acc += (t - s).double()
        return total
    d_whole, d_halves = delta([whole]), delta([h1, h2])
    for a, b, s0 in zip(d_whole, d_halves, start):
        scale = a.abs().max().item()
        assert scale > 0 or torch.equal(a, b)
        ulp = 1.2e-7 * (s0.abs().max().item() + scale)     # each update is read off fp32 parameters
        assert (a - b).abs().max().item() <= 2e-5 * scale + 8 * ulp


def test_sampler_is_keyed_by_position_not_by_batch(world):
    """The negatives of an interaction depend on (seed, epoch, dataset position) only: the same
    rows come out of the whole batch, of a half batch and of a re-run; another epoch differs."""
    cb, data, st = world
    m = _model(cb, data, 'sparse_adam')

    def negs(lo, hi, epoch):
        out = {'negs': None}
        m.fused_step(_batch(st, lo, hi), train=False, interactions=data, epoch=epoch, outputs=out)
        return out['negs']
    a = negs(0, BATCH, 5)
    assert a.shape == (BATCH, N_NEG) and int(a.min()) >= 0 and int(a.max()) < N_ITEMS
    assert torch.equal(a, negs(0, BATCH, 5))
    assert torch.equal(a[BATCH // 2:], negs(BATCH // 2, BATCH, 5))
    assert not torch.equal(a, negs(0, BATCH, 6))
    # uniform over the catalogue: every decile of item ids gets its share
    hist = torch.histc(a.float(), bins=10, min=0, max=N_ITEMS)
    assert (hist / a.numel() - 0.1).abs().max().item() < 0.003


def test_sparse_update_touches_only_gathered_rows(world):
    """torch.optim.SparseAdam semantics at full size: user rows outside the batch stay bit for
    bit and the user rows inside it move; item rows outside positives + sampled negatives stay
    bit for bit, and every positive and every selected negative moves. (At the first step a
    gathered row with an all-zero gradient still has zero moments, so it stays where it is.)"""
    cb, data, st = world
    m = _model(cb, data, 'sparse_adam')
    before = [t.clone() for t in _tables(m)]
    out = {'negs': None, 'sel': None}
    b = _batch(st, 0, BATCH)
    m.fused_step(b, optimizers=m.configure_optimizers(), interactions=data, epoch=1, outputs=out)
    after = _tables(m)
    in_batch = torch.zeros(N_USERS, dtype=torch.bool, device='cuda')
    in_batch[b.users.long()] = True
    changed_u = (after[0] != before[0]).any(dim=1)
    assert not bool((changed_u & ~in_batch).any())
    # (a sampled negative can coincide with the positive: that interaction's gradient is zero)
    assert float((changed_u & in_batch).sum()) >= 0.999 * float(in_batch.sum())
    gathered = torch.zeros(N_ITEMS, dtype=torch.bool, device='cuda')
    gathered[b.items.long()] = True
    gathered[out['negs'].flatten()] = True
    with_grad = torch.zeros(N_ITEMS, dtype=torch.bool, device='cuda')
    with_grad[b.items.long()] = True
    sel = out['sel'].long().view(-1, 1)
    with_grad[torch.gather(out['negs'], 1, sel).flatten()] = True
    changed_i = (after[1] != before[1]).any(dim=1)
    assert not bool((changed_i & ~gathered).any())
    assert float((changed_i & with_grad).sum()) >= 0.999 * float(with_grad.sum())
    # second step on another batch: moments are non-zero now, so EVERY gathered row moves
    mid = [t.clone() for t in after]
    out2 = {'negs': None}
    b2 = _batch(st, BATCH, 2 * BATCH)
    m.fused_step(b2, optimizers=m.configure_optimizers(), interactions=data, epoch=1,
                 outputs=out2)
    gathered2 = torch.zeros(N_ITEMS, dtype=torch.bool, device='cuda')
    gathered2[b2.items.long()] = True
    gathered2[out2['negs'].flatten()] = True
    changed2 = (_tables(m)[1] != mid[1]).any(dim=1)
    assert not bool((changed2 & ~gathered2).any())


def test_evaluation_paths_agree_at_full_item_count(world):
    """Exact fp32 path vs tensor-core path over all 26,744 items for 4,096 users: the same top-10
    ids and scores bit for bit, the same MAP@10 and MRR, and a repeated call returns the same bits."""
    cb, data, st = world
    from collie_b200 import metrics as M
    import scipy.sparse as sp
    m = _model(cb, data, 'sparse_adam')
    with torch.no_grad():
        for t in _tables(m):
            t.copy_(torch.randn_like(t) * 0.3)
    rng = np.random.default_rng(8)
    n_u = 4096
    tu = np.repeat(np.arange(n_u), 6)
    ti = rng.integers(0, N_ITEMS, len(tu))
    coo = sp.coo_matrix((np.ones(len(tu)), (tu, ti)), shape=(N_USERS, N_ITEMS))
    coo.sum_duplicates()
    indptr, items, nnz = M._targets_device(coo.tocsr(), np.arange(n_u), 'cuda')
    users = torch.arange(n_u, dtype=torch.int32, device='cuda')
    tabs = M._model_tables(m)

    def run(precision, tile):
        tot = {'mapk': 0.0, 'mrr': 0.0}
        keep = []
        for s in range(0, n_u, tile):
            e = min(s + tile, n_u)
            res = M._eval_tile('cuda', e - s, N_ITEMS, 10, indptr[s:e + 1], items,
                               {'topk', 'mapk', 'mrr'}, tables=tabs, user_ids=users[s:e],
                               precision=precision, tc_status=status if precision else None)
            for name in tot:
                tot[name] += M._per_user(name, res, nnz[s:e], N_ITEMS).sum().item()
            keep.append((res['topk_ids'].clone(), res['topk_scores'].clone()))
        return tot, keep
    status = torch.zeros(2, dtype=torch.int32, device='cuda')
    exact, k0 = run(0, 256)
    tc, k1 = run(1, n_u)
    # (users handed to the exact path inside the call: a few at most; none left over)
    assert int(status[1]) == 0 and int(status[0]) <= n_u // 100, status.tolist()
    for name in exact:
        np.testing.assert_allclose(tc[name], exact[name], rtol=1e-12)
    scores = k1[0][1]
    assert bool((scores[:, 1:] <= scores[:, :-1]).all())
    # ids and scores of the tensor-core path ARE the exact path's (random fp32 weights)
    assert torch.equal(torch.cat([a for a, _ in k0]), k1[0][0])
    assert torch.equal(torch.cat([b for _, b in k0]), k1[0][1])
    tc2, k2 = run(1, n_u)
    assert tc2 == tc and torch.equal(k2[0][0], k1[0][0]) and torch.equal(k2[0][1], k1[0][1])


def test_large_item_table_update_path_matches_oracle():
    """Item tables that do not stay in L2 (> 64 MB of parameter + moment rows) take the sorted
    touched-row list in front of k_update_rows: two steps against the oracle on a 60,000 x 128
    item table."""
    import collie_b200 as cb
    from oracle import collie_oracle as O
    rng = np.random.default_rng(33)
    n_users, n_items, dim, n_neg, batch = 3000, 60_000, 128, 3, 4096
    users = rng.integers(0, n_users, 3 * batch).astype(np.int32)
    items = rng.integers(0, n_items, 3 * batch).astype(np.int32)
    data = cb.Interactions(users=users, items=items, num_users=n_users, num_items=n_items,
                           num_negative_samples=n_neg, allow_missing_ids=True, seed=6,
                           remove_duplicate_user_item_pairs=False,
                           check_num_negative_samples_is_valid=False,
                           max_number_of_samples_to_consider=0)
    torch.manual_seed(2)
    model = cb.MatrixFactorizationModel(train=data, embedding_dim=dim, sparse=True, lr=1e-2,
                                        optimizer='sparse_adam', loss='bpr',
                                        lr_scheduler_func=None).cuda()
    ref = O.OracleMF(n_users, n_items, dim, sparse=True)
    names = ('user_embeddings', 'item_embeddings', 'user_biases', 'item_biases')
    with torch.no_grad():
        for name in names:
            getattr(model, name).weight.mul_(8.0)        # gradients well above fp32 noise
            getattr(ref, name).weight.copy_(getattr(model, name).weight.cpu())
    opts = model.configure_optimizers()
    ref_opts = O.make_optimizers(ref, 'sparse_adam', 'sgd', 1e-2, 1e-2)
    for step in range(2):
        sl = slice(step * batch, (step + 1) * batch)
        bu, bp = torch.from_numpy(users[sl]), torch.from_numpy(items[sl])
        bn = torch.from_numpy(rng.integers(0, n_items, (batch, n_neg)))
        loss = model.fused_step(((bu.cuda(), bp.cuda()), bn.cuda()), optimizers=opts).item()
        ref_loss = O.train_step(ref, ref_opts,
                                lambda: ref.implicit_loss(bu, bp, bn, 'adaptive_bpr'))
        np.testing.assert_allclose(loss, ref_loss, rtol=1e-5)
    from test_gpu_fullsize_oracle import _tables_close
    _tables_close(model, ref, 1e-2)
