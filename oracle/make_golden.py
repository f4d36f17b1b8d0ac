"""Generate ``tests/golden/*.npz`` by running the UNMODIFIED reference in the build container.

TEST INFRASTRUCTURE ONLY. Usage (build container only; ``/root/reference`` does not exist on the
GPU box, which is why the outputs are committed):

    PYTHONPATH=oracle/stubs:/root/reference python oracle/make_golden.py

The stubs under ``oracle/stubs`` stand in for ``pytorch_lightning`` / ``torchmetrics`` / ``fire``
(absent from this image); they contribute no arithmetic. Everything numerical below is computed by
``collie`` (v1.3.1) + torch (2.11.0 here; the reference pins 1.13.1).
"""
import os
import sys
from functools import partial

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), 'tests', 'golden')

import collie  # noqa: E402  (the reference)
from collie.interactions import ExplicitInteractions, Interactions  # noqa: E402
from collie.metrics import auc, evaluate_in_batches, mapk, mrr  # noqa: E402
from collie.model import MatrixFactorizationModel  # noqa: E402

assert collie.__file__.startswith('/root/reference'), collie.__file__


def synthetic_pairs(rng, n_users, n_items, n):
    flat = rng.choice(n_users * n_items, size=n, replace=False)
    return (flat // n_items).astype(np.int64), (flat % n_items).astype(np.int64)


def randomize(model, rng, scale=0.5):
    with torch.no_grad():
        for p in model.parameters():
            p.copy_(torch.from_numpy(rng.normal(0, scale, size=tuple(p.shape)).astype(np.float32)))


def state(model):
    return {
        'U': model.user_embeddings.weight.detach().numpy().copy(),
        'I': model.item_embeddings.weight.detach().numpy().copy(),
        'bu': model.user_biases.weight.detach().numpy().copy(),
        'bi': model.item_biases.weight.detach().numpy().copy(),
    }


def implicit_case(loss, n_neg, optimizer, sparse, seed, with_metadata=False, dim=16,
                  n_users=37, n_items=53, n_inter=400, batch=64, steps=3):
    rng = np.random.default_rng(seed)
    torch.manual_seed(seed)
    users, items = synthetic_pairs(rng, n_users, n_items, n_inter)
    train = Interactions(users=users, items=items, ratings=np.ones(n_inter),
                         num_negative_samples=n_neg, num_users=n_users, num_items=n_items,
                         allow_missing_ids=True, remove_duplicate_user_item_pairs=False,
                         check_num_negative_samples_is_valid=False)
    kwargs = {}
    meta = {}
    if with_metadata:
        meta = {'a': rng.integers(0, 3, size=n_items), 'b': rng.integers(0, 4, size=n_items)}
        kwargs = dict(metadata_for_loss={k: torch.from_numpy(v) for k, v in meta.items()},
                      metadata_for_loss_weights={'a': 0.2, 'b': 0.3})
    model = MatrixFactorizationModel(train=train, embedding_dim=dim, sparse=sparse, lr=1e-2,
                                     bias_lr=5e-2, optimizer=optimizer, bias_optimizer='sgd',
                                     loss=loss, lr_scheduler_func=None, **kwargs)
    randomize(model, rng)
    out = {'n_users': n_users, 'n_items': n_items, 'dim': dim, 'n_neg': n_neg,
           'loss_name': model.loss_function.__name__, 'lr': 1e-2, 'bias_lr': 5e-2}
    for k, v in meta.items():
        out['meta_' + k] = v
    for k, v in state(model).items():
        out['init_' + k] = v

    optimizers = model.configure_optimizers()
    if isinstance(optimizers, tuple):
        optimizers = optimizers[0]
    if not isinstance(optimizers, (list, tuple)):
        optimizers = [optimizers]
    model.train()
    for s in range(steps):
        bu_ = rng.integers(0, n_users, size=batch)
        bp_ = rng.integers(0, n_items, size=batch)
        bn_ = rng.integers(0, n_items, size=(batch, n_neg))
        batch_t = ((torch.from_numpy(bu_.astype(np.int32)), torch.from_numpy(bp_.astype(np.int32))),
                   torch.from_numpy(bn_))
        for opt in optimizers:
            opt.zero_grad()
        loss_t = model.calculate_loss(batch_t)
        if s == 0:
            with torch.no_grad():
                out['step0_pos_scores'] = model(batch_t[0][0].long(), batch_t[0][1].long()).numpy()
        loss_t.backward()
        if s == 0:
            for name, p in (('U', model.user_embeddings.weight), ('I', model.item_embeddings.weight),
                            ('bu', model.user_biases.weight), ('bi', model.item_biases.weight)):
                g = p.grad
                out['step0_grad_' + name] = (g.to_dense() if g.is_sparse else g).numpy().copy()
        for i, opt in enumerate(optimizers):
            model.optimizer_step(epoch=0, batch_idx=s, optimizer=opt, optimizer_idx=i)
        out[f'step{s}_users'] = bu_
        out[f'step{s}_pos'] = bp_
        out[f'step{s}_negs'] = bn_
        out[f'step{s}_loss'] = np.float64(loss_t.item())
    for k, v in state(model).items():
        out['final_' + k] = v
    out['steps'] = steps
    return out


def explicit_case(loss, optimizer, sparse, seed, y_range=None, dim=16, n_users=29, n_items=41,
                  n_inter=300, batch=48, steps=3):
    rng = np.random.default_rng(seed)
    users, items = synthetic_pairs(rng, n_users, n_items, n_inter)
    ratings = rng.integers(1, 6, size=n_inter).astype(np.float64)
    train = ExplicitInteractions(users=users, items=items, ratings=ratings, num_users=n_users,
                                 num_items=n_items, allow_missing_ids=True,
                                 remove_duplicate_user_item_pairs=False)
    model = MatrixFactorizationModel(train=train, embedding_dim=dim, sparse=sparse, lr=1e-2,
                                     bias_lr=5e-2, optimizer=optimizer, bias_optimizer='sgd',
                                     loss=loss, y_range=y_range, lr_scheduler_func=None)
    randomize(model, rng)
    out = {'n_users': n_users, 'n_items': n_items, 'dim': dim, 'lr': 1e-2, 'bias_lr': 5e-2,
           'y_range': np.array(y_range if y_range is not None else [np.nan, np.nan])}
    for k, v in state(model).items():
        out['init_' + k] = v
    optimizers = model.configure_optimizers()
    if not isinstance(optimizers, (list, tuple)):
        optimizers = [optimizers]
    for s in range(steps):
        bu_ = rng.integers(0, n_users, size=batch)
        bi_ = rng.integers(0, n_items, size=batch)
        br_ = rng.integers(1, 6, size=batch).astype(np.float32)
        batch_t = (torch.from_numpy(bu_.astype(np.int32)), torch.from_numpy(bi_.astype(np.int32)),
                   torch.from_numpy(br_))
        for opt in optimizers:
            opt.zero_grad()
        loss_t = model.calculate_loss(batch_t)
        loss_t.backward()
        if s == 0:
            for name, p in (('U', model.user_embeddings.weight), ('I', model.item_embeddings.weight),
                            ('bu', model.user_biases.weight), ('bi', model.item_biases.weight)):
                g = p.grad
                out['step0_grad_' + name] = (g.to_dense() if g.is_sparse else g).numpy().copy()
        for i, opt in enumerate(optimizers):
            model.optimizer_step(epoch=0, batch_idx=s, optimizer=opt, optimizer_idx=i)
        out[f'step{s}_users'], out[f'step{s}_items'], out[f'step{s}_ratings'] = bu_, bi_, br_
        out[f'step{s}_loss'] = np.float64(loss_t.item())
    for k, v in state(model).items():
        out['final_' + k] = v
    out['steps'] = steps
    return out


def metrics_case(seed, n_users=37, n_items=101, dim=12, n_test=260, exact_grid=False):
    rng = np.random.default_rng(seed)
    users, items = synthetic_pairs(rng, n_users, n_items, n_test)
    keep = users != 5                  # one user without test interactions
    users, items = users[keep], items[keep]
    test = Interactions(users=users, items=items, ratings=np.ones(len(users)), num_users=n_users,
                        num_items=n_items, allow_missing_ids=True,
                        remove_duplicate_user_item_pairs=False,
                        check_num_negative_samples_is_valid=False)
    model = MatrixFactorizationModel(train=test, embedding_dim=dim, lr_scheduler_func=None)
    if exact_grid:
        # exactly representable weights (multiples of 2^-2 in [-1, 1]): any summation order and any
        # bf16-split tensor-core product gives the same fp32 scores, with many genuine ties
        with torch.no_grad():
            for p in model.parameters():
                p.copy_(torch.from_numpy(
                    (rng.integers(-4, 5, size=tuple(p.shape)) / 4.0).astype(np.float32)))
    else:
        randomize(model, rng, scale=0.7)
    model.eval()
    out = {'n_users': n_users, 'n_items': n_items, 'dim': dim, 'test_users': users,
           'test_items': items}
    for k, v in state(model).items():
        out[k] = v
    for k_ in (1, 4, 10):
        for bs in (20, 7):
            res = evaluate_in_batches([mapk, mrr, auc], test, model, k=k_, batch_size=bs,
                                      verbose=False)
            out[f'k{k_}_bs{bs}'] = np.array(res, dtype=np.float64)
    with torch.no_grad():
        uu = torch.arange(n_users).repeat_interleave(n_items)
        ii = torch.arange(n_items).repeat(n_users)
        out['scores'] = model(uu, ii).view(n_users, n_items).numpy()
    return out


def similarity_case(seed, n_users=23, n_items=57, dim=20):
    """item_item_similarity / user_user_similarity (base_pipeline.py:783-859), in id order."""
    rng = np.random.default_rng(seed)
    users, items = synthetic_pairs(rng, n_users, n_items, 200)
    data = Interactions(users=users, items=items, ratings=np.ones(len(users)), num_users=n_users,
                        num_items=n_items, allow_missing_ids=True,
                        remove_duplicate_user_item_pairs=False,
                        check_num_negative_samples_is_valid=False)
    model = MatrixFactorizationModel(train=data, embedding_dim=dim, lr_scheduler_func=None)
    randomize(model, rng, scale=0.7)
    model.eval()
    out = {'n_users': n_users, 'n_items': n_items, 'dim': dim,
           'item_ids': np.array([0, 7, n_items - 1]), 'user_ids': np.array([0, 11, n_users - 1])}
    for k, v in state(model).items():
        out[k] = v
    for i in out['item_ids']:
        out[f'item_sim_{i}'] = model.item_item_similarity(int(i)).sort_index().values
    for u in out['user_ids']:
        out[f'user_sim_{u}'] = model.user_user_similarity(int(u)).sort_index().values
    return out


def split_case(seed, n_users=40, n_items=70, n=900):
    """random_split (exact entries) and stratified_split (per-user sizes) of the reference
    (collie/cross_validation.py:51-306)."""
    from collie.cross_validation import random_split, stratified_split
    rng = np.random.default_rng(seed)
    users, items = synthetic_pairs(rng, n_users, n_items, n)
    users[:3] = n_users - 1                      # a user with exactly three interactions
    keep = np.ones(n, dtype=bool)
    keep[3:] = users[3:] != n_users - 1
    users, items = users[keep], items[keep]
    data = Interactions(users=users, items=items, ratings=np.ones(len(users)), num_users=n_users,
                        num_items=n_items, allow_missing_ids=True,
                        remove_duplicate_user_item_pairs=False,
                        check_num_negative_samples_is_valid=False)
    out = {'n_users': n_users, 'n_items': n_items, 'users': data.mat.row, 'items': data.mat.col}
    for tag, (vp, tp) in {'a': (0.0, 0.2), 'b': (0.1, 0.25)}.items():
        parts = random_split(data, val_p=vp, test_p=tp, seed=seed + 1)
        for name, part in zip(('train', 'val', 'test') if vp > 0 else ('train', 'test'), parts):
            out[f'random_{tag}_{name}_users'] = part.mat.row
            out[f'random_{tag}_{name}_items'] = part.mat.col
        parts = stratified_split(data, val_p=vp, test_p=tp, seed=seed + 2, processes=0)
        for name, part in zip(('train', 'val', 'test') if vp > 0 else ('train', 'test'), parts):
            out[f'strat_{tag}_{name}_counts'] = np.bincount(part.mat.row, minlength=n_users)
    return out


def checkpoint_case(seed, n_users=37, n_items=53, dim=16):
    """A checkpoint WRITTEN BY THE REFERENCE (``save_model``, base_pipeline.py:880-900) after one
    training epoch, its tables, and its own predictions for a few pairs: what ``load_model_path=``
    of the B200 model must read back (tests/golden/reference_checkpoint.pth)."""
    from collie.model import CollieMinimalTrainer
    rng = np.random.default_rng(seed)
    torch.manual_seed(seed)
    users, items = synthetic_pairs(rng, n_users, n_items, 500)
    train = Interactions(users=users, items=items, ratings=np.ones(len(users)),
                         num_negative_samples=4, num_users=n_users, num_items=n_items,
                         allow_missing_ids=True, remove_duplicate_user_item_pairs=False,
                         check_num_negative_samples_is_valid=False)
    model = MatrixFactorizationModel(train=train, embedding_dim=dim, sparse=False, lr=1e-2,
                                     bias_lr=5e-2, optimizer='adam', bias_optimizer='sgd',
                                     loss='adaptive_hinge', lr_scheduler_func=None)
    trainer = CollieMinimalTrainer(model, max_epochs=1, gpus=0, logger=None, verbosity=0,
                                   enable_model_summary=False)
    trainer.fit(model)
    model.eval()
    model.save_model(os.path.join(OUT, 'reference_checkpoint.pth'))
    pu = rng.integers(0, n_users, 64)
    pi = rng.integers(0, n_items, 64)
    with torch.no_grad():
        preds = model(torch.from_numpy(pu), torch.from_numpy(pi)).numpy()
    out = {'pred_users': pu, 'pred_items': pi, 'preds': preds,
           'hparams_keys': np.array(sorted(model.hparams.keys())),
           'num_epochs_completed': int(model.hparams.num_epochs_completed)}
    for k, v in state(model).items():
        out[k] = v
    return out


def hybrid_case(seed, freeze, n_users=37, n_items=53, dim=12, n_meta=6, n_neg=4, batch=64,
                steps=3, meta_layers=None):
    """``HybridPretrainedModel`` of the reference (hybrid_pretrained_matrix_factorization.py):
    predictions, and losses + every parameter after ``steps`` iterations of the reference's own
    loop (calculate_loss -> backward -> Adam step) with injected negatives."""
    from collie.model import HybridPretrainedModel
    rng = np.random.default_rng(seed)
    torch.manual_seed(seed)
    users, items = synthetic_pairs(rng, n_users, n_items, 400)
    train = Interactions(users=users, items=items, ratings=np.ones(len(users)),
                         num_negative_samples=n_neg, num_users=n_users, num_items=n_items,
                         allow_missing_ids=True, remove_duplicate_user_item_pairs=False,
                         check_num_negative_samples_is_valid=False)
    mf = MatrixFactorizationModel(train=train, embedding_dim=dim, lr_scheduler_func=None)
    randomize(mf, rng, scale=0.5)
    meta = rng.normal(0, 1, size=(n_items, n_meta)).astype(np.float32)
    model = HybridPretrainedModel(train=train, item_metadata=meta, trained_model=mf,
                                  item_metadata_layers_dims=meta_layers,
                                  combined_layers_dims=[16, 8], freeze_embeddings=freeze,
                                  lr=1e-2, optimizer='adam', loss='adaptive_hinge',
                                  lr_scheduler_func=None)
    with torch.no_grad():
        for name, p in model.named_parameters():
            if 'layer' in name:
                p.copy_(torch.from_numpy(rng.normal(0, 0.4, size=tuple(p.shape)).astype(np.float32)))
    out = {'n_users': n_users, 'n_items': n_items, 'dim': dim, 'n_neg': n_neg, 'lr': 1e-2,
           'freeze': int(freeze), 'item_metadata': meta,
           'meta_layers': np.array(meta_layers or [], dtype=np.int64)}
    for k, v in model.state_dict().items():
        if '_trained_model' not in k:
            out['init_' + k] = v.detach().numpy().copy()
    model.eval()
    pu, pi = rng.integers(0, n_users, 80), rng.integers(0, n_items, 80)
    with torch.no_grad():
        out['pred_users'], out['pred_items'] = pu, pi
        out['preds'] = model(torch.from_numpy(pu), torch.from_numpy(pi)).numpy()
    model.train()
    opt = model.configure_optimizers()
    assert isinstance(opt, torch.optim.Optimizer)
    for s in range(steps):
        bu = rng.integers(0, n_users, batch)
        bp = rng.integers(0, n_items, batch)
        bn = rng.integers(0, n_items, (batch, n_neg))
        b = ((torch.from_numpy(bu), torch.from_numpy(bp)), torch.from_numpy(bn))
        opt.zero_grad()
        loss = model.calculate_loss(b)
        loss.backward()
        opt.step()
        out[f'step{s}_users'], out[f'step{s}_pos'], out[f'step{s}_negs'] = bu, bp, bn
        out[f'step{s}_loss'] = float(loss.detach())
    for k, v in model.state_dict().items():
        if '_trained_model' not in k:
            out['final_' + k] = v.detach().numpy().copy()
    return out


def cml_case(seed, optimizer='adam', sparse=False, loss='hinge', n_users=37, n_items=53, dim=12,
             n_neg=4, batch=64, steps=3):
    """``CollaborativeMetricLearningModel`` of the reference (collaborative_metric_learning.py:
    86-124: two embedding tables, prediction = pairwise_distance): predictions, and losses + both
    tables after ``steps`` iterations of the reference's own loop with injected negatives."""
    from collie.model import CollaborativeMetricLearningModel
    rng = np.random.default_rng(seed)
    torch.manual_seed(seed)
    users, items = synthetic_pairs(rng, n_users, n_items, 400)
    train = Interactions(users=users, items=items, ratings=np.ones(len(users)),
                         num_negative_samples=n_neg, num_users=n_users, num_items=n_items,
                         allow_missing_ids=True, remove_duplicate_user_item_pairs=False,
                         check_num_negative_samples_is_valid=False)
    model = CollaborativeMetricLearningModel(train=train, embedding_dim=dim, sparse=sparse, lr=1e-2,
                                             optimizer=optimizer, loss=loss, lr_scheduler_func=None)
    with torch.no_grad():
        for p in model.parameters():
            p.copy_(torch.from_numpy(rng.normal(0, 0.5, size=tuple(p.shape)).astype(np.float32)))
    out = {'n_users': n_users, 'n_items': n_items, 'dim': dim, 'n_neg': n_neg, 'lr': 1e-2}
    for k, v in model.state_dict().items():
        out['init_' + k] = v.detach().numpy().copy()
    model.eval()
    pu, pi = rng.integers(0, n_users, 80), rng.integers(0, n_items, 80)
    with torch.no_grad():
        out['pred_users'], out['pred_items'] = pu, pi
        out['preds'] = model(torch.from_numpy(pu), torch.from_numpy(pi)).numpy()
    model.train()
    opt = model.configure_optimizers()
    assert isinstance(opt, torch.optim.Optimizer)
    for s in range(steps):
        bu = rng.integers(0, n_users, batch)
        bp = rng.integers(0, n_items, batch)
        bn = rng.integers(0, n_items, (batch, n_neg))
        b = ((torch.from_numpy(bu), torch.from_numpy(bp)), torch.from_numpy(bn))
        opt.zero_grad()
        loss = model.calculate_loss(b)
        loss.backward()
        opt.step()
        out[f'step{s}_users'], out[f'step{s}_pos'], out[f'step{s}_negs'] = bu, bp, bn
        out[f'step{s}_loss'] = float(loss.detach())
    for k, v in model.state_dict().items():
        out['final_' + k] = v.detach().numpy().copy()
    return out


def main():
    os.makedirs(OUT, exist_ok=True)
    cases = {
        'implicit_bpr_n1_sparse_adam': implicit_case('bpr', 1, 'sparse_adam', True, 11),
        'implicit_hinge_n1_adam': implicit_case('hinge', 1, 'adam', False, 12),
        'implicit_adaptive_bpr_n5_sparse_adam': implicit_case('bpr', 5, 'sparse_adam', True, 13),
        'implicit_adaptive_hinge_n10_adam': implicit_case('adaptive', 10, 'adam', False, 14),
        'implicit_warp_n7_sparse_adam': implicit_case('warp', 7, 'sparse_adam', True, 15),
        'implicit_warp_n4_sgd': implicit_case('warp', 4, 'sgd', False, 16),
        'implicit_adaptive_hinge_meta_n5': implicit_case('adaptive_hinge', 5, 'sparse_adam', True,
                                                         17, with_metadata=True),
        'implicit_warp_meta_n6': implicit_case('warp', 6, 'adam', False, 18, with_metadata=True),
        'implicit_adaptive_bpr_d10_n5': implicit_case('bpr', 5, 'adam', False, 19, dim=10),
        'implicit_adaptive_bpr_d30_n3': implicit_case('bpr', 3, 'sparse_adam', True, 20, dim=30),
        'implicit_adaptive_bpr_d128_n5': implicit_case('bpr', 5, 'sparse_adam', True, 21, dim=128),
        'implicit_hinge_adagrad': implicit_case('hinge', 1, 'adagrad', False, 22),
        'explicit_mse_adam': explicit_case('mse', 'adam', False, 31),
        'explicit_mse_sparse_adam_d64': explicit_case('mse', 'sparse_adam', True, 32, dim=64),
        'explicit_mae_sgd': explicit_case('mae', 'sgd', False, 33),
        'explicit_mse_yrange': explicit_case('mse', 'adam', False, 34, y_range=(1.0, 5.0)),
        'metrics_random': metrics_case(41),
        'metrics_exact_grid': metrics_case(42, exact_grid=True),
        'metrics_d128': metrics_case(43, n_users=64, n_items=333, dim=128, n_test=700),
    }
    if len(sys.argv) > 1 and sys.argv[1] in ('similarity', 'checkpoint', 'hybrid', 'cml'):   # the cases added later
        cases = {}
    if len(sys.argv) > 1 and sys.argv[1] == 'cml':
        cases = {'cml_hinge_adam': cml_case(91),
                 'cml_adaptive_sparse_adam': cml_case(92, optimizer='sparse_adam', sparse=True,
                                                      loss='adaptive_hinge', dim=32),
                 'cml_bpr_sgd': cml_case(93, optimizer='sgd', loss='bpr', dim=7)}
        for name, arrays in cases.items():
            np.savez_compressed(os.path.join(OUT, name + '.npz'), **arrays)
            print(name, 'ok')
        return
    if len(sys.argv) > 1 and sys.argv[1] == 'hybrid':
        cases = {'hybrid_frozen': hybrid_case(81, True),
                 'hybrid_unfrozen': hybrid_case(82, False),
                 'hybrid_meta_layers': hybrid_case(83, False, meta_layers=[10, 7])}
        for name, arrays in cases.items():
            np.savez_compressed(os.path.join(OUT, name + '.npz'), **arrays)
            print(name, 'ok')
        return
    if len(sys.argv) > 1 and sys.argv[1] == 'checkpoint':
        cases = {'reference_checkpoint': checkpoint_case(71)}
        for name, arrays in cases.items():
            np.savez_compressed(os.path.join(OUT, name + '.npz'), **arrays)
            print(name, 'ok')
        return
    cases['splits'] = split_case(61)
    cases['similarity_random'] = similarity_case(51)
    cases['similarity_d128'] = similarity_case(52, n_users=40, n_items=90, dim=128)
    for name, arrays in cases.items():
        np.savez_compressed(os.path.join(OUT, name + '.npz'), **arrays)
        print(name, 'ok')
    print('torch', torch.__version__, 'collie', collie.__version__, file=sys.stderr)


if __name__ == '__main__':
    main()
