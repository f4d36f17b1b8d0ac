"""Multi-GPU NCCL tests: ``ShardedMatrixFactorization`` (users partitioned, item table row-sharded)
reproduces the single-GPU step on the union of the ranks' batches, for 2 / 4 / 8 ranks and for
every way the item side can be exchanged -- NCCL collectives, the owner-side kernel with plain
peer loads / stores, and the same kernel over NVSwitch multicast addresses (``multimem``).
Semantics under test: one optimizer step per unique row per GLOBAL batch
(collie/model/base/trainer.py:459-470 is the single-process form).

Needs >= world CUDA devices (``gpurun --gpus N``); every test launches its own ranks with
``torch.distributed.run`` the way the driver launches ``bench.py``."""
import os
import socket
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# tolerances of the single-GPU golden tests (tests/test_gpu_parity.py)
RTOL, ATOL = 1e-5, 1e-6

CASES = {
    # the BASELINE C3 recipe in small: 'bpr' + 5 negatives -> adaptive BPR, sparse Adam + SGD bias
    'bpr_n5_d128': dict(n_users=257, n_items=203, dim=128, n_neg=5, loss='bpr',
                        optimizer='sparse_adam', sparse=True, lr=1e-2),
    # the BASELINE C5 recipe in small: WARP + 10 negatives; D = 30 (the reference's default) runs
    # the generic 64-bit-vector kernels and the exchange kernel's V = 2 form
    'warp_n10_d30': dict(n_users=257, n_items=203, dim=30, n_neg=10, loss='warp',
                         optimizer='sparse_adam', sparse=True, lr=1e-2),
    # a catalogue far larger than a batch: most item rows are NOT gathered in a given step, some
    # only by one rank (the other cases touch every row every step)
    'sparse_touch_d32': dict(n_users=3001, n_items=5003, dim=32, n_neg=4, loss='bpr',
                             optimizer='sparse_adam', sparse=True, lr=1e-2),
    # the reference's DEFAULT optimizer: dense Adam moves every row every step, on every owner
    'hinge_n3_adam_d16': dict(n_users=120, n_items=75, dim=16, n_neg=3, loss='hinge',
                              optimizer='adam', sparse=False, lr=1e-2),
}
MODES = (('nccl', False), ('p2p', False), ('p2p', True))     # (exchange, multicast)
STEPS, GLOBAL_BATCH = 5, 640


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _close(got, exp, what):
    """Tables after several optimizer steps, sharded run vs single-GPU run.

    The two runs add the same fp32 contributions to an item-row gradient in different orders (per
    rank first, then over ranks), so a gradient element differs by a few ulp of its LARGEST term.
    Where the terms of an element nearly cancel, that is a large RELATIVE error of the sum, and Adam
    divides by sqrt(v): in the first steps the update is ~ lr * g / |g|, i.e. it follows the
    relative error of g, not the absolute one. An element whose gradient cancels to 1e-4 of its
    terms therefore moves by ~1e-4 * lr = 1e-6 more or less. So: the golden tolerance (1e-5 rel,
    1e-6 abs) for at least 98 % of the elements (measured: > 99.9 %), and 1e-4 rel + 1e-5 abs (1e-3 * lr) for all."""
    got, exp = got.detach().cpu().numpy(), exp.detach().cpu().numpy()
    err = np.abs(got - exp)
    tight = err <= ATOL + RTOL * np.abs(exp)
    assert tight.mean() >= 0.98, (what, float(tight.mean()), float(err.max()))
    np.testing.assert_allclose(got, exp, rtol=1e-4, atol=1e-5, err_msg=what)


def _check_case(name, case, mode, multicast, rank, world, dev, with_eval):
    import collie_b200 as cb
    from collie_b200 import distributed as D
    from collie_b200.interactions import DeviceBatch
    n_users, n_items, dim, n_neg = case['n_users'], case['n_items'], case['dim'], case['n_neg']
    rng = np.random.default_rng(7)                              # identical on every rank
    flat = rng.choice(n_users * n_items, size=STEPS * GLOBAL_BATCH, replace=False)
    users, items = flat // n_items, flat % n_items
    per_rank, padded = D.shard_rows(n_items, world)
    kw = dict(num_negative_samples=n_neg, allow_missing_ids=True, seed=5,
              check_num_negative_samples_is_valid=False, max_number_of_samples_to_consider=0)
    mkw = dict(embedding_dim=dim, sparse=case['sparse'], lr=case['lr'], bias_lr=3e-2,
               optimizer=case['optimizer'], bias_optimizer='sgd', loss=case['loss'],
               lr_scheduler_func=None)
    # ---- the single-GPU model on the whole data (every rank builds the same one)
    full = cb.Interactions(users=users, items=items, num_users=n_users, num_items=n_items, **kw)
    ref = cb.MatrixFactorizationModel(train=full, **mkw).to(dev)
    gen = torch.Generator().manual_seed(3)
    with torch.no_grad():
        for p in ref.parameters():
            p.copy_(torch.randn(p.shape, generator=gen) * 0.3)
    ref_opts = ref.configure_optimizers()
    # ---- this rank's shard
    lu, li, gpos = D.partition_interactions(users, items, rank, world)
    local = cb.Interactions(users=lu, items=li, num_users=D.num_local_users(n_users, rank, world),
                            num_items=padded, **kw)
    local.global_positions = gpos
    model = cb.MatrixFactorizationModel(train=local, **mkw).to(dev)
    mine = torch.arange(rank, n_users, world, device=dev)
    with torch.no_grad():
        model.user_embeddings.weight.copy_(ref.user_embeddings.weight[mine])
        model.user_biases.weight.copy_(ref.user_biases.weight[mine])
        model.item_embeddings.weight.zero_()
        model.item_biases.weight.zero_()
        if rank == 0:       # rank 0's item table is broadcast by the wrapper
            model.item_embeddings.weight[:n_items].copy_(ref.item_embeddings.weight)
            model.item_biases.weight[:n_items].copy_(ref.item_biases.weight)
    sharded = D.ShardedMatrixFactorization(model, n_items, exchange=mode, multicast=multicast)
    assert (sharded._p2p is not None) == (mode == 'p2p')
    opts = model.configure_optimizers()
    st_full, st_loc = full.device_state(dev), local.device_state(dev)
    for step in range(STEPS):
        # global batch: dataset positions [step*G, (step+1)*G); each rank takes its own users
        sel = np.arange(step * GLOBAL_BATCH, (step + 1) * GLOBAL_BATCH)
        idx = torch.from_numpy(sel).to(dev)
        gb = DeviceBatch(st_full['users'][idx], st_full['items'][idx], idx, None)
        ref_loss = ref.fused_step(gb, optimizers=ref_opts, interactions=full, epoch=step)
        lsel = torch.from_numpy(np.nonzero(np.isin(gpos, sel))[0]).to(dev)
        lb = DeviceBatch(st_loc['users'][lsel], st_loc['items'][lsel], st_loc['gpos'][lsel], None)
        loss = sharded.train_step(lb, opts, interactions=local, epoch=step)
        np.testing.assert_allclose(loss.item(), ref_loss.item(), rtol=1e-5,
                                   err_msg=f'loss of step {step}')
    if mode == 'p2p':
        assert sharded._p2p['multicast'] == multicast
    # same contributions summed in a different order (per rank first), then the optimizer rule
    I, bi = sharded.gather_item_table()
    _close(I, ref.item_embeddings.weight, 'item embeddings')
    _close(bi, ref.item_biases.weight, 'item biases')
    _close(model.user_embeddings.weight, ref.user_embeddings.weight[mine], 'user embeddings')
    _close(model.user_biases.weight, ref.user_biases.weight[mine], 'user biases')
    # the padding rows of the table never move
    assert not bool(sharded.I_full[n_items:].any()) and not bool(sharded.bi_full[n_items:].any())
    if not with_eval:
        return
    # user-sharded evaluation == single-GPU evaluation
    test_flat = rng.choice(n_users * n_items, size=900, replace=False)
    tu, ti = test_flat // n_items, test_flat % n_items
    test_full = cb.Interactions(users=tu, items=ti, num_users=n_users, num_items=n_items, **kw)
    exp = cb.evaluate_in_batches([cb.mapk, cb.mrr, cb.auc], test_full, ref, k=5, verbose=False)
    ltu, lti, _ = D.partition_interactions(tu, ti, rank, world)
    test_loc = cb.Interactions(users=ltu, items=lti,
                               num_users=D.num_local_users(n_users, rank, world),
                               num_items=n_items, **kw)
    # (the sharded model's tables agree with ``ref`` to 1e-5 only: evaluate the SAME tables)
    with torch.no_grad():
        model.user_embeddings.weight.copy_(ref.user_embeddings.weight[mine])
        model.user_biases.weight.copy_(ref.user_biases.weight[mine])
        sharded.I_full[:n_items].copy_(ref.item_embeddings.weight)
        sharded.bi_full[:n_items].copy_(ref.item_biases.weight)
    got = D.evaluate_sharded([cb.mapk, cb.mrr, cb.auc], test_loc, model, k=5)
    np.testing.assert_allclose(got, exp, rtol=1e-5)
    # item-sharded evaluation (every rank scores all users against its block of item rows,
    # partial top-k lists merged, counters summed) == single-GPU evaluation, both arithmetics
    for precision in (0, 1):
        got_i = D.evaluate_item_sharded([cb.mapk, cb.mrr], test_full, ref, k=5,
                                        precision=precision, tile_users=16)
        np.testing.assert_allclose(got_i, exp[:2], rtol=1e-5)
    only = D.evaluate_item_sharded([cb.mrr], test_full, ref, k=5)
    np.testing.assert_allclose(only, exp[1], rtol=1e-5)


def _rank_main(rank, world, port):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device('cuda', rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)
    done = []
    try:
        for name, case in CASES.items():
            for k, (mode, multicast) in enumerate(MODES):
                _check_case(name, case, mode, multicast, rank, world, dev,
                            with_eval=(name == 'bpr_n5_d128' and k == 1))
                done.append(f'{name}/{mode}{"+multicast" if multicast else ""}')
                if rank == 0:
                    print(f'[world {world}] ok: {done[-1]}', flush=True)
        torch.cuda.synchronize()
        dist.barrier()
        dist.destroy_process_group()
        return 'ok'
    except BaseException:  # pragma: no cover
        # leave at once: the other ranks may be waiting in a collective for this one, and the
        # launcher ends them as soon as one rank has failed
        import traceback
        print(f'rank {rank} of {world}: FAILED after {done}:\n' + traceback.format_exc(),
              flush=True)
        os._exit(1)


def _trainer_main(rank, world, port):
    """``CollieTrainer(model, gpus=world)`` under torch.distributed.run: trains over the ranks and
    leaves every rank's model with the same, trained, full tables."""
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device('cuda', rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)
    try:
        import collie_b200 as cb
        rng = np.random.default_rng(11)
        n_users, n_items = 301, 187
        flat = rng.choice(n_users * n_items, size=12000, replace=False)
        kw = dict(num_users=n_users, num_items=n_items, num_negative_samples=5,
                  allow_missing_ids=True, check_num_negative_samples_is_valid=False,
                  max_number_of_samples_to_consider=0, seed=9)
        train = cb.Interactions(users=flat[:10000] // n_items, items=flat[:10000] % n_items, **kw)
        val = cb.Interactions(users=flat[10000:] // n_items, items=flat[10000:] % n_items, **kw)
        torch.manual_seed(4)
        model = cb.MatrixFactorizationModel(
            train=cb.InteractionsDataLoader(train, batch_size=1024, shuffle=True),
            val=cb.InteractionsDataLoader(val, batch_size=1024, shuffle=False),
            embedding_dim=32, sparse=True, lr=3e-2, bias_lr=3e-2, optimizer='sparse_adam',
            loss='bpr', lr_scheduler_func=None)
        start = {k: v.clone() for k, v in model.state_dict().items()}
        before = cb.evaluate_in_batches([cb.mapk], val, model.cuda(), k=10, verbose=False)
        trainer = cb.CollieTrainer(model, max_epochs=4, gpus=world, verbosity=0,
                                   enable_model_summary=False, logger=None)
        trainer.fit(model)
        assert model.hparams.num_epochs_completed == 4
        sd = model.state_dict()
        for k in sd:
            assert sd[k].shape == start[k].shape
            # (user biases receive +g and -g from every implicit interaction: they never move)
            assert k == 'user_biases.weight' or not torch.equal(sd[k].cpu(), start[k].cpu()), k
            # every rank holds the same tables
            ref = sd[k].clone()
            dist.broadcast(ref, src=0)
            assert torch.equal(ref, sd[k]), k
        after = cb.evaluate_in_batches([cb.mapk], val, model, k=10, verbose=False)
        train_map = cb.evaluate_in_batches([cb.mapk], train, model, k=10, verbose=False)
        assert train_map > 5 * max(before, 1e-3), (before, after, train_map)   # it learned
        dist.barrier()
        dist.destroy_process_group()
        return f'ok (train MAP@10 {train_map:.4f}, val {before:.4f} -> {after:.4f})'
    except BaseException:  # pragma: no cover
        import traceback
        print(f'rank {rank} of {world}: FAILED:\n' + traceback.format_exc(), flush=True)
        os._exit(1)


@pytest.mark.parametrize('world', [2])
def test_trainer_gpus_n_trains_over_the_ranks(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f'needs {world} GPUs')
    import subprocess
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node',
           str(world), '--master-addr', '127.0.0.1', '--master-port', str(_free_port()),
           os.path.abspath(__file__), 'trainer']
    out = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                         timeout=300).stdout
    for r in range(world):
        assert f'rank {r} of {world}: ok' in out, out[-6000:]
    print(out[-1500:])


@pytest.mark.parametrize('world', [2, 4, 8])
def test_sharded_step_matches_single_gpu(world):
    """Launches ``world`` ranks the way the driver does (torch.distributed.run, one per GPU)."""
    if torch.cuda.device_count() < world:
        pytest.skip(f'needs {world} GPUs')
    import subprocess
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node',
           str(world), '--master-addr', '127.0.0.1', '--master-port', str(_free_port()),
           os.path.abspath(__file__)]
    out = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                         timeout=300).stdout
    for r in range(world):
        assert f'rank {r} of {world}: ok' in out, out[-6000:]
    print(out[-3000:])


if __name__ == '__main__':   # one rank of the test above
    import faulthandler
    faulthandler.dump_traceback_later(240, exit=True)
    main_fn = _trainer_main if 'trainer' in sys.argv[1:] else _rank_main
    res = main_fn(int(os.environ['RANK']), int(os.environ['WORLD_SIZE']),
                  int(os.environ['MASTER_PORT']))
    print(f"rank {os.environ['RANK']} of {os.environ['WORLD_SIZE']}: {res}", flush=True)
