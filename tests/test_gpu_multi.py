"""2-GPU NCCL test: ``ShardedMatrixFactorization`` (users partitioned, item table row-sharded)
reproduces the single-GPU step on the union of the ranks' batches. Needs >= 2 CUDA devices; run
with ``gpurun --gpus 2``. Launches its own two ranks."""
import os
import socket
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _rank_main(rank, world, port, ret):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    import collie_b200 as cb
    from collie_b200 import distributed as D
    from collie_b200.interactions import DeviceBatch
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dev = torch.device('cuda', rank)
    dist.init_process_group('nccl', rank=rank, world_size=world, device_id=dev)
    try:
        for mode in ('nccl', 'p2p'):
            _check_mode(mode, rank, world, dev, dist, cb, D, DeviceBatch)
        ret[rank] = 'ok'
    except Exception as e:  # pragma: no cover
        import traceback
        ret[rank] = traceback.format_exc()
    finally:
        dist.destroy_process_group()


def _check_mode(mode, rank, world, dev, dist, cb, D, DeviceBatch):
    """mode: how the item side is exchanged ('nccl' collectives or the 'p2p' owner-side kernel)."""
    if True:
        n_users, n_items, dim, n_neg, steps = 61, 45, 32, 4, 3     # 45 items -> padded to 46
        rng = np.random.default_rng(7)                              # identical on every rank
        flat = rng.choice(n_users * n_items, size=1500, replace=False)
        users, items = flat // n_items, flat % n_items
        per_rank, padded = D.shard_rows(n_items, world)
        kw = dict(num_negative_samples=n_neg, allow_missing_ids=True, seed=5,
                  check_num_negative_samples_is_valid=False, max_number_of_samples_to_consider=0)
        # ---- the single-GPU model on the whole data (every rank builds the same one)
        full = cb.Interactions(users=users, items=items, num_users=n_users, num_items=n_items, **kw)
        ref = cb.MatrixFactorizationModel(train=full, embedding_dim=dim, sparse=True, lr=1e-2,
                                          optimizer='sparse_adam', loss='bpr',
                                          lr_scheduler_func=None).to(dev)
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
        model = cb.MatrixFactorizationModel(train=local, embedding_dim=dim, sparse=True, lr=1e-2,
                                            optimizer='sparse_adam', loss='bpr',
                                            lr_scheduler_func=None).to(dev)
        mine = torch.arange(rank, n_users, world, device=dev)
        with torch.no_grad():
            model.user_embeddings.weight.copy_(ref.user_embeddings.weight[mine])
            model.user_biases.weight.copy_(ref.user_biases.weight[mine])
            model.item_embeddings.weight.zero_()
            model.item_biases.weight.zero_()
            if rank == 0:       # rank 0's item table is broadcast by the wrapper
                model.item_embeddings.weight[:n_items].copy_(ref.item_embeddings.weight)
                model.item_biases.weight[:n_items].copy_(ref.item_biases.weight)
        sharded = D.ShardedMatrixFactorization(model, n_items, exchange=mode)
        assert (sharded._p2p is not None) == (mode == 'p2p')
        opts = model.configure_optimizers()
        st_full, st_loc = full.device_state(dev), local.device_state(dev)
        for step in range(steps):
            # global batch: dataset positions [step*400, step*400+400); each rank takes its users
            sel = np.arange(step * 400, (step + 1) * 400)
            idx = torch.from_numpy(sel).to(dev)
            gb = DeviceBatch(st_full['users'][idx], st_full['items'][idx], idx, None)
            ref_loss = ref.fused_step(gb, optimizers=ref_opts, interactions=full, epoch=step)
            lsel = torch.from_numpy(np.nonzero(np.isin(gpos, sel))[0]).to(dev)
            lb = DeviceBatch(st_loc['users'][lsel], st_loc['items'][lsel], st_loc['gpos'][lsel], None)
            loss = sharded.train_step(lb, opts, interactions=local, epoch=step)
            np.testing.assert_allclose(loss.item(), ref_loss.item(), rtol=1e-5)
        # (fp32 sums of the same contributions in a different order, then Adam's normalisation)
        I, bi = sharded.gather_item_table()
        np.testing.assert_allclose(I.cpu().numpy(), ref.item_embeddings.weight.detach().cpu().numpy(),
                                   rtol=1e-4, atol=1e-5)
        np.testing.assert_allclose(bi.cpu().numpy(), ref.item_biases.weight.detach().cpu().numpy(),
                                   rtol=1e-4, atol=1e-5)
        np.testing.assert_allclose(model.user_embeddings.weight.detach().cpu().numpy(),
                                   ref.user_embeddings.weight[mine].detach().cpu().numpy(),
                                   rtol=1e-4, atol=1e-5)
        # user-sharded evaluation == single-GPU evaluation
        test_flat = rng.choice(n_users * n_items, size=300, replace=False)
        tu, ti = test_flat // n_items, test_flat % n_items
        test_full = cb.Interactions(users=tu, items=ti, num_users=n_users, num_items=n_items, **kw)
        exp = cb.evaluate_in_batches([cb.mapk, cb.mrr, cb.auc], test_full, ref, k=5, verbose=False)
        ltu, lti, _ = D.partition_interactions(tu, ti, rank, world)
        test_loc = cb.Interactions(users=ltu, items=lti,
                                   num_users=D.num_local_users(n_users, rank, world),
                                   num_items=n_items, **kw)
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


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason='needs 2 GPUs')
def test_sharded_step_matches_single_gpu():
    """Launches two ranks the way the driver does (torch.distributed.run, one per GPU)."""
    import subprocess
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2',
           '--master-addr', '127.0.0.1', '--master-port', str(_free_port()),
           os.path.abspath(__file__)]
    out = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                         timeout=280).stdout
    assert "0 {0: 'ok'}" in out and "1 {1: 'ok'}" in out, out[-4000:]


if __name__ == '__main__':   # one rank of the test above
    import faulthandler
    faulthandler.dump_traceback_later(200, exit=True)
    r = {}
    _rank_main(int(os.environ['RANK']), int(os.environ['WORLD_SIZE']),
               int(os.environ['MASTER_PORT']), r)
    print(os.environ['RANK'], r, flush=True)
