"""World-size-2 ``gloo`` tests of the multi-GPU host logic (no kernels run here): user partition,
row-shard geometry, the in-place all-gather / reduce-scatter plumbing and the top-k merge."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from collie_b200 import distributed as D


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, fn, ret):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        fn(rank, world)
        ret[rank] = 'ok'
    except Exception as e:  # pragma: no cover
        ret[rank] = repr(e)
    finally:
        dist.destroy_process_group()


def _run(fn, world=2):
    ret = mp.Manager().dict()
    port = _free_port()
    procs = [mp.Process(target=_worker, args=(r, world, port, fn, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
    assert dict(ret) == {r: 'ok' for r in range(world)}, dict(ret)


def test_partition_covers_every_interaction_once():
    rng = np.random.default_rng(0)
    users, items = rng.integers(0, 37, 500), rng.integers(0, 50, 500)
    for world in (1, 2, 3, 8):
        seen = np.zeros(500, dtype=int)
        for r in range(world):
            lu, li, pos = D.partition_interactions(users, items, r, world)
            assert (lu * world + r == users[pos]).all() and (li == items[pos]).all()
            assert lu.max(initial=-1) < D.num_local_users(37, r, world)
            seen[pos] += 1
        assert (seen == 1).all()
        assert sum(D.num_local_users(37, r, world) for r in range(world)) == 37
        per, padded = D.shard_rows(50, world)
        assert per * world == padded >= 50 and padded - 50 < world


def _collectives(rank, world):
    per, padded = D.shard_rows(7, world)            # 7 items over 2 ranks -> 4 + 4 (1 pad row)
    full = torch.full((padded, 3), -1.0)
    shard = full[rank * per:(rank + 1) * per]
    shard.copy_(torch.arange(per * 3, dtype=torch.float32).view(per, 3) + 100 * rank)
    D.all_gather_rows(full, shard)                   # in place: shard is a slice of full
    for r in range(world):
        exp = torch.arange(per * 3, dtype=torch.float32).view(per, 3) + 100 * r
        assert torch.equal(full[r * per:(r + 1) * per], exp)
    grads = torch.arange(padded * 3, dtype=torch.float32).view(padded, 3) * (rank + 1)
    out = torch.zeros(per, 3)
    D.reduce_scatter_sum(out, grads)
    total = torch.arange(padded * 3, dtype=torch.float32).view(padded, 3) * sum(range(1, world + 1))
    assert torch.equal(out, total[rank * per:(rank + 1) * per])
    flags = torch.tensor([1, 0, 0, 1, 0, 0, 1, 0], dtype=torch.int32) * (rank == 0) + \
        torch.tensor([0, 0, 1, 1, 0, 0, 0, 0], dtype=torch.int32) * (rank == 1)
    fo = torch.zeros(per, dtype=torch.int32)
    D.reduce_scatter_sum(fo, flags)
    assert fo.tolist() == ([1, 0, 1, 2] if rank == 0 else [0, 0, 1, 0])
    # ranks with different numbers of batches agree on the smaller one
    assert D.common_steps_per_epoch([0] * (5 + rank)) == 5
    # loss pieces add up to the global mean
    loss = torch.tensor(float(rank + 1)) / world
    dist.all_reduce(loss)
    assert abs(loss.item() - sum(range(1, world + 1)) / world) < 1e-6


def test_collectives_world_size_2():
    _run(_collectives, 2)


def test_merge_topk_canonical_order():
    from oracle import collie_oracle as O
    rng = np.random.default_rng(1)
    scores = (rng.integers(-3, 4, size=(9, 40)) / 2.0).astype(np.float32)   # many ties
    k, shards = 5, 4
    cand_s, cand_i = [], []
    for sh in range(shards):
        lo, hi = sh * 10, (sh + 1) * 10
        ids = O.topk_ids(scores[:, lo:hi], k) + lo
        cand_i.append(ids)
        cand_s.append(np.take_along_axis(scores, ids, axis=1))
    cand_i[1][:, -1] = -1                                                      # padding entry
    s, i = D.merge_topk(torch.from_numpy(np.concatenate(cand_s, 1)),
                        torch.from_numpy(np.concatenate(cand_i, 1).astype(np.int32)), k)
    masked = scores.copy()
    drop = O.topk_ids(scores[:, 10:20], k)[:, -1] + 10
    # the padded candidate never was among the global top-k unless it really belongs there;
    # compare against the oracle on the candidate set itself
    cand = np.concatenate(cand_i, 1)
    for r in range(scores.shape[0]):
        ids = cand[r][cand[r] >= 0]
        order = sorted(ids.tolist(), key=lambda j: (-scores[r, j], j))[:k]
        assert i[r].tolist() == order
        assert s[r].tolist() == [float(scores[r, j]) for j in order]
    del masked, drop


def test_ap_from_merged_topk_ids_equals_oracle_mapk():
    """The AP@k that item-sharded evaluation reads off the merged top-k ids (distributed.
    ap_from_topk_ids) = the oracle's per-user MAP@k (collie/metrics.py:144-198)."""
    import scipy.sparse as sp
    from oracle import collie_oracle as O
    rng = np.random.default_rng(4)
    n_users, n_items, k = 23, 57, 6
    scores = rng.normal(size=(n_users, n_items)).astype(np.float32)
    tu = rng.integers(0, n_users, 90)
    tu = tu[tu != 7]                                   # a user without targets
    ti = rng.integers(0, n_items, len(tu))
    coo = sp.coo_matrix((np.ones(len(tu)), (tu, ti)), shape=(n_users, n_items))
    coo.sum_duplicates()
    csr = coo.tocsr()
    csr.sort_indices()
    users = np.arange(n_users)
    top = torch.from_numpy(O.topk_ids(scores, k).astype(np.int32))
    top[3, -2:] = -1                                   # padding entries are never hits
    counts = torch.from_numpy(np.diff(csr.indptr).astype(np.int64))
    rows = torch.repeat_interleave(torch.arange(n_users), counts)
    keys = rows * n_items + torch.from_numpy(csr.indices.astype(np.int64))
    got = D.ap_from_topk_ids(top, torch.arange(n_users), keys, counts, n_items).numpy()
    exp = O.mapk_per_user(csr, users, scores, k)
    keep = np.ones(n_users, dtype=bool)
    keep[3] = False
    np.testing.assert_allclose(got[keep], exp[keep], rtol=1e-12)
    assert got[7] == 0.0
    sub = torch.tensor([5, 11])                        # explicit row numbers
    np.testing.assert_allclose(D.ap_from_topk_ids(top[sub], sub, keys, counts[sub], n_items).numpy(),
                               exp[[5, 11]], rtol=1e-12)
