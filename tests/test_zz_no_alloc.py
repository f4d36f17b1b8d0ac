"""The allocation-free evaluation entry point, `cmf_eval_tile_noalloc`, against `cmf_eval_tile`.

Some hosts of this library cannot give it scratch memory: they were handed the tables and the
output buffers by another runtime and own no device allocator, or their framework has already
reserved the whole GPU. For them the library carries a fixed 16 MiB scratch array of its own. This
test drives both entry points through the C ABI on the same inputs, twice in a row with different
tile shapes (the array keeps the first call's contents), and expects identical results: ids and
counters bit for bit, AP / AUC sums exactly equal (the same kernels run, only the scratch differs).
It lives in its own file, collected last, so a failure here cannot hide other tests' results.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _call(entry, tables, users, n_items, k, indptr, tgt_items, workspace=None):
    from collie_b200 import _abi
    dev = tables[0].device
    n = users.shape[0]
    a = _abi.cmf_eval_t()
    a.U, a.I, a.bu, a.bi = (t.data_ptr() for t in tables)
    a.D = tables[0].shape[1]
    a.user_ids, a.n_users_tile = users.data_ptr(), n
    a.item_offset, a.num_items_shard, a.k = 0, n_items, k
    a.tgt_indptr, a.tgt_items = indptr.data_ptr(), tgt_items.data_ptr()
    a.want_mrr = a.want_auc = 1
    out = {
        'topk_ids': torch.full((n, k), -7, dtype=torch.int32, device=dev),
        'topk_scores': torch.zeros((n, k), dtype=torch.float32, device=dev),
        'ap': torch.zeros(n, dtype=torch.float64, device=dev),
        'n_before_best': torch.zeros(n, dtype=torch.int32, device=dev),
        'best_score': torch.zeros(n, dtype=torch.float32, device=dev),
        'best_id': torch.zeros(n, dtype=torch.int32, device=dev),
        'auc_pairs': torch.zeros(n, dtype=torch.float64, device=dev),
    }
    for key, t in out.items():
        setattr(a, key, t.data_ptr())
    if workspace is not None:
        a.workspace, a.workspace_bytes = workspace.data_ptr(), workspace.numel()
    rc = getattr(_abi.lib(), entry)(_abi.C.byref(a), _abi.stream_ptr(dev))
    torch.cuda.synchronize()
    return rc, out


def _problem(seed, n_users, n_items, dim, n_tile):
    import scipy.sparse as sp
    from collie_b200 import metrics as M
    rng = np.random.default_rng(seed)
    tables = tuple(torch.from_numpy(rng.normal(0, 0.6, shape).astype(np.float32)).cuda()
                   for shape in ((n_users, dim), (n_items, dim), (n_users, 1), (n_items, 1)))
    flat = rng.choice(n_users * n_items, size=6 * n_users, replace=False)
    coo = sp.coo_matrix((np.ones(len(flat)), (flat // n_items, flat % n_items)),
                        shape=(n_users, n_items))
    tile = np.sort(rng.choice(n_users, size=n_tile, replace=False))
    indptr, tgt_items, _ = M._targets_device(coo.tocsr(), tile, 'cuda')
    return tables, torch.from_numpy(tile.astype(np.int32)).cuda(), indptr, tgt_items


def test_noalloc_entry_matches_workspace_entry_twice_in_a_row():
    from collie_b200 import _abi
    lib = _abi.lib()
    # two different shapes back to back: the second, smaller call runs over the first one's leftovers
    for seed, (n_users, n_items, dim, n_tile, k) in enumerate([(90, 700, 32, 64, 10),
                                                               (40, 333, 128, 17, 5)]):
        tables, users, indptr, tgt_items = _problem(seed, n_users, n_items, dim, n_tile)
        need = lib.cmf_eval_workspace_bytes(n_tile, n_items, dim, k, 0)
        ws = torch.empty(int(need), dtype=torch.uint8, device='cuda')
        rc_ref, ref = _call('cmf_eval_tile', tables, users, n_items, k, indptr, tgt_items, ws)
        rc_new, new = _call('cmf_eval_tile_noalloc', tables, users, n_items, k, indptr, tgt_items)
        assert rc_ref == 0, lib.cmf_last_error()
        assert rc_new == 0, lib.cmf_last_error()
        for key in ref:
            assert torch.equal(ref[key], new[key]), key
        assert int(ref['topk_ids'].min()) >= 0          # the outputs were really written


def test_noalloc_entry_refuses_tiles_beyond_its_scratch():
    from collie_b200 import _abi
    lib = _abi.lib()
    tables, users, indptr, tgt_items = _problem(3, 30, 50, 8, 4)
    # claim a shard far beyond 16 MiB of scratch: refused before anything is launched
    a = _abi.cmf_eval_t()
    a.U, a.I, a.bu, a.bi = (t.data_ptr() for t in tables)
    a.D, a.user_ids, a.n_users_tile = 8, users.data_ptr(), 4096
    a.num_items_shard, a.k = 100000, 10
    a.tgt_indptr, a.tgt_items = indptr.data_ptr(), tgt_items.data_ptr()
    rc = lib.cmf_eval_tile_noalloc(_abi.C.byref(a), None)
    assert rc == -3
    msg = lib.cmf_last_error().decode()
    assert 'built-in scratch' in msg and '16777216' in msg
