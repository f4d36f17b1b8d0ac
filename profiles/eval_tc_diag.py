"""Tensor-core evaluation at the bench shape: time per call, users handed to the exact path, for
both operand forms (1 bf16 term: top-k / MAP only; 3 terms: with the MRR counter) and two weight
scales. Usage (GPU box): python profiles/eval_tc_diag.py"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from collie_b200 import metrics as M  # noqa: E402
import scipy.sparse as sp  # noqa: E402

n_users, n_items, dim, k = 32768, 26744, 128, 10
rng = np.random.default_rng(0)
tu = np.repeat(np.arange(n_users), 8)
ti = rng.integers(0, n_items, len(tu))
coo = sp.coo_matrix((np.ones(len(tu)), (tu, ti)), shape=(n_users, n_items))
coo.sum_duplicates()
indptr, items, nnz = M._targets_device(coo.tocsr(), np.arange(n_users), 'cuda')
users = torch.arange(n_users, dtype=torch.int32, device='cuda')
from collie_b200 import _abi  # noqa: E402
if len(sys.argv) > 1:
    _abi.lib().cmf_set_tuning(5, int(sys.argv[1]))
if len(sys.argv) > 2:          # experiments: 1 = no ranking at all, 2 = masks but no inserts
    _abi.lib().cmf_set_tuning(3, int(sys.argv[2]))
if len(sys.argv) > 3:
    _abi.lib().cmf_set_tuning(7, int(sys.argv[3]))
out = {'list_len_override': sys.argv[1:] or None}
for label, scale, bscale in (('scaled_embedding_init', 1.0 / (2.5 * dim), 0.0),
                             ('trained_like', 0.3 / np.sqrt(dim), 0.05)):
    g = torch.Generator(device='cuda').manual_seed(1)
    tabs = (torch.randn(n_users, dim, device='cuda', generator=g) * scale,
            torch.randn(n_items, dim, device='cuda', generator=g) * scale,
            torch.randn(n_users, 1, device='cuda', generator=g) * bscale,
            torch.randn(n_items, 1, device='cuda', generator=g) * bscale)
    for want in (('topk', 'mapk'), ('topk', 'mapk', 'mrr')):
        status = torch.zeros(8, dtype=torch.int32, device='cuda')
        ws = M._eval_tile('cuda', n_users, n_items, k, indptr, items, set(want), tables=tabs,
                          user_ids=users, precision=1, tc_status=status)['_workspace']
        torch.cuda.synchronize()
        first = status[:2].tolist()
        counters = status[4:].view(torch.int64).tolist()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            res = M._eval_tile('cuda', n_users, n_items, k, indptr, items, set(want), tables=tabs,
                               user_ids=users, precision=1, workspace=ws, tc_status=status)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        out[f'{label}/{"+".join(want)}'] = {
            'ms': ms, 'users_per_s': n_users / ms * 1e3, 'status_first_call': first, 'inserts_per_row': counters[0] / n_users,
            'insert_rounds_per_group': counters[1] / (n_users / 32 * n_items / 32),
            'useful_tflops': 2.0 * n_users * n_items * dim / ms / 1e9}
print(json.dumps(out, indent=1))
