"""Exact-path evaluation at the bench shape (for ncu): one 512-user tile, all three metrics.
Usage (GPU box): [ncu ...] python profiles/exact_eval_diag.py"""
import os
import sys

import numpy as np
import scipy.sparse as sp
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from collie_b200 import metrics as M  # noqa: E402

n_users, n_items, dim, k = 512, 26744, 128, 10
rng = np.random.default_rng(0)
tu = np.repeat(np.arange(n_users), 8)
ti = rng.integers(0, n_items, len(tu))
coo = sp.coo_matrix((np.ones(len(tu)), (tu, ti)), shape=(n_users, n_items))
coo.sum_duplicates()
indptr, items, nnz = M._targets_device(coo.tocsr(), np.arange(n_users), 'cuda')
users = torch.arange(n_users, dtype=torch.int32, device='cuda')
g = torch.Generator(device='cuda').manual_seed(1)
tabs = (torch.randn(n_users, dim, device='cuda', generator=g) * 0.03,
        torch.randn(n_items, dim, device='cuda', generator=g) * 0.03,
        torch.randn(n_users, 1, device='cuda', generator=g) * 0.05,
        torch.randn(n_items, 1, device='cuda', generator=g) * 0.05)
ws = None
for want in ({'topk', 'mapk', 'mrr', 'auc'}, {'topk', 'mapk', 'mrr'}):
    for _ in range(3):
        ws = M._eval_tile('cuda', n_users, n_items, k, indptr, items, want, tables=tabs,
                          user_ids=users, precision=0, workspace=ws)['_workspace']
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        M._eval_tile('cuda', n_users, n_items, k, indptr, items, want, tables=tabs,
                     user_ids=users, precision=0, workspace=ws)
    e1.record()
    torch.cuda.synchronize()
    print(sorted(want), 'ms per 512-user tile:', e0.elapsed_time(e1) / 10)
