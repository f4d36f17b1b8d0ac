"""Device time of cmf_eval_tile (both precisions) on the bench shape; prints JSON lines.
Usage (GPU box): python profiles/time_eval.py [n_users_tile]"""
import json
import os
import sys

import numpy as np
import scipy.sparse as sp
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from collie_b200 import metrics as M  # noqa: E402

n_users, n_items, dim = 138493, 26744, 128
nt = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
rng = np.random.default_rng(0)
tabs = tuple(torch.randn(s, device='cuda') * 0.1
             for s in ((n_users, dim), (n_items, dim), (n_users, 1), (n_items, 1)))
tu = np.repeat(np.arange(nt), 8)
ti = rng.integers(0, n_items, nt * 8)
coo = sp.coo_matrix((np.ones(len(tu)), (tu, ti)), shape=(n_users, n_items))
coo.sum_duplicates()
indptr, items, nnz = M._targets_device(coo.tocsr(), np.arange(nt), 'cuda')
users = torch.arange(nt, dtype=torch.int32, device='cuda')
import collie_b200._abi as _abi
if len(sys.argv) > 2:
    _abi.lib().cmf_set_tuning(3, int(sys.argv[2]))
for prec, want, tile in ((1, {'mapk'}, nt), (1, {'mapk', 'mrr'}, nt), (0, {'mapk'}, 256),
                         (0, {'mapk', 'mrr', 'auc'}, 256)):
    ws = None
    def run():
        global ws
        for s in range(0, nt, tile):
            e = min(s + tile, nt)
            res = M._eval_tile('cuda', e - s, n_items, 10, indptr[s:e + 1], items, want,
                               tables=tabs, user_ids=users[s:e], precision=prec, workspace=ws,
                               reuse_item_prep=(s > 0))
            ws = res['_workspace']
    run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        run()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    flops = 2.0 * nt * n_items * dim
    print(json.dumps(dict(precision=prec, want=sorted(want), users=nt, tile=tile, ms=round(ms, 3),
                          users_per_s=round(nt / ms * 1e3), useful_tflops=round(flops / ms / 1e9, 1),
                          issued_tflops=round(3 * flops / ms / 1e9, 1) if prec else None)))
