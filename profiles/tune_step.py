"""Phase timings of the fused training step on the bench workload for several tuning settings.
Usage (GPU box): python profiles/tune_step.py [--batch B] > gpurun_out/tune.json"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import collie_b200 as cb  # noqa: E402
from collie_b200 import _abi  # noqa: E402
from collie_b200.interactions import DeviceBatch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=262144)
    ap.add_argument('--interactions', type=int, default=4_000_000)
    ap.add_argument('--groups', default='8,16,32')
    ap.add_argument('--chunks', default='0,4,8,16')
    ap.add_argument('--no-train', action='store_true')
    args = ap.parse_args()
    w = bench.WORKLOAD
    users, items = bench.synth_interactions(w['num_users'], w['num_items'], args.interactions,
                                            w['seed'])
    data = cb.Interactions(users=users.numpy(), items=items.numpy(), num_users=w['num_users'],
                           num_items=w['num_items'], num_negative_samples=w['n_neg'],
                           allow_missing_ids=True, remove_duplicate_user_item_pairs=False,
                           check_num_negative_samples_is_valid=False,
                           max_number_of_samples_to_consider=0, seed=w['seed'])
    model = cb.MatrixFactorizationModel(train=data, embedding_dim=w['dim'], sparse=True,
                                        optimizer='sparse_adam', loss='bpr',
                                        lr_scheduler_func=None).cuda()
    opts = model.configure_optimizers()
    st = data.device_state('cuda')
    perm = torch.randperm(args.interactions, device='cuda')
    u, i = st['users'][perm], st['items'][perm]
    B = args.batch
    nb = args.interactions // B
    lib = _abi.lib()
    C = _abi.C
    out = []
    for g in [int(x) for x in args.groups.split(',')]:
        for L in [int(x) for x in args.chunks.split(',')]:
            lib.cmf_set_tuning(1, g)
            lib.cmf_set_tuning(0, L)
            for k in range(3):
                a = (k % nb) * B
                model.fused_step(DeviceBatch(u[a:a + B], i[a:a + B], perm[a:a + B], None),
                                 optimizers=opts, interactions=data, train=not args.no_train)
            torch.cuda.synchronize()
            lib.cmf_profile_enable(1)
            for k in range(10):
                a = ((k + 3) % nb) * B
                model.fused_step(DeviceBatch(u[a:a + B], i[a:a + B], perm[a:a + B], None),
                                 optimizers=opts, interactions=data, train=not args.no_train)
            ms = (C.c_double * 4)()
            split = C.c_int32()
            n, ir, ur = C.c_int64(), C.c_int64(), C.c_int64()
            lib.cmf_profile_read(ms, C.byref(split), C.byref(n), C.byref(ir), C.byref(ur))
            lib.cmf_profile_enable(0)
            names = ['fwd', 'sort_wait', 'bwd', 'update'] if split.value else \
                ['sort', 'fused', '-', 'update']
            rec = dict(G=g, L=L, split=split.value, item_rows=ir.value, user_rows=ur.value)
            for k, nm in enumerate(names):
                rec[nm + '_ms'] = round(ms[k] / n.value, 4)
            rec['total_ms'] = round(sum(ms) / n.value, 4)
            out.append(rec)
            print(json.dumps(rec), flush=True)


if __name__ == '__main__':
    main()
