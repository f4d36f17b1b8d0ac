"""BASELINE config C5 shape across N ranks (10 M users x 2 M items, D = 128, WARP, 10 negatives,
262,144 interactions per rank and step): the sharded training step and a user-sharded tensor-core
evaluation against the full item table. Validation run, not a bench line. Usage (GPU box):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port 29513 profiles/c5_sharded.py [steps]"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import collie_b200 as cb  # noqa: E402
from collie_b200 import distributed as D  # noqa: E402

rank, world, local = (int(os.environ[k]) for k in ('RANK', 'WORLD_SIZE', 'LOCAL_RANK'))
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)
NU, NI, DIM, NNEG, B = 10_000_000, 2_000_000, 128, 10, 262_144
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 12
rng = np.random.default_rng(1)                     # the same global dataset on every rank
n = world * B * 4
users = rng.integers(0, NU, n).astype(np.int64)
items = rng.integers(0, NI, n).astype(np.int64)
lu, li, gpos = D.partition_interactions(users, items, rank, world)
per_rank, padded = D.shard_rows(NI, world)
data = cb.Interactions(users=lu, items=li, num_users=D.num_local_users(NU, rank, world),
                       num_items=padded, num_negative_samples=NNEG, allow_missing_ids=True, seed=3,
                       remove_duplicate_user_item_pairs=False,
                       check_num_negative_samples_is_valid=False,
                       max_number_of_samples_to_consider=0)
data.global_positions = gpos
model = cb.MatrixFactorizationModel(train=data, embedding_dim=DIM, sparse=True, lr=1e-3,
                                    optimizer='sparse_adam', bias_optimizer='sgd', loss='warp',
                                    lr_scheduler_func=None).to(dev)
sharded = D.ShardedMatrixFactorization(model, NI, max_batch=B + B // 4)
opts = model.configure_optimizers()
bsz = min(B, (data.num_interactions // 4) // 1024 * 1024)
loader = cb.InteractionsDataLoader(data, batch_size=bsz, shuffle=True, drop_last=True)


def batches():
    ep = 0
    while True:
        for db in loader.device_batches(dev, epoch=ep):
            yield db, ep
        ep += 1


it = batches()
for _ in range(3):
    db, ep = next(it)
    sharded.train_step(db, opts, interactions=data, epoch=ep, global_batch=world * bsz,
                       reduce_loss=False)
torch.cuda.synchronize()
dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(steps):
    db, ep = next(it)
    loss = sharded.train_step(db, opts, interactions=data, epoch=ep, global_batch=world * bsz,
                              reduce_loss=True)
e1.record()
torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device=dev)
dist.all_reduce(ms, op=dist.ReduceOp.MAX)

# user-sharded evaluation of 4,096 users per rank against all 2 M items (tensor-core path)
tu = np.repeat(np.arange(4096), 4)
ti = rng.integers(0, NI, len(tu))
test = cb.Interactions(users=tu, items=ti, num_users=D.num_local_users(NU, rank, world),
                       num_items=padded, allow_missing_ids=True,
                       remove_duplicate_user_item_pairs=False,
                       check_num_negative_samples_is_valid=False,
                       max_number_of_samples_to_consider=0)
torch.cuda.synchronize()
v0, v1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
v0.record()
metrics = D.evaluate_sharded([cb.mapk, cb.mrr], test, model, k=10, precision=1)
v1.record()
torch.cuda.synchronize()
if rank == 0:
    print(json.dumps({
        'shape': 'C5: 10M users x 2M items, D=128, warp, N=10', 'n_gpus': world,
        'batch_per_gpu': bsz, 'ms_per_step': float(ms.item()),
        'interactions_per_s': world * bsz / (float(ms.item()) * 1e-3), 'loss': float(loss),
        'exchange': 'p2p multicast=%s' % sharded._p2p.get('multicast') if sharded._p2p else 'nccl',
        'eval_users': 4096 * world, 'eval_ms': v0.elapsed_time(v1), 'map10_mrr': metrics,
        'mem_GiB': torch.cuda.max_memory_allocated() / 2 ** 30}), flush=True)
dist.barrier()
dist.destroy_process_group()
