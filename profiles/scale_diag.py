"""Where a sharded training step spends its device time: torch.profiler (CUPTI) kernel totals over
a few steps, rank 0. Usage (GPU box, N ranks):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port 29512 profiles/scale_diag.py [steps]"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import collie_b200 as cb  # noqa: E402
from collie_b200 import distributed as D  # noqa: E402

rank, world, local = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)
w = bench.WORKLOAD
B = w['batch_per_gpu']
users, items = bench.synth_interactions(w['num_users'], w['num_items'], w['interactions'], w['seed'])
lu, li, gpos = D.partition_interactions(users.numpy(), items.numpy(), rank, world)
per_rank, padded = D.shard_rows(w['num_items'], world)
data = cb.Interactions(users=lu, items=li, num_users=D.num_local_users(w['num_users'], rank, world),
                       num_items=padded, num_negative_samples=w['n_neg'], allow_missing_ids=True,
                       remove_duplicate_user_item_pairs=False,
                       check_num_negative_samples_is_valid=False,
                       max_number_of_samples_to_consider=0, seed=w['seed'])
data.global_positions = gpos
model = cb.MatrixFactorizationModel(train=data, embedding_dim=w['dim'], sparse=True, lr=w['lr'],
                                    bias_lr=w['bias_lr'], optimizer='sparse_adam',
                                    bias_optimizer='sgd', loss='bpr', lr_scheduler_func=None).to(dev)
sharded = D.ShardedMatrixFactorization(model, w['num_items'])
opts = model.configure_optimizers()
loader = cb.InteractionsDataLoader(data, batch_size=B, shuffle=True, drop_last=True)
it = loader.device_batches(dev, epoch=0)
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
for _ in range(5):
    sharded.train_step(next(it), opts, interactions=data, epoch=1, global_batch=world * B,
                       reduce_loss=False)
torch.cuda.synchronize()
dist.barrier()
from torch.profiler import profile, ProfilerActivity  # noqa: E402
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(steps):
        sharded.train_step(next(it), opts, interactions=data, epoch=1, global_batch=world * B,
                           reduce_loss=False)
    torch.cuda.synchronize()
if rank == 0:
    rows = [(e.key, e.device_time_total / steps, e.count / steps) for e in prof.key_averages()
            if e.device_time_total > 0 and e.device_type.name == 'CUDA']
    rows.sort(key=lambda r: -r[1])
    print(f'per step, rank 0 of {world} (us, launches):')
    for k, t, c in rows[:25]:
        print(f'{t:9.1f} {c:6.1f}  {k[:110]}')
    print(f'{sum(r[1] for r in rows):9.1f}  total device time per step')
# per-step device time over a fresh epoch (the first step carries the epoch plan)
import time  # noqa: E402
it = None
torch.cuda.synchronize()
dist.barrier()
evs = [torch.cuda.Event(enable_timing=True) for _ in range(37)]
t0 = time.perf_counter()
it = loader.device_batches(dev, epoch=2)
evs[0].record()
host = []
for i in range(36):
    h = time.perf_counter()
    b = next(it)
    h1 = time.perf_counter()
    sharded.train_step(b, opts, interactions=data, epoch=2, global_batch=world * B, reduce_loss=False)
    host.append((h1 - h, time.perf_counter() - h1))
    evs[i + 1].record()
torch.cuda.synchronize()
if rank == 0:
    ms = [evs[i].elapsed_time(evs[i + 1]) for i in range(36)]
    print('device ms per step:', ' '.join(f'{x:.3f}' for x in ms))
    print('host next(batch) ms:', ' '.join(f'{a * 1e3:.3f}' for a, _ in host))
    print('host train_step ms:', ' '.join(f'{b * 1e3:.3f}' for _, b in host))
dist.barrier()
dist.destroy_process_group()
