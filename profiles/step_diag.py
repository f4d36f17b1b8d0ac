"""Device activities of one single-GPU training step (torch.profiler / CUPTI): every kernel, memset
and copy with its time per step, plus the idle time between them. Usage: python profiles/step_diag.py"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import collie_b200 as cb  # noqa: E402

w = bench.WORKLOAD
dev = torch.device('cuda', 0)
B = w['batch_per_gpu']
users, items = bench.synth_interactions(w['num_users'], w['num_items'], w['interactions'], w['seed'])
data = cb.Interactions(users=users.numpy(), items=items.numpy(), num_users=w['num_users'],
                       num_items=w['num_items'], num_negative_samples=w['n_neg'],
                       allow_missing_ids=True, remove_duplicate_user_item_pairs=False,
                       check_num_negative_samples_is_valid=False,
                       max_number_of_samples_to_consider=0, seed=w['seed'])
model = cb.MatrixFactorizationModel(train=data, embedding_dim=w['dim'], sparse=True, lr=w['lr'],
                                    bias_lr=w['bias_lr'], optimizer='sparse_adam',
                                    bias_optimizer='sgd', loss='bpr', lr_scheduler_func=None).to(dev)
opts = model.configure_optimizers()
loader = cb.InteractionsDataLoader(data, batch_size=B, shuffle=True, drop_last=True)
it = loader.device_batches(dev, epoch=0)
for _ in range(5):
    model.fused_step(next(it), optimizers=opts, interactions=data, epoch=1)
torch.cuda.synchronize()
steps = 20
from torch.profiler import profile, ProfilerActivity  # noqa: E402
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(steps):
        model.fused_step(next(it), optimizers=opts, interactions=data, epoch=1)
    torch.cuda.synchronize()
evs = sorted((e for e in prof.events() if e.device_type.name == 'CUDA'),
             key=lambda e: e.time_range.start)
busy = sum(e.time_range.end - e.time_range.start for e in evs)
span = evs[-1].time_range.end - evs[0].time_range.start
rows = {}
for e in evs:
    r = rows.setdefault(e.name, [0.0, 0])
    r[0] += e.time_range.end - e.time_range.start
    r[1] += 1
print('per step (us, count):')
for k, (t, c) in sorted(rows.items(), key=lambda kv: -kv[1][0]):
    print(f'{t / steps:9.1f} {c / steps:5.1f}  {k[:100]}')
print(f'{busy / steps:9.1f}  busy; {span / steps:9.1f} wall per step; idle {(span - busy) / steps:.1f}')
one = [e for e in evs if e.time_range.start >= evs[len(evs) // 2].time_range.start][:12]
t0 = one[0].time_range.start
for e in one:
    print(f'  +{e.time_range.start - t0:8.1f} .. +{e.time_range.end - t0:8.1f}  {e.name[:80]}')
