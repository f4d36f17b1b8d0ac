"""Training-step throughput of HybridPretrainedModel at the BASELINE C4 shape (ML-20M shape, 32-dim
item metadata, MLP [2D + 32 -> 128 -> 64 -> 32 -> 1], adaptive hinge): the unfused loop
calculate_loss -> backward -> optimizer step over cmf_gather_rows / cmf_scatter_add_rows and
library GEMMs. Not a bench.py line (C4 is a parity-test configuration there).
Usage (GPU box): python profiles/hybrid_step.py [batch] [n_neg]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import collie_b200 as cb  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
N = int(sys.argv[2]) if len(sys.argv) > 2 else 5
n_users, n_items, dim, n_meta = 138_493, 26_744, 128, 32
rng = np.random.default_rng(0)
n = 2_000_000
users = rng.integers(0, n_users, n)
items = np.minimum((rng.pareto(1.2, n) * 50).astype(np.int64), n_items - 1)
data = cb.Interactions(users=users, items=items, num_users=n_users, num_items=n_items,
                       num_negative_samples=N, allow_missing_ids=True,
                       remove_duplicate_user_item_pairs=False,
                       check_num_negative_samples_is_valid=False,
                       max_number_of_samples_to_consider=0, seed=1)
mf = cb.MatrixFactorizationModel(train=data, embedding_dim=dim, lr_scheduler_func=None).cuda()
meta = torch.from_numpy(rng.normal(0, 1, (n_items, n_meta)).astype(np.float32))
out = {'batch': B, 'n_neg': N}
for freeze in (True, False):
    model = cb.HybridPretrainedModel(train=data, item_metadata=meta, trained_model=mf,
                                     combined_layers_dims=[128, 64, 32], freeze_embeddings=freeze,
                                     lr=1e-3, optimizer='adam', loss='adaptive_hinge',
                                     lr_scheduler_func=None).cuda()
    model.train()
    opt = model.configure_optimizers()
    loader = cb.InteractionsDataLoader(data, batch_size=B, shuffle=True, drop_last=True)
    batches = list(loader.device_batches('cuda', epoch=1))[:12]

    def step(db):
        opt.zero_grad()
        loss = model.calculate_loss(db)
        loss.backward()
        opt.step()
        return loss

    for db in batches[:3]:
        step(db)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for db in batches[3:]:
        loss = step(db)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / len(batches[3:])
    out['frozen' if freeze else 'unfrozen'] = {
        'ms_per_step': ms, 'interactions_per_s': B / ms * 1e3,
        'scored_pairs_per_s': B * (1 + N) / ms * 1e3, 'loss': float(loss.detach())}
print(json.dumps(out, indent=1))
