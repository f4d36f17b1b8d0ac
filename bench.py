#!/usr/bin/env python
"""Benchmark of the MF training hot path (BASELINE.json metric) -- see DESIGN.md "Measurement".

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Own arm: one "step" = one fused training step (on-device negative sampling + gather + scores +
loss + backward + sparse Adam / SGD updates) over one batch of the workload
  C3 = implicit MF, loss 'bpr' with 5 negatives (-> adaptive BPR), sparse Adam + SGD biases,
       embedding_dim 128, synthetic MovieLens-20M shape (138,493 users x 26,744 items, 20 M
       interactions), batch 262,144 interactions per GPU.
`value` is interactions/s with the batch already in HBM; `e2e` is the same loop fed from pinned
host memory (H2D of every batch, D2H of every loss) through the public API. The reference arm
(`--impl reference`) times the oracle's restatement of the reference's own PyTorch CPU path on the
host cores (the reference is pure Python and cannot travel to the GPU box; oracle/ is used here
only as the thing being timed for the baseline, never by the product path).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOAD = {
    'name': 'C3 implicit MF bpr(N=5)->adaptive_bpr, sparse_adam + sgd bias, D=128, '
            'synthetic ML-20M shape',
    'num_users': 138493, 'num_items': 26744, 'interactions': 20_000_000, 'dim': 128, 'n_neg': 5,
    'batch_per_gpu': 262144, 'lr': 1e-3, 'bias_lr': 1e-2, 'seed': 2,
}


def synth_interactions(num_users, num_items, n, seed):
    """MovieLens-20M-like skew: item popularity ~ 1/(rank+60) (top item ~0.3 % of interactions,
    like ML-20M's most-rated film), user activity log-normal. CPU generator -> same tensors for
    every arm and every rank."""
    g = torch.Generator().manual_seed(seed)
    item_w = 1.0 / (torch.arange(num_items, dtype=torch.float64) + 61.0)
    item_cdf = torch.cumsum(item_w / item_w.sum(), 0)
    user_w = torch.exp(torch.randn(num_users, generator=g, dtype=torch.float64))
    user_cdf = torch.cumsum(user_w / user_w.sum(), 0)
    users = torch.searchsorted(user_cdf, torch.rand(n, generator=g, dtype=torch.float64))
    items = torch.searchsorted(item_cdf, torch.rand(n, generator=g, dtype=torch.float64))
    # item ids are shuffled so popularity is not monotone in the id
    relabel = torch.randperm(num_items, generator=g)
    return (users.clamp_(max=num_users - 1).to(torch.int32),
            relabel[items.clamp_(max=num_items - 1)].to(torch.int32))


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler(object):
    """SM clock and throttle reasons sampled through NVML from a thread DURING the timed region."""

    def __init__(self, gpu_index):
        self.gpu_index = gpu_index
        self.samples, self.reasons = [], set()
        self.max_mhz = None
        self._stop = None
        self._thread = None

    def start(self):
        import threading
        try:
            import pynvml
            pynvml.nvmlInit()
            uuid = None
            try:
                uuid = torch.cuda.get_device_properties(self.gpu_index).uuid
            except Exception:
                pass
            h = None
            if uuid is not None:
                try:
                    h = pynvml.nvmlDeviceGetHandleByUUID(('GPU-' + str(uuid)).encode())
                except Exception:
                    h = None
            if h is None:
                h = pynvml.nvmlDeviceGetHandleByIndex(self.gpu_index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
        except Exception:
            return
        names = {'hw_slowdown': pynvml.nvmlClocksThrottleReasonHwSlowdown,
                 'hw_thermal_slowdown': pynvml.nvmlClocksThrottleReasonHwThermalSlowdown,
                 'sw_thermal_slowdown': pynvml.nvmlClocksThrottleReasonSwThermalSlowdown,
                 'sw_power_cap': pynvml.nvmlClocksThrottleReasonSwPowerCap}
        self._stop = threading.Event()

        def loop():
            while not self._stop.is_set():
                try:
                    self.samples.append(float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)))
                    mask = pynvml.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                    for n_, bit in names.items():
                        if mask & bit:
                            self.reasons.add(n_)
                except Exception:
                    pass
                self._stop.wait(0.002)
        self._thread = threading.Thread(target=loop, daemon=True)
        self._thread.start()

    def stop(self):
        if self._thread is not None:
            self._stop.set()
            self._thread.join(timeout=2)
        out = {'sm_mhz': None, 'sm_max_mhz': self.max_mhz, 'reasons': sorted(self.reasons),
               'samples': len(self.samples)}
        if self.samples:
            out['sm_mhz'] = float(np.median(self.samples))
        return out


# ------------------------------------------------------------------------------------------ CPU arm
def cpu_port_run(users, items, steps, warmup, batch, threads):
    """The reference's CPU training step, restated (oracle/collie_oracle.py): nn.Embedding(sparse)
    x4, forward for positives and for users.repeat(N) negatives, adaptive BPR, autograd backward,
    torch.optim.SparseAdam + SGD(bias); negatives by np.random.randint like the approximate loader
    (collie/interactions/datasets.py:439-443). Returns seconds for `steps` timed steps."""
    from oracle import collie_oracle as O
    w = WORKLOAD
    torch.set_num_threads(threads)
    model = O.OracleMF(w['num_users'], w['num_items'], w['dim'], sparse=True, seed=0)
    opts = O.make_optimizers(model, 'sparse_adam', 'sgd', w['lr'], w['bias_lr'])
    rng = np.random.RandomState(w['seed'])
    n = users.shape[0]
    t0 = None
    for s in range(warmup + steps):
        if s == warmup:
            t0 = time.perf_counter()
        a = (s * batch) % max(1, n - batch)
        bu, bp = users[a:a + batch], items[a:a + batch]
        negs = torch.from_numpy(rng.randint(0, w['num_items'], size=(batch, w['n_neg'])))
        O.train_step(model, opts, lambda: model.implicit_loss(bu, bp, negs, 'adaptive_bpr'))
    return time.perf_counter() - t0


def run_reference(args, rank, world):
    """Reference arm: the reference's CPU path (oracle port) on all host threads. Exactly K timed
    steps after W warm-up steps; each step is a bounded sample of the workload's batch, sized from
    one calibration step so that the whole run ends within a few minutes."""
    if rank != 0:
        return
    w = WORKLOAD
    threads = os.cpu_count() or 1
    full = w['batch_per_gpu']
    total_steps = args.steps + args.warmup
    calib = 32768
    cu, ci = synth_interactions(w['num_users'], w['num_items'], calib * 3, w['seed'])
    t_calib = cpu_port_run(cu, ci, 1, 1, calib, threads)          # seconds for one 32k step
    budget = 150.0
    batch = full
    while batch > 4096 and t_calib * (batch / calib) * total_steps > budget:
        batch //= 2
    users, items = synth_interactions(w['num_users'], w['num_items'], batch * (total_steps + 1),
                                      w['seed'])
    secs = cpu_port_run(users, items, args.steps, args.warmup, batch, threads)
    value = args.steps * batch / secs
    sample = (f'{args.steps} timed steps (+{args.warmup} warm-up) of {batch} interactions each '
              f'(workload batch is {full}); same tables, distributions and hyper-parameters; '
              'reference CPU path restated in torch (oracle port), all host threads')
    line = {
        'impl': 'reference', 'metric': 'MF train interactions/s (fwd+bwd+sparse Adam)',
        'value': value, 'unit': 'interactions/s', 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': 1e3 * secs / args.steps, 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': config_dict(args.gpus),
        'cpu_baseline': {'value': value, 'unit': 'interactions/s', 'cores': threads,
                         'kind': 'port', 'sample': sample},
        'e2e': {'value': value, 'unit': 'interactions/s', 'h2d_bytes_per_step': 0,
                'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


def config_dict(n_gpus):
    w = WORKLOAD
    return {'workload': w['name'], 'num_users': w['num_users'], 'num_items': w['num_items'],
            'interactions': w['interactions'], 'embedding_dim': w['dim'],
            'num_negative_samples': w['n_neg'], 'loss': 'bpr (adaptive_bpr, N=5)',
            'optimizer': 'sparse_adam + sgd bias', 'batch_per_gpu': w['batch_per_gpu'],
            'global_batch': w['batch_per_gpu'] * n_gpus,
            'parallelism': f'user-partitioned data parallel x{n_gpus}, item table row-sharded, '
                           'item-side exchange in one kernel over NVLink peer memory'
            if n_gpus > 1 else 'single GPU',
            'l2': 'tables + Adam state = 255 MB > 126 MB L2; every step reads a new batch',
            'timed_region': 'K training steps incl. the device-side epoch plan (shuffle + per-batch '
                            'user order) of the epoch they belong to'}


# ------------------------------------------------------------------------------------------ GPU arm
def run_ours(args, rank, world, local_rank):
    import torch.distributed as dist
    import collie_b200 as cb
    from collie_b200 import _abi
    w = WORKLOAD
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    B = w['batch_per_gpu']
    steps, warmup = args.steps, max(args.warmup, 3)

    if world > 1:
        from collie_b200 import distributed as cbd
        return cbd.bench_sharded(args, rank, world, local_rank, WORKLOAD, synth_interactions,
                                 config_dict, ClockSampler)

    users, items = synth_interactions(w['num_users'], w['num_items'], w['interactions'], w['seed'])
    data = cb.Interactions(users=users.numpy(), items=items.numpy(), num_users=w['num_users'],
                           num_items=w['num_items'], num_negative_samples=w['n_neg'],
                           allow_missing_ids=True, remove_duplicate_user_item_pairs=False,
                           check_num_negative_samples_is_valid=False,
                           max_number_of_samples_to_consider=0, seed=w['seed'])
    model = cb.MatrixFactorizationModel(train=data, embedding_dim=w['dim'], sparse=True,
                                        lr=w['lr'], bias_lr=w['bias_lr'], optimizer='sparse_adam',
                                        bias_optimizer='sgd', loss='bpr',
                                        lr_scheduler_func=None).to(dev)
    opts = model.configure_optimizers()
    n = data.num_interactions
    g = torch.Generator().manual_seed(w['seed'])
    perm = torch.randperm(n, generator=g)
    users_h = users[perm].contiguous().pin_memory()
    items_h = items[perm].contiguous().pin_memory()
    gidx_h = perm.contiguous().pin_memory()
    n_batches = n // B
    loader = cb.InteractionsDataLoader(data, batch_size=B, shuffle=True, drop_last=True)

    class Batches(object):
        """Batches the way the trainer gets them: the loader's device-side epoch plan (shuffle +
        per-batch user order), rebuilt whenever an epoch is exhausted -- inside the timed region."""
        def __init__(self):
            self.epoch, self.it = 0, None

        def next(self):
            while True:
                if self.it is None:
                    self.it = loader.device_batches(dev, epoch=self.epoch)
                    self.epoch += 1
                try:
                    return next(self.it), self.epoch
                except StopIteration:
                    self.it = None

    feed = Batches()

    def one_step():
        db, epoch = feed.next()
        return model.fused_step(db, optimizers=opts, interactions=data, epoch=epoch)

    # ---- device-resident throughput
    for i in range(warmup):
        one_step()
    feed.it = None          # the timed region pays for its own epoch plan
    torch.cuda.synchronize()
    launches0 = _abi.launch_count()
    clocks = ClockSampler(local_rank)
    clocks.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    ev0.record()
    for i in range(steps):
        loss = one_step()
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    clk = clocks.stop()
    launches = _abi.launch_count() - launches0
    value = steps * B / (ms * 1e-3)

    # ---- end to end: pinned host batches in, loss out, every step (double-buffered copies)
    # high priority: the small sort kernels of the next batch get SMs as soon as some free up,
    # instead of queueing behind the step's big grids
    copy_stream = torch.cuda.Stream(dev, priority=-1)
    main = torch.cuda.current_stream(dev)
    from collie_b200.interactions import BatchSorter
    slots = [dict(u=torch.empty(B, dtype=torch.int32, device=dev),
                  i=torch.empty(B, dtype=torch.int32, device=dev),
                  g=torch.empty(B, dtype=torch.int64, device=dev),
                  sort=BatchSorter(B, w['num_users'], dev),
                  ready=torch.cuda.Event(), free=torch.cuda.Event()) for _ in range(2)]
    loss_host = torch.empty(steps + warmup, dtype=torch.float32).pin_memory()

    def e2e_loop(count, base):
        for s_ in slots:
            s_['free'].record(main)
        for j in range(count):
            slot = slots[j % 2]
            a = ((base + j) % n_batches) * B
            with torch.cuda.stream(copy_stream):
                # copy stream: the batch comes in and is ordered by user while the previous
                # step is still running on the main stream
                copy_stream.wait_event(slot['free'])
                slot['u'].copy_(users_h[a:a + B], non_blocking=True)
                slot['i'].copy_(items_h[a:a + B], non_blocking=True)
                slot['g'].copy_(gidx_h[a:a + B], non_blocking=True)
                db = slot['sort'](slot['u'], slot['i'], slot['g'])
                slot['ready'].record(copy_stream)
            main.wait_event(slot['ready'])
            out = model.fused_step(db, optimizers=opts, interactions=data, epoch=1)
            loss_host[j:j + 1].copy_(out.view(1), non_blocking=True)
            slot['free'].record(main)

    e2e_loop(warmup, 0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    e2e_loop(steps, warmup)
    e1.record()
    torch.cuda.synchronize()
    e2e_ms = e0.elapsed_time(e1)
    e2e_value = steps * B / (e2e_ms * 1e-3)
    h2d = B * (4 + 4 + 8)

    # ---- roofline of the dominant kernel (k_fused_step), timed by events inside the library
    lib = _abi.lib()
    lib.cmf_profile_enable(1)
    prof_steps = min(steps, 20)
    outs = {'negs': None}
    last = None
    for i in range(prof_steps):          # the same batches as the timed loop: the loader's plan
        last, ep_ = feed.next()
        model.fused_step(last, optimizers=opts, interactions=data, epoch=ep_,
                         outputs=outs if i == prof_steps - 1 else None)
    C = _abi.C
    ms4 = (C.c_double * 4)()
    split = C.c_int32()
    n_steps, item_rows, user_rows = C.c_int64(), C.c_int64(), C.c_int64()
    lib.cmf_profile_read(ms4, C.byref(split), C.byref(n_steps), C.byref(item_rows),
                         C.byref(user_rows))
    lib.cmf_profile_enable(0)
    uq = int(torch.unique(last.users).numel())
    iq = int(torch.unique(torch.cat([outs['negs'].flatten(), last.items.long()])).numel())
    D, N, T = w['dim'], w['n_neg'], 4
    ns = max(1, n_steps.value)
    row = D * T
    # algorithmic bytes: gather B*(2+N) rows + biases + ids; update 6 streams per
    # unique row + 2 bias streams
    gather_bytes = B * (2 + N) * row + B * (2 + N) * T + B * 2 * 4
    update_row = 6 * row + 2 * T
    step_bytes = gather_bytes + (uq + iq) * update_row
    inplace_users = uq - int(user_rows.value)
    if split.value:
        phases = {'k_fwd_loss': ms4[0] / ns, 'sort_tail': ms4[1] / ns,
                  'k_bwd_update': ms4[2] / ns, 'k_update_rows': ms4[3] / ns}
        kernel, kernel_ms, kernel_bytes = 'k_fwd_loss', ms4[0] / ns, gather_bytes
    else:
        phases = {'sort': ms4[0] / ns, 'k_fused_step': ms4[1] / ns, 'k_update_rows': ms4[3] / ns}
        kernel, kernel_ms = 'k_fused_step', ms4[1] / ns
        kernel_bytes = gather_bytes + inplace_users * update_row
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
    except Exception:
        pass
    peak = float(peaks.get('hbm_gbs', 6650.0))
    achieved = kernel_bytes / (kernel_ms * 1e-3) / 1e9
    roofline = {
        'bound': 'hbm', 'kernel': kernel, 'achieved': achieved, 'peak': peak,
        'unit': 'GB/s', 'frac': achieved / peak,
        'peak_source': 'MEASURED_PEAKS.json hbm_gbs (of measured)' if peaks else
                       'fallback 6650 GB/s (of fallback)',
        'traffic': None,
        'algorithmic_bytes_per_launch': kernel_bytes, 'kernel_ms': kernel_ms,
        'phase_ms': phases,
        'step_algorithmic_bytes': step_bytes,
        'step_achieved_gbs': step_bytes / (ms / steps * 1e-3) / 1e9,
        'step_frac': step_bytes / (ms / steps * 1e-3) / 1e9 / peak,
        'unique_users': uq, 'unique_items': iq, 'rows_finished_by_update_kernel':
            {'items': int(item_rows.value), 'users': int(user_rows.value)},
        'update_phase_algorithmic_bytes': (uq + iq) * update_row,
    }
    # the forward kernel's rows come out of L2 (item table 13.7 MB, the step's user rows 46 MB),
    # so next to the HBM roofline: live probes of read bandwidth (cmf_read_probe, a persistent
    # grid of 16-byte loads) over an L2-resident buffer and over one far larger than L2
    def read_probe(nbytes, repeats):
        buf = torch.ones(nbytes // 4, dtype=torch.float32, device=dev)
        sink = torch.zeros(1, dtype=torch.float32, device=dev)
        call = lambda: _abi.check(lib.cmf_read_probe(_abi.ptr(buf), nbytes, repeats, _abi.ptr(sink),
                                                     _abi.stream_ptr(dev)), 'cmf_read_probe')
        call()
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record()
        call()
        p1.record()
        torch.cuda.synchronize()
        return nbytes * repeats / (p0.elapsed_time(p1) * 1e-3) / 1e9
    l2_gbs = read_probe(48 << 20, 200)
    hbm_read_gbs = read_probe(4 << 30, 2)
    # ... and of the kernel's own access pattern without its arithmetic: the same number of random
    # 512-byte rows of the item table, 16 loads in flight per lane (cmf_gather_probe)
    gather_gbs = None
    if D == 128:
        n_rows = B * (2 + N)
        idx = torch.randint(0, w['num_items'], (n_rows,), dtype=torch.int32, device=dev)
        sink = torch.zeros(1, dtype=torch.float32, device=dev)
        tab = model.item_embeddings.weight.data
        gp = lambda: _abi.check(lib.cmf_gather_probe(_abi.ptr(tab), _abi.ptr(idx), n_rows, 4, 1, 4,
                                                     _abi.ptr(sink), _abi.stream_ptr(dev)),
                                'cmf_gather_probe')
        gp()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        for _ in range(5):
            gp()
        g1.record()
        torch.cuda.synchronize()
        gather_gbs = 5 * n_rows * 512 / (g0.elapsed_time(g1) * 1e-3) / 1e9
    roofline['read_probe_gbs'] = {'l2_resident_48MB': l2_gbs, 'hbm_4GB': hbm_read_gbs,
                                  'gather_512B_rows_l2': gather_gbs}
    roofline['kernel_frac_of_l2_read_probe'] = achieved / l2_gbs
    if gather_gbs:
        roofline['kernel_frac_of_gather_probe'] = achieved / gather_gbs
    traffic_file = os.path.join(ROOT, 'profiles', 'traffic.json')
    if os.path.exists(traffic_file):
        try:
            roofline['traffic'] = json.load(open(traffic_file)).get(kernel)
        except Exception:
            pass

    # ---- secondary headline: evaluation users/s (MAP@10 through the tensor-core path; MRR + AUC
    # through the exact path), on a bounded user sample, timed on the device
    eval_info = None
    if not args.no_eval:
        n_eval_users = 32768
        tu = torch.arange(n_eval_users, dtype=torch.int32)
        ti = torch.from_numpy(np.random.default_rng(5).integers(0, w['num_items'],
                                                                 (n_eval_users, 8))).flatten()
        test = cb.Interactions(users=tu.repeat_interleave(8).numpy(), items=ti.numpy(),
                               num_users=w['num_users'], num_items=w['num_items'],
                               allow_missing_ids=True, check_num_negative_samples_is_valid=False,
                               remove_duplicate_user_item_pairs=False,
                               max_number_of_samples_to_consider=0)
        from collie_b200 import metrics as M
        tables = M._model_tables(model)
        test_users = np.unique(test.mat.row)
        indptr, tgt_items, nnz = M._targets_device(test.mat.tocsr(), test_users, dev)
        users_dev = torch.from_numpy(test_users.astype(np.int32)).to(dev)

        def tc_pass(ws=None):
            return M._eval_tile(dev, n_eval_users, w['num_items'], 10, indptr, tgt_items,
                                {'mapk', 'mrr'}, tables=tables, user_ids=users_dev, precision=1,
                                workspace=ws)
        ws = tc_pass()['_workspace']
        torch.cuda.synchronize()
        t0e, t1e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 5
        t0e.record()
        for _ in range(reps):
            res = tc_pass(ws)
        t1e.record()
        torch.cuda.synchronize()
        tc_ms = t0e.elapsed_time(t1e) / reps
        map10 = M._per_user('mapk', res, nnz, w['num_items']).mean().item()
        t0 = time.perf_counter()
        m_all = cb.evaluate_in_batches([cb.mapk, cb.mrr, cb.auc], test, model, k=10, verbose=False)
        torch.cuda.synchronize()
        api_exact_first_s = time.perf_counter() - t0
        t0 = time.perf_counter()
        m_all = cb.evaluate_in_batches([cb.mapk, cb.mrr, cb.auc], test, model, k=10, verbose=False)
        torch.cuda.synchronize()
        api_exact_s = time.perf_counter() - t0
        # wall clock of the public call, results on the host: the first call on a test set (derives
        # its device-side form) and a later one (every following epoch's validation)
        t0 = time.perf_counter()
        m_tc = cb.evaluate_in_batches([cb.mapk, cb.mrr], test, model, k=10, verbose=False,
                                      precision=1)
        torch.cuda.synchronize()
        api_tc_first_s = time.perf_counter() - t0
        t0 = time.perf_counter()
        for _ in range(3):
            m_tc = cb.evaluate_in_batches([cb.mapk, cb.mrr], test, model, k=10, verbose=False,
                                          precision=1)
        torch.cuda.synchronize()
        api_tc_s = (time.perf_counter() - t0) / 3
        flops = 2.0 * n_eval_users * w['num_items'] * w['dim']
        kp = ((3 * w['dim'] + 4 + 15) // 16) * 16       # K = 16 steps that are issued
        issued = 2.0 * n_eval_users * w['num_items'] * kp
        tpeak = float(peaks.get('bf16_tflops', 1590.0))
        eval_info = {
            'metric': 'eval users/s MAP@10', 'value': n_eval_users / (tc_ms * 1e-3),
            'unit': 'users/s', 'users': n_eval_users,
            'path': 'cmf_eval_tile precision=1, MAP@10 + MRR: tcgen05/TMA GEMM over bf16-split '
                    'operands (hi.hi + hi.lo + lo.hi), top-k + MRR counter in the epilogue; device '
                    'time of the whole call (operand split, exact target scores, GEMM + top-k, '
                    'merge)',
            'ms': tc_ms, 'map10': map10,
            'roofline': {'bound': 'tensor', 'achieved': issued / (tc_ms * 1e-3) / 1e12,
                         'useful_tflops': flops / (tc_ms * 1e-3) / 1e12, 'peak': tpeak,
                         'unit': 'TFLOP/s', 'frac': issued / (tc_ms * 1e-3) / 1e12 / tpeak,
                         'note': 'achieved counts the issued bf16 MMA flops (3 split terms + bias '
                                 'slots); useful = 2 * users * items * D'},
            'api_users_per_s': {'evaluate_in_batches([mapk, mrr], precision=1)':
                                n_eval_users / api_tc_s,
                                'evaluate_in_batches([mapk, mrr], precision=1), first call':
                                n_eval_users / api_tc_first_s,
                                'evaluate_in_batches([mapk, mrr, auc]) exact fp32':
                                n_eval_users / api_exact_s,
                                'evaluate_in_batches([mapk, mrr, auc]) exact fp32, first call':
                                n_eval_users / api_exact_first_s},
            'metrics_exact': m_all, 'metrics_tc': m_tc,
        }

    # ---- CPU baseline on a bounded sample (rank 0, N = 1 only)
    cpu = None
    if not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        cpu_steps, cpu_warm = 3, 1
        nb = B * (cpu_steps + cpu_warm + 1)
        secs = cpu_port_run(users[:nb], items[:nb], cpu_steps, cpu_warm, B, threads)
        cpu = {'value': cpu_steps * B / secs, 'unit': 'interactions/s', 'cores': threads,
               'kind': 'port',
               'sample': f'{cpu_steps} steps of batch {B} (+{cpu_warm} warm-up), same tables and '
                         'distributions; reference CPU path restated in torch (oracle port)'}

    line = {
        'metric': 'MF train interactions/s (fwd+bwd+sparse Adam)', 'value': value,
        'unit': 'interactions/s', 'n_gpus': 1, 'steps': steps, 'warmup': warmup,
        'ms_per_step': ms / steps, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic', 'config': config_dict(1),
        'roofline': roofline, 'cpu_baseline': cpu,
        'e2e': {'value': e2e_value, 'unit': 'interactions/s', 'h2d_bytes_per_step': h2d,
                'd2h_bytes_per_step': 4, 'ms_per_step': e2e_ms / steps},
        'gpu_launches': int(launches),
        'gpu_launches_note': 'own kernels per step: k_iota + k_fwd_loss + k_bwd_update + k_update_rows; '
                             'the CUB radix sort adds library launches that are not counted',
        'clocks': clk, 'final_loss': float(loss), 'eval': eval_info,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=50)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-eval', action='store_true')
    args = ap.parse_args()
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    # stdout carries the JSON line and nothing else: libraries that print from C (NCCL's version
    # banner) are sent to stderr by pointing fd 1 there and keeping the real stdout for the line
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), 'w')
    os.dup2(2, 1)
    sys.stdout = real_stdout
    if args.impl == 'reference':
        run_reference(args, rank, world)
        return
    if not torch.cuda.is_available():
        raise SystemExit('bench.py needs a CUDA device: collie_b200 has no CPU path')
    run_ours(args, rank, world, local_rank)


if __name__ == '__main__':
    main()
