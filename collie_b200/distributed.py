"""Multi-GPU matrix-factorization training: one process per GPU over ``torch.distributed``.

The reference has no multi-GPU code of its own; this is the B200 design of SURVEY 8(e):

* **users are partitioned** (``owner(u) = u % world``, local id ``u // world``): a rank trains on the
  interactions of its own users, so user rows, their Adam moments and biases are always local and
  are finished in place by the fused step -- no user-side communication at all;
* **the item table is row-sharded** (contiguous blocks of ``ceil(I / world)`` rows, padded): every
  rank owns one block of rows with its optimizer state and keeps a full-width working copy for
  its gathers. One step =
    1. the local fused step on the rank's batch (``skip_item_update``: item-row gradients, bias
       gradients and "touched" flags stay in the step's workspace; the loss and the gradients are
       scaled for the GLOBAL batch of ``world * B`` interactions),
    2. a fence (all ranks' accumulators are complete),
    3. ``cmf_item_exchange_update``: ONE kernel over peer memory in which the owner of a row sums
       the ranks' gradient rows out of their workspaces, applies the rule to its shard (SparseAdam
       moves every row some rank gathered), stores the new row into every rank's copy and clears
       what it consumed -- plain peer loads / stores for two ranks, NVSwitch multicast
       (``multimem.ld_reduce`` / ``multimem.st``) for more,
    4. a fence (all copies are complete before anybody's next forward).
  Tables and workspaces live in symmetric memory (``torch.distributed._symmetric_memory``); the
  fence is its signal-pad barrier. ``exchange='nccl'`` (and any backend without peer memory, e.g.
  gloo in the CPU tests) uses collectives instead: ``all_gather`` of the item rows, the local step,
  ``reduce_scatter`` of the three accumulators, the owner's ``cmf_optimizer_rows_masked``.
  With MovieLens-like catalogues every step touches nearly every item row, so a dense exchange is
  the right one (13.7 MB per direction per step at the 20M shape); a catalogue whose steps touch a
  small fraction of the rows would exchange de-duplicated row lists -- same ownership, same
  owner-side update.

Global semantics: one batch of ``world * B`` interactions per step, one optimizer step per unique
row. Negatives are keyed by the interaction's global dataset position, so they do not depend on the
number of GPUs. The collectives take any backend: NCCL on GPUs, gloo in the CPU tests (which
exercise the plumbing, not the kernels).
"""
import numpy as np
import torch
import torch.distributed as dist

from . import _abi


def owner_of_user(users, world):
    return users % world


def local_user_id(users, world):
    return users // world


def shard_rows(num_items, world):
    """Rows per rank and padded row count of the row-sharded item table."""
    per_rank = (int(num_items) + world - 1) // world
    return per_rank, per_rank * world


def partition_interactions(users, items, rank, world, extra=None):
    """Interactions of the users this rank owns, with local user ids and their global dataset
    positions (``gidx``, the RNG key)."""
    users = np.asarray(users)
    mine = np.nonzero(owner_of_user(users, world) == rank)[0]
    out = [local_user_id(users[mine], world), np.asarray(items)[mine], mine.astype(np.int64)]
    if extra is not None:
        out.append(np.asarray(extra)[mine])
    return tuple(out)


def num_local_users(num_users, rank, world):
    return (int(num_users) - rank + world - 1) // world


def all_gather_rows(full, shard, group=None):
    """``full[r * rows : (r + 1) * rows] = shard of rank r`` for every r (in place when ``shard`` is
    this rank's slice of ``full``)."""
    dist.all_gather_into_tensor(full, shard, group=group)
    return full


def reduce_scatter_sum(out_shard, full, group=None):
    """Owner r receives the sum over ranks of ``full[r * rows : (r + 1) * rows]``."""
    try:
        dist.reduce_scatter_tensor(out_shard, full, op=dist.ReduceOp.SUM, group=group)
    except (RuntimeError, NotImplementedError, ValueError):
        # backends without reduce_scatter (gloo): all_reduce + slice, same result
        tmp = full.clone()
        dist.all_reduce(tmp, op=dist.ReduceOp.SUM, group=group)
        rows = out_shard.shape[0]
        r = dist.get_rank(group)
        out_shard.copy_(tmp[r * rows:(r + 1) * rows])
    return out_shard


class ShardedMatrixFactorization(object):
    """A ``MatrixFactorizationModel`` trained across ``world`` GPUs (see the module docstring).

    ``model`` is this rank's local model: its user tables hold the rank's own users
    (``num_local_users``), its item tables are the full-width working copy (padded to
    ``shard_rows``); ``model.hparams.num_items`` is the padded count.
    """

    def __init__(self, model, true_num_items, group=None, exchange='auto', max_batch=None,
                 max_negatives=None, multicast='auto'):
        """``exchange``: ``'p2p'`` = the owner-side kernel over peer memory
        (``cmf_item_exchange_update``; needs NCCL ranks of one node and symmetric memory),
        ``'nccl'`` = all-gather / reduce-scatter collectives, ``'auto'`` = p2p when it can be set
        up. ``max_batch`` / ``max_negatives`` size the symmetric training workspace of the p2p
        mode (default: the first step's batch, over all ranks, plus 25 %). ``multicast``: whether
        the p2p kernel reduces and broadcasts through NVSwitch multicast addresses
        (``multimem.ld_reduce`` / ``multimem.st``): ``'auto'`` = for more than two ranks when the
        allocator provides them (env ``CMF_MULTICAST=0`` turns it off), ``True`` = required (raises
        when unavailable), ``False`` = plain peer loads / stores."""
        self.model = model
        self.group = group
        self.exchange = exchange
        self._max_batch, self._max_negatives = max_batch, max_negatives
        self._p2p = None
        import os
        # how the ranks rendezvous around the exchange kernel: 'kernel' = barriers inside the
        # kernel itself (signal pad in symmetric memory), on a side stream next to the step's
        # remaining user-row updates; 'symm' / 'nccl' = a fence before and after it on the main
        # stream (symmetric-memory barrier kernel / one-element all-reduce)
        self._fence_kind = os.environ.get('CMF_FENCE', 'kernel')
        if self._fence_kind not in ('kernel', 'symm', 'nccl'):
            raise ValueError("CMF_FENCE must be 'kernel', 'symm' or 'nccl'")
        if multicast not in ('auto', True, False):
            raise ValueError("multicast must be 'auto', True or False")
        if multicast == 'auto' and os.environ.get('CMF_MULTICAST') in ('0', '1'):
            multicast = os.environ['CMF_MULTICAST'] == '1'      # experiments / A-B runs
        self._multicast = multicast
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.per_rank, padded = shard_rows(true_num_items, self.world)
        if model.hparams.num_items != padded:
            raise ValueError(f'the local model must hold the padded item table ({padded} rows), '
                             f'not {model.hparams.num_items}')
        self.true_num_items = int(true_num_items)
        lo, hi = self.rank * self.per_rank, (self.rank + 1) * self.per_rank
        self.I_full = model.item_embeddings.weight.data
        self.bi_full = model.item_biases.weight.data
        self.I_shard = self.I_full[lo:hi]
        self.bi_shard = self.bi_full[lo:hi]
        self._state = None
        self._step = 0
        # set to a list to have every p2p step append four CUDA events (after the local step,
        # after the first fence, after the exchange kernel, after the second fence)
        self.timing = None
        # every rank starts from rank 0's item table (user tables are disjoint)
        dist.broadcast(self.I_full, src=self._global_rank(0), group=group)
        dist.broadcast(self.bi_full, src=self._global_rank(0), group=group)
        if exchange not in ('auto', 'p2p', 'nccl'):
            raise ValueError("exchange must be 'auto', 'p2p' or 'nccl'")
        if exchange != 'nccl':
            self._setup_tables_p2p()

    # ---- peer-memory exchange ---------------------------------------------------------------
    def _symm(self):
        import torch.distributed._symmetric_memory as symm
        return symm

    def _setup_tables_p2p(self):
        """Move the item table and biases into symmetric memory (every rank maps every rank's
        copy). Falls back to the collectives in 'auto' mode when that is not possible."""
        try:
            if dist.get_backend(self.group) != 'nccl' or not self.I_full.is_cuda:
                raise RuntimeError('the p2p exchange needs NCCL ranks on CUDA devices')
            if self.world > _abi.CMF_MAX_PEERS:
                raise RuntimeError(f'the p2p exchange takes at most {_abi.CMF_MAX_PEERS} ranks')
            symm = self._symm()
            grp = self.group if self.group is not None else dist.group.WORLD
            m = self.model
            I_sym = symm.empty(tuple(self.I_full.shape), dtype=torch.float32,
                               device=self.I_full.device)
            bi_sym = symm.empty(tuple(self.bi_full.shape), dtype=torch.float32,
                                device=self.I_full.device)
            hI, hb = symm.rendezvous(I_sym, grp), symm.rendezvous(bi_sym, grp)
            I_sym.copy_(self.I_full)
            bi_sym.copy_(self.bi_full)
            m.item_embeddings.weight.data = I_sym
            m.item_biases.weight.data = bi_sym
            self.I_full, self.bi_full = I_sym, bi_sym
            lo, hi = self.rank * self.per_rank, (self.rank + 1) * self.per_rank
            self.I_shard, self.bi_shard = I_sym[lo:hi], bi_sym[lo:hi]
            self._p2p = {'I': [int(p) for p in hI.buffer_ptrs],
                         'bi': [int(p) for p in hb.buffer_ptrs], 'handles': [hI, hb]}
            self._tick = torch.zeros(1, dtype=torch.float32, device=self.I_full.device)
            self._p2p['multicast'] = False
            # signal pad of the in-kernel barriers, the stream the exchange runs on and the events
            # that tie it to the main stream
            pad = symm.empty(int(_abi.lib().cmf_exchange_pad_bytes()) // 4, dtype=torch.int32,
                             device=self.I_full.device)
            pad.zero_()
            hp = symm.rendezvous(pad, grp)
            self._p2p.update(pad=pad, pad_ptrs=[int(p) for p in hp.buffer_ptrs])
            self._p2p['handles'].append(hp)
            dev = self.I_full.device
            # (default priority: the step's own kernels, launched first, are placed first; the
            # exchange kernel's CTAs mostly wait at its barriers)
            self._xstream = torch.cuda.Stream(dev)
            self._ev_grads, self._ev_done = torch.cuda.Event(), torch.cuda.Event()
            self._ev_grads.record(torch.cuda.current_stream(dev))
            self._ev_done.record(torch.cuda.current_stream(dev))
            torch.cuda.synchronize(dev)
            hp.barrier(channel=0)        # every rank's pad is zero before anybody signals
        except Exception as e:
            if self.exchange == 'p2p':
                raise _abi.CmfError(f'p2p exchange unavailable: {e}')
            self._p2p = None
        if not self._all_ranks(self._p2p is not None):
            self._p2p = None       # one mode for everybody: some rank could not set it up

    def _all_ranks(self, ok):
        """True when ``ok`` holds on every rank (so that the ranks take the same exchange path)."""
        t = torch.tensor([1 if ok else 0], dtype=torch.int32, device=self.I_full.device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=self.group)
        return bool(t.item())

    def _setup_workspace_p2p(self, batch, n_neg):
        """The training workspace in symmetric memory (collective: every rank calls it at its
        first step); records every rank's accumulator addresses."""
        m, h = self.model, self.model.hparams
        symm = self._symm()
        grp = self.group if self.group is not None else dist.group.WORLD
        dev = self.I_full.device
        want = torch.tensor([int(self._max_batch or int(batch * 1.25) + 1),
                             int(self._max_negatives or n_neg)], dtype=torch.int64, device=dev)
        dist.all_reduce(want, op=dist.ReduceOp.MAX, group=self.group)
        cap_b, cap_n = int(want[0].item()), int(want[1].item())
        need = torch.tensor([int(_abi.lib().cmf_train_workspace_bytes(
            cap_b, cap_n, h.num_users, h.num_items, h.embedding_dim))], dtype=torch.int64,
            device=dev)
        dist.all_reduce(need, op=dist.ReduceOp.MAX, group=self.group)
        ws = symm.empty(int(need.item()), dtype=torch.uint8, device=dev)
        hw = symm.rendezvous(ws, grp)
        _abi.check(_abi.lib().cmf_workspace_init(_abi.ptr(ws), ws.numel(), _abi.stream_ptr(dev)),
                   'cmf_workspace_init')
        m._workspace, m._workspace_B = ws, cap_b
        C = _abi.C
        GI, gbi, flag = C.c_void_p(), C.c_void_p(), C.c_void_p()
        _abi.check(_abi.lib().cmf_train_workspace_items(
            _abi.ptr(ws), h.num_users, h.num_items, h.embedding_dim, C.byref(GI), C.byref(gbi),
            C.byref(flag)), 'cmf_train_workspace_items')
        base = ws.data_ptr()
        # the layout depends on the local user count, which differs between ranks by one
        offs = torch.tensor([GI.value - base, gbi.value - base, flag.value - base],
                            dtype=torch.int64, device=dev)
        all_offs = [torch.zeros_like(offs) for _ in range(self.world)]
        dist.all_gather(all_offs, offs, group=self.group)
        bases = [int(p) for p in hw.buffer_ptrs]
        self._p2p.update(
            GI=[bases[q] + int(all_offs[q][0]) for q in range(self.world)],
            gbi=[bases[q] + int(all_offs[q][1]) for q in range(self.world)],
            flag=[bases[q] + int(all_offs[q][2]) for q in range(self.world)],
            cap=(cap_b, cap_n), ws=ws)
        self._p2p['handles'].append(hw)
        peers = _abi.cmf_peers_t()
        peers.rank, peers.world, peers.D = self.rank, self.world, int(h.embedding_dim)
        peers.rows_per_rank = self.per_rank
        for q in range(self.world):
            peers.I[q], peers.bi[q] = self._p2p['I'][q], self._p2p['bi'][q]
            peers.GI[q], peers.gbi[q] = self._p2p['GI'][q], self._p2p['gbi'][q]
            peers.flag[q] = self._p2p['flag'][q]
        # NVSwitch multicast views (NVLS), when the allocator made them and the accumulators sit
        # at the same offset on every rank
        hI, hb = self._p2p['handles'][0], self._p2p['handles'][1]
        mc = [int(getattr(h_, 'multicast_ptr', 0) or 0) for h_ in (hI, hb, hw)]
        same = all(bool(torch.equal(all_offs[q], offs)) for q in range(self.world))
        if self._multicast is True and not (all(mc) and same):
            raise RuntimeError('NVSwitch multicast addresses are not available for the symmetric '
                               'buffers (or the accumulators sit at different offsets)')
        # (measured: 2 ranks are ~2 % faster with plain peer loads / stores, 4 ranks ~2 % faster
        # with multicast)
        if (self._multicast is True or (self._multicast == 'auto' and self.world > 2)) \
                and all(mc) and same:
            peers.mc_I, peers.mc_bi = mc[0], mc[1]
            peers.mc_GI = mc[2] + int(offs[0])
            peers.mc_gbi = mc[2] + int(offs[1])
            peers.mc_flag = mc[2] + int(offs[2])
            self._p2p['multicast'] = True
        if self._fence_kind == 'kernel':
            for q in range(self.world):
                peers.pad[q] = self._p2p['pad_ptrs'][q]
            # this rank's own accumulators: item_flag, GI, gbi are the first three areas of the
            # workspace (cmf_train_workspace_items), cleared as one span
            lo = min(int(offs[0]), int(offs[1]), int(offs[2]))
            hi = int(offs[1]) + 4 * int(h.num_items)
            hi += (-(hi - lo)) % 16      # (the areas are 256-byte aligned: the slack is unused)
            peers.clear_ptr = base + lo
            peers.clear_bytes = hi - lo
        self._p2p['peers'] = peers

    def _fence(self):
        """Stream-ordered rendezvous of all ranks: nobody's later work starts before everybody's
        earlier work is done (a one-element all-reduce: each rank's result needs each rank's
        input)."""
        if self._fence_kind == 'symm':
            # the symmetric-memory signal-pad barrier (one small kernel, no NCCL launch)
            self._p2p['handles'][0].barrier(channel=0, timeout_ms=20000)
        else:
            dist.all_reduce(self._tick, group=self.group)

    def _train_step_p2p(self, batch, optimizers, main, bias, st, interactions, epoch, global_batch,
                        reduce_loss):
        m = self.model
        B = int(batch.users.shape[0])
        n_neg = int(getattr(interactions, 'num_negative_samples', 0) or 0)
        if 'peers' not in self._p2p:
            err = None
            try:
                self._setup_workspace_p2p(B, n_neg)
            except Exception as e:          # (checked on all ranks together below)
                err = e
            if not self._all_ranks(err is None):
                if self.exchange == 'p2p':
                    raise _abi.CmfError(f'p2p exchange unavailable: {err}')
                self._p2p = None            # collectives from here on, on every rank
                m._workspace = None
                return None
        cap_b, cap_n = self._p2p['cap']
        if B > cap_b or n_neg > cap_n or m._workspace is not self._p2p['ws']:
            raise _abi.CmfError(
                f'batch of {B} x {n_neg} negatives exceeds the symmetric training workspace '
                f'({cap_b} x {cap_n}); pass max_batch / max_negatives to '
                'ShardedMatrixFactorization')
        in_kernel = self._fence_kind == 'kernel'
        loss = m.fused_step(batch, optimizers=optimizers, interactions=interactions, epoch=epoch,
                            global_batch=int(global_batch), skip_item_update=True,
                            num_items_logical=self.true_num_items,
                            item_grads_event=self._ev_grads if in_kernel else None)
        if m._workspace is not self._p2p['ws']:
            raise _abi.CmfError('the model replaced its symmetric training workspace')
        self._step += 1
        g_main, g_bias = main.param_groups[0], bias.param_groups[0]
        o = _abi.make_opt(main.rule, g_main['lr'], self._step, g_main.get('betas', (0.9, 0.999)),
                          g_main.get('eps', 1e-8), g_main.get('weight_decay', 0.0))
        ob = _abi.make_opt(bias.rule, g_bias['lr'], self._step, g_bias.get('betas', (0.9, 0.999)),
                           g_bias.get('eps', 1e-8), g_bias.get('weight_decay', 0.0))
        peers = self._p2p['peers']
        dev = self.I_full.device
        ev = None
        if self.timing is not None:
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]

        def launch(stream_ptr):
            _abi.check(_abi.lib().cmf_item_exchange_update(
                _abi.C.byref(peers), _abi.ptr(st['m']), _abi.ptr(st['v']), _abi.ptr(st['bm']),
                _abi.ptr(st['bv']), o, ob, stream_ptr), 'cmf_item_exchange_update')

        if in_kernel:
            # ONE kernel, entry and exit barriers inside, on its own stream: it starts when this
            # rank's item-side accumulators are complete and runs next to the step's remaining
            # user-row updates; the next step's forward waits for it
            peers.epoch = self._step & 0xFFFFFFFF or 1
            main_stream = torch.cuda.current_stream(dev)
            with torch.cuda.stream(self._xstream):
                self._xstream.wait_event(self._ev_grads)
                if ev:
                    ev[0].record()
                    ev[1].record()
                launch(self._xstream.cuda_stream)
                if ev:
                    ev[2].record()
                    ev[3].record()
                self._ev_done.record(self._xstream)
            if reduce_loss:
                dist.all_reduce(loss, group=self.group)
            main_stream.wait_event(self._ev_done)
        else:
            if ev:
                ev[0].record()
            # every rank's accumulators are complete ...
            if reduce_loss:
                dist.all_reduce(loss, group=self.group)     # (doubles as the rendezvous)
            else:
                self._fence()
            if ev:
                ev[1].record()
            launch(_abi.stream_ptr(dev))
            if ev:
                ev[2].record()
            # ... and every rank's table copy is complete before anybody's next step reads it
            self._fence()
            if ev:
                ev[3].record()
        if ev:
            self.timing.append(ev)
        return loss

    def _global_rank(self, r):
        return dist.get_global_rank(self.group, r) if self.group is not None else r

    def _item_state(self, main, bias):
        if self._state is None:
            dev, D = self.I_full.device, self.I_full.shape[1]
            z = lambda *shape: torch.zeros(shape, dtype=torch.float32, device=dev)
            self._state = dict(
                m=z(self.per_rank, D) if main.rule != 'sgd' else None,
                v=z(self.per_rank, D) if main.rule in ('adam', 'sparse_adam') else None,
                bm=z(self.per_rank) if bias.rule != 'sgd' else None,
                bv=z(self.per_rank) if bias.rule in ('adam', 'sparse_adam') else None,
                g=z(self.per_rank, D), gb=z(self.per_rank),
                flag=torch.zeros(self.per_rank, dtype=torch.int32, device=dev))
        return self._state

    def _workspace_views(self):
        m, h = self.model, self.model.hparams
        cached = getattr(self, '_views', None)
        if cached is not None and cached[0] == m._workspace.data_ptr():
            return cached[1]
        C = _abi.C
        GI, gbi, flag = C.c_void_p(), C.c_void_p(), C.c_void_p()
        _abi.check(_abi.lib().cmf_train_workspace_items(
            _abi.ptr(m._workspace), h.num_users, h.num_items, h.embedding_dim, C.byref(GI),
            C.byref(gbi), C.byref(flag)), 'cmf_train_workspace_items')
        base = m._workspace.data_ptr()
        n, D = h.num_items, h.embedding_dim

        def view(p, dtype, count, shape):
            off = p.value - base
            return m._workspace[off:off + count * 4].view(dtype).view(shape)
        views = (view(GI, torch.float32, n * D, (n, D)), view(gbi, torch.float32, n, (n,)),
                 view(flag, torch.int32, n, (n,)))
        self._views = (base, views)
        return views

    def train_step(self, batch, optimizers, interactions=None, epoch=0, global_batch=None,
                   reduce_loss=True):
        """One global step; returns the global mean loss (0-d tensor, identical on every rank).
        Collective: every rank must call it the same number of times (``common_steps_per_epoch``).
        ``global_batch``: total interactions of this step over all ranks (asked of the ranks when
        ``None``, which costs a host round trip; equal-sized batches pass ``world * B``)."""
        m = self.model
        if isinstance(optimizers, tuple):
            optimizers = optimizers[0]
        if not isinstance(optimizers, (list, tuple)):
            optimizers = [optimizers]
        main, bias = optimizers[0], optimizers[1] if len(optimizers) > 1 else optimizers[0]
        st = self._item_state(main, bias)
        if global_batch is None:
            t = torch.tensor([int(batch.users.shape[0])], dtype=torch.int64,
                             device=self.I_full.device)
            dist.all_reduce(t, group=self.group)
            global_batch = int(t.item())
        if self._p2p is not None:
            done = self._train_step_p2p(batch, optimizers, main, bias, st, interactions, epoch,
                                        global_batch, reduce_loss)
            if done is not None:
                return done
        # 1. everyone gets the current item rows
        all_gather_rows(self.I_full, self.I_shard, self.group)
        all_gather_rows(self.bi_full.view(-1), self.bi_shard.reshape(-1), self.group)
        # 2. local fused step, item side left in the workspace
        loss = m.fused_step(batch, optimizers=optimizers, interactions=interactions, epoch=epoch,
                            global_batch=int(global_batch), skip_item_update=True,
                            num_items_logical=self.true_num_items)
        GI, gbi, flag = self._workspace_views()
        # 3. gradients and touched flags to the owners
        reduce_scatter_sum(st['g'], GI, self.group)
        reduce_scatter_sum(st['gb'], gbi, self.group)
        reduce_scatter_sum(st['flag'], flag, self.group)
        GI.zero_()
        gbi.zero_()
        flag.zero_()
        # 4. owner-side masked update of this rank's rows (written straight into the working copy)
        self._step += 1
        g_main, g_bias = main.param_groups[0], bias.param_groups[0]
        o = _abi.make_opt(main.rule, g_main['lr'], self._step, g_main.get('betas', (0.9, 0.999)),
                          g_main.get('eps', 1e-8), g_main.get('weight_decay', 0.0))
        ob = _abi.make_opt(bias.rule, g_bias['lr'], self._step, g_bias.get('betas', (0.9, 0.999)),
                           g_bias.get('eps', 1e-8), g_bias.get('weight_decay', 0.0))
        rc = _abi.lib().cmf_optimizer_rows_masked(
            _abi.ptr(self.I_shard), _abi.ptr(st['m']), _abi.ptr(st['v']), _abi.ptr(st['g']),
            _abi.ptr(self.bi_shard), _abi.ptr(st['bm']), _abi.ptr(st['bv']), _abi.ptr(st['gb']),
            self.per_rank, self.I_full.shape[1], _abi.ptr(st['flag']), o, ob,
            _abi.stream_ptr(self.I_full.device))
        _abi.check(rc, 'cmf_optimizer_rows_masked')
        if reduce_loss:
            dist.all_reduce(loss, group=self.group)
        return loss

    def gather_item_table(self):
        """The up-to-date, unpadded item table and biases on every rank (for checkpoints / eval)."""
        if self._p2p is not None:      # the owners already stored their rows everywhere
            return self.I_full[:self.true_num_items], self.bi_full[:self.true_num_items]
        all_gather_rows(self.I_full, self.I_shard, self.group)
        all_gather_rows(self.bi_full.view(-1), self.bi_shard.reshape(-1), self.group)
        return self.I_full[:self.true_num_items], self.bi_full[:self.true_num_items]


def common_steps_per_epoch(loader, group=None):
    """Steps every rank can take in one epoch: the minimum of ``len(loader)`` over the ranks. The
    user partition gives the ranks slightly different numbers of interactions, and
    ``ShardedMatrixFactorization.train_step`` is collective -- a rank that took one more step than
    the others would wait for them forever."""
    dev = torch.device('cuda', torch.cuda.current_device()) \
        if dist.get_backend(group) == 'nccl' else torch.device('cpu')
    t = torch.tensor([len(loader)], dtype=torch.int64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
    return int(t.item())


def evaluate_sharded(metric_list, local_test, model, k=10, group=None, **kwargs):
    """``evaluate_in_batches`` over user shards: every rank scores its own test users against the
    (gathered) item table; per-metric sums and user counts are all-reduced."""
    from .metrics import evaluate_in_batches
    cache = local_test.__dict__.setdefault('_cmf_n_test_users', {})
    key = (id(local_test.mat), local_test.mat.nnz)
    if cache.get('key') != key:
        cache.update(key=key, n=int(len(np.unique(local_test.mat.row))))
    n_local = cache['n']
    if n_local:
        res = evaluate_in_batches(metric_list, local_test, model, k=k, verbose=False, **kwargs)
        res = res if isinstance(res, list) else [res]
    else:
        res = [0.0] * len(metric_list)
    dev = next(model.parameters()).device
    t = torch.tensor([r * n_local for r in res] + [float(n_local)], dtype=torch.float64,
                     device=dev)
    dist.all_reduce(t, group=group)
    out = (t[:-1] / t[-1]).tolist()
    return out[0] if len(out) == 1 else out


def evaluate_item_sharded(metric_list, test_interactions, model, k=10, group=None, precision=0,
                          tile_users=None):
    """``evaluate_in_batches`` for ``mapk`` / ``mrr`` with the ITEM axis sharded (for
    catalogues too large to score on one GPU): every rank holds the same test users and scores
    them against its own contiguous block of item rows.

    Per user tile: (1) the targets' exact scores come from the shard that owns each target; the
    best target per user (score descending, id ascending) is agreed with two all-reduces;
    (2) one ``cmf_eval_tile`` per rank over its item block gives the local top-k and the number of
    local items ranked before that best target; (3) the local top-k lists are all-gathered and
    merged in canonical order (``merge_topk``), AP@k is read off the merged ids; the counters are
    summed by all-reduce. Returns what ``evaluate_in_batches`` returns, identical on every rank.
    ``model`` holds the full-width table (e.g. the working copy of ``ShardedMatrixFactorization``);
    only this rank's block of it is read.
    """
    from . import metrics as M
    names = [M._FUSED.get(m) for m in metric_list]
    if any(n not in ('mapk', 'mrr') for n in names):
        raise ValueError('evaluate_item_sharded computes mapk and mrr; use evaluate_sharded '
                         '(user-sharded) for auc or custom metrics')
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    tables = M._model_tables(model)
    dev = tables[0].device
    n_items = int(test_interactions.num_items)
    if 'mapk' in names:
        M._check_k(k, n_items)
    per_rank = (n_items + world - 1) // world
    lo = min(rank * per_rank, n_items)
    hi = min(lo + per_rank, n_items)
    n_loc = hi - lo
    if tile_users is None:
        # users per call: the tensor-core path keeps no score tile, so a call takes as many users
        # as the collectives of the merge are worth; the exact path scores into a workspace of
        # 12 bytes per user x item
        tile_users = 32768 if precision == 1 else max(1, min(4096, (1 << 30) // max(1, 12 * n_loc)))
    test_users = np.unique(test_interactions.mat.row)
    indptr, tgt_items, nnz = M._targets_all_users_device(test_interactions.mat.tocsr(), test_users,
                                                         dev)
    users_dev = torch.from_numpy(test_users.astype(np.int32)).to(dev)
    n_test = len(test_users)
    # ---- (1) best target per user over all shards
    counts = nnz.long()
    row_of = torch.repeat_interleave(torch.arange(n_test, device=dev), counts)
    mine = (tgt_items >= lo) & (tgt_items < hi)
    best_s = torch.full((n_test,), -float('inf'), dtype=torch.float32, device=dev)
    big = torch.iinfo(torch.int32).max
    best_id = torch.full((n_test,), big, dtype=torch.int32, device=dev)
    any_mine = bool(mine.any())
    if any_mine:
        with torch.no_grad():
            sc = model(users_dev.long()[row_of[mine]], tgt_items[mine].long()).float()
        best_s.scatter_reduce_(0, row_of[mine], sc, reduce='amax', include_self=True)
    dist.all_reduce(best_s, op=dist.ReduceOp.MAX, group=group)
    if any_mine:
        at_best = sc == best_s[row_of[mine]]
        best_id.scatter_reduce_(0, row_of[mine][at_best], tgt_items[mine][at_best].int(),
                                reduce='amin', include_self=True)
    dist.all_reduce(best_id, op=dist.ReduceOp.MIN, group=group)
    best_id = torch.where(best_id == big, torch.full_like(best_id, -1), best_id)
    return _item_sharded_tiles(M, names, k, group, world, precision, tile_users, tables, dev,
                               n_items, lo, n_loc, indptr, tgt_items, counts, row_of, users_dev,
                               n_test, best_s, best_id)


def _item_sharded_tiles(M, names, k, group, world, precision, tile_users, tables, dev, n_items,
                        lo, n_loc, indptr, tgt_items, counts, row_of, users_dev, n_test, best_s,
                        best_id):
    """Steps (2) and (3) of ``evaluate_item_sharded``, tile by tile."""
    want = {'topk'} | ({'mrr'} if 'mrr' in names else set())
    sums = {n_: torch.zeros((), dtype=torch.float64, device=dev) for n_ in names}
    status = torch.zeros(2, dtype=torch.int32, device=dev) if precision == 1 else None
    ws = None
    for s0 in range(0, n_test, tile_users):
        e0 = min(s0 + tile_users, n_test)
        nt = e0 - s0
        loc_ids = torch.full((nt, k), -1, dtype=torch.int32, device=dev)
        loc_s = torch.full((nt, k), -float('inf'), dtype=torch.float32, device=dev)
        before = torch.zeros(nt, dtype=torch.long, device=dev)
        if n_loc > 0:       # (more ranks than item blocks: nothing to score on the last ones)
            kk = min(k, n_loc)
            res = M._eval_tile(dev, nt, n_loc, kk, indptr[s0:e0 + 1], tgt_items, want,
                               tables=tables, user_ids=users_dev[s0:e0], precision=precision,
                               workspace=ws, item_offset=lo, reuse_item_prep=(s0 > 0),
                               tc_status=status,
                               ext_best=(best_s[s0:e0].contiguous(), best_id[s0:e0].contiguous()))
            ws = res['_workspace']
            loc_ids[:, :kk], loc_s[:, :kk] = res['topk_ids'], res['topk_scores']
            if 'mrr' in names:
                before = res['n_before_best'].clamp(min=0).long()
        if 'mapk' in names:
            # [shard][user][k] lists of all ranks, merged and scored by one kernel
            all_ids = torch.empty((world, nt, k), dtype=torch.int32, device=dev)
            all_s = torch.empty((world, nt, k), dtype=torch.float32, device=dev)
            dist.all_gather_into_tensor(all_ids, loc_ids, group=group)
            dist.all_gather_into_tensor(all_s, loc_s, group=group)
            ap = torch.empty(nt, dtype=torch.float64, device=dev)
            tile_indptr = indptr[s0:e0 + 1]
            with _abi.on_device(dev):
                _abi.check(_abi.lib().cmf_merge_topk(
                    all_ids.data_ptr(), all_s.data_ptr(), world, nt, k, tile_indptr.data_ptr(),
                    tgt_items.data_ptr(), None, None, ap.data_ptr(),
                    torch.cuda.current_stream(dev).cuda_stream))
            sums['mapk'] += ap.sum()
        if 'mrr' in names:
            dist.all_reduce(before, group=group)
            has = best_id[s0:e0] >= 0
            sums['mrr'] += torch.where(has, 1.0 / (before.double() + 1.0),
                                       torch.zeros(nt, dtype=torch.float64, device=dev)).sum()
    if status is not None:
        dist.all_reduce(status, group=group)
        if int(status[1].item()) != 0:      # (see evaluate_in_batches: the exact path takes over)
            return _item_sharded_tiles(M, names, k, group, world, 0, tile_users, tables, dev,
                                       n_items, lo, n_loc, indptr, tgt_items, counts, row_of,
                                       users_dev, n_test, best_s, best_id)
    out = [float(sums[n_] / n_test) for n_ in names]
    return out[0] if len(out) == 1 else out


def ap_from_topk_ids(top_ids, rows, tgt_keys, n_targets, n_items):
    """(torch statement of what ``cmf_merge_topk`` computes; the tests check the kernel against it)
    AP@k per user from ranked ids ``[n, k]`` (-1 = padding): ``tgt_keys`` = sorted
    ``row * n_items + item`` of every (user row, target) pair, ``rows`` = the user row of each line
    of ``top_ids``, ``n_targets`` = targets per user. Same formula as ``mapk``
    (collie/metrics.py:176-196): sum_j hit_j * precision@j / min(k, n_targets); 0 without targets."""
    k = top_ids.shape[1]
    q = rows.view(-1, 1).long() * n_items + top_ids.long()
    if len(tgt_keys):
        at = torch.searchsorted(tgt_keys, q.clamp(min=0)).clamp(max=len(tgt_keys) - 1)
        hit = (top_ids >= 0) & (tgt_keys[at] == q)
    else:
        hit = torch.zeros_like(top_ids, dtype=torch.bool)
    hit = hit.double()
    prec = torch.cumsum(hit, dim=1) / torch.arange(1, k + 1, device=top_ids.device,
                                                   dtype=torch.float64)
    denom = torch.minimum(n_targets.long(), torch.full_like(n_targets.long(), k)).double()
    return torch.where(denom > 0, (prec * hit).sum(dim=1) / denom.clamp(min=1),
                       torch.zeros_like(denom))


def merge_topk(scores, ids, k):
    """(torch statement of what ``cmf_merge_topk`` computes; the tests check the kernel against it)
    Merge per-shard top-k candidates ``[n, shards * k]`` into the global top-k in canonical order
    (score descending, item id ascending); ids < 0 are padding. Returns ``(scores, ids)``."""
    key = ids.long() + (ids < 0).long() * (1 << 40)          # padding last
    by_id = torch.argsort(key, dim=1, stable=True)
    s = torch.gather(scores, 1, by_id)
    i = torch.gather(ids, 1, by_id)
    s_cmp = torch.where(i < 0, torch.full_like(s, -float('inf')), s)
    # a stable sort by descending score keeps the ascending id order among equal scores
    top = torch.argsort(-s_cmp, dim=1, stable=True)[:, :k]
    return torch.gather(s, 1, top), torch.gather(i, 1, top)
