"""Interaction datasets and loaders with the reference's API, resident on the GPU.

Mirrors (same names, argument meaning and error behaviour):
  collie/interactions/datasets.py   BaseInteractions :17-193, Interactions :196-445,
                                    ExplicitInteractions :448-562
  collie/interactions/dataloaders.py InteractionsDataLoader :70-173,
                                    ApproximateNegativeSamplingInteractionsDataLoader :176-294
  collie/interactions/samplers.py   ApproximateNegativeSampler :11-64 (shuffled index slices)

What changes: the COO arrays live in HBM as int32 tensors, an epoch's shuffle is one device
permutation, a batch is a slice of it, and negatives come from the counter-based Philox stream
(``cmf_sample_negatives``; the fused training step draws the same stream in-kernel). HDF5 loaders
are out of scope (disk I/O, not this path).
"""
import collections
import textwrap
import warnings
from typing import Iterable, Optional, Union

import numpy as np
import torch
from scipy.sparse import coo_matrix

from . import _abi


def _check_all_ids_present(array, max_value, name):
    present = np.zeros(int(max_value), dtype=bool)
    arr = np.asarray(array)
    if arr.size and (arr.min() < 0 or arr.max() >= max_value):
        raise ValueError(f'``{name}`` contains ids outside ``[0, {max_value})``.')
    present[arr] = True
    if not present.all():
        raise ValueError(
            f'``{name}`` does not contain every integer in ``[0, {max_value})``; pass '
            '``allow_missing_ids=True`` to skip this check.')


# inputs of at least this many entries are de-duplicated on the device (when there is one)
_DEVICE_INGEST_MIN = 1 << 18


def _ingest_on_device(row, col, num_users, num_items, device=None):
    """``cmf_ingest_pairs``: positions kept (first occurrence of every pair, dataset order), the
    positions their values come from (last occurrence), and ``(device, indptr, csr_items)``."""
    dev = torch.device('cuda', torch.cuda.current_device()) if device is None else device
    n = int(len(row))
    u = torch.from_numpy(np.ascontiguousarray(row, dtype=np.int32)).to(dev)
    i = torch.from_numpy(np.ascontiguousarray(col, dtype=np.int32)).to(dev)
    keep_pos = torch.empty(n, dtype=torch.int32, device=dev)
    keep_val = torch.empty(n, dtype=torch.int32, device=dev)
    n_out = torch.zeros(1, dtype=torch.int64, device=dev)
    indptr = torch.empty(int(num_users) + 1, dtype=torch.int64, device=dev)
    csr_items = torch.empty(n, dtype=torch.int32, device=dev)
    ws = torch.empty(int(_abi.lib().cmf_ingest_workspace_bytes(n)), dtype=torch.uint8, device=dev)
    with _abi.on_device(dev):
        _abi.check(_abi.lib().cmf_ingest_pairs(
            _abi.ptr(u), _abi.ptr(i), n, int(num_users), int(num_items), _abi.ptr(keep_pos),
            _abi.ptr(keep_val), _abi.ptr(n_out), _abi.ptr(indptr), _abi.ptr(csr_items),
            _abi.ptr(ws), ws.numel(), _abi.stream_ptr(dev)), 'cmf_ingest_pairs')
    m = int(n_out.item())
    return (keep_pos[:m].cpu().numpy().astype(np.int64), keep_val[:m].cpu().numpy().astype(np.int64),
            (dev, indptr, csr_items[:m].clone()))


def _infer_num(num, array):
    if num == 'infer':
        return int(np.max(array)) + 1
    return int(num)


class BaseInteractions(torch.utils.data.Dataset):
    """COO wrapper shared by implicit and explicit data (datasets.py:17-193)."""

    def __init__(self, mat=None, users=None, items=None, ratings=None, allow_missing_ids=False,
                 remove_duplicate_user_item_pairs=True, num_users='infer', num_items='infer'):
        if mat is None:
            assert users is not None and items is not None, (
                'Either 1) ``mat`` or 2) both ``users`` or ``items`` must be non-null!')
            if len(users) != len(items):
                raise ValueError('Lengths of ``users`` and ``items`` must be equal.')
            users = np.asarray(users)
            items = np.asarray(items)
            num_users = _infer_num(num_users, users)
            num_items = _infer_num(num_items, items)
            if allow_missing_ids is False:
                _check_all_ids_present(users, num_users, 'users')
                _check_all_ids_present(items, num_items, 'items')
            if ratings is not None:
                if len(users) != len(ratings):
                    raise ValueError(
                        'Length of ``ratings`` must be equal to lengths of ``users``, ``items``.')
                data = np.asarray(ratings)
            else:
                data = np.ones(len(users))
            mat = coo_matrix((data, (users, items)), shape=(num_users, num_items))
        else:
            mat = coo_matrix(mat)
            if num_users == 'infer':
                num_users = mat.shape[0]
            if num_items == 'infer':
                num_items = mat.shape[1]
            if allow_missing_ids is False:
                _check_all_ids_present(mat.row, num_users, 'mat.shape[0]')
                _check_all_ids_present(mat.col, num_items, 'mat.shape[1]')

        self._ingest_csr = None
        if remove_duplicate_user_item_pairs and mat.nnz >= _DEVICE_INGEST_MIN \
                and torch.cuda.is_available():
            # large inputs: the same rule (below) applied by ``cmf_ingest_pairs`` on the device,
            # which also leaves the positives CSR there
            pos, val, csr = _ingest_on_device(mat.row, mat.col, mat.shape[0], mat.shape[1])
            if len(pos) != mat.nnz:
                mat = coo_matrix((mat.data[val], (mat.row[pos], mat.col[pos])), shape=mat.shape)
            self._ingest_csr = csr
        elif remove_duplicate_user_item_pairs:
            # a duplicated (user, item) pair stays at the position of its FIRST occurrence and
            # takes the value of its LAST one: the reference pushes the entries through a dict
            # (``dok_matrix._update`` + ``tocoo``, datasets.py:137-146), which keeps insertion
            # order and overwrites values
            key = mat.row.astype(np.int64) * int(mat.shape[1]) + mat.col.astype(np.int64)
            _, first = np.unique(key, return_index=True)
            if len(first) != len(key):
                _, last_rev = np.unique(key[::-1], return_index=True)
                last = len(key) - 1 - last_rev               # (same key order as ``first``)
                order = np.argsort(first, kind='stable')
                pos, val = first[order], last[order]
                mat = coo_matrix((mat.data[val], (mat.row[pos], mat.col[pos])), shape=mat.shape)

        self.mat = mat
        self.allow_missing_ids = allow_missing_ids
        self.remove_duplicate_user_item_pairs = remove_duplicate_user_item_pairs
        self.num_users = int(num_users)
        self.num_items = int(num_items)
        self.num_interactions = self.mat.nnz
        self.min_rating = self.mat.data.min() if self.mat.nnz else None
        self.max_rating = self.mat.data.max() if self.mat.nnz else None
        self._dev = {}
        # dataset positions in a larger (global) dataset this one was cut from; keys the RNG
        self.global_positions = None

    def __len__(self):
        return self.num_interactions

    def todense(self):
        return self.mat.todense()

    def toarray(self):
        return self.mat.toarray()

    def head(self, n=5):
        n = self._prep_head_tail_n(n)
        return self.mat.tocsr()[range(n), :].toarray()

    def tail(self, n=5):
        n = self._prep_head_tail_n(n)
        return self.mat.tocsr()[range(-n, 0), :].toarray()

    def _prep_head_tail_n(self, n):
        if n < 0:
            n = self.num_users + n
        return min(n, self.num_users)

    # ---- device residency -------------------------------------------------------------------
    def device_state(self, device):
        """int32 COO arrays (and ratings / positives CSR where relevant) on ``device``."""
        device = torch.device(device)
        if device.type != 'cuda':
            raise _abi.CmfError('collie_b200 datasets are consumed on a CUDA device only')
        key = str(device)
        if key not in self._dev:
            st = {
                'users': torch.from_numpy(self.mat.row.astype(np.int32)).to(device),
                'items': torch.from_numpy(self.mat.col.astype(np.int32)).to(device),
            }
            if self.global_positions is not None:
                st['gpos'] = torch.from_numpy(
                    np.asarray(self.global_positions, dtype=np.int64)).to(device)
            self._extra_device_state(st, device)
            self._dev[key] = st
        return self._dev[key]

    def _extra_device_state(self, st, device):
        pass

    def device_records_by_user(self, device):
        """The dataset ordered by user (equal users in dataset order), built once per device:
        ``rec[p] = (user, item, gidx low word, gidx high word)`` as int32 ``[n, 4]`` and the ratings
        alongside (or ``None``); ``gidx`` = the entry's dataset position (``global_positions`` of
        it when set). Input of ``cmf_epoch_plan_presorted``."""
        st = self.device_state(device)
        if 'rec_by_user' not in st:
            users = st['users']
            n = users.shape[0]
            order = torch.sort(users.long(), stable=True)[1]
            gidx = st['gpos'][order] if 'gpos' in st else order
            rec = torch.empty((n, 4), dtype=torch.int32, device=users.device)
            rec[:, 0] = users[order]
            rec[:, 1] = st['items'][order]
            rec[:, 2:] = gidx.contiguous().view(torch.int32).view(n, 2)   # little endian: low, high
            st['rec_by_user'] = rec
            st['ratings_by_user'] = st['ratings'][order].contiguous() if 'ratings' in st else None
        return st['rec_by_user'], st['ratings_by_user']


class Interactions(BaseInteractions):
    """Implicit interactions + negative sampling (datasets.py:196-445)."""

    def __init__(self, mat=None, users=None, items=None, ratings=None, num_negative_samples=10,
                 allow_missing_ids=False, remove_duplicate_user_item_pairs=True,
                 num_users='infer', num_items='infer', check_num_negative_samples_is_valid=True,
                 max_number_of_samples_to_consider=200, seed=None):
        if mat is None and ratings is not None and 0 in set(ratings):
            warnings.warn('``ratings`` contain ``0``s, which are ignored for implicit data.'
                          ' Filtering these rows out.')
            keep = np.asarray(ratings) != 0
            users = np.asarray(users)[keep]
            items = np.asarray(items)[keep]
            ratings = np.asarray(ratings)[keep]
        super().__init__(mat=mat, users=users, items=items, ratings=ratings,
                         allow_missing_ids=allow_missing_ids,
                         remove_duplicate_user_item_pairs=remove_duplicate_user_item_pairs,
                         num_users=num_users, num_items=num_items)
        if seed is None:
            seed = int(np.random.randint(0, 2 ** 31 - 1))
        self.num_negative_samples = num_negative_samples
        self.max_number_of_samples_to_consider = max_number_of_samples_to_consider
        self.check_num_negative_samples_is_valid = check_num_negative_samples_is_valid
        self.seed = seed
        self.epoch = 0  # advances the Philox counter so every pass draws fresh negatives

        assert self.num_negative_samples >= 1
        if (self.num_negative_samples >= self.max_number_of_samples_to_consider
                and self.max_number_of_samples_to_consider > 0):
            warnings.warn('``num_negative_samples > max_number_of_samples_to_consider``. '
                          'Approximate negative sampling will be used.')
        if self.check_num_negative_samples_is_valid:
            max_degree = int(np.bincount(self.mat.row, minlength=1).max()) if self.mat.nnz else 0
            limit = self.num_items - max_degree
            assert self.num_negative_samples < limit, (
                f'``num_negative_samples`` must be less than {limit}!')

    @property
    def exact_sampling(self):
        # the reference takes the rejection loop whenever the budget is positive (datasets.py:392)
        return self.max_number_of_samples_to_consider > 0

    def __repr__(self):
        return textwrap.dedent(f'''
            Interactions object with {self.num_interactions} interactions between {self.num_users}
            users and {self.num_items} items, returning {self.num_negative_samples} negative
            samples per interaction.
            ''').replace('\n', ' ').strip()

    def _extra_device_state(self, st, device):
        # positives CSR (sorted item ids per user) replaces the reference's set((row, col))
        if self._ingest_csr is not None and self._ingest_csr[0] == device:
            st['indptr'], st['csr_items'] = self._ingest_csr[1], self._ingest_csr[2]
            return
        csr = self.mat.tocsr()
        csr.sort_indices()
        st['indptr'] = torch.from_numpy(csr.indptr.astype(np.int64)).to(device)
        st['csr_items'] = torch.from_numpy(csr.indices.astype(np.int32)).to(device)

    def sample_negatives(self, users_dev, gidx_dev, epoch=None):
        """Negative ids ``[B, N]`` (int64, on device) for interactions at positions ``gidx_dev``."""
        st = self.device_state(users_dev.device)
        B = int(users_dev.numel())
        out = torch.empty((B, self.num_negative_samples), dtype=torch.int64,
                          device=users_dev.device)
        exact = self.exact_sampling
        rc = _abi.lib().cmf_sample_negatives(
            self.seed, self.epoch if epoch is None else int(epoch), _abi.ptr(gidx_dev), 0,
            _abi.ptr(users_dev), B, self.num_negative_samples, self.num_items,
            _abi.ptr(st['indptr']) if exact else None,
            _abi.ptr(st['csr_items']) if exact else None,
            int(self.max_number_of_samples_to_consider), _abi.ptr(out),
            _abi.stream_ptr(users_dev.device))
        _abi.check(rc, 'cmf_sample_negatives')
        return out

    def __getitem__(self, index):
        """``((user_id, item_id), negative_item_ids)`` like the reference (datasets.py:378-388).

        Needs a CUDA device: negatives are drawn by the device sampler.
        """
        if not torch.cuda.is_available():
            raise _abi.CmfError('Interactions.__getitem__ draws negatives on the GPU; no CUDA '
                                'device is available and collie_b200 has no CPU fallback')
        user_id = self.mat.row[index]
        item_id = self.mat.col[index]
        scalar = np.ndim(index) == 0
        idx = np.atleast_1d(np.asarray(index, dtype=np.int64))
        idx = np.where(idx < 0, idx + self.num_interactions, idx)
        dev = torch.device('cuda', torch.cuda.current_device())
        negs = self.sample_negatives(
            torch.from_numpy(np.atleast_1d(user_id).astype(np.int32)).to(dev),
            torch.from_numpy(idx).to(dev)).cpu().numpy()
        return (user_id, item_id), (negs[0] if scalar else negs)


class ExplicitInteractions(BaseInteractions):
    """Explicit ratings (datasets.py:448-562)."""

    def __init__(self, mat=None, users=None, items=None, ratings=None, allow_missing_ids=False,
                 remove_duplicate_user_item_pairs=True, num_users='infer', num_items='infer'):
        if mat is None and ratings is None:
            raise ValueError('Ratings must be provided to ``ExplicitInteractions`` with ``mat`` or'
                             ' ``ratings`` - both cannot be ``None``!')
        super().__init__(mat=mat, users=users, items=items, ratings=ratings,
                         allow_missing_ids=allow_missing_ids,
                         remove_duplicate_user_item_pairs=remove_duplicate_user_item_pairs,
                         num_users=num_users, num_items=num_items)

    @property
    def num_negative_samples(self):
        raise AttributeError('``num_negative_samples`` does not exist for explicit datasets.')

    def __repr__(self):
        return textwrap.dedent(f'''
            ExplicitInteractions object with {self.num_interactions} interactions between
            {self.num_users} users and {self.num_items} items, with minimum rating of
            {self.min_rating} and maximum rating of {self.max_rating}.
            ''').replace('\n', ' ').strip()

    def _extra_device_state(self, st, device):
        st['ratings'] = torch.from_numpy(self.mat.data.astype(np.float32)).to(device)

    def __getitem__(self, index):
        return self.mat.row[index], self.mat.col[index], self.mat.data[index]


_INTERACTIONS_KWARGS = ('num_negative_samples', 'allow_missing_ids',
                        'remove_duplicate_user_item_pairs', 'num_users', 'num_items',
                        'check_num_negative_samples_is_valid',
                        'max_number_of_samples_to_consider', 'seed')

DeviceBatch = collections.namedtuple('DeviceBatch', 'users items gidx ratings sorted',
                                     defaults=(False,))


class BatchSorter(object):
    """Orders device batches by user on whatever stream is current (``cmf_epoch_plan`` over one
    batch), so a host-fed pipeline can do it on its copy stream, next to the previous step, and hand
    ``fused_step`` a ``DeviceBatch`` with ``sorted=True`` (which then skips its own sort).
    Holds the output buffers and the scratch for batches of up to ``max_batch`` interactions."""

    def __init__(self, max_batch, num_users, device, with_ratings=False):
        self.max_batch, self.num_users = int(max_batch), int(num_users)
        self.device = torch.device(device)
        n = self.max_batch
        self.users = torch.empty(n, dtype=torch.int32, device=self.device)
        self.items = torch.empty(n, dtype=torch.int32, device=self.device)
        self.gidx = torch.empty(n, dtype=torch.int64, device=self.device)
        self.ratings = torch.empty(n, dtype=torch.float32, device=self.device) \
            if with_ratings else None
        need = int(_abi.lib().cmf_epoch_plan_workspace_bytes(n))
        self.ws = torch.empty(need, dtype=torch.uint8, device=self.device)

    def __call__(self, users, items, gidx, ratings=None):
        """``users`` / ``items`` int32, ``gidx`` int64 (dataset positions), on the device."""
        n = int(users.shape[0])
        if n > self.max_batch:
            raise ValueError(f'batch of {n} exceeds this sorter ({self.max_batch})')
        if (ratings is not None) != (self.ratings is not None):
            raise ValueError('ratings given to a sorter built without them, or the reverse')
        with torch.cuda.device(self.device):
            _abi.check(_abi.lib().cmf_epoch_plan(
                _abi.ptr(users), _abi.ptr(items), _abi.ptr(ratings), _abi.ptr(gidx), n, max(n, 1),
                self.num_users, 0, 1, 0, 0, _abi.ptr(self.users), _abi.ptr(self.items),
                _abi.ptr(self.ratings), _abi.ptr(self.gidx), _abi.ptr(self.ws), self.ws.numel(),
                _abi.stream_ptr(self.device)), 'cmf_epoch_plan')
        return DeviceBatch(self.users[:n], self.items[:n], self.gidx[:n],
                           self.ratings[:n] if ratings is not None else None, True)


class BaseInteractionsDataLoader(object):
    """Properties shared by the loaders (dataloaders.py:16-67)."""

    def __init__(self, interactions, batch_size=1024, shuffle=False, drop_last=False, **kwargs):
        self.interactions = interactions
        self.dataset = interactions
        self.batch_size = batch_size
        self.shuffle = shuffle
        self.drop_last = drop_last
        self.loader_kwargs = kwargs  # torch DataLoader kwargs (num_workers, ...) are accepted
        self._epoch = 0

    @property
    def num_users(self):
        return self.interactions.num_users

    @property
    def num_items(self):
        return self.interactions.num_items

    @property
    def num_negative_samples(self):
        return self.interactions.num_negative_samples

    @property
    def num_interactions(self):
        return self.interactions.num_interactions

    @property
    def mat(self):
        return self.interactions.mat

    def __len__(self):
        n, b = self.num_interactions, self.batch_size
        return n // b if self.drop_last else (n + b - 1) // b

    # ---- device-side epoch plan --------------------------------------------------------------
    def device_batches(self, device, epoch=None, sort_within_batch=True):
        """Yield ``DeviceBatch`` slices of one pass over the data, all tensors on ``device``
        (slices of ``epoch_plan``)."""
        plan = self.epoch_plan(device, epoch=epoch, sort_within_batch=sort_within_batch)
        n, b = plan.users.shape[0], self.batch_size
        for s in range(0, n, b):
            e = min(s + b, n)
            yield DeviceBatch(plan.users[s:e], plan.items[s:e], plan.gidx[s:e],
                              plan.ratings[s:e] if plan.ratings is not None else None,
                              plan.sorted)

    def epoch_plan(self, device, epoch=None, sort_within_batch=True):
        """One pass over the data as ONE ``DeviceBatch`` of contiguous arrays laid out batch after
        batch (``drop_last`` already applied): what ``device_batches`` slices and what
        ``fused_step(..., epoch_batch_size=...)`` consumes whole.

        Shuffled passes permute the dataset positions once on the device (the reference shuffles
        an index array on the host, samplers.py:51-60); ``gidx`` keeps the dataset position of each
        interaction, which keys the negative-sampling stream. With ``sort_within_batch`` every
        batch comes ordered by user (``DeviceBatch.sorted``), which changes nothing but the
        floating-point summation order of a step.
        """
        st = self.interactions.device_state(device)
        n = self.num_interactions
        if epoch is None:
            epoch = self._epoch
            self._epoch += 1
        b = self.batch_size
        device = torch.device(device)
        has_ratings = 'ratings' in st
        users = torch.empty(n, dtype=torch.int32, device=device)
        items = torch.empty(n, dtype=torch.int32, device=device)
        gidx = torch.empty(n, dtype=torch.int64, device=device)
        ratings = torch.empty(n, dtype=torch.float32, device=device) if has_ratings else None
        if n > 0:
            # the whole plan is one native call
            lib = _abi.lib()
            need = int(lib.cmf_epoch_plan_workspace_bytes(n))
            ws = self._plan_ws.get(str(device)) if hasattr(self, '_plan_ws') else None
            if ws is None or ws.numel() < need:
                ws = torch.empty(need, dtype=torch.uint8, device=device)
                self._plan_ws = {str(device): ws}
            seed = int(getattr(self.interactions, 'seed', 0) or 0) & (2 ** 64 - 1)
            with torch.cuda.device(device):
                if self.shuffle and sort_within_batch:
                    # fast path: the entries ordered by user once per dataset, then per epoch a
                    # keyed bijection picks every entry's batch and one stable radix partition
                    # lays the batches out (already ordered by user)
                    rec, rat = self.interactions.device_records_by_user(device)
                    _abi.check(lib.cmf_epoch_plan_presorted(
                        _abi.ptr(rec), _abi.ptr(rat), n, b, seed, int(epoch) & 0xFFFFFFFF,
                        _abi.ptr(users), _abi.ptr(items), _abi.ptr(ratings), _abi.ptr(gidx),
                        _abi.ptr(ws), ws.numel(), _abi.stream_ptr(device)),
                        'cmf_epoch_plan_presorted')
                else:
                    _abi.check(lib.cmf_epoch_plan(
                        _abi.ptr(st['users']), _abi.ptr(st['items']),
                        _abi.ptr(st['ratings']) if has_ratings else None,
                        _abi.ptr(st.get('gpos')), n, b, int(self.num_users),
                        int(bool(self.shuffle)), int(bool(sort_within_batch)), seed,
                        int(epoch) & 0xFFFFFFFF, _abi.ptr(users), _abi.ptr(items),
                        _abi.ptr(ratings), _abi.ptr(gidx), _abi.ptr(ws), ws.numel(),
                        _abi.stream_ptr(device)), 'cmf_epoch_plan')
        stop = (n // b) * b if self.drop_last else n
        return DeviceBatch(users[:stop], items[:stop], gidx[:stop],
                           ratings[:stop] if ratings is not None else None,
                           bool(sort_within_batch))

    def __iter__(self):
        """Reference-format batches on the current CUDA device:
        implicit ``((users, items), negatives[B, N])``, explicit ``(users, items, ratings)``."""
        if not torch.cuda.is_available():
            raise _abi.CmfError('collie_b200 loaders produce device-resident batches; no CUDA '
                                'device is available and there is no CPU fallback')
        dev = torch.device('cuda', torch.cuda.current_device())
        epoch = self._epoch
        self._epoch += 1
        for db in self.device_batches(dev, epoch=epoch, sort_within_batch=False):
            if db.ratings is not None:
                yield db.users, db.items, db.ratings
            else:
                negs = self.interactions.sample_negatives(db.users, db.gidx, epoch=epoch)
                yield (db.users, db.items), negs


class InteractionsDataLoader(BaseInteractionsDataLoader):
    """dataloaders.py:70-173."""

    def __init__(self, interactions=None, mat=None, users=None, items=None, ratings=None,
                 batch_size=1024, shuffle=False, **kwargs):
        if interactions is None:
            ikw = {k: v for k, v in kwargs.items() if k in _INTERACTIONS_KWARGS}
            kwargs = {k: v for k, v in kwargs.items() if k not in _INTERACTIONS_KWARGS}
            interactions = Interactions(mat=mat, users=users, items=items, ratings=ratings, **ikw)
        super().__init__(interactions=interactions, batch_size=batch_size, shuffle=shuffle,
                         **kwargs)

    def __repr__(self):
        if hasattr(self.interactions, 'num_negative_samples'):
            extra = f'{self.num_negative_samples} negative samples per implicit interaction in'
        else:
            extra = 'explicit,'
        return textwrap.dedent(f'''
            InteractionsDataLoader object with {self.num_interactions} interactions between
            {self.num_users} users and {self.num_items} items, returning
            {extra}
            {'shuffled' if self.shuffle else 'non-shuffled'} batches of size {self.batch_size}.
            ''').replace('\n', ' ').strip()


class ApproximateNegativeSamplingInteractionsDataLoader(BaseInteractionsDataLoader):
    """dataloaders.py:176-294: forces approximate (uniform) negatives."""

    def __init__(self, interactions=None, mat=None, users=None, items=None, ratings=None,
                 batch_size=1024, shuffle=False, **kwargs):
        if isinstance(interactions, ExplicitInteractions):
            raise ValueError('``ApproximateNegativeSamplingInteractionsDataLoader`` does not '
                             'support explicit data types!')
        if interactions is None:
            ikw = {k: v for k, v in kwargs.items() if k in _INTERACTIONS_KWARGS}
            kwargs = {k: v for k, v in kwargs.items() if k not in _INTERACTIONS_KWARGS}
            ikw['max_number_of_samples_to_consider'] = 0
            interactions = Interactions(mat=mat, users=users, items=items, ratings=ratings, **ikw)
        else:
            interactions.max_number_of_samples_to_consider = 0
        super().__init__(interactions=interactions, batch_size=batch_size, shuffle=shuffle,
                         **kwargs)

    def __repr__(self):
        return textwrap.dedent(f'''
            ApproximateNegativeSamplingInteractionsDataLoader object with {self.num_interactions}
            interactions between {self.num_users} users and {self.num_items} items, returning
            {self.num_negative_samples} negative samples per implicit interaction in
            {'shuffled' if self.shuffle else 'non-shuffled'} batches of size
            {self.batch_size}.
            ''').replace('\n', ' ').strip()
