"""``collie_b200.cross_validation`` against the reference's own splits (tests/golden/splits.npz, made
by running the unmodified reference: oracle/make_golden.py ``split_case``) and against the
properties the reference's tests check (tests/test_cross_validation.py there). Host-side only."""
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'golden', 'splits.npz')
CASES = {'a': (0.0, 0.2), 'b': (0.1, 0.25)}


def _data(g, explicit=False):
    import collie_b200 as cb
    kw = dict(users=g['users'], items=g['items'], num_users=int(g['n_users']),
              num_items=int(g['n_items']), allow_missing_ids=True,
              remove_duplicate_user_item_pairs=False)
    if explicit:
        rng = np.random.default_rng(0)
        return cb.ExplicitInteractions(ratings=rng.integers(1, 6, len(g['users'])).astype(float),
                                       **kw)
    return cb.Interactions(check_num_negative_samples_is_valid=False, seed=7,
                           max_number_of_samples_to_consider=0, **kw)


def _names(vp):
    return ('train', 'val', 'test') if vp > 0 else ('train', 'test')


@pytest.mark.parametrize('tag', sorted(CASES))
def test_random_split_equals_the_reference_entry_for_entry(tag):
    from collie_b200.cross_validation import random_split
    g = np.load(GOLDEN)
    vp, tp = CASES[tag]
    parts = random_split(_data(g), val_p=vp, test_p=tp, seed=62)
    assert len(parts) == len(_names(vp))
    for name, part in zip(_names(vp), parts):
        assert np.array_equal(part.mat.row, g[f'random_{tag}_{name}_users'])
        assert np.array_equal(part.mat.col, g[f'random_{tag}_{name}_items'])
        assert part.num_users == int(g['n_users']) and part.num_items == int(g['n_items'])
    assert parts[0].seed == 7                       # the parent's sampler settings are kept


@pytest.mark.parametrize('tag', sorted(CASES))
def test_stratified_split_gives_every_user_the_reference_sizes(tag):
    from collie_b200.cross_validation import stratified_split
    g = np.load(GOLDEN)
    vp, tp = CASES[tag]
    data = _data(g)
    parts = stratified_split(data, val_p=vp, test_p=tp, seed=63)
    n_users = int(g['n_users'])
    seen = set()
    for name, part in zip(_names(vp), parts):
        counts = np.bincount(part.mat.row, minlength=n_users)
        assert np.array_equal(counts, g[f'strat_{tag}_{name}_counts'])
        pairs = set(zip(part.mat.row.tolist(), part.mat.col.tolist()))
        assert len(pairs) == part.mat.nnz and not (pairs & seen)     # disjoint parts
        seen |= pairs
    assert seen == set(zip(g['users'].tolist(), g['items'].tolist()))   # nothing lost
    # every user with interactions appears in every part
    present = np.bincount(g['users'], minlength=n_users) > 0
    for part in parts:
        assert np.all((np.bincount(part.mat.row, minlength=n_users) > 0) == present)
    # reproducible for a seed, different for another one
    again = stratified_split(data, val_p=vp, test_p=tp, seed=63)
    assert np.array_equal(again[-1].mat.col, parts[-1].mat.col)
    other = stratified_split(data, val_p=vp, test_p=tp, seed=64)
    assert not np.array_equal(other[-1].mat.col, parts[-1].mat.col)


def test_stratified_split_users_with_one_interaction():
    import collie_b200 as cb
    from collie_b200.cross_validation import stratified_split
    users = np.array([0, 0, 0, 0, 1, 2, 2, 2])
    items = np.array([0, 1, 2, 3, 0, 1, 2, 3])
    data = cb.Interactions(users=users, items=items, num_users=3, num_items=4,
                           allow_missing_ids=True, check_num_negative_samples_is_valid=False)
    with pytest.raises(ValueError, match='users with a single interaction'):
        stratified_split(data, test_p=0.25, seed=1)
    train, test = stratified_split(data, test_p=0.25, seed=1, force_split=True)
    assert 1 in train.mat.row and 1 not in test.mat.row
    assert len(train) + len(test) == len(data) and len(test) == 2


def test_explicit_data_and_argument_checks():
    from collie_b200 import ExplicitInteractions
    from collie_b200.cross_validation import random_split, stratified_split
    g = np.load(GOLDEN)
    data = _data(g, explicit=True)
    train, test = random_split(data, test_p=0.3, seed=5)
    assert isinstance(train, ExplicitInteractions) and isinstance(test, ExplicitInteractions)
    assert len(train) + len(test) == len(data)
    lookup = {(u, i): r for u, i, r in zip(data.mat.row, data.mat.col, data.mat.data)}
    assert all(lookup[(u, i)] == r for u, i, r in zip(test.mat.row, test.mat.col, test.mat.data))
    tr, va, te = stratified_split(data, val_p=0.1, test_p=0.2, seed=5)
    assert len(tr) + len(va) + len(te) == len(data)
    for bad in (dict(val_p=-0.1), dict(test_p=1.0), dict(val_p=0.5, test_p=0.5),
                dict(val_p=0.0, test_p=0.0)):
        with pytest.raises(ValueError, match='must be in the range'):
            random_split(data, **bad)
