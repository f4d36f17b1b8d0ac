"""Train / validation / test splits of an ``Interactions`` object (collie/cross_validation.py).

Host-side data preparation, like in the reference, but vectorised: ``stratified_split`` there runs
one ``sklearn.train_test_split`` per user through joblib (``cross_validation.py:234-306``); here the
per-user draw is one sort of (user, random key) pairs over the whole dataset.

* ``random_split`` reproduces the reference's split bit for bit for the same ``seed`` (the same
  ``np.random.seed`` + ``np.random.shuffle`` + cut-offs, ``cross_validation.py:105-137``).
* ``stratified_split`` gives every user the reference's split SIZES (sklearn's rule for one
  class: ``ceil(test_p * n)`` test interactions, the rest train; a user with one interaction cannot
  be split); which of a user's interactions land in the test set comes from another random stream
  than sklearn's, so the sets differ while their distribution is the same.
"""
from datetime import datetime

import numpy as np
from scipy.sparse import coo_matrix

from .interactions import ExplicitInteractions, Interactions


def _get_random_seed():
    """collie/utils.py:24-26."""
    return int(datetime.now().replace(microsecond=0).timestamp())


def _validate_val_p_and_test_p(val_p, test_p):
    """cross_validation.py:309-317 (same messages)."""
    if val_p >= 1 or val_p < 0:
        raise ValueError('``val_p`` must be in the range [0, 1).')
    if test_p >= 1 or test_p < 0:
        raise ValueError('``test_p`` must be in the range [0, 1).')
    if val_p + test_p >= 1 or val_p + test_p <= 0:
        raise ValueError('The sum of ``val_p`` and ``test_p`` must be in the range (0, 1).')


def _subset_interactions(interactions, idxs):
    """cross_validation.py:18-48: the entries ``idxs`` as a new object of the same kind, with the
    construction-time checks off (the parent already passed them)."""
    idxs = np.asarray(idxs, dtype=np.int64)
    mat = coo_matrix((interactions.mat.data[idxs],
                      (interactions.mat.row[idxs], interactions.mat.col[idxs])),
                     shape=(interactions.num_users, interactions.num_items))
    if isinstance(interactions, Interactions):
        return Interactions(
            mat=mat, num_negative_samples=interactions.num_negative_samples,
            allow_missing_ids=True, remove_duplicate_user_item_pairs=False,
            num_users=interactions.num_users, num_items=interactions.num_items,
            check_num_negative_samples_is_valid=False,
            max_number_of_samples_to_consider=interactions.max_number_of_samples_to_consider,
            seed=interactions.seed)
    return ExplicitInteractions(mat=mat, allow_missing_ids=True,
                                remove_duplicate_user_item_pairs=False,
                                num_users=interactions.num_users,
                                num_items=interactions.num_items)


def random_split(interactions, val_p=0.0, test_p=0.2, processes=None, seed=None):
    """Random train / (validation) / test split (cross_validation.py:51-137). Users are not
    guaranteed to appear in every part. ``processes`` is accepted and unused, as in the reference."""
    _validate_val_p_and_test_p(val_p, test_p)
    if seed is None:
        seed = _get_random_seed()
    n = len(interactions)
    state = np.random.RandomState(seed)          # the stream np.random.seed(seed) would start
    order = np.arange(n)
    state.shuffle(order)
    validate_cutoff = int((1.0 - (val_p + test_p)) * n)
    test_cutoff = int((1.0 - test_p) * n)
    train = _subset_interactions(interactions, order[:validate_cutoff])
    test = _subset_interactions(interactions, order[test_cutoff:])
    if val_p > 0:
        return train, _subset_interactions(interactions, order[validate_cutoff:test_cutoff]), test
    return train, test


def stratified_split(interactions, val_p=0.0, test_p=0.2, processes=-1, seed=None,
                     force_split=False):
    """Split so that every user appears in every part (cross_validation.py:139-231): of a user's
    ``n`` interactions ``ceil(test_p * n)`` go to the test set (sklearn's rule, which the reference
    applies per user), then the rest is split again for the validation set with
    ``val_p / (1 - test_p)``. A user with a single interaction cannot be split: ``ValueError``
    unless ``force_split`` (then it stays in the training set). ``processes`` is accepted and
    unused (nothing here runs per user)."""
    _validate_val_p_and_test_p(val_p, test_p)
    if seed is None:
        seed = _get_random_seed()
    train, test = _stratified_split(interactions, test_p, seed, force_split)
    if val_p > 0:
        train, validate = _stratified_split(train, val_p / (1 - test_p), seed, force_split)
        return train, validate, test
    return train, test


def _stratified_split(interactions, test_p, seed, force_split):
    """One train / test cut with per-user sizes (cross_validation.py:234-306)."""
    users = np.asarray(interactions.mat.row, dtype=np.int64)
    n = len(users)
    rng = np.random.default_rng(seed)
    # a random order inside every user: sort by (user, random key)
    order = np.lexsort((rng.random(n), users))
    su = users[order]
    starts = np.flatnonzero(np.r_[True, su[1:] != su[:-1]]) if n else np.zeros(0, dtype=np.int64)
    counts = np.diff(np.r_[starts, n])
    # sklearn.model_selection.train_test_split sizes: n_test = ceil(test_size * n), n_train =
    # floor((1 - test_size) * n); an empty train set is an error there
    n_test = np.ceil(test_p * counts).astype(np.int64)
    n_train = np.floor((1.0 - test_p) * counts).astype(np.int64)
    unsplittable = n_train == 0
    if unsplittable.any():
        if not force_split:
            raise ValueError(
                'Unable to stratify split on users - the ``interactions`` object contains users'
                ' with a single interaction. Either set ``force_split = True`` to put all users'
                ' with a single interaction in the training set or run'
                ' ``collie.utils.remove_users_with_fewer_than_n_interactions`` first.')
        n_test = np.where(unsplittable, 0, n_test)
    rank_in_user = np.arange(n) - np.repeat(starts, counts)
    is_test_sorted = rank_in_user < np.repeat(n_test, counts)
    is_test = np.zeros(n, dtype=bool)
    is_test[order] = is_test_sorted
    idx = np.arange(n)
    return (_subset_interactions(interactions, idx[~is_test]),
            _subset_interactions(interactions, idx[is_test]))
