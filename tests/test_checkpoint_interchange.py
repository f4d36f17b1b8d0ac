"""Checkpoint interchange with the reference: a ``.pth`` written by the reference's
``save_model`` (collie/model/base/base_pipeline.py:880-900; golden file made by
``oracle/make_golden.py checkpoint``) is read by ``load_model_path=`` here, and a file written by
this package's ``save_model`` is read by the reference's own loader (:245-257). The CPU part needs
no kernels (state dicts and hyper-parameters); predictions of the loaded model are a GPU test."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, 'tests', 'golden')
KEYS = {'U': 'user_embeddings.weight', 'I': 'item_embeddings.weight',
        'bu': 'user_biases.weight', 'bi': 'item_biases.weight'}


def _golden():
    return np.load(os.path.join(GOLDEN, 'reference_checkpoint.npz'), allow_pickle=False)


def _hide_lightning(monkeypatch):
    """The golden file pickles the hyper-parameters as a ``pytorch_lightning`` class (the stub's,
    in the build container; Lightning's ``AttributeDict`` in a real installation): make sure no
    such package can be imported, like on a machine without Lightning."""
    for name in list(sys.modules):
        if name.split('.')[0] == 'pytorch_lightning':
            monkeypatch.delitem(sys.modules, name)
    monkeypatch.setattr(sys, 'path', [p for p in sys.path if 'oracle' not in p])


def test_reference_checkpoint_loads_without_lightning(monkeypatch):
    import collie_b200 as cb
    _hide_lightning(monkeypatch)
    g = _golden()
    model = cb.MatrixFactorizationModel(
        load_model_path=os.path.join(GOLDEN, 'reference_checkpoint.pth'))
    sd = model.state_dict()
    assert set(sd) == set(KEYS.values())
    for short, key in KEYS.items():
        assert sd[key].dtype == torch.float32 and tuple(sd[key].shape) == g[short].shape
        assert np.array_equal(sd[key].numpy(), g[short])          # bit for bit
    h = model.hparams
    assert set(g['hparams_keys'].tolist()) <= set(h.keys()) | {'load_model_path', 'map_location'}
    assert h.num_users == g['U'].shape[0] and h.num_items == g['I'].shape[0]
    assert h.embedding_dim == g['U'].shape[1]
    assert h.num_epochs_completed == int(g['num_epochs_completed']) == 1
    assert h.loss == 'adaptive_hinge' and h.optimizer == 'adam' and h._is_implicit is True
    assert not model.training                                     # loaded models start in eval()


@pytest.mark.gpu
def test_reference_checkpoint_predicts_like_the_reference():
    import collie_b200 as cb
    g = _golden()
    model = cb.MatrixFactorizationModel(
        load_model_path=os.path.join(GOLDEN, 'reference_checkpoint.pth')).cuda()
    with torch.no_grad():
        got = model(torch.from_numpy(g['pred_users']).cuda(), torch.from_numpy(g['pred_items']).cuda())
    np.testing.assert_allclose(got.cpu().numpy(), g['preds'], rtol=1e-5, atol=1e-7)


def test_reference_loads_a_checkpoint_written_here(tmp_path):
    """The other direction: the reference's ``load_model_path=`` (through ``oracle/_ref``) reads what
    ``save_model`` of this package wrote, and scores with it what the oracle scores."""
    from oracle import ref_runner as R
    if not R.available():
        pytest.skip('oracle/_ref is not built (run __graft_entry__.build() next to /root/reference)')
    import collie_b200 as cb
    from oracle import collie_oracle as O
    rng = np.random.default_rng(5)
    n_users, n_items, dim = 23, 31, 12
    flat = rng.choice(n_users * n_items, size=200, replace=False)
    data = cb.Interactions(users=flat // n_items, items=flat % n_items, num_users=n_users,
                           num_items=n_items, num_negative_samples=3, allow_missing_ids=True,
                           check_num_negative_samples_is_valid=False)
    model = cb.MatrixFactorizationModel(train=data, embedding_dim=dim, loss='bpr',
                                        optimizer='sparse_adam', sparse=True,
                                        lr_scheduler_func=None)
    with torch.no_grad():
        for p in model.parameters():
            p.copy_(torch.from_numpy(rng.normal(0, 0.4, tuple(p.shape)).astype(np.float32)))
    path = str(tmp_path / 'ours.pth')
    model.save_model(path)
    R.import_reference()
    from collie.model import MatrixFactorizationModel as RefMF
    ref = RefMF(load_model_path=path)
    for key, value in model.state_dict().items():
        assert torch.equal(ref.state_dict()[key], value), key
    assert ref.hparams.embedding_dim == dim and ref.hparams.num_users == n_users
    users = torch.from_numpy(rng.integers(0, n_users, 50))
    items = torch.from_numpy(rng.integers(0, n_items, 50))
    with torch.no_grad():
        preds = ref(users, items)
    sd = model.state_dict()
    exp = O.mf_forward(sd['user_embeddings.weight'], sd['item_embeddings.weight'],
                       sd['user_biases.weight'], sd['item_biases.weight'], users, items)
    np.testing.assert_allclose(preds.numpy(), exp.numpy(), rtol=1e-6, atol=1e-7)
