"""scipy >= 1.13 dropped ``dok_matrix._update`` used at collie/interactions/datasets.py:141."""
try:
    import scipy.sparse as _sp

    if not hasattr(_sp.dok_matrix, '_update'):
        def _update(self, data):
            items = data.items() if hasattr(data, 'items') else data
            for k, v in items:
                self[k] = v
        _sp.dok_matrix._update = _update
except Exception:  # pragma: no cover
    pass
