"""Minimal stand-in for ``torchmetrics`` (pinned 0.11.4 by the reference, absent here).

TEST INFRASTRUCTURE ONLY. ``functional.auroc`` is implemented with
``sklearn.metrics.roc_auc_score``, the yardstick the reference's own test pins ``auc`` to
(``/root/reference/tests/test_metrics.py:140-154``).
"""
from . import functional  # noqa: F401


class Metric(object):
    pass
