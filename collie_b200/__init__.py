"""collie_b200: the matrix-factorization training + evaluation hot path of ShopRunner/collie,
rebuilt for NVIDIA B200 (sm_100a) behind Collie's own API for that path.

    from collie_b200 import (Interactions, MatrixFactorizationModel, CollieTrainer,
                             evaluate_in_batches, mapk, mrr, auc)

All arithmetic runs in ``lib/libcollie_b200.so`` (``include/collie_b200.h``); there is no CPU or
PyTorch fallback -- a missing library or a non-CUDA tensor raises.
"""
from . import _abi
from . import cross_validation  # noqa: F401  (random_split, stratified_split)
from .interactions import (ApproximateNegativeSamplingInteractionsDataLoader,
                           ExplicitInteractions, Interactions, InteractionsDataLoader)
from .metrics import (MeanAbsoluteError, MeanSquaredError, auc, evaluate_in_batches,
                      explicit_evaluate_in_batches, get_preds, mapk, mrr)
from .hybrid import HybridPretrainedModel
from .cml import CollaborativeMetricLearningModel
from . import loss  # noqa: F401  (bpr_loss, hinge_loss, adaptive_*, warp_loss)
from .model import (MatrixFactorizationModel, RowOptimizer, ScaledEmbedding, ZeroEmbedding)
from .trainer import CollieMinimalTrainer, CollieTrainer

__version__ = '0.2.0'

__all__ = [
    'ApproximateNegativeSamplingInteractionsDataLoader', 'CollaborativeMetricLearningModel',
    'CollieMinimalTrainer', 'CollieTrainer',
    'ExplicitInteractions', 'HybridPretrainedModel', 'Interactions', 'InteractionsDataLoader',
    'MatrixFactorizationModel',
    'MeanAbsoluteError', 'MeanSquaredError', 'explicit_evaluate_in_batches',
    'RowOptimizer', 'ScaledEmbedding', 'ZeroEmbedding', 'auc', 'evaluate_in_batches', 'get_preds',
    'mapk', 'mrr',
]
