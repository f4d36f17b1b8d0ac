"""ctypes binding of ``libcollie_b200.so`` (declared in ``include/collie_b200.h``).

There is no fallback: if the library is missing, or a call fails, an exception is raised. Nothing
here (or anywhere under ``collie_b200/``) imports ``oracle/``.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, 'lib', 'libcollie_b200.so')

CMF_MAX_METADATA = 8

LOSS_IDS = {'bpr': 0, 'hinge': 1, 'adaptive_bpr': 2, 'adaptive_hinge': 3, 'warp': 4, 'mse': 5,
            'mae': 6}
OPT_IDS = {'sparse_adam': 0, 'sgd': 1, 'adam': 2, 'adagrad': 3}


class CmfError(RuntimeError):
    pass


class cmf_opt_t(C.Structure):
    _fields_ = [('rule', C.c_int32), ('reserved', C.c_int32), ('lr', C.c_double),
                ('beta1', C.c_double), ('beta2', C.c_double), ('eps', C.c_double),
                ('weight_decay', C.c_double), ('step', C.c_int64)]


CMF_MAX_PEERS = 8


class cmf_peers_t(C.Structure):
    _fields_ = [('rank', C.c_int32), ('world', C.c_int32), ('D', C.c_int32),
                ('reserved', C.c_int32), ('rows_per_rank', C.c_int64),
                ('I', C.c_void_p * CMF_MAX_PEERS), ('bi', C.c_void_p * CMF_MAX_PEERS),
                ('GI', C.c_void_p * CMF_MAX_PEERS), ('gbi', C.c_void_p * CMF_MAX_PEERS),
                ('flag', C.c_void_p * CMF_MAX_PEERS),
                ('mc_I', C.c_void_p), ('mc_bi', C.c_void_p), ('mc_GI', C.c_void_p),
                ('mc_gbi', C.c_void_p), ('mc_flag', C.c_void_p),
                ('pad', C.c_void_p * CMF_MAX_PEERS), ('epoch', C.c_uint32),
                ('reserved2', C.c_uint32), ('clear_ptr', C.c_void_p),
                ('clear_bytes', C.c_size_t)]


class cmf_step_t(C.Structure):
    _fields_ = [
        ('U', C.c_void_p), ('I', C.c_void_p), ('bu', C.c_void_p), ('bi', C.c_void_p),
        ('U_m', C.c_void_p), ('U_v', C.c_void_p), ('I_m', C.c_void_p), ('I_v', C.c_void_p),
        ('bu_m', C.c_void_p), ('bu_v', C.c_void_p), ('bi_m', C.c_void_p), ('bi_v', C.c_void_p),
        ('num_users', C.c_int64), ('num_items', C.c_int64), ('D', C.c_int32),
        ('users', C.c_void_p), ('items', C.c_void_p), ('ratings', C.c_void_p),
        ('negs_in', C.c_void_p), ('gidx', C.c_void_p), ('gidx_base', C.c_int64),
        ('B', C.c_int64), ('N', C.c_int32), ('users_sorted', C.c_int32),
        ('seed', C.c_uint64), ('epoch', C.c_uint32), ('csr_indptr', C.c_void_p),
        ('csr_items', C.c_void_p), ('max_tries', C.c_int32),
        ('loss', C.c_int32), ('has_y_range', C.c_int32), ('y_lo', C.c_float), ('y_hi', C.c_float),
        ('n_meta', C.c_int32), ('meta', C.c_void_p * CMF_MAX_METADATA),
        ('meta_w', C.c_float * CMF_MAX_METADATA),
        ('emb_opt', cmf_opt_t), ('bias_opt', cmf_opt_t), ('train', C.c_int32),
        ('loss_out', C.c_void_p), ('negs_out', C.c_void_p), ('sel_out', C.c_void_p),
        ('pos_scores_out', C.c_void_p), ('neg_scores_out', C.c_void_p),
        ('workspace', C.c_void_p), ('workspace_bytes', C.c_size_t),
        ('B_global', C.c_int64), ('skip_item_update', C.c_int32), ('num_items_logical', C.c_int32),
        ('item_grads_ready_event', C.c_void_p), ('dropout_p', C.c_float),
        ('deterministic', C.c_int32),
    ]


class cmf_eval_t(C.Structure):
    _fields_ = [
        ('U', C.c_void_p), ('I', C.c_void_p), ('bu', C.c_void_p), ('bi', C.c_void_p),
        ('D', C.c_int32), ('user_ids', C.c_void_p), ('n_users_tile', C.c_int64),
        ('item_offset', C.c_int64), ('num_items_shard', C.c_int64), ('k', C.c_int32),
        ('tgt_indptr', C.c_void_p), ('tgt_items', C.c_void_p),
        ('want_mrr', C.c_int32), ('want_auc', C.c_int32), ('precision', C.c_int32),
        ('ext_best_score', C.c_void_p), ('ext_best_id', C.c_void_p),
        ('topk_ids', C.c_void_p), ('topk_scores', C.c_void_p), ('ap', C.c_void_p),
        ('n_before_best', C.c_void_p), ('best_score', C.c_void_p), ('best_id', C.c_void_p),
        ('auc_pairs', C.c_void_p), ('workspace', C.c_void_p), ('workspace_bytes', C.c_size_t),
        ('scores_in', C.c_void_p), ('reuse_item_prep', C.c_int32), ('tc_terms', C.c_int32),
        ('tc_status', C.c_void_p), ('tc_overflow_rows', C.c_void_p), ('tc_row_base', C.c_int64),
        ('tc_overflow_cap', C.c_int32), ('reserved4_', C.c_int32),
    ]


# every symbol include/collie_b200.h declares: name -> (restype, argtypes)
SIGNATURES = {
    'cmf_train_workspace_bytes': (C.c_size_t, [C.c_int64, C.c_int32, C.c_int64, C.c_int64,
                                               C.c_int32]),
    'cmf_workspace_init': (C.c_int, [C.c_void_p, C.c_size_t, C.c_void_p]),
    'cmf_train_workspace_items': (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int32,
                                            C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                            C.POINTER(C.c_void_p)]),
    'cmf_optimizer_rows_masked': (C.c_int, [C.c_void_p] * 8 + [C.c_int64, C.c_int32, C.c_void_p,
                                                              C.POINTER(cmf_opt_t),
                                                              C.POINTER(cmf_opt_t), C.c_void_p]),
    'cmf_train_step': (C.c_int, [C.POINTER(cmf_step_t), C.c_void_p]),
    'cmf_train_step_implicit': (C.c_int, [C.POINTER(cmf_step_t), C.c_void_p]),
    'cmf_train_step_explicit': (C.c_int, [C.POINTER(cmf_step_t), C.c_void_p]),
    'cmf_forward_scores': (C.c_int, [C.c_void_p] * 4 + [C.c_int32, C.c_void_p, C.c_void_p,
                                                       C.c_int32, C.c_int64,
                                                       C.POINTER(C.c_float), C.c_void_p,
                                                       C.c_void_p]),
    'cmf_backward_scores': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                      C.c_int32, C.c_int64, C.c_void_p, C.c_void_p,
                                      C.POINTER(C.c_float), C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p]),
    'cmf_row_similarity': (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_void_p,
                                    C.c_void_p]),
    'cmf_sample_negatives': (C.c_int, [C.c_uint64, C.c_uint32, C.c_void_p, C.c_int64, C.c_void_p,
                                       C.c_int64, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p,
                                       C.c_int32, C.c_void_p, C.c_void_p]),
    'cmf_train_epoch': (C.c_int, [C.POINTER(cmf_step_t), C.c_int64, C.c_void_p, C.c_void_p]),
    'cmf_epoch_plan_workspace_bytes': (C.c_size_t, [C.c_int64]),
    'cmf_epoch_plan': (C.c_int, [C.c_void_p] * 4 + [C.c_int64, C.c_int64, C.c_int64, C.c_int32,
                                               C.c_int32, C.c_uint64, C.c_uint32] +
                       [C.c_void_p] * 5 + [C.c_size_t, C.c_void_p]),
    'cmf_epoch_plan_presorted': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64,
                                          C.c_uint64, C.c_uint32] + [C.c_void_p] * 5 +
                                 [C.c_size_t, C.c_void_p]),
    'cmf_optimizer_rows': (C.c_int, [C.c_void_p] * 4 + [C.c_int64, C.c_int32, C.c_void_p,
                                                       C.c_int64, C.POINTER(cmf_opt_t), C.c_int32,
                                                       C.c_void_p]),
    'cmf_item_exchange_update': (C.c_int, [C.POINTER(cmf_peers_t)] + [C.c_void_p] * 4 +
                                 [C.POINTER(cmf_opt_t), C.POINTER(cmf_opt_t), C.c_void_p]),
    'cmf_exchange_pad_bytes': (C.c_size_t, []),
    'cmf_eval_workspace_bytes': (C.c_size_t, [C.c_int64, C.c_int64, C.c_int32, C.c_int32,
                                              C.c_int32]),
    'cmf_eval_tile': (C.c_int, [C.POINTER(cmf_eval_t), C.c_void_p]),
    'cmf_eval_tc_info': (C.c_int, [C.c_int32, C.POINTER(C.c_double)]),
    'cmf_eval_tile_noalloc': (C.c_int, [C.POINTER(cmf_eval_t), C.c_void_p]),
    'cmf_score_tile': (C.c_int, [C.c_void_p] * 4 + [C.c_int32, C.c_void_p, C.c_int64, C.c_int64,
                                                   C.c_void_p, C.c_void_p]),
    'cmf_ingest_workspace_bytes': (C.c_size_t, [C.c_int64]),
    'cmf_ingest_pairs': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p, C.c_size_t, C.c_void_p]),
    'cmf_pairwise_distance': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                        C.c_int32, C.c_int64, C.c_void_p, C.c_void_p]),
    'cmf_pairwise_distance_backward': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p,
                                                 C.c_void_p, C.c_int32, C.c_int64, C.c_void_p,
                                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'cmf_merge_topk': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_int32,
                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                 C.c_void_p]),
    'cmf_gather_rows': (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_int32,
                                  C.c_int64, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]),
    'cmf_scatter_add_rows': (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_int32,
                                       C.c_int64, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]),
    'cmf_read_probe': (C.c_int, [C.c_void_p, C.c_size_t, C.c_int32, C.c_void_p, C.c_void_p]),
    'cmf_gather_probe': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_int32,
                                  C.c_int32, C.c_void_p, C.c_void_p]),
    'cmf_red_probe': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]),
    'cmf_set_tuning': (C.c_int, [C.c_int32, C.c_int32]),
    'cmf_profile_enable': (C.c_int, [C.c_int32]),
    'cmf_profile_read': (C.c_int, [C.POINTER(C.c_double), C.POINTER(C.c_int32)]
                         + [C.POINTER(C.c_int64)] * 3),
    'cmf_last_error': (C.c_char_p, []),
    'cmf_version': (C.c_int, []),
    'cmf_launch_count': (C.c_int64, []),
}

_lib = None


def lib():
    """Load the library once. Raises ``CmfError`` (never falls back) when it is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise CmfError(
                f'{LIB_PATH} is missing: build it with `make` (or __graft_entry__.build()). '
                'collie_b200 has no CPU or PyTorch fallback.')
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def check(code, what=''):
    if code != 0:
        msg = lib().cmf_last_error().decode('utf-8', 'replace')
        # a few reference error conditions are ValueErrors there; keep the type
        if code == -1:
            raise ValueError(msg or f'{what} failed with code {code}')
        raise CmfError(msg or f'{what} failed with code {code}')


def ptr(t):
    """Device pointer of a tensor (``None`` -> NULL)."""
    if t is None:
        return None
    return C.c_void_p(t.data_ptr())


def stream_ptr(device=None):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def require_cuda(t, name):
    if not t.is_cuda:
        raise CmfError(f'{name} must live on a CUDA device: collie_b200 has no CPU path '
                       f'(got a tensor on {t.device})')
    if not t.is_contiguous():
        raise CmfError(f'{name} must be contiguous')
    return t


def launch_count():
    return int(lib().cmf_launch_count())


def make_opt(rule, lr, step, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
    o = cmf_opt_t()
    o.rule = OPT_IDS[rule] if isinstance(rule, str) else int(rule)
    o.lr = float(lr)
    o.beta1, o.beta2 = float(betas[0]), float(betas[1])
    o.eps = float(eps)
    o.weight_decay = float(weight_decay)
    o.step = int(step)
    return o


def on_device(device):
    """Context that makes ``device`` the current CUDA device. The library keeps per-device state
    (side streams, SM counts, the scratch of ``cmf_eval_tile_noalloc``) keyed by the CURRENT
    device, so every call is made with the device that owns its tensors current."""
    import contextlib
    import torch
    device = torch.device(device)
    return torch.cuda.device(device) if device.type == 'cuda' else contextlib.nullcontext()
