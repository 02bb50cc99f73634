"""Upstream module name kept: ``from evaluation_functions import compute_amn, ...``."""
from fedrec_b200.utils.metrics import (  # noqa: F401
    dcg_score, ndcg_score, mrr_score, compute_amn, evaluation_split, roc_auc_score)
