"""Per-task evaluation metrics with the reference's names and definitions (fs_mol/utils/metrics.py:20-195).  The reference
calls scikit-learn; these are the same quantities written in numpy (tests/test_host_logic.py checks them against
scikit-learn), so the evaluation step after the GPU path has no dependency beyond numpy."""
import dataclasses
from dataclasses import dataclass
from typing import Dict, List, Sequence, Tuple

import numpy as np


@dataclass(frozen=True)
class BinaryEvalMetrics:            # metrics.py:20-30
    size: int
    acc: float
    balanced_acc: float
    f1: float
    prec: float
    recall: float
    roc_auc: float
    avg_precision: float
    kappa: float


@dataclass(frozen=True)
class NumericEvalMetrics:           # metrics.py:107-112
    size: int
    mse: float
    mae: float
    r2: float


def _ratio(num: float, den: float, zero_division: float) -> float:
    return float(num / den) if den > 0 else float(zero_division)


def _binary_clf_curve(labels: np.ndarray, scores: np.ndarray):
    """(fps, tps) at every distinct score threshold, descending -- scikit-learn's `_binary_clf_curve`, same operations in
    the same order (the CSV files of the evaluation harness are compared byte for byte with the reference's, whose metrics
    are scikit-learn calls: fs_mol/utils/metrics.py:7-18)."""
    y = np.asarray(labels).astype(bool).astype(np.float64).reshape(-1)
    s = np.asarray(scores, dtype=np.float64).reshape(-1)
    order = np.argsort(s, kind="mergesort")[::-1]
    y, s = y[order], s[order]
    threshold_idxs = np.r_[np.where(np.diff(s))[0], y.size - 1]
    tps = np.cumsum(y, dtype=np.float64)[threshold_idxs]
    fps = 1 + threshold_idxs - tps
    return fps, tps


def roc_auc_score(labels: np.ndarray, scores: np.ndarray) -> float:
    """Area under the ROC curve as scikit-learn computes it: roc_curve (collinear points dropped) + trapezoidal rule."""
    fps, tps = _binary_clf_curve(labels, scores)
    if len(fps) > 2:
        keep = np.where(np.r_[True, np.logical_or(np.diff(fps, 2), np.diff(tps, 2)), True])[0]
        fps, tps = fps[keep], tps[keep]
    tps, fps = np.r_[0, tps], np.r_[0, fps]
    fpr, tpr = fps / fps[-1], tps / tps[-1]
    return float((np.diff(fpr) * (tpr[1:] + tpr[:-1]) / 2.0).sum())


def average_precision_score(labels: np.ndarray, scores: np.ndarray) -> float:
    """sum_k (R_k - R_{k-1}) P_k over the distinct score thresholds, in scikit-learn's order of operations
    (precision_recall_curve reversed to increasing thresholds, then -sum(diff(recall) * precision[:-1]))."""
    fps, tps = _binary_clf_curve(labels, scores)
    ps = tps + fps
    precision = np.zeros_like(tps)
    np.divide(tps, ps, out=precision, where=(ps != 0))
    recall = np.ones_like(tps) if tps[-1] == 0 else tps / tps[-1]
    precision, recall = np.hstack((precision[::-1], 1)), np.hstack((recall[::-1], 0))
    return float(max(0.0, -np.sum(np.diff(recall) * precision[:-1])))


def compute_binary_task_metrics(predictions: Sequence[float], labels: Sequence[float]) -> BinaryEvalMetrics:
    """metrics.py:38-58: predictions are probabilities, thresholded at 0.5; zero_division=1 for f1 / precision / recall;
    roc_auc is 0.0 for a single-class label set."""
    p = np.asarray(predictions, dtype=np.float64).reshape(-1)
    y = np.asarray(labels).reshape(-1).astype(bool)
    yhat = p >= 0.5
    tp = float(np.sum(yhat & y)); tn = float(np.sum(~yhat & ~y))
    fp = float(np.sum(yhat & ~y)); fn = float(np.sum(~yhat & y))
    n = float(len(y))
    n_pos, n_neg = tp + fn, tn + fp
    prec = _ratio(tp, tp + fp, 1.0)
    rec = _ratio(tp, tp + fn, 1.0)
    f1 = _ratio(2.0 * tp, 2.0 * tp + fp + fn, 1.0)
    # balanced accuracy = mean recall over the classes that are present
    recalls = [r for r, c in ((tp / n_pos if n_pos else 0.0, n_pos), (tn / n_neg if n_neg else 0.0, n_neg)) if c > 0]
    p_o = (tp + tn) / n
    p_e = ((tp + fp) * n_pos + (tn + fn) * n_neg) / (n * n)
    kappa = (p_o - p_e) / (1.0 - p_e) if p_e != 1.0 else float("nan")          # cohen_kappa_score: 0/0 on one class
    single_class = n_pos == n or n_pos == 0
    return BinaryEvalMetrics(size=int(n), acc=float(p_o), balanced_acc=float(np.mean(recalls)), f1=f1, prec=prec, recall=rec,
                             roc_auc=0.0 if single_class else roc_auc_score(y, p),
                             avg_precision=average_precision_score(y, p), kappa=float(kappa))


def r2_score_os(y_true, y_pred, y_train_mean: float = 0.0) -> float:
    """Out-of-sample R2 against the support-set mean (0 after the support standardisation), metrics.py:120-131."""
    y_true, y_pred = np.asarray(y_true, dtype=np.float64), np.asarray(y_pred, dtype=np.float64)
    assert len(y_true) == len(y_pred)
    den = ((y_true - y_train_mean) ** 2).sum()
    assert den != 0
    return float(1.0 - ((y_true - y_pred) ** 2).sum() / den)


def compute_numeric_task_metrics(predictions: Sequence[float], labels: Sequence[float]) -> NumericEvalMetrics:
    p, y = np.asarray(predictions, dtype=np.float64).reshape(-1), np.asarray(labels, dtype=np.float64).reshape(-1)
    assert len(p) == len(y)
    return NumericEvalMetrics(size=len(p), mse=float(np.mean((y - p) ** 2)), mae=float(np.mean(np.abs(y - p))), r2=r2_score_os(y, p))


def _avg_list(results: List, cls) -> Dict[str, Tuple[float, float]]:
    out = {}
    for f in dataclasses.fields(cls):
        v = [getattr(r, f.name) for r in results]
        out[f.name] = (np.mean(v), np.std(v))
    return out


def avg_task_metrics_list(results: List[BinaryEvalMetrics]) -> Dict[str, Tuple[float, float]]:
    return _avg_list(results, BinaryEvalMetrics)


def avg_task_numeric_metrics_list(results: List[NumericEvalMetrics]) -> Dict[str, Tuple[float, float]]:
    return _avg_list(results, NumericEvalMetrics)


def _avg_over_tasks(task_results: Dict[str, List], cls) -> Dict[str, Tuple[float, float]]:
    """mean / std ACROSS tasks of the per-task means (within-task stds dropped), metrics.py:61-80, 143-162."""
    per_task = {t: _avg_list(r, cls) for t, r in task_results.items()}
    out = {}
    for f in dataclasses.fields(cls):
        v = [x[f.name][0] for x in per_task.values()]
        out[f.name] = (np.mean(v), np.std(v))
    return out


def avg_metrics_over_tasks(task_results: Dict[str, List[BinaryEvalMetrics]]) -> Dict[str, Tuple[float, float]]:
    return _avg_over_tasks(task_results, BinaryEvalMetrics)


def avg_numeric_metrics_over_tasks(task_results: Dict[str, List[NumericEvalMetrics]]) -> Dict[str, Tuple[float, float]]:
    return _avg_over_tasks(task_results, NumericEvalMetrics)
