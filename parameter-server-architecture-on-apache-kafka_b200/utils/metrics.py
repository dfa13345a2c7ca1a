"""Weighted F1 and accuracy from a confusion matrix.

Reference: ``Metrics.java:15-24`` uses Spark's ``MulticlassClassificationEvaluator``
with ``metricName="f1"`` — the support-weighted mean of per-class F1 over the
classes that occur as *true* labels — and ``"accuracy"``.  The evaluation
kernel (``csrc/ps_eval.cu``) produces the confusion matrix and finalises the two
numbers with exactly this arithmetic; sklearn's ``f1_score(average="weighted")``
is the test oracle.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass
class Metrics:
    f1: float
    accuracy: float


def confusion_matrix(labels: np.ndarray, predictions: np.ndarray, num_class_rows: int) -> np.ndarray:
    """cm[true, predicted] counts."""
    cm = np.zeros((num_class_rows, num_class_rows), dtype=np.int64)
    np.add.at(cm, (np.asarray(labels, dtype=np.int64), np.asarray(predictions, dtype=np.int64)), 1)
    return cm


def metrics_from_confusion(cm: np.ndarray) -> Metrics:
    cm = np.asarray(cm, dtype=np.float64)
    total = cm.sum()
    if total == 0:
        return Metrics(0.0, 0.0)
    tp = np.diag(cm)
    support = cm.sum(axis=1)            # true-label counts
    predicted = cm.sum(axis=0)
    with np.errstate(divide="ignore", invalid="ignore"):
        precision = np.where(predicted > 0, tp / predicted, 0.0)
        recall = np.where(support > 0, tp / support, 0.0)
        f1 = np.where(precision + recall > 0, 2 * precision * recall / (precision + recall), 0.0)
    weighted = float((f1 * support).sum() / total)
    return Metrics(weighted, float(tp.sum() / total))


def evaluate(labels: np.ndarray, predictions: np.ndarray, num_class_rows: int) -> Metrics:
    return metrics_from_confusion(confusion_matrix(labels, predictions, num_class_rows))
