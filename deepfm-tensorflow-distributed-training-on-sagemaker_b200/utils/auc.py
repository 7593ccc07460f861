"""Streaming AUC with `tf.metrics.auc` defaults (reference: PS:249-251 / HVD:241-243).

TF-1 defaults: num_thresholds=200, ROC curve, trapezoidal summation. Thresholds are
[-eps, 1/(T-1), ..., (T-2)/(T-1), 1+eps] with eps=1e-7; a prediction counts as positive at
threshold t when pred > t.  TP/FP/TN/FN accumulate across batches (and across ranks: the four
[200] vectors are summed with one small all-reduce at eval end).  `tpr = (tp+e)/(tp+fn+e)`,
`fpr = fp/(fp+tn+e)`, e=1e-6.

Bucketed form used here and by the `auc_hist` kernel: bucket(pred) = number of thresholds
strictly below pred; per-class histograms over 201 buckets give the confusion counts by suffix sums.
"""
from __future__ import annotations

import numpy as np
import torch

NUM_THRESHOLDS = 200
_KEPS = 1e-7
_EPS = 1e-6


def thresholds(num: int = NUM_THRESHOLDS) -> np.ndarray:
    t = [(i + 1) * 1.0 / (num - 1) for i in range(num - 2)]
    return np.asarray([0.0 - _KEPS] + t + [1.0 + _KEPS], dtype=np.float32)


class StreamingAUC:
    def __init__(self, num_thresholds: int = NUM_THRESHOLDS):
        self.T = num_thresholds
        self.thr = torch.from_numpy(thresholds(num_thresholds))
        # hist[c, k]: number of class-c samples with exactly k thresholds strictly below the prediction
        self.hist = torch.zeros(2, num_thresholds + 1, dtype=torch.float64)

    def update(self, labels: torch.Tensor, probs: torch.Tensor) -> None:
        labels = labels.detach().reshape(-1).cpu().float()
        probs = probs.detach().reshape(-1).cpu().float()
        k = torch.bucketize(probs, self.thr, right=False)  # #thr < prob  (pred > thr)
        pos = labels > 0.5
        self.hist[1] += torch.bincount(k[pos], minlength=self.T + 1).double()
        self.hist[0] += torch.bincount(k[~pos], minlength=self.T + 1).double()

    def update_hist(self, hist: torch.Tensor) -> None:
        """Accumulate a [2, T+1] histogram produced by the `auc_hist` kernel."""
        self.hist += hist.detach().cpu().double()

    def result(self) -> float:
        return auc_from_hist(self.hist)


def auc_from_hist(hist: torch.Tensor) -> float:
    h = hist.double()
    T = h.shape[1] - 1
    # predicted positive at threshold i  <=>  bucket k >= i+1
    suffix = torch.flip(torch.cumsum(torch.flip(h, [1]), 1), [1])  # suffix[c, j] = sum_{k>=j}
    tp = suffix[1, 1:T + 1]
    fp = suffix[0, 1:T + 1]
    fn = h[1].sum() - tp
    tn = h[0].sum() - fp
    tpr = (tp + _EPS) / (tp + fn + _EPS)
    fpr = fp / (fp + tn + _EPS)
    return float(((fpr[:-1] - fpr[1:]) * (tpr[:-1] + tpr[1:]) / 2.0).sum())


def exact_auc(labels: np.ndarray, probs: np.ndarray) -> float:
    """Sort-based (Mann-Whitney) AUC with tie handling; test oracle for the bucketed metric."""
    labels = np.asarray(labels).reshape(-1) > 0.5
    probs = np.asarray(probs, dtype=np.float64).reshape(-1)
    order = np.argsort(probs, kind="mergesort")
    ranks = np.empty(len(probs), dtype=np.float64)
    sp = probs[order]
    i = 0
    while i < len(sp):
        j = i
        while j + 1 < len(sp) and sp[j + 1] == sp[i]:
            j += 1
        ranks[order[i:j + 1]] = 0.5 * (i + j) + 1.0
        i = j + 1
    npos, nneg = labels.sum(), (~labels).sum()
    if npos == 0 or nneg == 0:
        return 0.0
    return float((ranks[labels].sum() - npos * (npos + 1) / 2.0) / (npos * nneg))
