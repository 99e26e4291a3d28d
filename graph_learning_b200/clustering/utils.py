"""Label initialisation helpers (host side; same interface as the reference's
``graph_learn/clustering/utils.py``: ``init_labels`` :9, ``one_hot`` :50).  Like the reference these
defer to scikit-learn -- they are not on the accelerated path."""
from __future__ import annotations

import numpy as np
from sklearn.cluster import k_means, kmeans_plusplus


def init_labels(x, n_clusters: int, init_params: str, random_state) -> np.ndarray:
    """First estimate of the cluster assignments: 'random' (stratified shuffle), 'kmeans' or
    'k-means++'.  Raises ``ValueError`` for anything else."""
    n_samples = x.shape[0]
    if init_params == "random":
        # class c fills slots [c*q, (c+1)*q), the remainder stays 0 (as the reference does)
        labels = _stratified(n_samples, n_clusters)
        random_state.shuffle(labels)
        return labels
    if init_params == "kmeans":
        return k_means(x, n_clusters, init="random", n_init=1, random_state=random_state)[1]
    if init_params == "k-means++":
        return kmeans_plusplus(x, n_clusters, random_state=random_state)
    raise ValueError(f"Invalid init_params: {init_params}")


def _stratified(n_samples: int, n_clusters: int) -> np.ndarray:
    labels = np.zeros(n_samples, dtype=np.int64)
    per = n_samples // n_clusters
    for c in range(n_clusters):
        labels[c * per:(c + 1) * per] = c
    return labels


def one_hot(labels, n_labels: int | None = None) -> np.ndarray:
    """(n_samples,) integer labels -> (n_samples, n_labels) indicator matrix."""
    labels = np.asarray(labels)
    n_labels = n_labels or int(labels.max()) + 1
    out = np.zeros((labels.shape[0], n_labels))
    out[np.arange(labels.shape[0]), labels] = 1
    return out
