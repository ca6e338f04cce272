"""Hyper-detection: embedding-history cosine test + PCA/DBSCAN on embedding deltas, with rollback.

Reference parity: ``/root/reference/src/Utils.py:391-416`` (``cosine``), ``:419-436``
(``DBSCAN_phase2``) and the driver at ``server.py:496-536``: per selected client the current
embedding is tested against the history (deque, maxlen 5 — the config's ``cosine-search`` is read
but ignored, Q12) from round >= 18, appended, and ``all_embeddings.npy`` is rewritten; from round
18 on, clients flagged by *both* tests are removed from the selection and the hyper-network is
rolled back to the pre-round snapshot.  The embeddings are 8-dimensional, so this stays on the
host (numpy / sklearn) exactly like the reference.
"""
from __future__ import annotations

from collections import deque
from typing import Deque, Dict, List, Sequence

import numpy as np


def cosine_anomaly(previous_embeddings: np.ndarray, current_embedding: np.ndarray, k: float = 2.0) -> bool:
    """True when cos(current, mean normalised history) < mu - k*sigma of the history's own cosines."""
    hist = np.asarray(previous_embeddings, dtype=np.float64)
    cur = np.asarray(current_embedding, dtype=np.float64)
    hist_n = hist / np.linalg.norm(hist, axis=1, keepdims=True)
    cur_n = cur / np.linalg.norm(cur, axis=1, keepdims=True)
    centre = hist_n.mean(axis=0)
    cnorm = np.linalg.norm(centre)
    sim = (cur_n @ centre) / (np.linalg.norm(cur_n) * cnorm)
    hist_sim = (hist * centre).sum(axis=1) / (np.linalg.norm(hist, axis=1) * cnorm)
    mu, sigma = hist_sim.mean(), max(hist_sim.std(), 1e-6)
    return bool(np.all(sim < mu - k * sigma))


def dbscan_outliers(before: Sequence[np.ndarray], after: Sequence[np.ndarray], selected: Sequence[int],
                    n_components: int = 3, eps: float = 0.008, min_samples: int = 3) -> List[int]:
    """PCA(n_components) of the per-client embedding deltas, DBSCAN, label -1 => outlier client."""
    from sklearn.cluster import DBSCAN
    from sklearn.decomposition import PCA
    delta = np.array([after[c] for c in selected]) - np.array([before[c] for c in selected])
    delta = delta.reshape(len(selected), -1)
    nc = min(n_components, delta.shape[0], delta.shape[1])
    proj = PCA(n_components=nc).fit_transform(delta)
    labels = DBSCAN(eps=eps, min_samples=min_samples).fit(proj).labels_
    return [selected[i] for i in np.where(labels == -1)[0]]


class EmbeddingMonitor:
    """State + policy of ``server.py:496-536``."""

    def __init__(self, total_clients: int, history: int = 5, start_round: int = 18,
                 n_components: int = 3, eps: float = 0.007, min_samples: int = 3,
                 save_path: str = "all_embeddings.npy"):
        self.hist: List[Deque[np.ndarray]] = [deque(maxlen=history) for _ in range(total_clients)]
        self.start_round = start_round
        self.n_components, self.eps, self.min_samples = n_components, eps, min_samples
        self.save_path = save_path

    def observe(self, round_no: int, selected: Sequence[int], embeddings: Dict[int, np.ndarray]) -> List[int]:
        """Feed this round's embeddings (``[1, E]`` each); returns the clients to remove."""
        flagged: List[int] = []
        for c in selected:
            cur = np.asarray(embeddings[c], dtype=np.float32).reshape(1, -1)
            if round_no >= self.start_round and len(self.hist[c]) > 0:
                if cosine_anomaly(np.vstack(self.hist[c]), cur):
                    flagged.append(c)
            self.hist[c].append(cur)
        if self.save_path:
            np.save(self.save_path, np.array([list(dq) for dq in self.hist], dtype=object))
        if round_no < self.start_round:
            return []
        ok = [c for c in selected if len(self.hist[c]) >= 2]
        if len(ok) < 2:
            return []
        before = {c: self.hist[c][-2] for c in ok}
        after = {c: self.hist[c][-1] for c in ok}
        outliers = dbscan_outliers(before, after, ok, self.n_components, self.eps, self.min_samples)
        return sorted(set(flagged) & set(outliers))
