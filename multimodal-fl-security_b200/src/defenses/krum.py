"""Krum / Multi-Krum -- mirrors /root/reference/src/defenses/krum.py:14-237.

The O(n^2) Python loop of fp32 ``torch.norm(fi - fj).item()`` (krum.py:92-97) becomes one
centred FP64 Gram matrix (K4) plus a small distance/score kernel (K5).  Scores are sums
of UN-squared distances, as the reference computes them (krum.py:95,129), and the final
ranking is the same host-side ``np.argsort`` call (krum.py:174) so tie order follows
numpy exactly.
"""
from typing import Any, Dict, List

import numpy as np
import torch

from flagg import engine
from .base_defense import BaseDefense


class KrumDefense(BaseDefense):
    def __init__(self, defense_config: Dict[str, Any]):
        super().__init__(defense_config)
        self.num_malicious = defense_config.get('num_malicious', 1)
        self.multi_k = defense_config.get('multi_k', 1)
        self.selected_clients = []
        self.rejected_clients = []
        self.client_scores = []

    def _flatten_update(self, update: List[torch.Tensor]) -> torch.Tensor:
        return torch.cat([t.flatten().float() for t in update])

    def _distances_and_scores(self, u: engine.Updates, f: int):
        return engine.krum_distances_scores(u, f, group=self._group, want_D=False)

    def _compute_distances(self, client_updates) -> np.ndarray:
        """Pairwise L2 distance matrix, float64 [n, n] (krum.py:73-99)."""
        u = engine.Updates(client_updates)
        n = u.n
        if n >= 3:      # scores with one neighbour come for free and route non-finite rows to the exact path
            D, _ = engine.krum_distances_scores(u, n - 3, group=self._group, want_D=True)
        else:
            D, _ = engine.krum_scores(engine.sqdist(u, group=self._group), 0, want_D=True, want_scores=False, squared=True)
        u.close()
        return D

    def aggregate(self, client_updates, num_examples):
        n = len(client_updates)
        f = self.num_malicious
        if n < 2 * f + 3:                                   # krum.py:153-157
            raise ValueError(
                f"Krum requires n >= 2f + 3. Got n={n}, f={f}. "
                f"Need at least {2*f + 3} clients."
            )
        u = engine.Updates(client_updates)
        _, scores = self._distances_and_scores(u, f)
        self.client_scores = [np.float64(s) for s in scores]
        sorted_indices = np.argsort(self.client_scores)      # krum.py:174
        self.selected_clients = sorted_indices[:self.multi_k].tolist()
        self.rejected_clients = sorted_indices[self.multi_k:].tolist()
        if self.multi_k == 1:
            sel = client_updates[self.selected_clients[0]]   # the client's own list object (krum.py:178-181)
            if self._outputs is not None:                    # epilogue form: the selected row lands in the model
                engine.segments_transform(engine.T_SCALE, 1.0, sel, None, u.check_outputs(self._outputs))
            u.close()
            return sel
        # krum.py:187-190: sum(selected, in score order) / multi_k
        flat = engine.row_combine(u, self.selected_clients, None, None, np.float32(self.multi_k), outputs=self._outputs)
        dt = [d if d.is_floating_point else torch.float32 for d in u.in_dtypes]
        out = self._emit(u, flat, dt)
        u.close()      # drops the device copy of host inputs now, not at the next GC pass
        return out

    def detect_malicious(self, client_updates, num_examples) -> List[int]:
        if not self.client_scores:
            self.aggregate(client_updates, num_examples)
        return self.rejected_clients

    def get_metrics(self) -> Dict[str, Any]:
        return {
            'defense_type': 'krum',
            'num_malicious_assumed': self.num_malicious,
            'multi_k': self.multi_k,
            'selected_clients': self.selected_clients,
            'rejected_clients': self.rejected_clients,
            'client_scores': self.client_scores,
        }

    def __repr__(self) -> str:
        return f"KrumDefense(f={self.num_malicious}, k={self.multi_k})"


class MultiKrumDefense(KrumDefense):
    def __init__(self, defense_config: Dict[str, Any]):
        if 'multi_k' not in defense_config:                  # mutates the caller's dict, as krum.py:234-235
            defense_config['multi_k'] = defense_config.get('default_k', 3)
        super().__init__(defense_config)
