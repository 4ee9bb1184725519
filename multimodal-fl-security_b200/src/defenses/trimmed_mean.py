"""Trimmed mean / median / geometric median -- mirrors
/root/reference/src/defenses/trimmed_mean.py:14-272.

Median and trimmed mean are one launch of K1 (column-wise order statistics) over the
un-packed update lists; the geometric median's Weiszfeld loop is composed from K1
(initial point), K2-style distances and K3 (weighted combine).
"""
from typing import Any, Dict, List

import numpy as np
import torch

from flagg import engine, MEDIAN, TRIMMED_MEAN
from .base_defense import BaseDefense



class TrimmedMeanDefense(BaseDefense):
    def __init__(self, defense_config: Dict[str, Any]):
        super().__init__(defense_config)
        self.trim_ratio = defense_config.get('trim_ratio', 0.1)
        self.num_trimmed_per_end = 0

    def aggregate(self, client_updates, num_examples):
        n = len(client_updates)
        self.num_trimmed_per_end = max(1, int(n * self.trim_ratio))     # trimmed_mean.py:66
        if n - 2 * self.num_trimmed_per_end < 1:                        # :69-72
            return self._coordinate_wise_median(client_updates)
        u = engine.Updates(client_updates)
        # .float() outputs (trimmed_mean.py:79,160)
        return self._emit(u, engine.colwise_select(u, TRIMMED_MEAN, self.num_trimmed_per_end, outputs=self._outputs))

    def _coordinate_wise_median(self, client_updates):
        u = engine.Updates(client_updates)
        return self._emit(u, engine.colwise_select(u, MEDIAN, outputs=self._outputs))

    def get_metrics(self) -> Dict[str, Any]:
        return {'defense_type': 'trimmed_mean', 'trim_ratio': self.trim_ratio,
                'num_trimmed_per_end': self.num_trimmed_per_end}

    def __repr__(self) -> str:
        return f"TrimmedMeanDefense(trim_ratio={self.trim_ratio})"


class MedianDefense(BaseDefense):
    def __init__(self, defense_config: Dict[str, Any] = None):
        super().__init__(defense_config or {})

    def aggregate(self, client_updates, num_examples):
        u = engine.Updates(client_updates)
        return self._emit(u, engine.colwise_select(u, MEDIAN, outputs=self._outputs))

    def get_metrics(self) -> Dict[str, Any]:
        return {'defense_type': 'median'}

    def __repr__(self) -> str:
        return "MedianDefense()"


class GeometricMedianDefense(BaseDefense):
    """Weiszfeld iterations (trimmed_mean.py:225-265): start at the coordinate-wise
    median, then x <- sum_i(u_i / ||u_i - x||) / sum_i(1 / ||u_i - x||)."""

    def __init__(self, defense_config: Dict[str, Any] = None):
        super().__init__(defense_config or {})
        self.max_iters = defense_config.get('max_iters', 100) if defense_config else 100
        self.tolerance = defense_config.get('tolerance', 1e-5) if defense_config else 1e-5
        self.num_iters = 0

    def aggregate(self, client_updates, num_examples):
        u0 = engine.Updates(client_updates)
        # the whole loop runs on the device (engine.weiszfeld): distances, weights, combine and the
        # stopping norm; with 'process_group' the distances and the change are whole-vector values
        # on every rank, so all ranks stop at the same iteration (trimmed_mean.py:254-260)
        current, self.num_iters = engine.weiszfeld(u0, self.max_iters, self.tolerance, group=self._group)
        if self._outputs is not None:
            u0.scatter(current, u0.check_outputs(self._outputs))
            current = None
        out = self._emit(u0, current)
        u0.close()
        return out

    def get_metrics(self) -> Dict[str, Any]:
        return {'defense_type': 'geometric_median', 'max_iters': self.max_iters, 'num_iters': self.num_iters}
