"""DP-SGD, gradient clipping, norm bounding -- mirrors
/root/reference/src/defenses/differential_privacy.py:15-349.

All three are "row norm -> per-row factor -> num_examples-weighted mean": K2 gives every
row's L2^2 and Linf in one pass over the matrix, the factors are formed on the host in
the reference's own precision, K3 applies them in the reference's operation order.
"""
from typing import Any, Dict, List, Tuple

import numpy as np
import torch

from flagg import engine
from .base_defense import BaseDefense, _f32


def _l2_norms32(st) -> np.ndarray:
    """torch.norm(flat) as an fp32 value per client."""
    with np.errstate(all='ignore'):
        return np.sqrt(st['sumsq']).astype(np.float32)


def _out_dtypes(u):
    return [d if d.is_floating_point else torch.float32 for d in u.in_dtypes]


class DPSGDDefense(BaseDefense):
    def __init__(self, defense_config: Dict[str, Any]):
        super().__init__(defense_config)
        self.clip_norm = defense_config.get('clip_norm', 10.0)
        self.noise_multiplier = defense_config.get('noise_multiplier', 0.005)
        self.target_epsilon = defense_config.get('target_epsilon', 8.0)
        self.target_delta = defense_config.get('target_delta', 1e-5)
        self.rounds_completed = 0
        self.privacy_spent = 0.0

    def _scales(self, norms32: np.ndarray):
        """differential_privacy.py:90-92: ``norm > clip`` compares in fp32 and the scale is
        the 0-dim fp32 tensor ``clip / norm`` == norm.reciprocal() * clip (two roundings)."""
        clip32 = _f32(self.clip_norm)
        out = []
        with np.errstate(all='ignore'):
            for nv in norms32:
                out.append(np.float32(np.float32(1.0) / nv) * clip32 if nv > clip32 else np.float32(1.0))
        return out

    def clip_gradient(self, update: List[torch.Tensor]) -> Tuple[List[torch.Tensor], float]:
        u = engine.Updates([update])
        norm32 = _l2_norms32(engine.row_stats(u, group=self._group))[0]
        if norm32 > _f32(self.clip_norm):
            flat = engine.row_combine(u, [0], self._scales([norm32]), None, 1.0)
            return u.split(flat, to_host=u.on_host, dtypes=_out_dtypes(u)), float(norm32)
        return update, float(norm32)

    def add_noise(self, update: List[torch.Tensor], num_clients: int) -> List[torch.Tensor]:
        # Gaussian noise stays torch.randn_like on the output tensors (:112-117): the RNG
        # stream is the caller's, not something a custom kernel could reproduce.
        noise_std = self.clip_norm * self.noise_multiplier / num_clients
        return [p + torch.randn_like(p) * noise_std for p in update]

    def aggregate(self, client_updates, num_examples):
        u = engine.Updates(client_updates)
        n = u.n
        # K2 -> (all-reduce) -> clip factors on the device (:90-92, fp32) -> K3, one read-back at the end
        flat, _, _, _ = engine.weighted_round(u, engine.W_DPSGD, num_examples, p0=self.clip_norm, group=self._group,
                                              outputs=self._outputs)
        aggregated = self._emit(u, flat, _out_dtypes(u))
        if flat is None:      # epilogue form: the noise of :112-117 lands in the model tensors too (same RNG calls)
            noise_std = self.clip_norm * self.noise_multiplier / n
            for p in aggregated:
                p.add_(torch.randn_like(p) * noise_std)
            noisy = aggregated
        else:
            noisy = self.add_noise(aggregated, n)
        self.rounds_completed += 1
        self._update_privacy_spent(n)
        return noisy

    def _update_privacy_spent(self, num_clients: int):
        self.privacy_spent += np.sqrt(2 * np.log(1 / self.target_delta)) / self.noise_multiplier   # :175-176

    def get_privacy_spent(self) -> float:
        return self.privacy_spent

    def is_budget_exhausted(self) -> bool:
        return self.privacy_spent >= self.target_epsilon

    def get_metrics(self) -> Dict[str, Any]:
        return {'defense_type': 'dp_sgd', 'clip_norm': self.clip_norm, 'noise_multiplier': self.noise_multiplier,
                'rounds_completed': self.rounds_completed, 'privacy_spent': self.privacy_spent,
                'target_epsilon': self.target_epsilon}

    def __repr__(self) -> str:
        return (f"DPSGDDefense(clip={self.clip_norm}, noise={self.noise_multiplier}, "
                f"eps={self.privacy_spent:.2f}/{self.target_epsilon})")


class GradientClippingDefense(BaseDefense):
    def __init__(self, defense_config: Dict[str, Any]):
        super().__init__(defense_config)
        self.clip_norm = defense_config.get('clip_norm', 1.0)
        self.clip_type = defense_config.get('clip_type', 'l2')
        self.clipped_count = 0
        self.original_norms = []

    def _norms(self, u) -> List[float]:
        st = engine.row_stats(u, group=self._group)
        if self.clip_type == 'linf':                          # :229-230
            return [float(v) for v in st['linf']]
        return [float(v) for v in _l2_norms32(st)]            # :227-228, :231-232

    def _compute_norm(self, update: List[torch.Tensor]) -> float:
        return self._norms(engine.Updates([update]))[0]

    def _scale(self, norm: float):
        """:242-245 -- ``clip_norm / norm`` is a Python float here, cast to fp32 by the multiply."""
        return _f32(self.clip_norm / norm) if norm > self.clip_norm else np.float32(1.0)

    def clip_update(self, update: List[torch.Tensor]) -> List[torch.Tensor]:
        u = engine.Updates([update])
        norm = self._norms(u)[0]
        self.original_norms.append(norm)
        if norm > self.clip_norm:
            self.clipped_count += 1
            flat = engine.row_combine(u, [0], [self._scale(norm)], None, 1.0)
            return u.split(flat, to_host=u.on_host, dtypes=_out_dtypes(u))
        return update

    def aggregate(self, client_updates, num_examples):
        self.clipped_count = 0
        self.original_norms = []
        u = engine.Updates(client_updates)
        # K2 -> (all-reduce) -> clip factors on the device (:242-245: Python float clip / norm, then fp32) -> K3
        flat, norms, _, clipped = engine.weighted_round(u, engine.W_CLIP, num_examples, p0=self.clip_norm,
                                                        use_linf=self.clip_type == 'linf', group=self._group, outputs=self._outputs)
        self.original_norms = norms.tolist()
        self.clipped_count = clipped
        return self._emit(u, flat, _out_dtypes(u))

    def get_metrics(self) -> Dict[str, Any]:
        return {'defense_type': 'gradient_clipping', 'clip_norm': self.clip_norm, 'clip_type': self.clip_type,
                'clipped_count': self.clipped_count, 'original_norms': self.original_norms}


class NormBoundingDefense(BaseDefense):
    def __init__(self, defense_config: Dict[str, Any]):
        super().__init__(defense_config)
        self.max_norm = defense_config.get('max_norm', 10.0)
        self.min_norm = defense_config.get('min_norm', 0.0)
        self.rejected_clients = []

    def aggregate(self, client_updates, num_examples):
        self.rejected_clients = []
        u = engine.Updates(client_updates)
        # K2 -> (all-reduce) -> keep mask min <= norm <= max (:313; all rejected -> all used, :319-322) and the kept
        # clients' example total on the device -> K3 over the kept rows
        flat, norms, _, _ = engine.weighted_round(u, engine.W_NORMBOUND, num_examples, p0=self.max_norm, p1=self.min_norm,
                                                  group=self._group, outputs=self._outputs)
        self.rejected_clients = [i for i, nv in enumerate(norms.tolist()) if not (self.min_norm <= nv <= self.max_norm)]
        return self._emit(u, flat, _out_dtypes(u))

    def detect_malicious(self, client_updates, num_examples) -> List[int]:
        return self.rejected_clients

    def get_metrics(self) -> Dict[str, Any]:
        return {'defense_type': 'norm_bounding', 'max_norm': self.max_norm, 'min_norm': self.min_norm,
                'rejected_clients': self.rejected_clients}
