"""Model-poisoning attacks -- mirrors /root/reference/src/attacks/model_poisoning.py:15-284.

Each ``poison_update`` is one launch of ``flagg_segments_transform`` over the client's P tensors
in the reference's own fp32 operation order (bit-identical results).  ``poison_update(...,
out=local_update)`` transforms the rows of a resident update matrix in place, so an
attack -> defense round never leaves the GPU.  Random perturbations stay
``torch.randn_like`` (the RNG stream is the caller's), as the DP noise does.
"""
from typing import Any, Dict, List, Optional

import numpy as np
import torch

from flagg import engine
from .base_attack import BaseAttack


class ModelReplacementAttack(BaseAttack):
    def __init__(self, attack_config: Dict[str, Any]):
        super().__init__(attack_config)
        self.scale_factor = attack_config.get('scale_factor', 10.0)
        self.num_malicious = attack_config.get('num_malicious', 1)
        self.boost_factor = attack_config.get('boost_factor', None)

    def poison_update(self, local_update, global_model, num_clients, out=None):
        if self.boost_factor is not None:                                   # :63-67
            scale = self.boost_factor
        else:
            scale = (num_clients / self.num_malicious) * self.scale_factor
        # old + (new - old) * scale, three fp32 roundings (:74-77)
        return engine.segments_transform(engine.T_REPLACE, scale, local_update, global_model, out)

    def get_metrics(self) -> Dict[str, Any]:
        return {'attack_type': 'model_replacement', 'scale_factor': self.scale_factor,
                'num_malicious': self.num_malicious}


class AdaptiveKrumAttack(BaseAttack):
    def __init__(self, attack_config: Dict[str, Any]):
        super().__init__(attack_config)
        self.perturbation_scale = attack_config.get('perturbation_scale', 0.1)
        self.target_direction = attack_config.get('target_direction', None)
        self.num_malicious = attack_config.get('num_malicious', 1)
        self.estimated_benign_center = None

    def estimate_benign_center(self, benign_updates: List[List[torch.Tensor]]):
        """Mean of the benign rows, Python sum() order (:125-138) = K3 with unit weights."""
        if not benign_updates:
            return None
        u = engine.Updates(benign_updates)
        flat = engine.row_combine(u, None, None, None, np.float32(u.n))
        dt = [d if d.is_floating_point else torch.float32 for d in u.in_dtypes]
        out = u.split(flat, to_host=u.on_host, dtypes=dt)
        u.close()
        return out

    def poison_update(self, local_update, global_model, num_clients, benign_updates=None, out=None):
        if benign_updates:
            self.estimated_benign_center = self.estimate_benign_center(benign_updates)
        if self.estimated_benign_center is None:
            return self._simple_perturbation(local_update, global_model)
        if self.target_direction is not None:                               # :176-177
            return engine.segments_transform(engine.T_AXPY, self.perturbation_scale, self.estimated_benign_center,
                                             self.target_direction, out)
        # random perturbation (:179-180): the caller's torch RNG stream
        return [c.clone() + torch.randn_like(c) * self.perturbation_scale for c in self.estimated_benign_center]

    def _simple_perturbation(self, local_update, global_model):
        return [p + torch.randn_like(p) * self.perturbation_scale for p in local_update]      # :193-197

    def get_metrics(self) -> Dict[str, Any]:
        return {'attack_type': 'adaptive_krum', 'perturbation_scale': self.perturbation_scale,
                'num_malicious': self.num_malicious}


class ScalingAttack(BaseAttack):
    def __init__(self, attack_config: Dict[str, Any]):
        super().__init__(attack_config)
        self.scale = attack_config.get('scale', 100.0)

    def poison_update(self, local_update, global_model, num_clients, out=None):
        return engine.segments_transform(engine.T_SCALE, self.scale, local_update, None, out)      # :236

    def get_metrics(self) -> Dict[str, Any]:
        return {'attack_type': 'scaling', 'scale': self.scale}


class InnerProductManipulationAttack(BaseAttack):
    def __init__(self, attack_config: Dict[str, Any]):
        super().__init__(attack_config)
        self.epsilon = attack_config.get('epsilon', 0.1)

    def poison_update(self, local_update, global_model, num_clients, benign_mean: Optional[List[torch.Tensor]] = None,
                      out=None):
        if benign_mean is None:                                             # :271-273
            return engine.segments_transform(engine.T_SCALE, -1.0, local_update, None, out)
        # -epsilon * sign(benign_mean) (:276-281)
        return engine.segments_transform(engine.T_SIGN, -self.epsilon, benign_mean, None, out)

    def get_metrics(self) -> Dict[str, Any]:
        return {'attack_type': 'ipm', 'epsilon': self.epsilon}
