"""BaseDefense / NoDefense -- mirrors /root/reference/src/defenses/base_defense.py:13-97.

Engine knobs ride in the config dict as optional extra keys (ignored by the reference):
  'output_params'  list of P contiguous fp32 device tensors (e.g. list(global_model.parameters())): the
                   aggregation kernel writes its result straight into them and aggregate() returns them,
                   which turns the per-tensor ``param.copy_(new_val)`` loop of run_experiments.py:255-259
                   into a self-copy.  Device inputs only.
  'process_group'  a flagg.Comm (native NCCL communicator) or a torch.distributed group: the parameter dimension is sharded over its
                   ranks, each rank passes only its column shard, partial Gram matrices /
                   norms / dots are all-reduced (NCCL) before any selection.
"""
from abc import ABC, abstractmethod
from typing import Any, Dict, List

import numpy as np
import torch

from flagg import engine


def _f32(v) -> np.float32:
    """A Python scalar as torch applies it to an fp32 tensor."""
    return np.float32(float(v))


class BaseDefense(ABC):
    def __init__(self, defense_config: Dict[str, Any]):
        self.config = defense_config
        self.name = self.__class__.__name__

    @property
    def _group(self):
        return self.config.get('process_group') if isinstance(self.config, dict) else None

    @property
    def _outputs(self):
        return self.config.get('output_params') if isinstance(self.config, dict) else None

    def _emit(self, u: engine.Updates, flat, dtypes=None):
        """The aggregate as the reference returns it: P tensors shaped like the inputs -- the caller's
        own 'output_params' when the kernel wrote into them (flat is None then)."""
        if flat is None:
            return list(self._outputs)
        return u.split(flat, to_host=u.on_host, dtypes=dtypes)

    @abstractmethod
    def aggregate(self, client_updates: List[List[torch.Tensor]], num_examples: List[int]) -> List[torch.Tensor]:
        pass

    def detect_malicious(self, client_updates, num_examples) -> List[int]:
        return []

    def get_metrics(self) -> Dict[str, Any]:
        return {}

    def __repr__(self) -> str:
        return f"{self.name}(config={self.config})"

    # shared by FedAvg, clipping, norm bounding: num_examples-weighted mean in the
    # reference's operation order (base_defense.py:88-95) via K3
    def _weighted_mean(self, u: engine.Updates, rows, num_examples, scale=None):
        w_ne = [_f32(int(num_examples[i])) for i in rows]
        total = sum(int(num_examples[i]) for i in rows)
        w_sc = None if scale is None else [scale[i] for i in rows]
        flat = engine.row_combine(u, rows, w_sc, w_ne, _f32(total), outputs=self._outputs)
        return flat


class NoDefense(BaseDefense):
    """Plain FedAvg (base_defense.py:74-97)."""

    def __init__(self, defense_config: Dict[str, Any] = None):
        super().__init__(defense_config or {})

    def aggregate(self, client_updates, num_examples):
        u = engine.Updates(client_updates)
        flat = self._weighted_mean(u, list(range(u.n)), num_examples)
        # int * tensor / int keeps floating input dtypes (fp64 in -> fp64 out)
        dt = [d if d.is_floating_point else torch.float32 for d in u.in_dtypes]
        out = self._emit(u, flat, dt)
        u.close()
        return out
