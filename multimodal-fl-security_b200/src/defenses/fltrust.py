"""FLTrust -- mirrors /root/reference/src/defenses/fltrust.py:17-308.

Trust scores and the renormalisation both need whole-row reductions, so the update
matrix is read twice: K2 produces every ||x_i||^2, x_i.g0 and ||g0||^2 in one pass
(replacing n flatten+dot+2 norms, fltrust.py:170-176, and n more flatten+norms,
:205-209), K3 applies ts_i * (x_i * ||g0||/||x_i||) / sum(ts) in the reference's
operation order (:262-268).  The server-gradient step (:93-150) is model training, not
aggregation, and stays in PyTorch.
"""
from typing import Any, Dict, List, Optional

import numpy as np
import torch
import torch.nn as nn
import torch.optim as optim
from torch.utils.data import DataLoader, Dataset, Subset

from flagg import engine
from flagg.sharding import shard_bounds
from .base_defense import BaseDefense


class FLTrustDefense(BaseDefense):
    def __init__(self, defense_config: Dict[str, Any]):
        super().__init__(defense_config)
        self.root_dataset_size = defense_config.get('root_dataset_size', 100)
        self.learning_rate = defense_config.get('learning_rate', 0.01)
        self.local_epochs = defense_config.get('local_epochs', 1)
        self.batch_size = defense_config.get('batch_size', 32)
        self.root_dataset = None
        self.model = None
        self.device = defense_config.get('device', 'cpu')
        self.trust_scores = []
        self.server_gradient = None

    def set_root_dataset(self, dataset: Dataset) -> None:
        if len(dataset) > self.root_dataset_size:            # fltrust.py:75-83
            idx = np.random.choice(len(dataset), size=self.root_dataset_size, replace=False)
            self.root_dataset = Subset(dataset, idx.tolist())
        else:
            self.root_dataset = dataset

    def set_model(self, model: nn.Module) -> None:
        self.model = model.to(self.device)

    def _flatten(self, tensors: List[torch.Tensor]) -> torch.Tensor:
        return torch.cat([t.flatten().float() for t in tensors])

    def _compute_server_gradient(self, global_params: List[torch.Tensor]) -> List[torch.Tensor]:
        """One short SGD run on the root dataset; g0 = theta_after - theta_before
        (fltrust.py:93-150).  Plain PyTorch: this is training, not the aggregation path."""
        if self.root_dataset is None:
            raise ValueError("Root dataset not set. Call set_root_dataset() first.")
        if self.model is None:
            raise ValueError("Model not set. Call set_model() first.")
        sd = self.model.state_dict()
        for i, key in enumerate(sd.keys()):
            sd[key] = global_params[i].clone()
        self.model.load_state_dict(sd)
        before = [p.clone() for p in global_params]
        self.model.train()
        loss_fn = nn.CrossEntropyLoss()
        opt = optim.SGD(self.model.parameters(), lr=self.learning_rate, momentum=0.9)
        loader = DataLoader(self.root_dataset, batch_size=self.batch_size, shuffle=True)
        for _ in range(self.local_epochs):
            for images, labels in loader:
                images, labels = images.to(self.device), labels.to(self.device)
                opt.zero_grad()
                loss_fn(self.model(images), labels).backward()
                opt.step()
        after = [p.data.clone() for p in self.model.parameters()]
        self.server_gradient = [after[i] - before[i] for i in range(len(before))]
        return self.server_gradient

    # kept for API parity; the engine computes all of these at once in aggregate()
    def _compute_trust_score(self, client_update, server_gradient) -> float:
        u = engine.Updates([client_update])
        st = engine.row_stats(u, self._flatten(server_gradient), group=self._group)
        return self._trust_from_stats(st)[0][0]

    @staticmethod
    def _trust_from_stats(st):
        """fp32 arithmetic of fltrust.py:174-186 and :208-214 on the K2 reductions."""
        f32 = np.float32
        with np.errstate(all='ignore'):
            sn = f32(np.sqrt(st['g0_sumsq']))
            cn = np.sqrt(np.asarray(st['sumsq'], dtype=np.float64)).astype(f32)
            q = np.asarray(st['dot'], dtype=np.float64).astype(f32) / (cn * sn)          # fp32 / fp32
            tiny = cn < 1e-10
            # Python's max(0.0, q) keeps 0.0 unless q > 0.0 (so NaN -> 0.0), fltrust.py:186
            trust = np.where(~tiny & (sn >= 1e-10) & (q > 0), q, f32(0.0)).astype(np.float64)
            scale = np.where(tiny, f32(1.0), sn / cn).astype(f32)
        return trust.tolist(), list(scale)

    def aggregate(self, client_updates, num_examples, global_params: Optional[List[torch.Tensor]] = None):
        if global_params is None:
            raise ValueError("global_params must be provided for FLTrust")     # fltrust.py:235-238
        server_gradient = self._compute_server_gradient(global_params)
        u = engine.Updates(client_updates)
        g0 = self._flatten(server_gradient)
        sharded = self._group is not None and g0.numel() != u.d
        if sharded:
            # parameter-sharded call: the clients' lists hold this rank's columns only (flagg.sharding),
            # g0 comes from the whole model -- cut it to the same column range
            rank, world = engine.group_rank_world(self._group)
            lo, hi = shard_bounds(int(g0.numel()), world)[rank]
            if hi - lo != u.d:
                raise ValueError(f"sharded FLTrust: this rank's updates hold {u.d} columns, its shard of g0 is [{lo}, {hi})")
            g0 = g0[lo:hi]
        # K2 -> all-reduce -> trust scores / renormalisation on the device -> K3: no host round trip in the middle
        # (the fp32 arithmetic of fltrust.py:174-186, 208-214 and Python's double-precision sum(trust), :256)
        flat, info, _, no_trust = engine.weighted_round(u, engine.W_FLTRUST, None, g0, group=self._group, outputs=self._outputs)
        self.trust_scores = info.tolist()
        if no_trust:                                                            # sum(trust) < 1e-10 -> g0 (:258-260)
            if self._outputs is not None:
                u.scatter(g0.to(u.device), u.check_outputs(self._outputs))
                return list(self._outputs)
            if sharded:
                return u.split(g0.to(u.device), to_host=u.on_host)
            return server_gradient
        dt = [d if d.is_floating_point else torch.float32 for d in u.in_dtypes]
        out = self._emit(u, flat, dt)
        u.close()
        return out

    def detect_malicious(self, client_updates, num_examples, threshold: float = 0.1) -> List[int]:
        if not self.trust_scores:
            return []
        return [i for i, s in enumerate(self.trust_scores) if s < threshold]

    def get_metrics(self) -> Dict[str, Any]:
        return {
            'defense_type': 'fltrust',
            'root_dataset_size': self.root_dataset_size,
            'trust_scores': self.trust_scores,
            'avg_trust': np.mean(self.trust_scores) if self.trust_scores else 0.0,
        }

    def __repr__(self) -> str:
        return f"FLTrustDefense(root_size={self.root_dataset_size})"
