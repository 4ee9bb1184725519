"""BaseAttack / NoAttack -- mirrors /root/reference/src/attacks/base_attack.py (interface only)."""
from abc import ABC, abstractmethod
from typing import Any, Dict, List

import torch


class BaseAttack(ABC):
    def __init__(self, attack_config: Dict[str, Any]):
        self.config = attack_config
        self.name = self.__class__.__name__

    def poison_data(self, dataset):
        return dataset

    @abstractmethod
    def poison_update(self, local_update: List[torch.Tensor], global_model: List[torch.Tensor],
                      num_clients: int) -> List[torch.Tensor]:
        pass

    def is_data_poisoning(self) -> bool:
        return False

    def is_model_poisoning(self) -> bool:
        return True

    def get_metrics(self) -> Dict[str, Any]:
        return {}

    def __repr__(self) -> str:
        return f"{self.name}(config={self.config})"


class NoAttack(BaseAttack):
    def poison_update(self, local_update, global_model, num_clients):
        return local_update

    def is_model_poisoning(self) -> bool:
        return False
