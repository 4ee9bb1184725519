"""Model-poisoning attacks of the reference as device row transforms
(/root/reference/src/attacks/__init__.py:1-58 lists more attack types: the data-poisoning ones
-- label flip, backdoor -- act on datasets, not on the update matrix, and are out of this
engine's scope; SURVEY.md section 8f-3 names the four below)."""
from .base_attack import BaseAttack, NoAttack
from .model_poisoning import (ModelReplacementAttack, AdaptiveKrumAttack, ScalingAttack,
                              InnerProductManipulationAttack)

__all__ = ['BaseAttack', 'NoAttack', 'ModelReplacementAttack', 'AdaptiveKrumAttack', 'ScalingAttack',
           'InnerProductManipulationAttack']

_ATTACKS = {
    'none': NoAttack,
    'model_replacement': ModelReplacementAttack,
    'adaptive_krum': AdaptiveKrumAttack,
    'scaling': ScalingAttack,
    'ipm': InnerProductManipulationAttack,
}


def get_attack(attack_type: str, attack_config: dict):
    """Factory by name (src/attacks/__init__.py:31-58), model-poisoning entries only."""
    if attack_type not in _ATTACKS:
        raise ValueError(f"Unknown attack type: {attack_type}. Available: {list(_ATTACKS.keys())}")
    return _ATTACKS[attack_type](attack_config)
