"""Drop-in mirror of the reference's ``src.defenses`` package
(/root/reference/src/defenses/__init__.py:1-59): same class names, constructor dicts,
attributes, exceptions and factory, computed by hand-written sm_100a kernels through
libflagg.so.  Put ``multimodal-fl-security_b200/`` on sys.path in place of the
reference's repo root and ``from src.defenses import get_defense`` keeps working."""
from .base_defense import BaseDefense, NoDefense
from .krum import KrumDefense, MultiKrumDefense
from .trimmed_mean import TrimmedMeanDefense, MedianDefense, GeometricMedianDefense
from .fltrust import FLTrustDefense
from .differential_privacy import DPSGDDefense, GradientClippingDefense, NormBoundingDefense

__all__ = [
    'BaseDefense', 'NoDefense', 'KrumDefense', 'MultiKrumDefense', 'TrimmedMeanDefense', 'MedianDefense',
    'GeometricMedianDefense', 'FLTrustDefense', 'DPSGDDefense', 'GradientClippingDefense', 'NormBoundingDefense',
]

_DEFENSES = {
    'none': NoDefense, 'fedavg': NoDefense,
    'krum': KrumDefense, 'multi_krum': MultiKrumDefense,
    'trimmed_mean': TrimmedMeanDefense, 'median': MedianDefense, 'geometric_median': GeometricMedianDefense,
    'fltrust': FLTrustDefense,
    'dp_sgd': DPSGDDefense, 'gradient_clipping': GradientClippingDefense, 'norm_bounding': NormBoundingDefense,
}


def get_defense(defense_type: str, defense_config: dict):
    """Factory by name; same 11 names and the same ValueError as the reference
    (src/defenses/__init__.py:28-59)."""
    if defense_type not in _DEFENSES:
        raise ValueError(f"Unknown defense type: {defense_type}. Available: {list(_DEFENSES.keys())}")
    return _DEFENSES[defense_type](defense_config)
